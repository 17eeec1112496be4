// ref_stubs/libmesh/elem.h — STUB (see libmesh_common.h in this directory): what the reference's shape-function, quadrature
// and FEMap sources ask of an element — its type, p-level, nodes, mapping type and default order.  has_affine_map() is a flag
// the caller sets: the geometric test itself (Hex27::has_affine_map ..., src/geom/cell_*.C) is part of the element class
// hierarchy, which cannot be compiled stand-alone; both branches of FEMap::compute_map can be forced and compared.
#ifndef REF_STUB_ELEM_H
#define REF_STUB_ELEM_H
#include "libmesh/enum_elem_type.h"
#include "libmesh/enum_order.h"
#include "libmesh/point.h"
namespace libMesh {
// DECLARATIONS of what is taken from the element classes (include/geom/cell_hex8.h ...): the side-node tables (num_sides,
// nodes_per_side, side_nodes_map) and has_affine_map().  The DEFINITIONS — tables and function bodies — are the reference's own,
// cut out of src/geom/cell_*.C at build time (oracle/ref_excerpt.py, set "fe") and compiled in ref_shim.cpp.
class Elem;
#define REF_CELL_STUB(Name, NS, NPS)                                                                                      \
  struct Name {                                                                                                            \
    static const int num_sides = NS, nodes_per_side = NPS;                                                                 \
    static const unsigned int side_nodes_map[num_sides][nodes_per_side];                                                   \
    static constexpr Real affine_tol = TOLERANCE * TOLERANCE; /* Elem::affine_tol, include/geom/elem.h:2070 */             \
    const Elem* _e = nullptr;                                                                                              \
    const Point& point(const unsigned int i) const;                                                                        \
    bool has_affine_map() const; /* DECLARATION; the definition is the reference's, cut out of src/geom/cell_*.C */        \
  };
REF_CELL_STUB(Hex8, 6, 4)
REF_CELL_STUB(Hex20, 6, 8)
REF_CELL_STUB(Hex27, 6, 9)
REF_CELL_STUB(Tet10, 4, 6)
REF_CELL_STUB(Prism18, 5, 9)

class Elem {
public:
  // Elem::type_to_n_nodes_map (src/geom/elem.C), for the types this shim is asked for
  struct NNodesMap {
    unsigned int operator[](ElemType t) const {
      switch (t) {
        case EDGE2: return 2; case EDGE3: return 3; case TRI3: return 3; case TRI6: return 6; case QUAD4: return 4; case QUAD8: return 8; case QUAD9: return 9;
        case TET4: return 4; case TET10: return 10; case HEX8: return 8; case HEX20: return 20; case HEX27: return 27; case PRISM6: return 6; case PRISM18: return 18;
        default: return invalid_uint;
      }
    }
  };
  static constexpr NNodesMap type_to_n_nodes_map{};
  explicit Elem(ElemType t = INVALID_ELEM, unsigned int p = 0) : _t(t), _p(p) {}
  virtual ~Elem() = default;
  ElemType type() const { return _t; }
  unsigned int p_level() const { return _p; }
  unsigned short dim() const { return (unsigned short)(_t <= EDGE4 ? 1 : (_t <= QUAD9 ? 2 : 3)); }  // the types this shim is asked for
  bool runtime_topology() const { return false; }
  dof_id_type id() const { return _id; }
  void set_id(dof_id_type i) { _id = i; }
  // nodes
  void set_nodes(const double* xyz, unsigned int n) { _nodes.clear(); for (unsigned int i = 0; i < n; ++i) _nodes.emplace_back(Point(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), i); }
  unsigned int n_nodes() const { return (unsigned int)_nodes.size(); }
  const Point& point(unsigned int i) const { return _nodes[i]; }
  const Node* node_ptr(unsigned int i) const { return &_nodes[i]; }
  const Node& node_ref(unsigned int i) const { return _nodes[i]; }
  IntRange<unsigned short> node_index_range() const { return IntRange<unsigned short>(0, (unsigned short)_nodes.size()); }
  // mapping
  ElemMappingType mapping_type() const { return LAGRANGE_MAP; }
  unsigned char mapping_data() const { return 0; }
  Order default_order() const { return (_t == HEX8 || _t == QUAD4 || _t == TET4 || _t == PRISM6 || _t == TRI3 || _t == EDGE2) ? FIRST : SECOND; }  // cell_*.h
  bool is_linear() const { return _t == TET4 || _t == TRI3 || _t == EDGE2; }  // elem.h: constant Jacobian by type
  bool infinite() const { return false; }
  // Elem::has_affine_map: the element class's own test (Hex27::has_affine_map ...), unless a caller forces a branch of compute_map
  bool has_affine_map() const;
  void force_affine(bool a) { _affine = a ? 1 : 0; }
  void auto_affine() { _affine = -1; }
  unsigned int local_singular_node(const Point&, Real = 0.) const { return invalid_uint; }
  bool on_reference_element(const Point&, Real = 0.) const { return true; }
  Real hmax() const { return 1.; }
  Point master_point(unsigned int) const { return Point(); }  // only inverse_map (never called here) asks
  // sides: Elem::n_sides / local_side_node / side_type / build_side_ptr over the side_nodes_map tables above
  unsigned int n_sides() const { return (_t == HEX8 || _t == HEX20 || _t == HEX27) ? 6u : (_t == TET10 ? 4u : (_t == PRISM18 ? 5u : 0u)); }
  IntRange<unsigned short> side_index_range() const { return IntRange<unsigned short>(0, (unsigned short)n_sides()); }
  ElemType side_type(unsigned int s) const {
    switch (_t) {
      case HEX8: return QUAD4; case HEX20: return QUAD8; case HEX27: return QUAD9; case TET10: return TRI6;
      case PRISM18: return (s == 0 || s == 4) ? TRI6 : QUAD9;
      default: return INVALID_ELEM;
    }
  }
  unsigned int local_side_node(unsigned int s, unsigned int i) const {
    switch (_t) {
      case HEX8: return Hex8::side_nodes_map[s][i]; case HEX20: return Hex20::side_nodes_map[s][i]; case HEX27: return Hex27::side_nodes_map[s][i];
      case TET10: return Tet10::side_nodes_map[s][i]; case PRISM18: return Prism18::side_nodes_map[s][i];
      default: return invalid_uint;
    }
  }
  std::unique_ptr<Elem> build_side_ptr(unsigned int s) const {
    const ElemType st = side_type(s);
    auto side = std::make_unique<Elem>(st, _p);
    const unsigned int n = type_to_n_nodes_map[st];
    std::vector<double> c(3 * (std::size_t)n);
    for (unsigned int i = 0; i < n; ++i) for (int d = 0; d < 3; ++d) c[3 * i + d] = _nodes[local_side_node(s, i)](d);
    side->set_nodes(c.data(), n);
    return side;
  }
  bool is_node_on_side(unsigned int n, unsigned int s) const {
    const unsigned int ns = type_to_n_nodes_map[side_type(s)];
    for (unsigned int i = 0; i < ns; ++i) if (local_side_node(s, i) == n) return true;
    return false;
  }
  // edges: named by main() of systems_of_equations_ex6.C only (node / edge boundary ids), which is compiled but not run
  IntRange<unsigned short> edge_index_range() const { return IntRange<unsigned short>(0, 0); }
  bool is_edge_on_side(unsigned int, unsigned int) const { libmesh_not_implemented(); }
  // neighbours and boundary ids, as far as an assembly loop asks: is the side on the boundary, and with which id
  void set_side_boundary_ids(const short* ids, unsigned int n) { _side_bid.assign(ids, ids + n); }
  short side_boundary_id(unsigned int s) const { return s < _side_bid.size() ? _side_bid[s] : (short)-1; }
  const Elem* interior_parent() const { return nullptr; }                  // the 1D / 2D branches of compute_face_map ask (never run here)
  dof_id_type node_id(unsigned int i) const { return _nodes[i].id(); }
  const Elem* neighbor_ptr(unsigned int s) const { return side_boundary_id(s) >= 0 ? nullptr : this; }  // non-null = "has a neighbour"
  void print_info(std::ostream& os) const { os << get_info(); }
  std::string get_info() const { return "Elem (stub) type " + std::to_string((int)_t); }
private:
  ElemType _t;
  unsigned int _p;
  int _affine = -1;
  dof_id_type _id = 0;
  std::vector<short> _side_bid;
  std::vector<Node> _nodes;
};
#define REF_CELL_POINT(Name) inline const Point& Name::point(const unsigned int i) const { return _e->point(i); }
REF_CELL_POINT(Hex8) REF_CELL_POINT(Hex20) REF_CELL_POINT(Hex27) REF_CELL_POINT(Tet10) REF_CELL_POINT(Prism18)
inline bool Elem::has_affine_map() const {
  if (_affine >= 0) return _affine != 0;
  switch (_t) {
    case HEX8: { Hex8 c; c._e = this; return c.has_affine_map(); }
    case HEX20: { Hex20 c; c._e = this; return c.has_affine_map(); }
    case HEX27: { Hex27 c; c._e = this; return c.has_affine_map(); }
    case TET10: { Tet10 c; c._e = this; return c.has_affine_map(); }
    case PRISM18: { Prism18 c; c._e = this; return c.has_affine_map(); }
    default: return false;  // sides and the like: the general branch is always right
  }
}
}  // namespace libMesh
#endif
