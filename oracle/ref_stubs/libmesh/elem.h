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
class Elem {
public:
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
  bool has_affine_map() const { return _affine; }
  void force_affine(bool a) { _affine = a; }
  unsigned int local_singular_node(const Point&, Real = 0.) const { return invalid_uint; }
  bool on_reference_element(const Point&, Real = 0.) const { return true; }
  Real hmax() const { return 1.; }
  Point master_point(unsigned int) const { return Point(); }  // only inverse_map (never called here) asks
  void print_info(std::ostream& os) const { os << get_info(); }
  std::string get_info() const { return "Elem (stub) type " + std::to_string((int)_t); }
private:
  ElemType _t;
  unsigned int _p;
  bool _affine = false;
  dof_id_type _id = 0;
  std::vector<Node> _nodes;
};
}  // namespace libMesh
#endif
