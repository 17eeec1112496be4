// ref_mesh/ref_mesh_world.h — STUB mesh world (test infrastructure, written for this repository; not reference code).
//
// Purpose: run the reference's own MeshTools::Generation::build_cube (src/mesh/mesh_generation.C:317-1938, with its idx() helpers
// :81-157) — the DEFINITIONS are cut out of the file where it lies at build time (oracle/ref_excerpt.py, sets "mesh_idx" /
// "mesh"; the file as a whole pulls in every element class, the sphere / extrusion / Delaunay generators and MeshRefinement) —
// against a mesh that only RECORDS what the generator does: points with their ids, elements with their node pointers, side and
// node boundary ids.  tests/test_oracle_ref_mesh.py compares the record with the oracle's build_cube restatement (orc_mesh.cpp)
// before renumbering: lattice indexing, node coordinates (the reference's own scaling arithmetic), connectivity, boundary sides,
// and the 24-tet split that precedes all_second_order for TET10.  prepare_for_use() here is ONE step of the reference's: its own
// ReplicatedMesh::renumber_nodes_and_elements (src/mesh/replicated_mesh.C:700-857, excerpted likewise) — the node renumbering in
// order of first touch and the removal of orphaned lattice nodes that give a build_cube mesh its final numbering; the record can
// be read before and after it.  For TET10 the reference's own UnstructuredMesh::all_second_order_range with its helpers
// all_increased_order_range / transfer_elem / map_hi_order_node (src/mesh/unstructured_mesh.C:60-587, :1731-1795),
// Elem::second_order_equivalent_type (src/geom/elem.C) and Tet10's mid-edge table raise the TET4 split.  On the finished mesh the reference's own DofMap::distribute_local_dofs_var_major
// (src/base/dof_map.C:1417-1535) numbers the dofs of one group of LAGRANGE variables, and its own LinearPartitioner::partition_range
// (src/partitioning/linear_partitioner.C:36-117) assigns the elements to ranks.  What this world does NOT do: neighbour search, point
// locators, ghosting — nothing the path reads.
#ifndef REF_MESH_WORLD_H
#define REF_MESH_WORLD_H
#include <algorithm>
#include <array>
#include <map>
#include <unordered_map>
#include <unordered_set>

#include "libmesh/libmesh_common.h"
#include "libmesh/enum_elem_type.h"  // the reference's own enums
#include "libmesh/enum_order.h"
#include "libmesh/enum_to_string.h"
#include "libmesh/point.h"

#define LIBMESH_ENABLE_UNIQUE_ID 1

namespace libMesh {
typedef short boundary_id_type;
typedef unsigned int processor_id_type;
typedef uint64_t unique_id_type;
#define timpi_parallel_only(comm) ((void)0)

// the side-node table build_side_ptr of the HEX27 base elements needs (include/geom/cell_hex27.h); DEFINITION excerpted from
// src/geom/cell_hex27.C
struct Hex27 { static const int num_sides = 6, nodes_per_side = 9; static const unsigned int side_nodes_map[num_sides][nodes_per_side]; };
// the mid-edge nodes of a TET10 and the vertices they sit between (include/geom/cell_tet10.h:277); DEFINITION excerpted from
// src/geom/cell_tet10.C
struct Tet10 { static const unsigned short int _second_order_adjacent_vertices[6][2]; };

class MNode : public Point {  // include/geom/node.h: a point with an id
public:
  MNode(const Point& p, dof_id_type id) : Point(p), _id(id) {}
  dof_id_type id() const { return _id; }
  void set_id(dof_id_type i) { _id = i; }
  // DofObject (include/base/dof_object.h), as far as DofMap::distribute_local_dofs_var_major asks: one system, one variable group
  // of LAGRANGE variables whose order is the element's (one component per node), one processor
  static const dof_id_type invalid_id = static_cast<dof_id_type>(-1);
  unsigned int n_comp_group(const unsigned int, const unsigned int) const { return 1; }
  static const processor_id_type invalid_processor_id = static_cast<processor_id_type>(-1);
  processor_id_type processor_id() const { return _pid; }
  processor_id_type& processor_id() { return _pid; }
  void set_unique_id(unique_id_type) {}
  dof_id_type vg_dof_base(const unsigned int, const unsigned int) const { return _dof_base; }
  void set_vg_dof_base(const unsigned int, const unsigned int, const dof_id_type db) { _dof_base = db; }
private:
  dof_id_type _id;
  dof_id_type _dof_base = invalid_id;
  processor_id_type _pid = 0;
};
typedef MNode DofObject;
typedef unsigned short subdomain_id_type;
#define Node MNode  /* ref_stubs' point.h already names a value-type Node for the FE library; this world's nodes are shared by pointer */

class Elem {
public:
  explicit Elem(ElemType t = INVALID_ELEM) : _t(t), _nodes(type_n_nodes(t), nullptr) {}
  // Elem::type_to_n_nodes_map / n_vertices / default_order (src/geom/elem.C, include/geom/cell_*.h), for the types build_cube makes
  static unsigned int type_n_nodes(ElemType t) {
    switch (t) {
      case NODEELEM: return 1; case EDGE2: return 2; case EDGE3: return 3; case EDGE4: return 4; case TRI3: return 3; case TRI6: return 6; case TRI7: return 7;
      case QUAD4: return 4; case QUAD8: return 8; case QUAD9: return 9; case TET4: return 4; case TET10: return 10; case TET14: return 14;
      case HEX8: return 8; case HEX20: return 20; case HEX27: return 27; case PRISM6: return 6; case PRISM15: return 15; case PRISM18: return 18;
      case PRISM20: return 20; case PRISM21: return 21; case PYRAMID5: return 5; case PYRAMID13: return 13; case PYRAMID14: return 14; case PYRAMID18: return 18;
      default: return 0;
    }
  }
  unsigned int n_vertices() const {
    switch (_t) {
      case TET4: case TET10: case TET14: return 4; case HEX8: case HEX20: case HEX27: return 8;
      case PRISM6: case PRISM15: case PRISM18: case PRISM20: case PRISM21: return 6; case PYRAMID5: case PYRAMID13: case PYRAMID14: case PYRAMID18: return 5;
      case QUAD4: case QUAD8: case QUAD9: return 4; case TRI3: case TRI6: case TRI7: return 3; case EDGE2: case EDGE3: case EDGE4: return 2;
      default: return n_nodes();
    }
  }
  Order default_order() const { return n_vertices() == n_nodes() ? FIRST : SECOND; }
  // Tet10::n_second_order_adjacent_vertices / second_order_adjacent_vertex (include/geom/cell_tet10.h:193-203, src/geom/cell_tet10.C:412-419)
  unsigned int n_second_order_adjacent_vertices(const unsigned int) const { libmesh_error_msg_if(_t != TET10, "TET10 only in this stub"); return 2; }
  unsigned short int second_order_adjacent_vertex(const unsigned int n, const unsigned int v) const {
    libmesh_error_msg_if(_t != TET10, "TET10 only in this stub");
    return Tet10::_second_order_adjacent_vertices[n - this->n_vertices()][v];
  }
  // DECLARATION (include/geom/elem.h); the definition is the reference's, src/geom/elem.C:3168-3290
  static ElemType second_order_equivalent_type(const ElemType et, const bool full_ordered = true);
  dof_id_type node_id(unsigned int i) const { return _nodes.at(i)->id(); }
  unsigned int level() const { return 0; }
  static const processor_id_type invalid_processor_id = static_cast<processor_id_type>(-1);
  // no neighbour, interior-parent or extra-integer data exists before prepare_for_use on a generated mesh
  Elem* neighbor_ptr(unsigned int) const { return nullptr; }
  unsigned int which_neighbor_am_i(const Elem*) const { return invalid_uint; }
  void set_neighbor(unsigned int, Elem*) {}
  Elem* interior_parent() const { return nullptr; }
  void set_interior_parent(Elem*) {}
  unsigned int n_extra_integers() const { return 0; }
  void add_extra_integers(unsigned int) {}
  void set_extra_integer(unsigned int, dof_id_type) {}
  dof_id_type get_extra_integer(unsigned int) const { return 0; }
  void inherit_data_from(const Elem& o) { _pid = o._pid; }
  unique_id_type unique_id() const { return 0; }
  void set_unique_id(unique_id_type) {}
  virtual ~Elem() = default;
  static std::unique_ptr<Elem> build(const ElemType t) { return std::make_unique<Elem>(t); }
  static std::unique_ptr<Elem> build_with_id(const ElemType t, dof_id_type id) { auto e = std::make_unique<Elem>(t); e->_id = id; return e; }
  ElemType type() const { return _t; }
  dof_id_type id() const { return _id; }
  void set_id(dof_id_type i) { _id = i; }
  dof_id_type& set_id() { return _id; }
  void set_node(unsigned int i, Node* n) { if (_nodes.size() <= i) _nodes.resize(i + 1, nullptr); _nodes[i] = n; }  // (polygons grow)
  Node* node_ptr(unsigned int i) const { return _nodes.at(i); }
  unsigned int n_nodes() const { return (unsigned int)_nodes.size(); }
  Node& node_ref(unsigned int i) const { return *_nodes.at(i); }
  subdomain_id_type subdomain_id() const { return 0; }
  processor_id_type& processor_id() { return _pid; }  // DofObject::processor_id
  processor_id_type processor_id() const { return _pid; }
  unsigned int n_comp_group(const unsigned int, const unsigned int) const { return 0; }  // LAGRANGE: no element dofs
  dof_id_type vg_dof_base(const unsigned int, const unsigned int) const { return (dof_id_type)-1; }
  void set_vg_dof_base(const unsigned int, const unsigned int, const dof_id_type) { libmesh_not_implemented(); }
  // Elem::node_ref_range: the element's nodes by reference
  struct NodeRefRange {
    struct iterator {
      Node* const* p;
      Node& operator*() const { return **p; }
      iterator& operator++() { ++p; return *this; }
      bool operator!=(const iterator& o) const { return p != o.p; }
    };
    Node* const* b; Node* const* e;
    iterator begin() const { return {b}; }
    iterator end() const { return {e}; }
  };
  NodeRefRange node_ref_range() const { return {_nodes.data(), _nodes.data() + _nodes.size()}; }
  IntRange<unsigned short> side_index_range() const {
    const unsigned short n = (_t == HEX8 || _t == HEX20 || _t == HEX27) ? 6 : ((_t == TET4 || _t == TET10) ? 4 : ((_t == PRISM6 || _t == PRISM18) ? 5 : 0));
    return IntRange<unsigned short>(0, n);
  }
  // Hex27::build_side_ptr: a QUAD9 over the nodes of Hex27::side_nodes_map[s]
  void build_side_ptr(std::unique_ptr<Elem>& side, const unsigned int s) {
    libmesh_error_msg_if(_t != HEX27, "build_side_ptr: HEX27 only in this stub");
    side = std::make_unique<Elem>(QUAD9);
    for (unsigned int i = 0; i < 9; ++i) side->set_node(i, _nodes[Hex27::side_nodes_map[s][i]]);
  }
private:
  ElemType _t;
  dof_id_type _id = 0;
  processor_id_type _pid = 0;
  std::vector<Node*> _nodes;
};
static Elem* const remote_elem = reinterpret_cast<Elem*>(sizeof(void*));  // include/geom/remote_elem.h: a sentinel never dereferenced
class Polygon : public Elem { public: using Elem::Elem; };
class C0Polygon : public Polygon { public: explicit C0Polygon(unsigned int) : Polygon(C0POLYGON) {} };
class C0Polyhedron : public Elem {
public:
  C0Polyhedron(const std::vector<std::shared_ptr<Polygon>>&, std::unique_ptr<Node>&) : Elem(C0POLYHEDRON) { libmesh_not_implemented(); }
};

class BoundaryInfo {
public:
  static const boundary_id_type invalid_id = -123;  // include/mesh/boundary_info.h
  void add_side(const Elem* e, const unsigned short int side, const boundary_id_type id) { sides.push_back({e, side, id}); }
  void add_node(const Node* n, const boundary_id_type id) { nodes.push_back({n, id}); }
  void boundary_ids(const Elem* const e, const unsigned short int side, std::vector<boundary_id_type>& out) const {
    out.clear();
    for (const auto& s : sides) if (s.e == e && s.side == side) out.push_back(s.id);
  }
  void remove(const Elem* e) { for (auto& s : sides) if (s.e == e) s.e = nullptr; }
  // BoundaryInfo::copy_boundary_ids (src/mesh/boundary_info.C): the side ids of one element given to another
  void copy_boundary_ids(const BoundaryInfo& old_bi, const Elem* const old_elem, const Elem* const new_elem) {
    const std::size_t n = old_bi.sides.size();
    for (std::size_t i = 0; i < n; ++i) if (old_bi.sides[i].e == old_elem) { const Side s = old_bi.sides[i]; sides.push_back({new_elem, s.side, s.id}); }
  }
  void remove(const Node* n) { for (auto& s : nodes) if (s.n == n) s.n = nullptr; }
  std::string& sideset_name(boundary_id_type id) { return _ss[id]; }
  std::string& nodeset_name(boundary_id_type id) { return _ns[id]; }
  struct Side { const Elem* e; unsigned short side; boundary_id_type id; };
  struct NodeId { const Node* n; boundary_id_type id; };
  std::vector<Side> sides;
  std::vector<NodeId> nodes;
private:
  std::map<boundary_id_type, std::string> _ss, _ns;
};

// include/utils/simple_range.h as_range
template <typename It> struct SimpleRange { It b, e; It begin() const { return b; } It end() const { return e; } };
template <typename It> SimpleRange<It> as_range(It b, It e) { return {b, e}; }

// storage as in ReplicatedMesh (include/mesh/replicated_mesh.h): _nodes[id], _elements[id], null where deleted
class UnstructuredMesh {
public:
  virtual ~UnstructuredMesh() { clear(); }
  // as far as LinearPartitioner::partition_range asks: a serial (replicated, one process) mesh
  typedef std::vector<Elem*>::iterator element_iterator;
  bool is_serial() const { return true; }
  struct Comm {  // one process
    template <typename S> void set_union(S&) const {}
    template <typename T> void min(T&) const {}
    template <typename T> void max(T&) const {}
    template <typename T> bool verify(const T&) const { return true; }
  };
  Comm comm() const { return Comm(); }
  Elem* query_elem_ptr(const dof_id_type i) { return i < _elements.size() ? _elements[i] : nullptr; }
  // as far as all_increased_order_range (src/mesh/unstructured_mesh.C:322-587) asks
  element_iterator elements_begin() { return _elements.begin(); }
  element_iterator elements_end() { return _elements.end(); }
  virtual void update_parallel_id_counts() {}
  unique_id_type parallel_max_unique_id() const { return 0; }
  void set_next_unique_id(unique_id_type) {}
  bool is_replicated() const { return true; }
  bool is_prepared() const { return false; }
  bool allow_find_neighbors() const { return _find_neighbors; }
  void allow_find_neighbors(bool b) { _find_neighbors = b; }
  processor_id_type processor_id() const { return 0; }
  processor_id_type n_processors() const { return 1; }
  dof_id_type n_nodes() const { return _n_nodes; }
  const Point& point(const dof_id_type i) const { return *_nodes.at(i); }
  void own_node(Node&) {}
  Elem* insert_elem(std::unique_ptr<Elem> e) {  // ReplicatedMesh::insert_elem: replaces the element of the same id
    const dof_id_type i = e->id();
    if (_elements.size() <= i) _elements.resize(i + 1, nullptr);
    if (_elements[i]) this->delete_elem(_elements[i]);  // replicated_mesh.C:358
    _elements[i] = e.release();
    return _elements[i];
  }
  // DECLARATION (include/mesh/unstructured_mesh.h:147); the definition is the reference's, src/mesh/unstructured_mesh.C:1731-1795
  void all_second_order_range(const SimpleRange<element_iterator>& range, const bool full_ordered = true);
  bool _find_neighbors = true;
  void clear() {
    for (Node* n : _nodes) delete n;
    for (Elem* e : _elements) delete e;
    _nodes.clear(); _elements.clear(); _bi = BoundaryInfo(); _n_nodes = 0;
    second_order_requests = 0; prepared = false; have_snapshot = false;
  }
  BoundaryInfo& get_boundary_info() { return _bi; }
  void set_mesh_dimension(unsigned char d) { _dim = d; }
  void set_spatial_dimension(unsigned char) {}
  unsigned int mesh_dimension() const { return _dim; }
  void reserve_elem(dof_id_type n) { reserved_elem = n; }
  void reserve_nodes(dof_id_type n) { reserved_nodes = n; }
  Node* add_point(const Point& p, const dof_id_type id = (dof_id_type)-1, const processor_id_type = 0) {  // ReplicatedMesh::add_point: at _nodes[id], or appended
    const dof_id_type i = id == (dof_id_type)-1 ? (dof_id_type)_nodes.size() : id;
    if (_nodes.size() <= i) _nodes.resize(i + 1, nullptr);
    libmesh_error_msg_if(_nodes[i], "node id " << i << " is taken");
    _nodes[i] = new Node(p, i);
    ++_n_nodes;
    return _nodes[i];
  }
  Node* add_node(std::unique_ptr<Node> n) { _nodes.push_back(n.release()); ++_n_nodes; return _nodes.back(); }
  Node* node_ptr(const dof_id_type i) { return _nodes.at(i); }
  Elem* add_elem(std::unique_ptr<Elem> e) {  // ReplicatedMesh::add_elem: at _elements[id]
    const dof_id_type i = e->id();
    if (_elements.size() <= i) _elements.resize(i + 1, nullptr);
    libmesh_error_msg_if(_elements[i], "element id " << i << " is taken");
    _elements[i] = e.release();
    return _elements[i];
  }
  dof_id_type n_elem() const { dof_id_type n = 0; for (const Elem* e : _elements) n += e != nullptr; return n; }
  std::vector<Node*> node_ptr_range() { std::vector<Node*> v; for (Node* n : _nodes) if (n) v.push_back(n); return v; }
  std::vector<Elem*> element_ptr_range() { std::vector<Elem*> v; for (Elem* e : _elements) if (e) v.push_back(e); return v; }
  std::vector<Elem*> active_local_element_ptr_range() { return element_ptr_range(); }  // one level, one processor
  std::vector<Node*> local_node_ptr_range() { return node_ptr_range(); }
  void delete_elem(Elem* e) { _bi.remove(e); _elements.at(e->id()) = nullptr; delete e; }  // ReplicatedMesh::delete_elem (replicated_mesh.C:391-435)
  // MeshBase::all_second_order (src/mesh/mesh_base.C): the whole mesh
  void all_second_order(const bool full_ordered = true) {
    ++second_order_requests;
    snapshot_before();
    this->all_second_order_range({this->elements_begin(), this->elements_end()}, full_ordered);
  }
  // the record as it is now, once: kept for the tests (what the generator built before any renumbering or order raising)
  void snapshot_before();
  bool have_snapshot = false;
  std::vector<double> xyz_before;
  std::vector<uint32_t> conn_before;
  void all_complete_order() { libmesh_not_implemented(); }
  virtual void prepare_for_use() { prepared = true; }
  int second_order_requests = 0;
  bool prepared = false;
  dof_id_type reserved_elem = 0, reserved_nodes = 0;
  std::vector<Node*> _nodes;
  std::vector<Elem*> _elements;
  dof_id_type _n_nodes = 0;
protected:
  BoundaryInfo _bi;
  unsigned char _dim = 0;
};
class ReplicatedMesh : public UnstructuredMesh {
public:
  // DECLARATION (include/mesh/replicated_mesh.h); the definition is the reference's, src/mesh/replicated_mesh.C:700-857
  void renumber_nodes_and_elements();
  // MeshBase::prepare_for_use, as far as this world goes: the numbering step, once the generator is done.  The record before it
  // is kept for the tests.  (A mesh that still waits for all_second_order is left alone.)
  void prepare_for_use() override {
    prepared = true;
    snapshot_before();
    this->renumber_nodes_and_elements();
  }
  void update_parallel_id_counts() override {}
  bool _skip_renumber_nodes_and_elements = false;
  std::map<const Node*, int> _constraint_rows;
  struct { bool has_removed_orphaned_nodes = false; } _preparation;
};

typedef UnstructuredMesh MeshBase;
// named by the partial-range branch of all_increased_order_range (never taken: the whole mesh is raised)
struct GhostingFunctor { typedef std::map<const Elem*, const int*> map_type; };
class GhostPointNeighbors {
public:
  explicit GhostPointNeighbors(const MeshBase&) {}
  void operator()(MeshBase::element_iterator, MeshBase::element_iterator, processor_id_type, GhostingFunctor::map_type&) { libmesh_not_implemented(); }
};
struct MeshCommunication { void make_nodes_parallel_consistent(MeshBase&) {} };

// include/base/dof_map.h, variable.h: one system with one group of `n` LAGRANGE variables of the same type (what
// System::add_variable makes of consecutive variables of one FEType, identify_variable_groups = true)
enum FEFamilyStub { LAGRANGE_FAMILY = 0, SCALAR = 31 };
struct FETypeStub { FEFamilyStub family = LAGRANGE_FAMILY; };
class VariableGroup {
public:
  explicit VariableGroup(unsigned int n) : _n(n) {}
  unsigned int n_variables() const { return _n; }
  const FETypeStub& type() const { return _type; }
  bool active_on_subdomain(subdomain_id_type) const { return true; }
private:
  unsigned int _n;
  FETypeStub _type;
};
class DofMap {
public:
  explicit DofMap(unsigned int n_vars) : _group(n_vars) {}
  unsigned int sys_number() const { return 0; }
  unsigned int n_variable_groups() const { return 1; }
  const VariableGroup& variable_group(unsigned int) const { return _group; }
  unsigned int processor_id() const { return 0; }
  void distribute_scalar_dofs(dof_id_type&) {}
  // DECLARATION (include/base/dof_map.h); the definition is the reference's, src/base/dof_map.C:1417-1535
  void distribute_local_dofs_var_major(dof_id_type& next_free_dof, MeshBase& mesh,
                                       const std::map<const Node*, std::set<subdomain_id_type>>& constraining_subdomains);
private:
  VariableGroup _group;
};

// include/partitioning/linear_partitioner.h: DECLARATION; the definition of partition_range is the reference's,
// src/partitioning/linear_partitioner.C:36-117
class LinearPartitioner {
public:
  void partition_range(MeshBase& mesh, MeshBase::element_iterator it, MeshBase::element_iterator end, const unsigned int n);
  void single_partition_range(MeshBase::element_iterator it, MeshBase::element_iterator end) { for (; it != end; ++it) (*it)->processor_id() = 0; }  // partitioner.C
};

template <typename Output> class FunctionBase { public: virtual ~FunctionBase() = default; };
namespace MeshTools {
namespace Modification { inline void redistribute(UnstructuredMesh&, const FunctionBase<Real>&) { libmesh_not_implemented(); } }
namespace Generation {
// DECLARATION (include/mesh/mesh_generation.h:70-86); the definition is the reference's
void build_cube(UnstructuredMesh& mesh, const unsigned int nx = 0, const unsigned int ny = 0, const unsigned int nz = 0, const Real xmin = 0., const Real xmax = 1.,
                const Real ymin = 0., const Real ymax = 1., const Real zmin = 0., const Real zmax = 1., const ElemType type = INVALID_ELEM,
                const bool gauss_lobatto_grid = false);
namespace Private {
// the Gauss-Lobatto grid option of build_cube (mesh_generation.C:165-308) is not on the path: named, never constructed with true
class GaussLobattoRedistributionFunction : public FunctionBase<Real> {
public:
  GaussLobattoRedistributionFunction(unsigned int, Real, Real, unsigned int = 0, Real = 0, Real = 0, unsigned int = 0, Real = 0, Real = 0) { libmesh_not_implemented(); }
};
}  // namespace Private
}  // namespace Generation
}  // namespace MeshTools
}  // namespace libMesh
#endif
