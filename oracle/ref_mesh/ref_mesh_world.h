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
// be read before and after it.  On the finished mesh the reference's own DofMap::distribute_local_dofs_var_major
// (src/base/dof_map.C:1417-1535) numbers the dofs of one group of LAGRANGE variables, and its own LinearPartitioner::partition_range
// (src/partitioning/linear_partitioner.C:36-117) assigns the elements to ranks.  What this world does NOT do: UnstructuredMesh::all_second_order (recorded as a request; a TET10
// mesh is left as the TET4 split, unrenumbered), neighbours, partitioning.
#ifndef REF_MESH_WORLD_H
#define REF_MESH_WORLD_H
#include <array>
#include <map>
#include <unordered_set>

#include "libmesh/libmesh_common.h"
#include "libmesh/enum_elem_type.h"  // the reference's own enum
#include "libmesh/enum_to_string.h"
#include "libmesh/point.h"

#define LIBMESH_ENABLE_UNIQUE_ID 1

namespace libMesh {
typedef short boundary_id_type;

// the side-node table build_side_ptr of the HEX27 base elements needs (include/geom/cell_hex27.h); DEFINITION excerpted from
// src/geom/cell_hex27.C
struct Hex27 { static const int num_sides = 6, nodes_per_side = 9; static const unsigned int side_nodes_map[num_sides][nodes_per_side]; };

class MNode : public Point {  // include/geom/node.h: a point with an id
public:
  MNode(const Point& p, dof_id_type id) : Point(p), _id(id) {}
  dof_id_type id() const { return _id; }
  void set_id(dof_id_type i) { _id = i; }
  // DofObject (include/base/dof_object.h), as far as DofMap::distribute_local_dofs_var_major asks: one system, one variable group
  // of LAGRANGE variables whose order is the element's (one component per node), one processor
  static const dof_id_type invalid_id = static_cast<dof_id_type>(-1);
  unsigned int n_comp_group(const unsigned int, const unsigned int) const { return 1; }
  unsigned int processor_id() const { return 0; }
  dof_id_type vg_dof_base(const unsigned int, const unsigned int) const { return _dof_base; }
  void set_vg_dof_base(const unsigned int, const unsigned int, const dof_id_type db) { _dof_base = db; }
private:
  dof_id_type _id;
  dof_id_type _dof_base = invalid_id;
};
typedef MNode DofObject;
typedef unsigned short subdomain_id_type;
#define Node MNode  /* ref_stubs' point.h already names a value-type Node for the FE library; this world's nodes are shared by pointer */

class Elem {
public:
  explicit Elem(ElemType t = INVALID_ELEM) : _t(t) {}
  virtual ~Elem() = default;
  static std::unique_ptr<Elem> build(const ElemType t) { return std::make_unique<Elem>(t); }
  static std::unique_ptr<Elem> build_with_id(const ElemType t, dof_id_type id) { auto e = std::make_unique<Elem>(t); e->_id = id; return e; }
  ElemType type() const { return _t; }
  dof_id_type id() const { return _id; }
  void set_id(dof_id_type i) { _id = i; }
  dof_id_type& set_id() { return _id; }
  void set_node(unsigned int i, Node* n) { if (_nodes.size() <= i) _nodes.resize(i + 1, nullptr); _nodes[i] = n; }
  Node* node_ptr(unsigned int i) const { return _nodes.at(i); }
  unsigned int n_nodes() const { return (unsigned int)_nodes.size(); }
  Node& node_ref(unsigned int i) const { return *_nodes.at(i); }
  subdomain_id_type subdomain_id() const { return 0; }
  unsigned int& processor_id() { return _pid; }  // DofObject::processor_id
  unsigned int processor_id() const { return _pid; }
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
  IntRange<unsigned short> side_index_range() const { return IntRange<unsigned short>(0, (unsigned short)(_t == HEX27 ? 6 : 0)); }
  // Hex27::build_side_ptr: a QUAD9 over the nodes of Hex27::side_nodes_map[s]
  void build_side_ptr(std::unique_ptr<Elem>& side, const unsigned int s) {
    libmesh_error_msg_if(_t != HEX27, "build_side_ptr: HEX27 only in this stub");
    side = std::make_unique<Elem>(QUAD9);
    for (unsigned int i = 0; i < 9; ++i) side->set_node(i, _nodes[Hex27::side_nodes_map[s][i]]);
  }
private:
  ElemType _t;
  dof_id_type _id = 0;
  unsigned int _pid = 0;
  std::vector<Node*> _nodes;
};
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
  struct Comm { template <typename S> void set_union(S&) const {} };
  Comm comm() const { return Comm(); }
  Elem* query_elem_ptr(const dof_id_type i) { return i < _elements.size() ? _elements[i] : nullptr; }
  void clear() {
    for (Node* n : _nodes) delete n;
    for (Elem* e : _elements) delete e;
    _nodes.clear(); _elements.clear(); _bi = BoundaryInfo(); _n_nodes = 0;
    second_order_requests = 0; prepared = false;
  }
  BoundaryInfo& get_boundary_info() { return _bi; }
  void set_mesh_dimension(unsigned char d) { _dim = d; }
  void set_spatial_dimension(unsigned char) {}
  unsigned int mesh_dimension() const { return _dim; }
  void reserve_elem(dof_id_type n) { reserved_elem = n; }
  void reserve_nodes(dof_id_type n) { reserved_nodes = n; }
  Node* add_point(const Point& p, const dof_id_type id = (dof_id_type)-1) {  // ReplicatedMesh::add_point: at _nodes[id], or appended
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
  void delete_elem(Elem* e) { _elements.at(e->id()) = nullptr; delete e; }  // ReplicatedMesh::delete_elem
  void all_second_order(const bool = true) { ++second_order_requests; }     // not done here (see the header comment)
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
    if (!second_order_requests) this->renumber_nodes_and_elements();
  }
  void snapshot_before();
  std::vector<double> xyz_before;
  std::vector<uint32_t> conn_before;
  void update_parallel_id_counts() {}
  bool _skip_renumber_nodes_and_elements = false;
  std::map<const Node*, int> _constraint_rows;
  struct { bool has_removed_orphaned_nodes = false; } _preparation;
};

typedef UnstructuredMesh MeshBase;

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

typedef unsigned int processor_id_type;
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
