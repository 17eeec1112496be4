// ref_mesh_shim.cpp — the reference's own MeshTools::Generation::build_cube over the recording mesh of ref_mesh/ref_mesh_world.h:
// oracle/_ref/libref_mesh.so.  Test infrastructure (see that header); nothing under libmesh_b200/ may touch it.
#include "ref_mesh_world.h"

namespace libMesh {
namespace MeshTools { namespace Generation { namespace Private {
#include "ref_mesh_idx.inc"  // unsigned int idx(type, nx, i, j) and idx(type, nx, ny, i, j, k), mesh_generation.C:81-157
} } }
namespace {
#include "ref_mesh_so_excerpts.inc"  // map_hi_order_node, transfer_elem, all_increased_order_range: unstructured_mesh.C:60-587
}
#include "ref_mesh_excerpts.inc"  // void MeshTools::Generation::build_cube(...), mesh_generation.C:317-1938; Hex27::side_nodes_map;
                                  // void ReplicatedMesh::renumber_nodes_and_elements(), replicated_mesh.C:700-857;
                                  // void DofMap::distribute_local_dofs_var_major(...), dof_map.C:1417-1535;
                                  // void LinearPartitioner::partition_range(...), linear_partitioner.C:36-117
void UnstructuredMesh::snapshot_before() {
  if (have_snapshot) return;
  have_snapshot = true;
  xyz_before.clear();
  conn_before.clear();
  for (const Node* n : _nodes) for (int d = 0; d < 3; ++d) xyz_before.push_back(n ? (*n)(d) : 0.);
  for (const Elem* e : _elements) if (e) for (unsigned int a = 0; a < e->n_nodes(); ++a) conn_before.push_back(e->node_ptr(a)->id());
}
}  // namespace libMesh

using namespace libMesh;

namespace {
ReplicatedMesh g_mesh;
std::vector<Elem*> g_active;
}

extern "C" {

// build_cube(nx, ny, nz, bounds, type): returns 0 and leaves the record to the getters below, or -1 with the message
int ref_build_cube(unsigned int nx, unsigned int ny, unsigned int nz, const double* b, int elem_type, char* err, int err_len) {
  try {
    MeshTools::Generation::build_cube(g_mesh, nx, ny, nz, b[0], b[1], b[2], b[3], b[4], b[5], static_cast<ElemType>(elem_type));
    g_active = g_mesh.element_ptr_range();
    return 0;
  } catch (const std::exception& e) {
    if (err && err_len > 0) {
      int n = 0;
      for (const char* c = e.what(); *c && n < err_len - 1; ++c) err[n++] = *c;
      err[n] = 0;
    }
    return -1;
  }
}
// sizes: n_nodes, n_elem (active), nodes per element, number of boundary sides, all_second_order requests, prepare_for_use called
void ref_mesh_sizes(int64_t* out) {
  out[0] = (int64_t)g_mesh._nodes.size();
  out[1] = (int64_t)g_active.size();
  out[2] = g_active.empty() ? 0 : (int64_t)g_active[0]->n_nodes();
  int64_t ns = 0;
  for (const auto& s : g_mesh.get_boundary_info().sides) ns += s.e != nullptr;
  out[3] = ns;
  out[4] = g_mesh.second_order_requests;
  out[5] = g_mesh.prepared ? 1 : 0;
}
// DofMap::distribute_local_dofs_var_major on the mesh of the last ref_build_cube, for one group of n_vars LAGRANGE variables whose
// order is the element's: dof_base[n_nodes] (the dof of variable v at a node is dof_base + v, DofObject::dof_number with one
// component per variable); returns n_dofs, or -1
int64_t ref_distribute_dofs(unsigned int n_vars, uint32_t* dof_base, char* err, int err_len) {
  try {
    for (Node* n : g_mesh._nodes) if (n) n->set_vg_dof_base(0, 0, DofObject::invalid_id);
    DofMap dof_map(n_vars);
    dof_id_type next_free_dof = 0;
    dof_map.distribute_local_dofs_var_major(next_free_dof, g_mesh, {});
    for (std::size_t i = 0; i < g_mesh._nodes.size(); ++i) dof_base[i] = g_mesh._nodes[i] ? g_mesh._nodes[i]->vg_dof_base(0, 0) : (uint32_t)-1;
    return (int64_t)next_free_dof;
  } catch (const std::exception& e) {
    if (err && err_len > 0) {
      int n = 0;
      for (const char* c = e.what(); *c && n < err_len - 1; ++c) err[n++] = *c;
      err[n] = 0;
    }
    return -1;
  }
}

// LinearPartitioner::partition_range over all elements of the mesh of the last ref_build_cube: elem_pid[n_elem]
int ref_partition_linear(unsigned int n_parts, int32_t* elem_pid, char* err, int err_len) {
  try {
    LinearPartitioner lp;
    lp.partition_range(g_mesh, g_mesh._elements.begin(), g_mesh._elements.end(), n_parts);
    for (std::size_t e = 0; e < g_mesh._elements.size(); ++e) elem_pid[e] = (int32_t)g_mesh._elements[e]->processor_id();
    return 0;
  } catch (const std::exception& e) {
    if (err && err_len > 0) {
      int n = 0;
      for (const char* c = e.what(); *c && n < err_len - 1; ++c) err[n++] = *c;
      err[n] = 0;
    }
    return -1;
  }
}

// the record as it was when prepare_for_use was called (before the renumbering): sizes (lattice nodes, connectivity entries), arrays
void ref_mesh_before_sizes(int64_t* out) { out[0] = (int64_t)g_mesh.xyz_before.size() / 3; out[1] = (int64_t)g_mesh.conn_before.size(); }
void ref_mesh_before(double* xyz, uint32_t* conn) {
  std::copy(g_mesh.xyz_before.begin(), g_mesh.xyz_before.end(), xyz);
  std::copy(g_mesh.conn_before.begin(), g_mesh.conn_before.end(), conn);
}
// xyz[n_nodes][3]; conn[n_elem][npe] (node ids) in the order the elements were added; elem ids; boundary sides (position of the
// element in that order, side, id) in the order they were added
void ref_mesh_get(double* xyz, uint32_t* conn, uint32_t* elem_id, uint32_t* side_elem, uint8_t* side_no, int16_t* side_bid) {
  for (std::size_t n = 0; n < g_mesh._nodes.size(); ++n) for (int d = 0; d < 3; ++d) xyz[3 * n + d] = g_mesh._nodes[n] ? (*g_mesh._nodes[n])(d) : 0.;
  std::map<const Elem*, uint32_t> pos;
  const std::size_t npe = g_active.empty() ? 0 : g_active[0]->n_nodes();
  for (std::size_t e = 0; e < g_active.size(); ++e) {
    pos[g_active[e]] = (uint32_t)e;
    elem_id[e] = g_active[e]->id();
    for (std::size_t a = 0; a < npe; ++a) conn[e * npe + a] = g_active[e]->node_ptr((unsigned int)a)->id();
  }
  std::size_t k = 0;
  for (const auto& s : g_mesh.get_boundary_info().sides) {
    if (!s.e) continue;
    side_elem[k] = pos.at(s.e);
    side_no[k] = (uint8_t)s.side;
    side_bid[k] = s.id;
    ++k;
  }
}

}  // extern "C"
