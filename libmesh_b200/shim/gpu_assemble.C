// gpu_assemble.C — see gpu_assemble.h.  Flattens what the reference reaches through pointers (Elem -> Node ->
// DofObject, src/fe/fe_map.C:1361-1366, src/base/dof_map.C:2201-2314) into the arrays libfeasm takes, once per mesh,
// then runs the assembly on the GPU and hands the RHS back to the system's NumericVector.
//
// Several ranks (one process per GPU, mesh.comm().size() > 1): every rank flattens its active_local elements; rows are
// numbered owned first (dof - first_dof()), then the ghost rows — dofs of local elements that another rank owns — in
// ascending global id.  The context joins an NCCL communicator bootstrapped over the libMesh communicator
// (GpuSparseMatrix::ensure_communicator), builds its local pattern on the device and synchronises it with the peers
// (feasm_build_exchange: what SparsityPattern::Build::parallel_sync does, sparsity_pattern.C:446-575); the owned rows of
// the result are checked against the Graph libMesh handed to update_sparsity_pattern.  An assembly is then
// feasm_assemble_begin -> feasm_exchange_begin -> feasm_assemble_end -> feasm_exchange_end: the contributions to rows of
// other ranks travel while the owned rows are reduced, the role of the stash in MatAssemblyBegin/End and of
// DistributedVector::close (distributed_vector.C:399-470).
#include "gpu_assemble.h"

#include <algorithm>
#include <cstring>
#include <unordered_map>

namespace feasm_libmesh {
using namespace libMesh;

namespace {
int feasm_type(ElemType t) {
  switch (t) {
    case HEX8: return FEASM_HEX8;
    case HEX27: return FEASM_HEX27;
    case TET10: return FEASM_TET10;
    case HEX20: return FEASM_HEX20;
    case PRISM18: return FEASM_PRISM18;
    default: libmesh_error_msg("libfeasm supports HEX8, HEX20, HEX27, TET10 and PRISM18 elements; got ElemType " << (int)t);
  }
}

// FNV-1a over the bit patterns of the node coordinates: a moved mesh is uploaded again, an unchanged one is not
uint64_t checksum(const std::vector<double>& v) {
  uint64_t h = 1469598103934665603ull;
  for (double d : v) { uint64_t u; std::memcpy(&u, &d, 8); h = (h ^ u) * 1099511628211ull; }
  return h;
}
}  // namespace

void assemble(EquationSystems& es, const std::string& system_name, const feasm_params& params, const SideLoad* side_load,
              const SidePenalty* side_penalty) {
  LinearImplicitSystem& system = es.get_system<LinearImplicitSystem>(system_name);
  auto* A = dynamic_cast<GpuSparseMatrix<Number>*>(system.matrix);
  if (!A) libmesh_error_msg("feasm_libmesh::assemble: system.matrix is not a GpuSparseMatrix; add it with "
                            "system.add_matrix<GpuSparseMatrix>(\"System Matrix\") before EquationSystems::init()");
  const unsigned n_ranks = A->n_processors();
  std::lock_guard<std::mutex> lock(A->mutex());
  const MeshBase& mesh = system.get_mesh();
  DofMap& dof_map = system.get_dof_map();
  feasm_ctx* ctx = A->ctx();
  const dof_id_type first = dof_map.first_dof(), n_local = dof_map.end_dof() - first;

  // ---- flatten the mesh and the dof map.  What was uploaded is remembered IN the matrix (GpuSparseMatrix::uploaded();
  // reset by clear() and update_sparsity_pattern(), so a System::reinit or a new matrix starts over).  The node
  // coordinates are read on every call, as the reference reads elem->point() in every assemble, and uploaded again when
  // they moved; connectivity and dof numbering are flattened again when the mesh, its size or the dof ranges changed.
  std::size_t n_elem = 0;
  for (const Elem* e : mesh.active_local_element_ptr_range()) { (void)e; ++n_elem; }
  if (!n_elem) libmesh_error_msg("no local elements");
  auto& up = A->uploaded();
  const bool reflatten = up.mesh != &mesh || up.n_elem != n_elem || up.n_dofs != dof_map.n_dofs() || up.n_local != n_local || up.first != first;
  {
    std::unordered_map<dof_id_type, uint32_t> local_node;
    std::vector<double> xyz;
    std::vector<uint32_t> conn, elem_dofs;
    std::vector<dof_id_type> di;
    int type = 0, nd = 0;
    // ghost rows: the dofs of local elements outside [first_dof, end_dof), ascending (empty on one rank)
    std::vector<dof_id_type> ghosts;
    if (n_ranks > 1) {
      if (reflatten) {
        for (const Elem* elem : mesh.active_local_element_ptr_range()) {
          dof_map.dof_indices(elem, di);
          for (dof_id_type d : di) if (d < first || d >= first + n_local) ghosts.push_back(d);
        }
        std::sort(ghosts.begin(), ghosts.end());
        ghosts.erase(std::unique(ghosts.begin(), ghosts.end()), ghosts.end());
        A->set_ghosts(ghosts);
      } else {
        ghosts = A->ghosts();
      }
    }
    for (const Elem* elem : mesh.active_local_element_ptr_range()) {
      const int t = feasm_type(elem->type());
      if (type && t != type) libmesh_error_msg("libfeasm assembles one element type per context");
      type = t;
      for (unsigned a = 0; a < elem->n_nodes(); ++a) {  // element-local node order, as FEMap::compute_map reads it
        const Node* node = elem->node_ptr(a);
        auto it = local_node.find(node->id());
        if (it == local_node.end()) {
          it = local_node.emplace(node->id(), (uint32_t)local_node.size()).first;
          xyz.push_back((*node)(0)); xyz.push_back((*node)(1)); xyz.push_back((*node)(2));
        }
        if (reflatten) conn.push_back(it->second);
      }
      if (!reflatten) continue;
      dof_map.dof_indices(elem, di);  // variable-major (src/base/dof_map.C:2280-2308)
      if (nd && (int)di.size() != nd) libmesh_error_msg("libfeasm needs the same number of dofs on every element");
      nd = (int)di.size();
      for (dof_id_type d : di) {
        if (n_ranks == 1 && (d < first || d >= first + n_local)) libmesh_error_msg("dof outside the local range on a one-rank run");
        elem_dofs.push_back(A->local_row(d));
      }
    }
    const uint64_t sum = checksum(xyz);
    if (reflatten) {
      A->check(feasm_set_mesh(ctx, type, (int64_t)n_elem, (int64_t)local_node.size(), xyz.data(), conn.data()));
      A->pattern_invalidated();
      if (n_ranks == 1) {
        A->check(feasm_set_dofs(ctx, nd / type, n_local, n_local, elem_dofs.data(), nullptr));
        A->send_pattern();
      } else {
        std::vector<uint32_t> gid(n_local + ghosts.size());
        for (dof_id_type r = 0; r < n_local; ++r) gid[r] = first + r;
        std::copy(ghosts.begin(), ghosts.end(), gid.begin() + n_local);
        A->check(feasm_set_dofs(ctx, nd / type, (int64_t)gid.size(), n_local, elem_dofs.data(), gid.data()));
        std::vector<uint64_t> first_dof(n_ranks + 1);
        for (unsigned p = 0; p < n_ranks; ++p) first_dof[p] = dof_map.first_dof(p);
        first_dof[n_ranks] = dof_map.end_dof(n_ranks - 1);
        A->ensure_communicator();
        A->check(feasm_build_pattern(ctx));
        A->check(feasm_build_exchange(ctx, first_dof.data()));  // collective
        A->check(feasm_build_halo(ctx, first_dof.data()));      // collective: the products / solves after the assembly
        A->adopt_device_pattern();
      }
      up.mesh = &mesh; up.n_elem = n_elem; up.n_dofs = dof_map.n_dofs(); up.n_local = n_local; up.first = first;
    } else if (sum != up.xyz_sum) {
      A->check(feasm_update_coordinates(ctx, xyz.data()));
    }
    up.xyz_sum = sum;
  }

  // ---- loaded sides: the `if (elem->neighbor_ptr(side) == nullptr) ... has_boundary_id(elem, side, id)` test of
  // systems_of_equations_ex6.C:243-251, evaluated once per call into an (element, side) list
  {
    std::vector<uint32_t> side_elems;
    std::vector<uint8_t> side_ids;
    if (side_load) {
      const BoundaryInfo& bi = mesh.get_boundary_info();
      uint32_t le = 0;  // index in active_local order = the element index handed to feasm_set_mesh
      for (const Elem* elem : mesh.active_local_element_ptr_range()) {
        for (auto side : elem->side_index_range())
          if (elem->neighbor_ptr(side) == nullptr && bi.has_boundary_id(elem, side, side_load->id)) {
            side_elems.push_back(le);
            side_ids.push_back((uint8_t)side);
          }
        ++le;
      }
    }
    A->check(feasm_set_side_loads(ctx, (int64_t)side_elems.size(), side_elems.data(), side_ids.data(),
                                  side_load ? side_load->g.data() : nullptr, side_load ? (int)side_load->g.size() : 0));
  }

  // ---- penalty sides: the `if (elem->neighbor_ptr(side) == nullptr)` loop of introduction_ex3.C:403-455 ----
  {
    std::vector<uint32_t> pen_elems;
    std::vector<uint8_t> pen_sides;
    if (side_penalty) {
      uint32_t le = 0;
      for (const Elem* elem : mesh.active_local_element_ptr_range()) {
        for (auto side : elem->side_index_range())
          if (elem->neighbor_ptr(side) == nullptr) {
            pen_elems.push_back(le);
            pen_sides.push_back((uint8_t)side);
          }
        ++le;
      }
    }
    A->check(feasm_set_side_penalty(ctx, (int64_t)pen_elems.size(), pen_elems.data(), pen_sides.data(),
                                    side_penalty ? side_penalty->penalty : 0., side_penalty ? side_penalty->kind : 0,
                                    side_penalty ? side_penalty->value : 0.));
  }

  // ---- Dirichlet constraints: DofMap::is_constrained_dof + primal constraint values (empty constraint rows) ----
  std::vector<uint32_t> con_rows;
  std::vector<double> con_vals;
  const DofConstraintValueMap& vals = dof_map.get_primal_constraint_values();
  auto constrained = [&](dof_id_type d, uint32_t row) {
    if (!dof_map.is_constrained_dof(d)) return;
    const auto it = vals.find(d);
    con_rows.push_back(row);
    con_vals.push_back(it == vals.end() ? 0. : it->second);
  };
  for (dof_id_type d = first; d < first + n_local; ++d) constrained(d, d - first);
  // the element-level constraint needs the ghost dofs of the local elements too (libMesh keeps their constraint rows on
  // this rank as well: DofMap::scatter_constraints, dof_map_constraints.C)
  for (std::size_t g = 0; g < A->ghosts().size(); ++g) constrained(A->ghosts()[g], (uint32_t)(n_local + g));
  // Hanging nodes and other non-empty constraint rows (DofMap::constraint_rows_begin, dof_map.h:1164): the same list with
  // its rows; elements that touch such a dof are widened on the device as heterogeneously_constrain_element_matrix_and_vector
  // widens them (feasm_set_constraint_rows).  Without any, the Dirichlet list goes the short way.
  bool any_row = false;
  for (auto it = dof_map.constraint_rows_begin(); it != dof_map.constraint_rows_end() && !any_row; ++it) any_row = !it->second.empty();
  if (any_row) {
    if (n_ranks > 1) libmesh_error_msg("feasm_libmesh::assemble: non-empty constraint rows on several ranks are not supported yet");
    std::vector<uint64_t> row_ptr(1, 0);
    std::vector<uint32_t> cols;
    std::vector<double> coefs;
    auto it = dof_map.constraint_rows_begin();
    for (std::size_t k = 0; k < con_rows.size(); ++k) {  // con_rows ascends with the dof id, as the constraint map does
      while (it != dof_map.constraint_rows_end() && it->first < con_rows[k] + first) ++it;
      if (it != dof_map.constraint_rows_end() && it->first == con_rows[k] + first)
        for (const auto& item : it->second) { cols.push_back(item.first - first); coefs.push_back(item.second); }
      row_ptr.push_back(cols.size());
    }
    A->check(feasm_set_constraint_rows(ctx, (int64_t)con_rows.size(), con_rows.data(), row_ptr.data(), cols.data(), coefs.data(), con_vals.data()));
  } else {
    A->check(feasm_set_constraints(ctx, (int64_t)con_rows.size(), con_rows.data(), con_vals.data()));
  }

  // ---- the hot path -----------------------------------------------------------------------------------------------
  if (n_ranks == 1) {
    A->check(feasm_assemble(ctx, &params));
  } else {
    A->check(feasm_assemble_begin(ctx, &params));  // element phase (interface elements first), ghost rows reduced and packed
    A->check(feasm_exchange_begin(ctx));           // one grouped ncclSend / ncclRecv per neighbour, on the communication stream
    A->check(feasm_assemble_end(ctx));             // the other elements and the owned rows, under the transfer
    A->check(feasm_exchange_end(ctx));             // the peers' contributions added in rank order
  }
  A->invalidate_host_copy();
  if (auto* grhs = dynamic_cast<GpuNumericVector<Number>*>(system.rhs)) {
    // route (2) of SURVEY 8b: the RHS never leaves the device
    A->check(feasm_synchronize(ctx));  // throws on a degenerate map like FEMap does
    grhs->assign_device(feasm_device_rhs(ctx), n_local);
  } else {
    // route (1): any NumericVector, through NumericVector::operator=(const std::vector<T>&), numeric_vector.h:295
    std::vector<Number> rhs(n_local + A->ghosts().size());
    A->check(feasm_get_values(ctx, nullptr, rhs.data()));  // synchronises; throws on a degenerate map like FEMap does
    rhs.resize(n_local);                                   // the owned rows; ghost rows were added into their owners
    *system.rhs = rhs;
  }
  system.rhs->close();
}

void assemble_poisson_ex4(EquationSystems& es, const std::string& name) {
  feasm_params p{};
  p.problem = FEASM_POISSON; p.quadrature_order = 5; p.forcing = FEASM_FORCING_EX4_FD; p.constrain = FEASM_CONSTRAIN_HETEROGENEOUS;
  assemble(es, name, p);
}
void assemble_laplace(EquationSystems& es, const std::string& name) {
  feasm_params p{};
  p.problem = FEASM_POISSON; p.quadrature_order = 5; p.forcing = FEASM_FORCING_NONE; p.constrain = FEASM_CONSTRAIN_HETEROGENEOUS;
  assemble(es, name, p);
}
void assemble_elasticity_ex6(EquationSystems& es, const std::string& name) {
  feasm_params p{};
  p.problem = FEASM_ELASTICITY; p.constrain = FEASM_CONSTRAIN_HOMOGENEOUS;
  const SideLoad traction{2 /* BOUNDARY_ID_MAX_X */, {0., 0., -1.}};
  assemble(es, name, p, &traction);
}
}  // namespace feasm_libmesh

// explicit instantiation: every virtual of the adaptor is compiled in this translation unit
template class libMesh::GpuSparseMatrix<libMesh::Number>;
template class libMesh::GpuNumericVector<libMesh::Number>;
