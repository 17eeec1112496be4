// shim_test_ranks.C — the libMesh adaptor on SEVERAL ranks (one process per GPU): N processes of this program, started by
// tests/test_gpu_shim.py with FEASM_MOCK_RANK / FEASM_MOCK_SIZE / FEASM_MOCK_DIR (the mock communicator's stand-in for MPI),
// each with its own GPU.  Every process holds the whole (replicated) mock mesh, partitioned by element id slabs
// (LinearPartitioner), with libMesh's per-rank contiguous dof numbering (feasm_distribute_dofs).  The hook assembles the
// local elements, NCCL carries the ghost rows; the owned rows of every rank are compared with a one-context assembly of
// the whole mesh through the C ABI on the same GPU.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <set>

#include "gpu_assemble.h"

using namespace libMesh;

#define REQUIRE(c) do { if (!(c)) { std::printf("rank %u FAILED %s:%d: %s\n", (unsigned)comm.rank(), __FILE__, __LINE__, #c); return 1; } } while (0)
#define CK(call) do { int rc__ = (call); if (rc__) { std::printf("C ABI error %d at %s:%d\n", rc__, __FILE__, __LINE__); return 1; } } while (0)

static int run(const Parallel::Communicator& comm, int n_vars) {
  const unsigned rank = comm.rank(), world = comm.size();
  const int nx = 4, ny = 3, nz = 2, npe = 27, nd = npe * n_vars;
  const double bounds_p[6] = {-1, 1, -1, 1, -1, 1}, bounds_e[6] = {0, 1, 0, 0.3, 0, 0.1};
  const double* bounds = n_vars == 1 ? bounds_p : bounds_e;
  int64_t n_elem = 0, n_nodes = 0;
  CK(feasm_cube_sizes(FEASM_HEX27, nx, ny, nz, &n_elem, &n_nodes));
  std::vector<double> xyz(3 * n_nodes);
  std::vector<uint32_t> conn(npe * n_elem), base(n_nodes), first(world + 1);
  std::vector<uint8_t> mask(n_nodes);
  std::vector<int32_t> npid(n_nodes), epid(n_elem);
  CK(feasm_build_cube(FEASM_HEX27, nx, ny, nz, bounds, xyz.data(), conn.data(), mask.data()));
  const int64_t blk = std::max<int64_t>(n_elem / world, 1);
  for (int64_t e = 0; e < n_elem; ++e) epid[e] = e / blk < (int64_t)world ? (int32_t)(e / blk) : 0;  // linear_partitioner.C:60-75
  const int64_t n_dofs = feasm_distribute_dofs(npe, n_elem, n_nodes, conn.data(), n_vars, (int)world, epid.data(), base.data(), first.data(), npid.data());
  REQUIRE(n_dofs == n_nodes * n_vars);

  std::vector<Node> nodes(n_nodes);
  for (int64_t i = 0; i < n_nodes; ++i) { nodes[i]._id = (dof_id_type)i; for (int d = 0; d < 3; ++d) nodes[i].c[d] = xyz[3 * i + d]; }
  std::vector<Elem> elems(n_elem);
  MeshBase mesh;
  DofMap dof_map;
  dof_map._n_dofs = (dof_id_type)n_dofs; dof_map._nd = (unsigned)nd; dof_map._rank = rank;
  dof_map._first.assign(first.begin(), first.end());
  std::vector<uint32_t> load_elems;
  std::vector<uint8_t> load_sides;
  for (int64_t e = 0; e < n_elem; ++e) {
    elems[e]._type = HEX27; elems[e]._id = (dof_id_type)e; elems[e]._pid = (processor_id_type)epid[e];
    for (int a = 0; a < npe; ++a) elems[e]._nodes.push_back(&nodes[conn[npe * e + a]]);
    for (int v = 0; v < n_vars; ++v)
      for (int a = 0; a < npe; ++a) dof_map._elem_dofs.push_back(base[conn[npe * e + a]] + v);  // variable-major, dof_map.C:2280-2308
    if ((unsigned)epid[e] == rank) mesh._local.push_back(&elems[e]);                              // active_local_element_ptr_range
    const int i = (int)(e % nx), j = (int)((e / nx) % ny), k = (int)(e / (nx * ny));
    elems[e]._neighbors = {k > 0 ? &elems[e - nx * ny] : nullptr, j > 0 ? &elems[e - nx] : nullptr, i < nx - 1 ? &elems[e + 1] : nullptr,
                           j < ny - 1 ? &elems[e + nx] : nullptr, i > 0 ? &elems[e - 1] : nullptr, k < nz - 1 ? &elems[e + nx * ny] : nullptr};
    if (i == nx - 1) {
      mesh._bi._sides.insert(std::make_tuple((const Elem*)&elems[e], (unsigned short)2, (boundary_id_type)2));
      load_elems.push_back((uint32_t)e); load_sides.push_back(2);
    }
  }
  const double hp = 1.5707963267948966;
  for (int64_t i = 0; i < n_nodes; ++i) {
    if (n_vars == 1 && mask[i]) dof_map._vals[base[i]] = std::cos(hp * xyz[3 * i]) * std::sin(hp * xyz[3 * i + 1]) * std::cos(hp * xyz[3 * i + 2]);
    if (n_vars == 3 && (mask[i] & 0x10))
      for (int v = 0; v < 3; ++v) dof_map._vals[base[i] + v] = 0.;  // clamped at x-min (boundary id 4), systems_of_equations_ex6.C:330-345
  }
  // the Graph libMesh hands to update_sparsity_pattern on this rank: its OWN rows, already united with what the other ranks'
  // elements contribute (SparsityPattern::Build::parallel_sync) — here computed from all elements of the replicated mesh
  const dof_id_type f0 = first[rank], f1 = first[rank + 1], n_local = f1 - f0;
  std::vector<std::set<dof_id_type>> rows(n_dofs);
  for (int64_t e = 0; e < n_elem; ++e)
    for (int a = 0; a < nd; ++a)
      for (int b = 0; b < nd; ++b) rows[dof_map._elem_dofs[nd * e + a]].insert(dof_map._elem_dofs[nd * e + b]);
  SparsityPattern::Graph graph(n_local);
  for (dof_id_type r = 0; r < n_local; ++r) graph[r].assign(rows[f0 + r].begin(), rows[f0 + r].end());

  GpuSparseMatrix<Number> A(comm);
  NumericVector<Number> rhs(comm);
  LinearImplicitSystem system;
  system._mesh = &mesh; system._dof = &dof_map; system.matrix = &A; system.rhs = &rhs;
  EquationSystems es;
  es._sys = &system;
  A.attach_dof_map(dof_map);
  A.update_sparsity_pattern(graph);
  A.init();
  REQUIRE(A.m() == (numeric_index_type)n_dofs && A.row_start() == f0 && A.row_stop() == f1);
  auto hook = n_vars == 1 ? feasm_libmesh::assemble_poisson_ex4 : feasm_libmesh::assemble_elasticity_ex6;
  A.zero();
  hook(es, "sys");   // the merged device pattern is checked against `graph` inside (adopt_device_pattern)
  A.close();

  // ---- the whole mesh on one context, straight through the C ABI (same GPU) ----------------------------------------
  feasm_ctx* ctx = nullptr;
  CK(feasm_create((int)rank, &ctx));
  CK(feasm_set_mesh(ctx, FEASM_HEX27, n_elem, n_nodes, xyz.data(), conn.data()));
  CK(feasm_set_dofs(ctx, n_vars, n_dofs, n_dofs, dof_map._elem_dofs.data(), nullptr));
  std::vector<uint32_t> crow;
  std::vector<double> cval;
  for (const auto& kv : dof_map._vals) { crow.push_back(kv.first); cval.push_back(kv.second); }
  CK(feasm_set_constraints(ctx, (int64_t)crow.size(), crow.data(), cval.data()));
  CK(feasm_build_pattern(ctx));
  feasm_params p{};
  if (n_vars == 1) { p.problem = FEASM_POISSON; p.quadrature_order = 5; p.forcing = FEASM_FORCING_EX4_FD; p.constrain = FEASM_CONSTRAIN_HETEROGENEOUS; }
  else {
    p.problem = FEASM_ELASTICITY; p.constrain = FEASM_CONSTRAIN_HOMOGENEOUS;
    const double g[3] = {0., 0., -1.};
    CK(feasm_set_side_loads(ctx, (int64_t)load_elems.size(), load_elems.data(), load_sides.data(), g, 3));
  }
  CK(feasm_assemble(ctx, &p));
  std::vector<uint64_t> rp(n_dofs + 1);
  std::vector<uint32_t> ci((std::size_t)feasm_nnz(ctx));
  CK(feasm_get_pattern(ctx, rp.data(), ci.data()));
  std::vector<double> v(ci.size()), b(n_dofs);
  CK(feasm_get_values(ctx, v.data(), b.data()));

  // ---- owned rows: same columns; same bits where no other rank contributes, 1e-12 where the peers' partial sums are added
  std::vector<char> shared_row(n_dofs, 0);   // rows with contributions from more than one rank
  {
    std::vector<int> owner_seen(n_dofs, -1);
    for (int64_t e = 0; e < n_elem; ++e)
      for (int a = 0; a < nd; ++a) {
        const dof_id_type d = dof_map._elem_dofs[nd * e + a];
        if (owner_seen[d] >= 0 && owner_seen[d] != epid[e]) shared_row[d] = 1;
        owner_seen[d] = epid[e];
      }
  }
  REQUIRE(rhs._v.size() == n_local);
  std::size_t n_shared = 0, n_bitwise = 0;
  double worst = 0., linf = 0.;
  for (dof_id_type r = f0; r < f1; ++r) {
    std::vector<numeric_index_type> idx;
    std::vector<Number> val;
    A.get_row(r, idx, val);
    REQUIRE(idx.size() == rp[r + 1] - rp[r] && std::equal(idx.begin(), idx.end(), ci.begin() + rp[r]));
    double rowmax = 0., rowsum = 0.;
    for (std::size_t k = 0; k < idx.size(); ++k) rowmax = std::max(rowmax, std::fabs(v[rp[r] + k]));
    for (std::size_t k = 0; k < idx.size(); ++k) {
      const double ref = v[rp[r] + k];
      rowsum += std::fabs(val[k]);
      if (shared_row[r]) worst = std::max(worst, std::fabs(val[k] - ref) / std::max(std::max(std::fabs(ref), std::fabs(val[k])), 1e-2 * rowmax));
      else REQUIRE(val[k] == ref);
    }
    linf = std::max(linf, rowsum);
    if (shared_row[r]) { ++n_shared; REQUIRE(std::fabs(rhs._v[r - f0] - b[r]) <= 1e-12 * std::max(std::fabs(b[r]), 1e-300) || std::fabs(rhs._v[r - f0] - b[r]) < 1e-15); }
    else { ++n_bitwise; REQUIRE(rhs._v[r - f0] == b[r]); }
    REQUIRE(A(r, r) == val[std::lower_bound(idx.begin(), idx.end(), r) - idx.begin()]);
  }
  REQUIRE(worst <= 1e-12);
  REQUIRE(world == 1 || rank != 0 || n_shared > 0);  // shared rows belong to the lowest touching rank (node.C:99-102)
  // linfty_norm is a global quantity (max over the ranks' rows)
  double linf_all = 0.;
  for (int64_t r = 0; r < n_dofs; ++r) { double s = 0.; for (uint64_t k = rp[r]; k < rp[r + 1]; ++k) s += std::fabs(v[k]); linf_all = std::max(linf_all, s); }
  REQUIRE(std::fabs(A.linfty_norm() - linf_all) <= 1e-12 * linf_all);

  // ---- a second assembly (nothing is flattened again) gives the same bits; so does a fresh matrix --------------------
  const std::vector<Number> first_values = A.host_values();
  hook(es, "sys");
  A.close();
  REQUIRE(A.host_values() == first_values);

  // ---- route (2): the right-hand side stays on the device in a distributed GpuNumericVector -------------------------
  {
    GpuNumericVector<Number> grhs(comm, (numeric_index_type)n_dofs, n_local);
    REQUIRE(grhs.size() == (numeric_index_type)n_dofs && grhs.local_size() == n_local && grhs.first_local_index() == f0 && grhs.last_local_index() == f1);
    system.rhs = &grhs;
    hook(es, "sys");
    A.close();
    system.rhs = &rhs;
    std::vector<Number> whole;
    grhs.localize(whole);                       // gathers: every rank holds the whole vector
    REQUIRE(whole.size() == (std::size_t)n_dofs);
    double bmax = 0., s = 0., l1 = 0., dot = 0.;
    for (int64_t i = 0; i < n_dofs; ++i) { bmax = std::max(bmax, std::fabs(b[i])); s += b[i]; l1 += std::fabs(b[i]); dot += b[i] * b[i]; }
    for (int64_t i = 0; i < n_dofs; ++i) REQUIRE(std::fabs(whole[i] - b[i]) <= 1e-12 * bmax);
    for (dof_id_type r = f0; r < f1; ++r) REQUIRE(whole[r] == rhs._v[r - f0]);
    auto near = [](double x, double y) { return std::fabs(x - y) <= 1e-11 * (std::fabs(y) + 1e-300) || std::fabs(x - y) < 1e-13; };
    REQUIRE(near(grhs.sum(), s) && near(grhs.l1_norm(), l1) && near(grhs.dot(grhs), dot) && near(grhs.l2_norm(), std::sqrt(dot)));
    REQUIRE(std::fabs(grhs.linfty_norm() - bmax) <= 1e-12 * bmax);
    REQUIRE(n_local == 0 || grhs(f0) == whole[f0]);
    auto g2 = grhs.clone();
    g2->scale(2.0);
    g2->add(-1.0, grhs);
    REQUIRE(near(g2->dot(grhs), dot));
    GpuNumericVector<Number> g3(comm, (numeric_index_type)n_dofs, n_local);
    g3 = whole;                                 // from the whole vector: the local part is kept
    REQUIRE(near(g3.dot(grhs), dot));
    // the distributed matrix used on the device: y += A x with x's ghost entries fetched from their owners (feasm_build_halo
    // was set up by the hook), against the product with the one-context matrix
    {
      std::vector<Number> xw(n_dofs), yref(n_dofs, 0.), yw;
      for (int64_t i = 0; i < n_dofs; ++i) xw[i] = std::cos(0.37 * (double)i) + 0.1;
      double ymax = 0.;
      for (int64_t r = 0; r < n_dofs; ++r) { for (uint64_t k = rp[r]; k < rp[r + 1]; ++k) yref[r] += v[k] * xw[ci[k]]; ymax = std::max(ymax, std::fabs(yref[r])); }
      GpuNumericVector<Number> gx(comm, (numeric_index_type)n_dofs, n_local), gy(comm, (numeric_index_type)n_dofs, n_local);
      gx = xw;
      gy.zero();
      gy.add_vector(gx, A);
      gy.localize(yw);
      for (int64_t i = 0; i < n_dofs; ++i) REQUIRE(std::fabs(yw[i] - yref[i]) <= 1e-13 * ymax);
      // LinearSolver::solve on the distributed system: all ranks agree on the iteration count; solution = the one-context solve
      std::vector<Number> xs, xone(n_dofs);
      const auto res = A.solve_cg(xs, 1e-11, 4000);
      int its1 = 0;
      double rel1 = 0.;
      const int rc1 = feasm_cg_solve(ctx, 1e-11, 4000, nullptr, xone.data(), &its1, &rel1);
      REQUIRE(rc1 == FEASM_OK && res.second <= 1e-11 && xs.size() == n_local);
      unsigned its_max = res.first, its_min = res.first;
      comm.max(its_max); comm.min(its_min);
      REQUIRE(its_max == its_min);
      double xmax = 0.;
      for (int64_t i = 0; i < n_dofs; ++i) xmax = std::max(xmax, std::fabs(xone[i]));
      for (dof_id_type r = f0; r < f1; ++r) REQUIRE(std::fabs(xs[r - f0] - xone[r]) <= 1e-8 * xmax);
    }
    if (world > 1 && n_local < (dof_id_type)n_dofs) {
      bool threw = false;
      try { (void)grhs(f0 > 0 ? 0 : f1); } catch (const LogicError&) { threw = true; }
      REQUIRE(threw);                           // an entry of another rank is not readable here
    }
  }
  std::printf("rank %u of %u, %d variable(s): %zu local elements, rows [%u, %u) + %zu ghost rows; %zu rows bitwise equal to the one-context assembly, "
              "%zu rows with peer contributions within %.1e\n", rank, world, n_vars, mesh._local.size(), f0, f1, A.ghosts().size(), n_bitwise, n_shared, worst);
  comm.barrier();
  A.clear();
  feasm_destroy(ctx);
  return 0;
}

int main() {
  Parallel::Communicator comm;
  try {
    if (run(comm, 1)) return 1;
    if (run(comm, 3)) return 1;
  } catch (const std::exception& e) {
    std::printf("rank %u: exception: %s\n", (unsigned)comm.rank(), e.what());
    return 1;
  }
  std::printf("rank %u: SHIM RANKS OK\n", (unsigned)comm.rank());
  return 0;
}
