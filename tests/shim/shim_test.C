// shim_test.C — runs the libMesh adaptor (GpuSparseMatrix + the batched assemble hook) end to end against the mock
// classes of libmesh_mock.h on a real B200, and checks it against a second, direct drive of the C ABI with the same
// inputs: what travels through the adaptor (mesh flattening, dof_indices, Graph hand-over, constraints, RHS hand-back,
// the SparseMatrix virtuals) must change nothing.  Built by shim/Makefile, run by tests/test_gpu_shim.py.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <set>
#include <sstream>

#include "compare_mode.h"
#include "gpu_assemble.h"

using namespace libMesh;

#define REQUIRE(c) do { if (!(c)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)
#define CK(call) do { int rc__ = (call); if (rc__) { std::printf("C ABI error %d at %s:%d\n", rc__, __FILE__, __LINE__); return 1; } } while (0)

int main() {
  // ---- a 3x2x2 HEX27 mesh on [-1,1]^3 from the host generators of the C ABI (MeshTools::Generation::build_cube) ----
  const int nx = 3, ny = 2, nz = 2;
  const double bounds[6] = {-1, 1, -1, 1, -1, 1};
  int64_t n_elem = 0, n_nodes = 0;
  CK(feasm_cube_sizes(FEASM_HEX27, nx, ny, nz, &n_elem, &n_nodes));
  std::vector<double> xyz(3 * n_nodes);
  std::vector<uint32_t> conn(27 * n_elem), base(n_nodes), first(2);
  std::vector<uint8_t> mask(n_nodes);
  std::vector<int32_t> npid(n_nodes);
  CK(feasm_build_cube(FEASM_HEX27, nx, ny, nz, bounds, xyz.data(), conn.data(), mask.data()));
  const int64_t n_dofs = feasm_distribute_dofs(27, n_elem, n_nodes, conn.data(), 1, 1, nullptr, base.data(), first.data(), npid.data());
  REQUIRE(n_dofs == n_nodes);

  std::vector<Node> nodes(n_nodes);
  for (int64_t i = 0; i < n_nodes; ++i) { nodes[i]._id = (dof_id_type)i; for (int d = 0; d < 3; ++d) nodes[i].c[d] = xyz[3 * i + d]; }
  std::vector<Elem> elems(n_elem);
  MeshBase mesh;
  std::vector<uint32_t> load_elems;  // elements whose side 2 (x-max) lies on the box face: boundary id 2
  std::vector<uint8_t> load_sides;
  DofMap dof_map;
  dof_map._n_dofs = (dof_id_type)n_dofs; dof_map._nd = 27;
  for (int64_t e = 0; e < n_elem; ++e) {
    elems[e]._type = HEX27; elems[e]._id = (dof_id_type)e; elems[e]._pid = 0;
    for (int a = 0; a < 27; ++a) {
      elems[e]._nodes.push_back(&nodes[conn[27 * e + a]]);
      dof_map._elem_dofs.push_back(base[conn[27 * e + a]]);
    }
    mesh._local.push_back(&elems[e]);
    // neighbours in build_cube's k-j-i element order; sides 0..5 = z-min, y-min, x-max, y-max, x-min, z-max
    const int i = (int)(e % nx), j = (int)((e / nx) % ny), k = (int)(e / (nx * ny));
    elems[e]._neighbors = {k > 0 ? &elems[e - nx * ny] : nullptr, j > 0 ? &elems[e - nx] : nullptr, i < nx - 1 ? &elems[e + 1] : nullptr,
                           j < ny - 1 ? &elems[e + nx] : nullptr, i > 0 ? &elems[e - 1] : nullptr, k < nz - 1 ? &elems[e + nx * ny] : nullptr};
    if (i == nx - 1) {
      mesh._bi._sides.insert(std::make_tuple((const Elem*)&elems[e], (unsigned short)2, (boundary_id_type)2));
      load_elems.push_back((uint32_t)e); load_sides.push_back(2);
    }
  }
  // Dirichlet dofs on the whole boundary with the exact solution of ex4 (DirichletBoundary, introduction_ex4.C:242-274)
  const double hp = 1.5707963267948966;
  for (int64_t i = 0; i < n_nodes; ++i)
    if (mask[i]) dof_map._vals[base[i]] = std::cos(hp * xyz[3 * i]) * std::sin(hp * xyz[3 * i + 1]) * std::cos(hp * xyz[3 * i + 2]);
  // SparsityPattern::Build: sorted unique union of the element dof sets
  std::vector<std::set<dof_id_type>> rows(n_dofs);
  for (int64_t e = 0; e < n_elem; ++e)
    for (int a = 0; a < 27; ++a)
      for (int b = 0; b < 27; ++b) rows[dof_map._elem_dofs[27 * e + a]].insert(dof_map._elem_dofs[27 * e + b]);
  SparsityPattern::Graph graph(n_dofs);
  for (int64_t r = 0; r < n_dofs; ++r) graph[r].assign(rows[r].begin(), rows[r].end());

  // ---- the libMesh side: what EquationSystems::init() and System::solve() do with the adaptor ----------------------
  Parallel::Communicator comm;
  GpuSparseMatrix<Number> A(comm);
  NumericVector<Number> rhs(comm);
  LinearImplicitSystem system;
  system._mesh = &mesh; system._dof = &dof_map; system.matrix = &A; system.rhs = &rhs;
  EquationSystems es;
  es._sys = &system;
  A.attach_dof_map(dof_map);             // DofMap::attach_matrix, dof_map.C:240-287
  REQUIRE(A.need_full_sparsity_pattern());
  A.update_sparsity_pattern(graph);
  A.init();
  REQUIRE(A.initialized() && A.m() == (numeric_index_type)n_dofs && A.row_stop() == (numeric_index_type)n_dofs);
  A.zero();                              // ImplicitSystem::assemble, implicit_system.C:101-105
  feasm_libmesh::assemble_poisson_ex4(es, "Poisson");
  A.close();
  REQUIRE(A.closed());

  // ---- the same problem straight through the C ABI ---------------------------------------------------------------
  feasm_ctx* ctx = nullptr;
  CK(feasm_create(0, &ctx));
  CK(feasm_set_mesh(ctx, FEASM_HEX27, n_elem, n_nodes, xyz.data(), conn.data()));
  CK(feasm_set_dofs(ctx, 1, n_dofs, n_dofs, dof_map._elem_dofs.data(), nullptr));
  std::vector<uint32_t> crow;
  std::vector<double> cval;
  for (const auto& kv : dof_map._vals) { crow.push_back(kv.first); cval.push_back(kv.second); }
  CK(feasm_set_constraints(ctx, (int64_t)crow.size(), crow.data(), cval.data()));
  CK(feasm_set_pattern(ctx, A.row_ptr().data(), A.col_idx().data()));
  feasm_params p{};
  p.problem = FEASM_POISSON; p.quadrature_order = 5; p.forcing = FEASM_FORCING_EX4_FD; p.constrain = FEASM_CONSTRAIN_HETEROGENEOUS;
  CK(feasm_assemble(ctx, &p));
  std::vector<double> v(A.col_idx().size()), b(n_dofs);
  CK(feasm_get_values(ctx, v.data(), b.data()));

  // the adaptor changed nothing: same bits in the matrix (operator(), get_row, host_values) and in the RHS
  REQUIRE(rhs._v.size() == (std::size_t)n_dofs);
  for (int64_t r = 0; r < n_dofs; ++r) REQUIRE(rhs._v[r] == b[r]);
  REQUIRE(A.host_values() == v);
  double l1 = 0., linf = 0.;
  std::vector<double> colsum(n_dofs, 0.);
  for (int64_t r = 0; r < n_dofs; ++r) {
    std::vector<numeric_index_type> idx;
    std::vector<Number> val;
    A.get_row((numeric_index_type)r, idx, val);
    REQUIRE(idx == graph[r]);                                   // ascending columns (sparse_matrix_test.h:88-149)
    double rowsum = 0.;
    for (std::size_t k = 0; k < idx.size(); ++k) {
      REQUIRE(val[k] == v[A.row_ptr()[r] + k]);
      REQUIRE(A((numeric_index_type)r, idx[k]) == val[k]);
      rowsum += std::fabs(val[k]); colsum[idx[k]] += std::fabs(val[k]);
    }
    linf = std::max(linf, rowsum);
  }
  for (double c : colsum) l1 = std::max(l1, c);
  REQUIRE(std::fabs(A.l1_norm() - l1) <= 1e-12 * l1 && std::fabs(A.linfty_norm() - linf) <= 1e-12 * linf);
  REQUIRE(A(0, (numeric_index_type)n_dofs - 1) == 0.);          // outside the pattern reads as zero
  // a constrained dof: diagonal = number of touching elements, rhs = that count x g
  {
    const dof_id_type d = dof_map._vals.begin()->first;
    int touching = 0;
    for (dof_id_type x : dof_map._elem_dofs) touching += x == d;
    REQUIRE(A(d, d) == (double)touching && std::fabs(rhs._v[d] - touching * dof_map._vals[d]) <= 1e-14);
  }
  NumericVector<Number> diag(comm);
  A.get_diagonal(diag);
  for (int64_t r = 0; r < n_dofs; ++r) REQUIRE(diag._v[r] == A((numeric_index_type)r, (numeric_index_type)r));

  // ---- the single-entry virtuals ---------------------------------------------------------------------------------
  const numeric_index_type i0 = graph[5][0], j0 = graph[5].size() > 1 ? graph[5][1] : graph[5][0];
  const double before = A(5, j0);
  A.add(5, j0, 0.25);
  A.close();
  REQUIRE(A(5, j0) == before + 0.25);
  (void)i0;
  DenseMatrix<Number> dm;
  dm._m = dm._n = 2; dm._val = {1., 2., 3., 4.};
  std::vector<numeric_index_type> di = {graph[7][0], graph[7][1]};
  const double a00 = A(di[0], di[0]), a01 = A(di[0], di[1]), a10 = A(di[1], di[0]), a11 = A(di[1], di[1]);
  A.add_matrix(dm, di);
  A.close();
  REQUIRE(A(di[0], di[0]) == a00 + 1. && A(di[0], di[1]) == a01 + 2. && A(di[1], di[0]) == a10 + 3. && A(di[1], di[1]) == a11 + 4.);
  bool threw = false;
  try { A.add(0, (numeric_index_type)n_dofs - 1, 1.0); A.close(); } catch (const LogicError&) { threw = true; }
  REQUIRE(threw);                                               // non-zero outside the pattern is an error ...
  A.add(0, (numeric_index_type)n_dofs - 1, 0.0);                // ... zero is allowed (sparse_matrix.h:300-303)
  A.close();
  std::ostringstream os;
  A.print_personal(os);
  REQUIRE(!os.str().empty());

  // ---- after assembly: product and CG solve on the device matrix -----------------------------------------------
  A.zero();
  feasm_libmesh::assemble_poisson_ex4(es, "Poisson");
  A.close();
  std::vector<Number> x(n_dofs, 1.0), y;
  A.multiply(y, x);
  for (int64_t r = 0; r < n_dofs; ++r) {
    double s = 0.;
    for (uint64_t k = A.row_ptr()[r]; k < A.row_ptr()[r + 1]; ++k) s += v[k];
    REQUIRE(std::fabs(y[r] - s) <= 1e-12 * (std::fabs(s) + 1.));
  }
  std::vector<Number> u;
  const auto res = A.solve_cg(u, 1e-12, 2000);
  REQUIRE(res.second <= 1e-12 && res.first > 0);
  double err = 0.;
  for (int64_t i = 0; i < n_nodes; ++i)
    err = std::max(err, std::fabs(u[base[i]] - std::cos(hp * xyz[3 * i]) * std::sin(hp * xyz[3 * i + 1]) * std::cos(hp * xyz[3 * i + 2])));
  REQUIRE(err < 5e-3);                                          // coarse Q2 mesh: discretisation error only

  // ---- GpuNumericVector: route (2), the vector algebra against the host mock vector, a CG written on the interface ----
  {
    GpuNumericVector<Number> grhs(comm, (numeric_index_type)n_dofs);
    system.rhs = &grhs;
    A.zero();
    feasm_libmesh::assemble_poisson_ex4(es, "Poisson");   // RHS handed over on the device
    A.close();
    std::vector<Number> hb;
    grhs.localize(hb);
    REQUIRE(hb == b);
    system.rhs = &rhs;

    std::vector<Number> ha(n_dofs), hc(n_dofs);
    for (int64_t i = 0; i < n_dofs; ++i) { ha[i] = std::sin(0.37 * i) + 1.5; hc[i] = std::cos(0.11 * i) * 2.0 - 0.3; }
    NumericVector<Number> ra(comm), rc(comm);
    ra = ha; rc = hc;
    GpuNumericVector<Number> ga(comm, (numeric_index_type)n_dofs), gc(comm, (numeric_index_type)n_dofs);
    ga = ha; gc = hc;
    auto close_to = [](double x, double y) { return std::fabs(x - y) <= 1e-12 * (std::fabs(y) + 1.); };
    auto same = [&](const GpuNumericVector<Number>& g, const NumericVector<Number>& r) {
      std::vector<Number> h;
      g.localize(h);
      for (std::size_t i = 0; i < h.size(); ++i) if (!close_to(h[i], r._v[i])) return false;
      return true;
    };
    REQUIRE(ga.size() == (numeric_index_type)n_dofs && ga.local_size() == ga.size() && ga.initialized());
    REQUIRE(close_to(ga.dot(gc), ra.dot(rc)) && close_to(ga.l2_norm(), ra.l2_norm()) && close_to(ga.l1_norm(), ra.l1_norm()));
    REQUIRE(ga.linfty_norm() == ra.linfty_norm() && gc.min() == rc.min() && gc.max() == rc.max() && close_to(gc.sum(), rc.sum()));
    REQUIRE(close_to(ga.dot(rc), ra.dot(rc)));                       // a host vector as the second operand
    ga.add(2.5, gc); ra.add(2.5, rc); REQUIRE(same(ga, ra));
    ga += gc; ra += rc; ga -= gc; ra -= rc; REQUIRE(same(ga, ra));
    ga.scale(-0.75); ra.scale(-0.75); ga.add(0.125); ra.add(0.125); REQUIRE(same(ga, ra));
    ga.abs(); ra.abs(); REQUIRE(same(ga, ra));
    ga *= gc; ra *= rc; REQUIRE(same(ga, ra));
    gc.add(10.); rc.add(10.); ga /= gc; ra /= rc; REQUIRE(same(ga, ra));
    gc.reciprocal(); rc.reciprocal(); REQUIRE(same(gc, rc));
    ga.pointwise_mult(gc, gc); ra.pointwise_mult(rc, rc); REQUIRE(same(ga, ra));
    ga.pointwise_divide(gc, ga); ra.pointwise_divide(rc, ra); REQUIRE(same(ga, ra));
    ga.set(3, 42.); ga.add(3, 0.5); ga.close(); REQUIRE(ga(3) == 42.5);
    auto gcl = ga.clone();
    REQUIRE(gcl->dot(*gcl) == ga.dot(ga));
    auto gz = ga.zero_clone();
    REQUIRE(gz->linfty_norm() == 0. && gz->size() == ga.size());
    ga = 7.0; REQUIRE(ga.min() == 7.0 && ga.max() == 7.0);
    ga = rc; REQUIRE(same(ga, rc));                                   // assignment from a host NumericVector
    std::vector<Number> sub;
    ga.localize(sub, std::vector<numeric_index_type>{5, 0, 9});
    REQUIRE(sub.size() == 3 && sub[1] == ga(0));

    // y = A x through the vector interface (what SparseMatrix::vector_mult resolves to), then CG on the interface
    GpuNumericVector<Number> gx(comm, (numeric_index_type)n_dofs), gy(comm, (numeric_index_type)n_dofs);
    gx = 1.0;
    gy.add_vector(gx, A);
    std::vector<Number> hy;
    gy.localize(hy);
    for (int64_t r = 0; r < n_dofs; ++r) REQUIRE(close_to(hy[r], y[r]));
    GpuNumericVector<Number> xs(comm, (numeric_index_type)n_dofs), rr(comm, (numeric_index_type)n_dofs), pp(comm, (numeric_index_type)n_dofs),
        qq(comm, (numeric_index_type)n_dofs);
    rr = grhs; pp = rr;
    double rho = rr.dot(rr);
    const double stop = 1e-24 * rho;
    int its = 0;
    while (rho > stop && its < 2000) {
      qq.zero(); qq.add_vector(pp, A);
      const double alpha = rho / pp.dot(qq);
      xs.add(alpha, pp); rr.add(-alpha, qq);
      const double rho_new = rr.dot(rr);
      pp.scale(rho_new / rho); pp.add(rr);
      rho = rho_new; ++its;
    }
    std::vector<Number> hx;
    xs.localize(hx);
    double diff = 0.;
    for (int64_t i = 0; i < n_dofs; ++i) diff = std::max(diff, std::fabs(hx[i] - u[i]));
    REQUIRE(its > 0 && its < 2000 && diff < 1e-9);
  }

  // ---- a side load through the adaptor (BoundaryInfo + neighbor_ptr -> feasm_set_side_loads) vs the C ABI directly ----
  {
    const feasm_libmesh::SideLoad flux{2, {1.7}};
    A.zero();
    feasm_libmesh::assemble(es, "Poisson", p, &flux);
    A.close();
    const double g1 = 1.7;
    CK(feasm_set_side_loads(ctx, (int64_t)load_elems.size(), load_elems.data(), load_sides.data(), &g1, 1));
    CK(feasm_assemble(ctx, &p));
    std::vector<double> b2(n_dofs);
    CK(feasm_get_values(ctx, nullptr, b2.data()));
    for (int64_t r = 0; r < n_dofs; ++r) REQUIRE(rhs._v[r] == b2[r]);
    // every node of the loaded face is a Dirichlet node here, so the flux shows in the element vector before the
    // constraint (where the reference adds it, ex6 :240-257), not in the assembled rows
    const int64_t le = load_elems[0];
    std::vector<double> ke(27 * 27), fe1(27), fe0(27);
    CK(feasm_element_matrices(ctx, &p, le, le + 1, 0, ke.data(), fe1.data()));
    CK(feasm_set_side_loads(ctx, 0, nullptr, nullptr, nullptr, 1));
    CK(feasm_element_matrices(ctx, &p, le, le + 1, 0, ke.data(), fe0.data()));
    feasm_libmesh::assemble(es, "Poisson", p);                      // the adaptor: a call without a load clears the list
    double moved = 0.;
    for (int a = 0; a < 27; ++a) moved += fe1[a] - fe0[a];
    REQUIRE(std::fabs(moved - g1 * (2. / ny) * (2. / nz)) < 1e-12);  // = g * area of the x-max side (sum_a phi_a = 1)
    for (int64_t r = 0; r < n_dofs; ++r) REQUIRE(rhs._v[r] == b[r]);
  }

  // ---- the ex3 penalty loop through the adaptor (every side without a neighbour) vs the C ABI directly ----
  {
    std::vector<uint32_t> pe;
    std::vector<uint8_t> ps;
    for (int64_t e = 0; e < n_elem; ++e)
      for (unsigned s = 0; s < 6; ++s)
        if (elems[e]._neighbors[s] == nullptr) { pe.push_back((uint32_t)e); ps.push_back((uint8_t)s); }
    REQUIRE(pe.size() == (size_t)(2 * (nx * ny + ny * nz + nx * nz)));
    const feasm_libmesh::SidePenalty pen{1.e10, FEASM_PENALTY_EX4_EXACT, 0.};
    A.zero();
    feasm_libmesh::assemble(es, "Poisson", p, nullptr, &pen);
    A.close();
    CK(feasm_set_side_penalty(ctx, (int64_t)pe.size(), pe.data(), ps.data(), 1.e10, FEASM_PENALTY_EX4_EXACT, 0.));
    CK(feasm_assemble(ctx, &p));
    std::vector<double> b3(n_dofs);
    CK(feasm_get_values(ctx, nullptr, b3.data()));
    double moved = 0.;
    for (int64_t r = 0; r < n_dofs; ++r) { REQUIRE(rhs._v[r] == b3[r]); moved = std::max(moved, std::fabs(b3[r] - b[r])); }
    REQUIRE(moved == 0.);  // every boundary dof is a Dirichlet row here: the constraint pass overwrites the penalised rows
    std::vector<double> k1(27 * 27), f1(27), k0(27 * 27), f0(27);
    CK(feasm_element_matrices(ctx, &p, 0, 1, 0, k1.data(), f1.data()));
    CK(feasm_set_side_penalty(ctx, 0, nullptr, nullptr, 0., 0, 0.));
    CK(feasm_element_matrices(ctx, &p, 0, 1, 0, k0.data(), f0.data()));
    double dk = 0.;
    for (int i = 0; i < 27 * 27; ++i) dk += k1[i] - k0[i];
    // element 0 has three boundary sides of area (2/ny)(2/nz), (2/nx)(2/nz), (2/nx)(2/ny): sum_ij phi_i phi_j = 1
    const double area = (2. / ny) * (2. / nz) + (2. / nx) * (2. / nz) + (2. / nx) * (2. / ny);
    REQUIRE(std::fabs(dk - 1.e10 * area) < 1e-3 * 1.e10 * 1e-9);
    feasm_libmesh::assemble(es, "Poisson", p);  // and a call without the term clears the adaptor's list
    for (int64_t r = 0; r < n_dofs; ++r) REQUIRE(rhs._v[r] == b[r]);
  }

  // ---- the rest of the SparseMatrix surface: clone, set, add(a, X), get_transpose, add_vector_transpose -------------------
  {
    auto Cl = A.clone();                                             // sparse_matrix.h:226: same values, its own storage
    auto* C2 = static_cast<GpuSparseMatrix<Number>*>(Cl.get());
    REQUIRE(C2->host_values() == A.host_values());
    const numeric_index_type r5 = 5, c5 = graph[5][0];
    const double a5 = A(r5, c5);
    C2->set(r5, c5, 123.5);                                          // :295
    C2->close();
    REQUIRE((*C2)(r5, c5) == 123.5 && A(r5, c5) == a5);              // the clone is not a view
    C2->set(r5, c5, a5);
    C2->add(-1., A);                                                 // :344: C2 = C2 - A = 0
    C2->close();
    for (double v : C2->host_values()) REQUIRE(v == 0.);
    C2->add(2., A);
    C2->close();
    for (std::size_t k = 0; k < A.host_values().size(); ++k) REQUIRE(C2->host_values()[k] == 2. * A.host_values()[k]);
    // make A unsymmetric in one slot, then transpose
    const numeric_index_type ri = 7, cj = graph[7][0] == 7 ? graph[7][1] : graph[7][0];
    A.add(ri, cj, 0.375);
    A.close();
    GpuSparseMatrix<Number> At(comm);
    A.get_transpose(At);                                             // :598
    for (int64_t r = 0; r < n_dofs; ++r) {
      std::vector<numeric_index_type> idx;
      std::vector<Number> val;
      At.get_row((numeric_index_type)r, idx, val);
      for (std::size_t k = 0; k < idx.size(); ++k) REQUIRE(val[k] == A(idx[k], (numeric_index_type)r));
      for (std::size_t k = 1; k < idx.size(); ++k) REQUIRE(idx[k - 1] < idx[k]);
    }
    REQUIRE(At(cj, ri) == A(ri, cj) && At(ri, cj) == A(cj, ri));
    // y += A^T v against y += At v
    GpuNumericVector<Number> gv2(comm, (numeric_index_type)n_dofs), y1(comm, (numeric_index_type)n_dofs), y2(comm, (numeric_index_type)n_dofs);
    std::vector<Number> hv(n_dofs);
    for (int64_t i = 0; i < n_dofs; ++i) hv[i] = std::sin(0.3 * i) + 0.1;
    gv2 = hv;
    y1.add_vector_transpose(gv2, A);
    y2.add_vector(gv2, At);
    std::vector<Number> h1, h2;
    y1.localize(h1); y2.localize(h2);
    REQUIRE(h1 == h2);
    A.add(ri, cj, -0.375);
    A.close();
  }
  // ---- SparseMatrix::init + the per-element virtuals WITHOUT the batched hook (tests/numerics/sparse_matrix_test.h:88-149) ---
  {
    GpuSparseMatrix<Number> B(comm);
    B.attach_dof_map(dof_map);
    B.update_sparsity_pattern(graph);
    B.init();
    DenseMatrix<Number> d3;
    d3._m = d3._n = 3; d3._val = {1., 2., 3., 4., 5., 6., 7., 8., 9.};
    std::vector<numeric_index_type> dofs3 = {dof_map._elem_dofs[0], dof_map._elem_dofs[1], dof_map._elem_dofs[2]};
    B.add_matrix(d3, dofs3);
    B.add_matrix(d3, dofs3);
    B.close();
    for (int a = 0; a < 3; ++a)
      for (int c = 0; c < 3; ++c) REQUIRE(B(dofs3[a], dofs3[c]) == 2. * d3._val[3 * a + c]);
    std::vector<numeric_index_type> idx;
    std::vector<Number> val;
    B.get_row(dofs3[1], idx, val);
    REQUIRE(idx == graph[dofs3[1]]);
    double row_sum = 0.;
    for (double v : val) row_sum += v;
    REQUIRE(row_sum == 2. * (4. + 5. + 6.));
    B.zero();
    B.close();
    REQUIRE(B.linfty_norm() == 0.);
  }
  // ---- a non-empty constraint row through the hook (the hanging-node algebra: the element is widened by the constraining
  // dofs, C^T K C) vs the C ABI driven directly with the same rows ---------------------------------------------------
  {
    std::vector<dof_id_type> inner;
    for (int64_t i = 0; i < n_nodes && inner.size() < 3; ++i) if (!mask[i]) inner.push_back(base[i]);
    REQUIRE(inner.size() == 3);
    const dof_id_type slave = inner[2];
    dof_map._rows[slave] = {{inner[0], 0.5}, {inner[1], 0.5}};
    std::vector<uint32_t> gd;
    std::vector<uint64_t> grp(1, 0);
    std::vector<uint32_t> gc;
    std::vector<double> gw, gv;
    for (dof_id_type d = 0; d < (dof_id_type)n_dofs; ++d)
      if (dof_map.is_constrained_dof(d)) {
        gd.push_back(d);
        if (d == slave) { gc.push_back(inner[0]); gc.push_back(inner[1]); gw.push_back(0.5); gw.push_back(0.5); }
        grp.push_back(gc.size());
        const auto v = dof_map._vals.find(d);
        gv.push_back(v == dof_map._vals.end() ? 0. : v->second);
      }
    feasm_ctx* c2 = nullptr;
    CK(feasm_create(0, &c2));
    CK(feasm_set_mesh(c2, FEASM_HEX27, n_elem, n_nodes, xyz.data(), conn.data()));
    CK(feasm_set_dofs(c2, 1, n_dofs, n_dofs, dof_map._elem_dofs.data(), nullptr));
    CK(feasm_set_constraint_rows(c2, (int64_t)gd.size(), gd.data(), grp.data(), gc.data(), gw.data(), gv.data()));
    CK(feasm_build_pattern(c2));  // with the couplings DofMap::find_connected_dofs adds: the Graph libMesh would hand over
    std::vector<uint64_t> rp2(n_dofs + 1);
    std::vector<uint32_t> ci2((std::size_t)feasm_nnz(c2));
    CK(feasm_get_pattern(c2, rp2.data(), ci2.data()));
    REQUIRE(ci2.size() > A.col_idx().size());
    CK(feasm_assemble(c2, &p));
    std::vector<double> v2(ci2.size()), b2(n_dofs);
    CK(feasm_get_values(c2, v2.data(), b2.data()));
    SparsityPattern::Graph graph2(n_dofs);
    for (int64_t r = 0; r < n_dofs; ++r) graph2[r].assign(ci2.begin() + rp2[r], ci2.begin() + rp2[r + 1]);
    GpuSparseMatrix<Number> Ah(comm);
    Ah.attach_dof_map(dof_map);
    Ah.update_sparsity_pattern(graph2);
    Ah.init();
    system.matrix = &Ah;
    feasm_libmesh::assemble_poisson_ex4(es, "Poisson");
    Ah.close();
    REQUIRE(Ah.host_values() == v2);
    for (int64_t r = 0; r < n_dofs; ++r) REQUIRE(rhs._v[r] == b2[r]);
    // the constrained row: its diagonal counts the touching elements, its masters carry -coef times that count, rhs 0
    const double t = Ah(slave, slave);
    REQUIRE(t >= 1. && Ah(slave, inner[0]) == -0.5 * t && Ah(slave, inner[1]) == -0.5 * t && rhs._v[slave] == 0.);
    system.matrix = &A;
    dof_map._rows.clear();
    feasm_libmesh::assemble_poisson_ex4(es, "Poisson");  // back to Dirichlet rows only
    REQUIRE(A.host_values() == v);
    feasm_destroy(c2);
  }

  // ---- a moved mesh: the hook reads the node coordinates on every call (stale flatten cache) ---------------
  {
    feasm_libmesh::assemble_poisson_ex4(es, "Poisson");
    const std::vector<Number> v_before = A.host_values();
    std::vector<double> x0;
    for (auto& nd : nodes) { x0.push_back(nd.c[0]); nd.c[0] *= 1.25; }  // stretch in x: every element matrix changes
    feasm_libmesh::assemble_poisson_ex4(es, "Poisson");
    REQUIRE(A.host_values() != v_before);
    for (std::size_t i = 0; i < nodes.size(); ++i) nodes[i].c[0] = x0[i];
    feasm_libmesh::assemble_poisson_ex4(es, "Poisson");
    REQUIRE(A.host_values() == v_before);
    // System::reinit path: clear() drops the context, the next hook call uploads everything again
    A.clear();
    A.attach_dof_map(dof_map);
    A.update_sparsity_pattern(graph);
    A.init();
    feasm_libmesh::assemble_poisson_ex4(es, "Poisson");
    REQUIRE(A.host_values() == v_before);
  }

  // ---- compare mode (compare_mode.h), here with the hook on both sides: exercises the comparison a libMesh box runs against
  // the reference's own assemble function
  {
    GpuSparseMatrix<Number> A2(comm);
    A2.attach_dof_map(dof_map);
    A2.update_sparsity_pattern(graph);
    A2.init();
    NumericVector<Number> b1(comm), b2(comm);
    b1.init((numeric_index_type)n_dofs); b2.init((numeric_index_type)n_dofs);
    auto rep = feasm_libmesh::compare_assembly(es, "Poisson", feasm_libmesh::assemble_poisson_ex4, A, b1, feasm_libmesh::assemble_poisson_ex4, A2, b2);
    REQUIRE(rep.passes() && rep.pattern_equal && rep.strict_failures == 0 && rep.entries == A.col_idx().size());
    A2.add(5, graph[5][0], 1e-9);  // a planted difference is found
    A2.close();
    rep = feasm_libmesh::compare_systems(A, b1, A2, b2);
    REQUIRE(!rep.passes() && rep.strict_failures == 1);
    rep.print("compare mode self-check (one planted difference)");
  }

  auto Z = A.zero_clone();
  REQUIRE(Z->m() == A.m());
  A.clear();
  REQUIRE(!A.initialized());
  feasm_destroy(ctx);
  std::printf("SHIM TEST OK: %lld elements, %lld dofs, CG %u iterations, nodal error %.2e\n", (long long)n_elem, (long long)n_dofs, res.first, err);
  return 0;
}
