// ref_app_shim.cpp — the stub world of ref_app/ref_world.h (definitions) and the C entry points of oracle/_ref/libref_app.so.
// Test infrastructure: see ref_world.h for what in that library is the reference's own code and what is glue.  Nothing under
// libmesh_b200/ may touch it.
#include <chrono>

#include "ref_world.h"
#include "libmesh/quadrature_gauss.h"

namespace libMesh {

// ---- the reference's own function bodies, excerpted at build time by oracle/Makefile (awk, from the first line of each
// definition to its closing brace) out of src/base/dof_map_constraints.C and src/fe/h1_fe_transformation.C into the build
// intermediate _ref/gen/ref_excerpts.inc (removed after the compile; never committed) -------------------------------------------------
#include "ref_excerpts.inc"
template class H1FETransformation<Real>;

RefWorldData& ref_world() {
  static RefWorldData w;
  return w;
}

namespace MeshTools { namespace Generation {
void build_cube(UnstructuredMesh& mesh, const unsigned int nx, const unsigned int ny, const unsigned int nz, const Real xmin, const Real xmax, const Real ymin,
                const Real ymax, const Real zmin, const Real zmax, const ElemType type, const bool gauss_lobatto_grid) {
  const RefWorldData& w = ref_world();
  const double asked[6] = {xmin, xmax, ymin, ymax, zmin, zmax};
  libmesh_error_msg_if(gauss_lobatto_grid, "gauss_lobatto_grid is not loaded");
  libmesh_error_msg_if((int)type != w.elem_type || nx != w.nx || ny != w.ny || nz != w.nz || std::memcmp(asked, w.bounds, sizeof(asked)) != 0,
                       "build_cube(" << nx << ", " << ny << ", " << nz << ", type " << (int)type << ") is not the mesh that was loaded ("
                                     << w.nx << ", " << w.ny << ", " << w.nz << ", type " << w.elem_type << ")");
  const uint64_t ne = w.n_elem();
  mesh._elems.clear();
  mesh._ptrs.clear();
  mesh._elems.reserve(ne);
  std::vector<double> c((std::size_t)w.npe * 3);
  for (uint64_t e = 0; e < ne; ++e) {
    for (int a = 0; a < w.npe; ++a) std::memcpy(&c[3 * a], &w.xyz[3 * (std::size_t)w.conn[e * w.npe + a]], 3 * sizeof(double));
    mesh._elems.emplace_back(type, 0);
    mesh._elems.back().set_nodes(c.data(), (unsigned int)w.npe);
    if (w.affine[e] != 255) mesh._elems.back().force_affine(w.affine[e] != 0);  // 255: the element class's own has_affine_map decides
    mesh._elems.back().set_id((dof_id_type)e);
    if (!w.side_bid.empty()) mesh._elems.back().set_side_boundary_ids(&w.side_bid[e * w.n_sides], (unsigned int)w.n_sides);
  }
  for (const Elem& e : mesh._elems) mesh._ptrs.push_back(&e);
}
} }

// ---- DofMap ---------------------------------------------------------------------------------------------------------------------------
void DofMap::dof_indices(const Elem* const elem, std::vector<dof_id_type>& di) const {
  const RefWorldData& w = ref_world();
  di.clear();
  for (int v = 0; v < w.n_vars; ++v)
    for (int a = 0; a < w.npe; ++a) di.push_back(w.node_dof[(std::size_t)v * w.n_nodes() + w.conn[(std::size_t)elem->id() * w.npe + a]]);
}
void DofMap::dof_indices(const Elem* const elem, std::vector<dof_id_type>& di, const unsigned int vn) const {
  const RefWorldData& w = ref_world();
  di.clear();
  for (int a = 0; a < w.npe; ++a) di.push_back(w.node_dof[(std::size_t)vn * w.n_nodes() + w.conn[(std::size_t)elem->id() * w.npe + a]]);
}
void DofMap::create_dof_constraints() {
  RefWorldData& w = ref_world();
  _dof_constraints.clear();
  _primal_constraint_values.clear();
  const uint64_t nn = w.n_nodes();
  for (const DirichletBoundary& d : _dirichlet) {
    unsigned int mask = 0;
    for (boundary_id_type b : d.b) if (b >= 0 && b < 8) mask |= 1u << b;
    DenseVector<Number> out((unsigned int)w.n_vars);
    for (uint64_t n = 0; n < nn; ++n) {
      if (!(w.node_bmask[n] & mask)) continue;
      out.zero();
      (*d.f)(Point(w.xyz[3 * n], w.xyz[3 * n + 1], w.xyz[3 * n + 2]), 0., out);
      for (std::size_t i = 0; i < d.variables.size(); ++i) {
        const unsigned int var = d.variables[i];
        const dof_id_type dof = w.node_dof[(std::size_t)var * nn + n];
        const Number g = out(d.f_system_order ? var : (unsigned int)i);
        _dof_constraints[dof] = DofConstraintRow();           // a Dirichlet constraint: an empty row ...
        if (g != 0.) _primal_constraint_values[dof] = g;      // ... and its value (DofMap::add_constraint_row keeps non-zero values only)
        else _primal_constraint_values.erase(dof);
      }
    }
  }
  // rows loaded as such (ref_world_set_constraint_rows)
  for (std::size_t c = 0; c < w.crow_dof.size(); ++c) {
    DofConstraintRow row;
    for (uint64_t k = w.crow_ptr[c]; k < w.crow_ptr[c + 1]; ++k) row[w.crow_col[k]] = w.crow_coef[k];
    _dof_constraints[w.crow_dof[c]] = row;
    if (w.crow_val[c] != 0.) _primal_constraint_values[w.crow_dof[c]] = w.crow_val[c];
    else _primal_constraint_values.erase(w.crow_dof[c]);
  }
}

// ---- FEBase ---------------------------------------------------------------------------------------------------------------------------
template <> FEGenericBase<Real>::FEGenericBase(unsigned int d, const FEType& t) : dim(d), fe_type(t), _fe_trans(std::make_unique<H1FETransformation<Real>>()) {
  // the getters switch the calculations on (fe_map.h:216-380), as FEAbstract::determine_calculations does for get_dphi()
  _fe_map.get_xyz(); _fe_map.get_JxW(); _fe_map.get_jacobian();
  _fe_map.get_dxidx(); _fe_map.get_dxidy(); _fe_map.get_dxidz(); _fe_map.get_detadx(); _fe_map.get_detady(); _fe_map.get_detadz();
  _fe_map.get_dzetadx(); _fe_map.get_dzetady(); _fe_map.get_dzetadz();
  _fe_map.get_normals();
}
template <> std::unique_ptr<FEGenericBase<Real>> FEGenericBase<Real>::build(const unsigned int dim, const FEType& type) {
  libmesh_error_msg_if(dim != 3 || type.family != LAGRANGE, "this stub builds FE<3, LAGRANGE> only");
  return std::make_unique<FEGenericBase<Real>>(dim, type);
}
// init_shape_functions (fe.C:372-600): phi and its reference derivatives at the given points
template <> void FEGenericBase<Real>::init_shape_functions(const std::vector<Point>& qp, const Elem* elem) {
  const unsigned int ns = elem->n_nodes();
  const std::size_t nq = qp.size();
  for (auto* t : {&phi, &dphidxi, &dphideta, &dphidzeta, &dphidx, &dphidy, &dphidz}) t->assign(ns, std::vector<Real>(nq, 0.));
  dphi.assign(ns, std::vector<RealGradient>(nq));
  FE<3, LAGRANGE>::all_shapes(elem, fe_type.order, qp, phi, true);
  std::vector<std::vector<Real>>* comps[3] = {&dphidxi, &dphideta, &dphidzeta};
  FE<3, LAGRANGE>::all_shape_derivs(elem, fe_type.order, qp, comps, true);
}
template <> void FEGenericBase<Real>::reinit(const Elem* elem) {
  libmesh_error_msg_if(!qrule, "no quadrature rule attached");
  libmesh_error_msg_if((Order)fe_type.order != elem->default_order(), "the variable's order is not the element's");
  qrule->init(*elem);                                         // fe.C:222
  const std::vector<Point>& qp = qrule->get_points();
  if (_elem_type != elem->type()) {                           // fe.C:226-240: tables when the element type changes
    _elem_type = elem->type();
    _fe_map.init_reference_to_physical_map<3>(qp, elem);
    init_shape_functions(qp, elem);
  }
  _fe_map.compute_map(dim, qrule->get_weights(), elem, /*calculate_d2phi=*/false);  // fe.C:340
  _fe_trans->map_dphi(dim, elem, qp, *this, dphi, dphidx, dphidy, dphidz);          // compute_shape_functions, fe_base.C
}
template <> void FEGenericBase<Real>::reinit(const Elem* elem, const unsigned int s) {
  libmesh_error_msg_if(!qrule, "no quadrature rule attached");
  const std::unique_ptr<const Elem> side(elem->build_side_ptr(s));                  // fe_boundary.C:137
  qrule->init(*side, elem->p_level());                                              // :172
  if (_elem_type != elem->type() || side->type() != last_side) {                    // :176-190
    _elem_type = elem->type();
    last_side = side->type();
    _fe_map.init_face_shape_functions<3>(qrule->get_points(), side.get());
  }
  _fe_map.compute_face_map(3, qrule->get_weights(), side.get());                    // :204
  const std::vector<Real> JxW_int(_fe_map.get_JxW());                               // :211
  // side_map, :347-400: the side's points in the parent's reference space
  const std::vector<Point>& reference_side_points = qrule->get_points();
  std::vector<Point> qp(reference_side_points.size());
  for (Point& p : qp) p.zero();
  std::vector<Point> refspace_nodes;
  FEAbstract::get_refspace_nodes(elem->type(), refspace_nodes);
  const std::vector<std::vector<Real>>& psi_map = _fe_map.get_psi();
  for (std::size_t i = 0; i < psi_map.size(); ++i) {
    const Point& side_node = refspace_nodes[elem->local_side_node(s, (unsigned int)i)];
    for (std::size_t p = 0; p < qp.size(); ++p) qp[p].add_scaled(side_node, psi_map[i][p]);
  }
  _side_ref_points = qp;
  // reinit(elem, &qp), fe.C:206-216 and :333-338 (no weights: 1)
  _fe_map.init_reference_to_physical_map<3>(qp, elem);
  init_shape_functions(qp, elem);
  std::vector<Real> dummy_weights(qp.size(), 1.);
  _fe_map.compute_map(dim, dummy_weights, elem, /*calculate_d2phi=*/false);
  _fe_trans->map_dphi(dim, elem, qp, *this, dphi, dphidx, dphidy, dphidz);
  _fe_map.get_JxW() = JxW_int;                                                      // :237
}

// ---- EquationSystems ------------------------------------------------------------------------------------------------------------------
void EquationSystems::init() {
  RefWorldData& w = ref_world();
  for (auto& s : _systems) {
    if (s->n_vars() != (unsigned int)w.n_vars && s->get_dof_map()._dirichlet.empty()) continue;  // an auxiliary system: not loaded
    libmesh_error_msg_if(s->n_vars() != (unsigned int)w.n_vars, "system " << s->name() << " has " << s->n_vars() << " variables, " << w.n_vars << " were loaded");
    s->get_dof_map().create_dof_constraints();
    w.con_dof.clear();
    w.con_val.clear();
    const DofMap& dm = s->get_dof_map();
    for (const auto& row : dm._dof_constraints) {
      w.con_dof.push_back(row.first);
      const auto it = dm._primal_constraint_values.find(row.first);
      w.con_val.push_back(it == dm._primal_constraint_values.end() ? 0. : it->second);
    }
  }
  w.values.assign(w.col_idx.size(), 0.);
  w.touched.assign(w.col_idx.size(), 0);
  w.rhs.assign(w.n_dofs, 0.);
}

}  // namespace libMesh

using namespace libMesh;
// examples/introduction/introduction_ex4/introduction_ex4.C, compiled with -Dmain=ref_ex4_main
int ref_ex4_main(int argc, char** argv);
void assemble_poisson(EquationSystems& es, const std::string& system_name);
void exact_solution_wrapper(DenseVector<Number>& output, const Point& p, const Real);

namespace {
int fail(const std::exception& e, char* err, int err_len) {
  if (err && err_len > 0) {
    int n = 0;
    for (const char* c = e.what(); *c && n < err_len - 1; ++c) err[n++] = *c;
    err[n] = 0;
  }
  return -1;
}
}  // namespace

extern "C" {

// the mesh build_cube will be asked for, the dof numbering and the sparsity pattern (the oracle's integer structures)
int ref_world_load(int elem_type, int npe, unsigned int nx, unsigned int ny, unsigned int nz, const double* bounds, uint64_t n_nodes, const double* xyz,
                   uint64_t n_elem, const uint32_t* conn, const uint8_t* affine, const uint8_t* node_bmask, int n_vars, uint64_t n_dofs,
                   const uint32_t* node_dof, const uint64_t* row_ptr, const uint32_t* col_idx) {
  RefWorldData& w = ref_world();
  w = RefWorldData();
  w.elem_type = elem_type; w.npe = npe; w.nx = nx; w.ny = ny; w.nz = nz;
  std::memcpy(w.bounds, bounds, sizeof(w.bounds));
  w.xyz.assign(xyz, xyz + 3 * n_nodes);
  w.conn.assign(conn, conn + n_elem * npe);
  w.affine.assign(affine, affine + n_elem);
  w.node_bmask.assign(node_bmask, node_bmask + n_nodes);
  w.n_vars = n_vars; w.n_dofs = n_dofs;
  w.node_dof.assign(node_dof, node_dof + (uint64_t)n_vars * n_nodes);
  w.row_ptr.assign(row_ptr, row_ptr + n_dofs + 1);
  w.col_idx.assign(col_idx, col_idx + row_ptr[n_dofs]);
  return 0;
}

// BoundaryInfo's ids of the boundary sides: bid[n_elem][n_sides], -1 on interior sides (what Elem::neighbor_ptr(side) == nullptr and
// BoundaryInfo::has_boundary_id(elem, side, id) answer from)
int ref_world_set_sides(int n_sides, const int16_t* bid) {
  RefWorldData& w = ref_world();
  w.n_sides = n_sides;
  w.side_bid.assign(bid, bid + w.n_elem() * (uint64_t)n_sides);
  return 0;
}

// constraint rows given as such: hanging nodes (and Dirichlet rows as empty rows with a value), CSR over the constrained dofs
int ref_world_set_constraint_rows(int64_t n, const uint32_t* dofs, const uint64_t* row_ptr, const uint32_t* cols, const double* coefs, const double* values) {
  RefWorldData& w = ref_world();
  w.crow_dof.assign(dofs, dofs + n);
  w.crow_ptr.assign(row_ptr, row_ptr + n + 1);
  w.crow_col.assign(cols, cols + row_ptr[n]);
  w.crow_coef.assign(coefs, coefs + row_ptr[n]);
  w.crow_val.assign(values, values + n);
  return 0;
}

// main() of introduction_ex4.C with the given command line ("-d 3 -n 4 ..."); returns its return value, or -1 with the message
int ref_ex4_run(int argc, const char* const* argv, char* err, int err_len) {
  try {
    std::vector<std::string> store(argv, argv + argc);
    std::vector<char*> av;
    for (auto& s : store) av.push_back(s.data());
    return ref_ex4_main(argc, av.data());
  } catch (const std::exception& e) { return fail(e, err, err_len); }
}

// assemble_poisson of introduction_ex4.C called the way its main() sets it up (:190-274), but on whatever mesh was loaded — the
// element types main() does not ask build_cube for (TET10, HEX20, PRISM18), skewed or locally refined meshes: a "Poisson" system
// with one LAGRANGE variable of `order`, DirichletBoundary with exact_solution_wrapper on the boundary ids of `boundary_mask`.
int ref_ex4_assemble(int order, unsigned int boundary_mask, char* err, int err_len) {
  try {
    const RefWorldData& w = ref_world();
    const char* none[1] = {"ref_ex4_assemble"};
    LibMeshInit init(1, none);
    Mesh mesh(init.comm());
    MeshTools::Generation::build_cube(mesh, w.nx, w.ny, w.nz, w.bounds[0], w.bounds[1], w.bounds[2], w.bounds[3], w.bounds[4], w.bounds[5],
                                      static_cast<ElemType>(w.elem_type));
    EquationSystems equation_systems(mesh);
    LinearImplicitSystem& system = equation_systems.add_system<LinearImplicitSystem>("Poisson");
    const unsigned int u_var = system.add_variable("u", static_cast<Order>(order), LAGRANGE);
    system.attach_assemble_function(assemble_poisson);
    std::set<boundary_id_type> boundary_ids;
    for (int b = 0; b < 8; ++b) if (boundary_mask & (1u << b)) boundary_ids.insert((boundary_id_type)b);
    AnalyticFunction<> exact_solution_object(exact_solution_wrapper);
    DirichletBoundary dirichlet_bc(boundary_ids, {u_var}, exact_solution_object);
    system.get_dof_map().add_dirichlet_boundary(dirichlet_bc);
    equation_systems.init();
    const auto t0 = std::chrono::steady_clock::now();
    system.solve();
    ref_world().solve_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return 0;
  } catch (const std::exception& e) { return fail(e, err, err_len); }
}

// wall-clock seconds of the last system.solve() (= zero + the example's assembly function) of ref_ex4_assemble
double ref_world_solve_seconds(void) { return ref_world().solve_seconds; }

// entries of the loaded sparsity pattern that no add_matrix call of the run added into (an entry outside the pattern throws): 0 means
// the pattern is exactly the set of (row, column) pairs the example's element loop touches
int64_t ref_world_untouched_entries(void) {
  int64_t n = 0;
  for (uint8_t t : ref_world().touched) n += !t;
  return n;
}

// what the run left behind: matrix values in the loaded pattern's order, the right-hand side, the Dirichlet constraints
int64_t ref_world_results(double* values, double* rhs, uint32_t* con_dof, double* con_val) {
  const RefWorldData& w = ref_world();
  if (values) std::memcpy(values, w.values.data(), sizeof(double) * w.values.size());
  if (rhs) std::memcpy(rhs, w.rhs.data(), sizeof(double) * w.rhs.size());
  if (con_dof) std::memcpy(con_dof, w.con_dof.data(), sizeof(uint32_t) * w.con_dof.size());
  if (con_val) std::memcpy(con_val, w.con_val.data(), sizeof(double) * w.con_val.size());
  return (int64_t)w.con_dof.size();
}

// FE::reinit(elem, side) on one element given by its node coordinates, with QGauss(2, qorder) on the side: the number of side points
// nqf (<= cap), ref[nqf][3] (the points in the parent's reference space), xyz[nqf][3], JxW[nqf], normals[nqf][3], phi[n][nqf].
int ref_side_reinit(int elem_type, int n_nodes, const double* node_xyz, int affine, int side, int fe_order, int qorder, int cap, double* ref, double* xyz,
                    double* JxW, double* normals, double* phi, char* err, int err_len) {
  try {
    Elem elem(static_cast<ElemType>(elem_type), 0);
    elem.set_nodes(node_xyz, (unsigned int)n_nodes);
    if (affine >= 0) elem.force_affine(affine != 0);
    std::unique_ptr<FEBase> fe_face(FEBase::build(3, FEType(fe_order, LAGRANGE)));
    QGauss qface(2, static_cast<Order>(qorder));
    fe_face->attach_quadrature_rule(&qface);
    fe_face->reinit(&elem, (unsigned int)side);
    const int nqf = (int)qface.n_points();
    libmesh_error_msg_if(nqf > cap, nqf << " side points, room for " << cap);
    for (int p = 0; p < nqf; ++p) {
      JxW[p] = fe_face->get_JxW()[p];
      for (int d = 0; d < 3; ++d) {
        ref[3 * p + d] = fe_face->side_reference_points()[p](d);
        xyz[3 * p + d] = fe_face->get_xyz()[p](d);
        normals[3 * p + d] = fe_face->get_normals()[p](d);
      }
      for (int i = 0; i < n_nodes; ++i) phi[(std::size_t)i * nqf + p] = fe_face->get_phi()[i][p];
    }
    return nqf;
  } catch (const std::exception& e) { return fail(e, err, err_len); }
}

// DofMap::constrain_element_matrix_and_vector (hetero = 0) / heterogeneously_constrain_element_matrix_and_vector (hetero = 1) of the
// reference on one element: K [n][n], F [n], dofs [n]; the constraints as rows con_dof[c] = sum_k coef * row_dof + con_val[c]
// (CSR over c).  The element's dof set may grow (constraining dofs that are not the element's are appended): returns the new size
// m (<= cap) with K_out [m][m], F_out [m], dofs_out [m]; -1 when the reference code throws.
int ref_constrain_element(int n, const double* K, const double* F, const uint32_t* dofs, int n_con, const uint32_t* con_dof, const uint64_t* con_ptr,
                          const uint32_t* row_dof, const double* row_coef, const double* con_val, int hetero, int asymmetric, int cap,
                          uint32_t* dofs_out, double* K_out, double* F_out, char* err, int err_len) {
  try {
    DofMap dm;
    for (int c = 0; c < n_con; ++c) {
      DofMap::DofConstraintRow row;
      for (uint64_t k = con_ptr[c]; k < con_ptr[c + 1]; ++k) row[row_dof[k]] = row_coef[k];
      dm._dof_constraints[con_dof[c]] = row;
      if (con_val && con_val[c] != 0.) dm._primal_constraint_values[con_dof[c]] = con_val[c];
    }
    DenseMatrix<Number> Ke(n, n);
    DenseVector<Number> Fe(n);
    std::vector<dof_id_type> di(dofs, dofs + n);
    for (int i = 0; i < n; ++i) {
      Fe(i) = F[i];
      for (int j = 0; j < n; ++j) Ke(i, j) = K[(std::size_t)i * n + j];
    }
    if (hetero) dm.heterogeneously_constrain_element_matrix_and_vector(Ke, Fe, di, asymmetric != 0);
    else dm.constrain_element_matrix_and_vector(Ke, Fe, di, asymmetric != 0);
    const int m = (int)di.size();
    libmesh_error_msg_if(m > cap, "the constrained element has " << m << " dofs, room for " << cap);
    libmesh_error_msg_if((int)Ke.m() != m || (int)Ke.n() != m || (int)Fe.size() != m, "constrained sizes disagree");
    for (int i = 0; i < m; ++i) {
      dofs_out[i] = di[i];
      F_out[i] = Fe(i);
      for (int j = 0; j < m; ++j) K_out[(std::size_t)i * m + j] = Ke(i, j);
    }
    return m;
  } catch (const std::exception& e) { return fail(e, err, err_len); }
}

}  // extern "C"
