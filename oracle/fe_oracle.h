/* fe_oracle.h — C interface of the CPU ORACLE.
 *
 * TEST INFRASTRUCTURE ONLY.  This library is a plain-C++ restatement of libMesh's finite-element
 * assembly path (FEMap::compute_map + FE::reinit -> Ke/Fe -> DofMap::dof_indices ->
 * SparseMatrix::add_matrix / NumericVector::add_vector).  It is the checker for the CUDA product in
 * libmesh_b200/ and the timed CPU baseline of bench.py.  Nothing in the product may include, link or
 * call it; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Parity status: PINNED AGAINST OUTPUTS OF THE REFERENCE ITSELF, RUN HERE, for every floating-point number of the path.
 * libMesh as a whole cannot be compiled in the build container (TIMPI / MetaPhysicL submodules are empty, libmesh_config.h
 * is generated), but the reference files that hold the arithmetic of the path can, where they lie, against stub declarations
 * (oracle/Makefile -> oracle/_ref/; oracle/ref_stubs/, oracle/ref_app/):
 *   libref_fe.so   src/fe/fe_lagrange_shape_{2,3}D.C, include/fe/fe_lagrange_shape_1D.h, src/quadrature/quadrature*.C,
 *                  src/fe/fe_map.C, the face maps of src/fe/fe_boundary.C, the examples' exact_solution.C: this oracle's Gauss
 *                  rules, shape tables, FE::reinit and FE::reinit(elem, side) output are BIT-IDENTICAL on every element type
 *                  of the path, affine and general branch (tests/test_oracle_ref.py, golden tables reference_fe_*.npz);
 *   libref_app.so  the examples themselves — introduction_ex4.C (main and assemble_poisson), systems_of_equations_ex6.C
 *                  (LinearElasticity::assemble) — over the reference's DenseMatrix sources and its DofMap constraint
 *                  functions: the GLOBAL matrix and right-hand side they produce are BIT-IDENTICAL to orc_assemble's, config 1
 *                  (20^3 HEX27) at full size included (tests/test_oracle_ref_app.py; golden outputs reference_ex{4,6}_runs.npz,
 *                  which tests/test_golden_ref_app.py and the GPU suite compare with on machines without /root/reference).
 *   libref_mesh.so the integer side — MeshTools::Generation::build_cube, UnstructuredMesh::all_second_order_range (TET10),
 *                  ReplicatedMesh::renumber_nodes_and_elements,
 *                  DofMap::distribute_local_dofs_var_major, LinearPartitioner::partition_range — over a recording mesh: this
 *                  oracle's mesh (coordinates bitwise), node numbering, dof numbering and element shards ARE the reference's
 *                  (tests/test_oracle_ref_mesh.py); Hex8 ... Prism18::has_affine_map likewise (libref_fe.so), and the sparsity
 *                  pattern is exactly the set of entries the reference's element loop adds into (libref_app.so).
 * Still only restated (glue): the call order inside FE::reinit, dof_indices as a lookup, DirichletBoundary on LAGRANGE
 * variables as nodal evaluation — held by the reference's unit-test known answers
 * (tests/fe/fe_test.h, tests/numerics/sparse_matrix_test.h, tests/systems/systems_test.C, tests/base/dof_map_test.C,
 * tests/mesh/mesh_generation_test.C) and its published run facts, restated in tests/test_oracle_*.py.
 *
 * Every function cites the reference file:line it follows (paths relative to /root/reference).
 */
#ifndef FE_ORACLE_H
#define FE_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* element types (values = nodes per element) */
#define ORC_HEX8 8
#define ORC_HEX27 27
#define ORC_TET10 10
#define ORC_HEX20 20
#define ORC_PRISM18 18

/* problems */
#define ORC_POISSON_EX4 1    /* examples/introduction/introduction_ex4/introduction_ex4.C:313-563 */
#define ORC_ELASTICITY_EX6 2 /* examples/systems_of_equations/systems_of_equations_ex6/systems_of_equations_ex6.C:156-264; the traction
                              * term :240-257 through orc_system_set_side_loads */
#define ORC_LAPLACE 3        /* Poisson with f = 0 (config 4) */

/* forcing selection for ORC_POISSON_EX4 */
#define ORC_FORCING_FD 0        /* the finite-difference Laplacian of introduction_ex4.C:500-537 */
#define ORC_FORCING_TABULATED 1 /* f at quadrature points supplied by the caller, [n_elem][nq] */

typedef struct orc_mesh orc_mesh;
typedef struct orc_system orc_system;

/* ---- quadrature / shape tables (src/quadrature, src/fe) ---- */
/* qorder: libMesh Order of the Gauss rule (3 = THIRD, 5 = FIFTH).  Returns nq; fills up to cap entries. */
int orc_quadrature(int elem_type, int qorder, int cap, double* points /*nq*3*/, double* weights /*nq*/);
/* reference shape functions at arbitrary reference points: phi[n][np], dphidxi/eta/zeta[n][np] */
int orc_shapes(int elem_type, int np, const double* ref_points, double* phi, double* dxi, double* deta, double* dzeta);
/* FE::reinit on one element with explicit reference points + weights (FE::reinit(elem, pts, wts)):
 * outputs xyz[np][3], JxW[np], phi[n][np], dphi[n][np][3]; returns 1 if the affine branch was taken, 0 if
 * general, <0 on a degenerate map. */
int orc_fe_reinit(int elem_type, const double* node_xyz /*n*3*/, int np, const double* ref_points,
                  const double* weights, double* xyz, double* JxW, double* phi, double* dphi);
int orc_has_affine_map(int elem_type, const double* node_xyz);
/* FE::reinit(elem, side) with the side's own Gauss rule (src/fe/fe_boundary.C:109-240), HEX8/HEX27: returns the number
 * of side quadrature points nqf (<= cap) and fills ref_points[nqf][3] (the points in the parent's reference space,
 * FE::side_map), xyz[nqf][3], JxW[nqf], normals[nqf][3], phi[n][nqf] (any may be NULL); < 0 on error. */
int orc_side_reinit(int elem_type, const double* node_xyz /*n*3*/, int side, int qorder, int cap, double* ref_points,
                    double* xyz, double* JxW, double* normals, double* phi);

/* ---- mesh (src/mesh/mesh_generation.C build_cube) ---- */
orc_mesh* orc_mesh_build_cube(int elem_type, int nx, int ny, int nz, double xmin, double xmax, double ymin,
                              double ymax, double zmin, double zmax, int renumber_nodes);
orc_mesh* orc_mesh_from_arrays(int elem_type, int64_t n_elem, int64_t n_nodes, const double* xyz, const uint32_t* conn,
                               const uint8_t* boundary_mask /* may be NULL */);
void orc_mesh_free(orc_mesh*);
int64_t orc_mesh_n_elem(const orc_mesh*);
int64_t orc_mesh_n_nodes(const orc_mesh*);
int orc_mesh_nodes_per_elem(const orc_mesh*);
void orc_mesh_get(const orc_mesh*, double* xyz /*n_nodes*3*/, uint32_t* conn /*n_elem*npe*/,
                  uint8_t* node_boundary_mask /*n_nodes, bit b = on boundary id b*/);
/* tests/fe/fe_test.h:40-64 SkewFunc: x += 0.1*y ... ; kind 1 = seeded interior perturbation */
void orc_mesh_skew(orc_mesh*, int kind, double amplitude, uint64_t seed);
/* (element, side) pairs carrying one boundary id (BoundaryInfo after build_cube); returns the count (fills <= cap) */
int64_t orc_mesh_boundary_sides(const orc_mesh*, int boundary_id, int64_t cap, uint32_t* elems, uint8_t* sides);
/* LinearPartitioner (src/partitioning/linear_partitioner.C:36-80) */
void orc_mesh_partition_linear(orc_mesh*, int n_parts);
void orc_mesh_set_partition(orc_mesh*, int n_parts, const int32_t* elem_pid);
void orc_mesh_get_partition(const orc_mesh*, int32_t* elem_pid, int32_t* node_pid);

/* ---- system: dofs, constraints, sparsity, assembly ---- */
orc_system* orc_system_create(const orc_mesh*, int n_vars);
void orc_system_free(orc_system*);
int64_t orc_system_n_dofs(const orc_system*);
int orc_system_dofs_per_elem(const orc_system*);
void orc_system_get_dofs(const orc_system*, uint32_t* elem_dofs /*n_elem*nd*/, uint32_t* node_dof_base /*n_nodes*/,
                         uint32_t* first_dof /*n_parts+1*/);
/* Dirichlet on the boundary ids in id_mask; kind 0: g = exact_solution (ex4), kind 1: g = 0 (ex6).  Returns count. */
int64_t orc_system_add_dirichlet(orc_system*, uint32_t id_mask, int kind);
void orc_system_get_constraints(const orc_system*, uint32_t* dofs, double* values);
/* Side loads, systems_of_equations_ex6.C:240-257: Fe_var[i](a) += JxW_face[qp] * (g[i] * phi_face[a][qp]) on the listed
 * sides (QGauss(dim-1) of the volume rule's order), added after the volume terms and before the element constraint.
 * g has one component per variable.  n = 0 clears. */
void orc_system_set_side_loads(orc_system*, int64_t n, const uint32_t* elems, const uint8_t* sides, const double* g);
/* Penalty sides, introduction_ex3.C:386-455: Ke(i,j) += JxW_face*penalty*phi_i*phi_j, Fe(i) += JxW_face*penalty*value*phi_i;
 * kind 0: value constant, kind 1: value = exact_solution at the side quadrature point.  n = 0 clears. */
void orc_system_set_side_penalty(orc_system*, int64_t n, const uint32_t* elems, const uint8_t* sides, double penalty, int kind, double value);
/* SparsityPattern::Build (src/base/sparsity_pattern.C:96-363): returns nnz */
int64_t orc_system_build_sparsity(orc_system*);
void orc_system_get_sparsity(const orc_system*, uint64_t* row_ptr /*n_dofs+1*/, uint32_t* col_idx /*nnz*/);
/* element matrices for elements [e0,e1): Ke [ne][nd*nd] row-major, Fe [ne][nd]; constrained!=0 applies the
 * (heterogeneous for ex4, homogeneous for ex6) element constraint.  qorder 0 = default of the problem. */
int orc_element_matrices(const orc_system*, int problem, int qorder, int forcing, const double* f_tab,
                         int64_t e0, int64_t e1, int constrained, double* Ke, double* Fe);
/* forcing f(x_q) of ex4 at every quadrature point: f[n_elem][nq], xyz_q[n_elem][nq][3] (either may be NULL) */
int orc_forcing_at_qps(const orc_system*, int qorder, double* f, double* xyz_q);
/* full assembly into CSR values (nnz) and rhs (n_dofs), zeroed first.  n_threads 1 = the reference's serial
 * loop; >1 = OpenMP over elements with atomic adds.  phase_seconds[4] = elem init, Ke, Fe, matrix insertion
 * (filled for n_threads==1 only).  Returns 0, or <0 on error (degenerate map, missing sparsity entry). */
int orc_assemble(const orc_system*, int problem, int qorder, int forcing, const double* f_tab, int n_threads,
                 double* values, double* rhs, double* phase_seconds);
/* the same loop restricted to elements [e0,e1), all work done and discarded into the same arrays (baseline timing) */
int orc_assemble_range(const orc_system*, int problem, int qorder, int forcing, const double* f_tab, int n_threads,
                       int64_t e0, int64_t e1, double* values, double* rhs);
double orc_exact_solution(double x, double y, double z);
const char* orc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
