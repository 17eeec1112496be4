/* feasm.h — C ABI of libfeasm, the B200-native finite-element assembly path for libMesh.
 *
 * libfeasm replaces ONE path of libMesh: the per-element loop a user's assemble function runs
 * (examples/introduction/introduction_ex4/introduction_ex4.C:407-558):
 *     DofMap::dof_indices  ->  FE::reinit / FEMap::compute_map  ->  Ke/Fe contraction  ->
 *     DofMap::(heterogeneously_)constrain_element_matrix_and_vector  ->
 *     SparseMatrix::add_matrix / NumericVector::add_vector.
 * Everything here is plain C: opaque context, plain pointers and sizes, int return codes (0 = ok,
 * negative = error, message via feasm_last_error).  Host arrays are copied at the call; the context owns
 * all device memory.  One context per GPU; calls on one context must be serialised by the caller (the
 * C++ adaptor in libmesh_b200/shim/ holds a mutex, mirroring PetscMatrix's _petsc_matrix_mutex,
 * src/numerics/petsc_matrix.C:747).  All floating point is IEEE FP64 (libMesh Real = double,
 * include/base/libmesh_common.h:144); dof ids are uint32 (dof_id_type, include/base/id_types.h:66-76);
 * CSR offsets are uint64 so a 180 GB value array stays addressable.
 *
 * There is no CPU fallback: every entry point that computes fails with FEASM_ERR_CUDA when no sm_100
 * device is usable.
 */
#ifndef FEASM_H
#define FEASM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEASM_VERSION 1

/* element types; the value is the node count (libMesh ElemType HEX8 / HEX20 / HEX27 / TET10 / PRISM18) */
#define FEASM_HEX8 8
#define FEASM_HEX27 27
#define FEASM_TET10 10
#define FEASM_PRISM18 18 /* triangular prism, TRI6 x EDGE3 (cell_prism18.C, fe_lagrange_shape_3D.C:1077-1094); volume terms only */
#define FEASM_HEX20 20 /* serendipity hex (cell_hex20.C, fe_lagrange_shape_3D.C:848-925); QUAD8 sides (fe_lagrange_shape_2D.C:434-471) */

/* problems (the Ke/Fe contraction that runs per element) */
#define FEASM_POISSON 1    /* introduction_ex4.C:464-537: Ke = int grad phi_i . grad phi_j, Fe = int f phi_i */
#define FEASM_ELASTICITY 2 /* systems_of_equations_ex6.C:220-237: 3-variable isotropic elasticity + body force */

/* forcing f(x_q) for FEASM_POISSON */
#define FEASM_FORCING_NONE 0      /* Laplace: Fe = 0 before constraints */
#define FEASM_FORCING_EX4_FD 1    /* introduction_ex4.C:500-537, the FD Laplacian of exact_solution, on device */
#define FEASM_FORCING_TABULATED 2 /* caller-supplied f[n_elem][nq] (the analogue of user code evaluating f) */

/* constraint flavour */
#define FEASM_CONSTRAIN_HETEROGENEOUS 0 /* dof_map_constraints.C:2518-2614 (ex4) */
#define FEASM_CONSTRAIN_HOMOGENEOUS 1   /* dof_map_constraints.C:2439-2514 (ex6) */

/* error codes */
#define FEASM_OK 0
#define FEASM_ERR_ARG -1
#define FEASM_ERR_CUDA -2
#define FEASM_ERR_STATE -3
#define FEASM_ERR_DEGENERATE -4 /* det J <= 0 somewhere: libMesh's DegenerateMap, fe_map.C:467-506 */
#define FEASM_ERR_PATTERN -5    /* add to an entry outside the sparsity pattern (sparse_matrix.h:300-303) */
#define FEASM_ERR_NOMEM -6
#define FEASM_ERR_BREAKDOWN -7 /* feasm_cg_solve: p.Ap <= 0 */
#define FEASM_ERR_COMM -8      /* NCCL missing or an NCCL call failed */
#define FEASM_NOT_CONVERGED 1  /* feasm_cg_solve: max_its reached above the tolerance (x, iterations, residual are valid) */

typedef struct feasm_ctx feasm_ctx;

typedef struct feasm_params {
  int problem;           /* FEASM_POISSON | FEASM_ELASTICITY */
  int quadrature_order;  /* libMesh Order of the QGauss rule: 3 (THIRD) or 5 (FIFTH); 0 = the example's default */
  int forcing;           /* FEASM_FORCING_* (Poisson only) */
  int constrain;         /* FEASM_CONSTRAIN_* */
  const double* f_host;  /* FEASM_FORCING_TABULATED: host array [n_elem][nq], copied at the call, or NULL ... */
  const double* f_device;/* ... in which case this device pointer is read in place */
  double young_modulus;  /* elasticity; 0 = the example's E = 1 */
  double poisson_ratio;  /* elasticity; 0 = the example's nu = 0.3 */
  double body_force[3];  /* elasticity; all zero = the example's (0,0,-1) */
  int affine_tensor;     /* Poisson: 1 = elements with an affine map (Elem::has_affine_map) take the reference-tensor
                            contraction Ke = sum_cc' G_cc' N_cc' (same Ke up to rounding, 1/5 of the FP64 work); the others
                            still go through the quadrature loop.  0 = the reference's algorithm for every element */
} feasm_params;

/* ---- lifetime ------------------------------------------------------------------------------------ */
int feasm_create(int device, feasm_ctx** out);
void feasm_destroy(feasm_ctx* ctx);
const char* feasm_last_error(const feasm_ctx* ctx); /* ctx may be NULL: error of the last failed feasm_create */
int feasm_version(void);

/* ---- inputs: what the host libMesh already holds ----------------------------------------------- */
/* Flattened MeshBase: node coordinates (Node = Point, 3 doubles) and element->node connectivity in libMesh's
 * element-local node order (Elem::node_ptr(i)); replaces the Elem->Node pointer chase of FEMap::compute_map
 * (src/fe/fe_map.C:1361-1366). */
int feasm_set_mesh(feasm_ctx* ctx, int elem_type, int64_t n_elem, int64_t n_nodes, const double* xyz,
                   const uint32_t* conn);
/* Only the coordinates change (moving mesh / e2e timing): re-upload xyz. */
int feasm_update_coordinates(feasm_ctx* ctx, const double* xyz);
/* Result of DofMap::dof_indices(elem, di) for every local element (src/base/dof_map.C:2201-2314), variable-major:
 * elem_dofs[e][v*npe + a].  Dofs are LOCAL row indices 0..n_rows-1 of this context's matrix; global_ids (may be
 * NULL = identity) maps them to libMesh's global dof ids, which is what the pattern's col_idx holds.
 * Rows [0, n_owned) are the rows this rank owns ([first_dof, end_dof) in libMesh); rows [n_owned, n_rows) are
 * ghost rows whose contributions are shipped to their owner (DistributedVector::close semantics,
 * src/numerics/distributed_vector.C:399-470). */
int feasm_set_dofs(feasm_ctx* ctx, int n_vars, int64_t n_rows, int64_t n_owned, const uint32_t* elem_dofs,
                   const uint32_t* global_ids);
/* DofMap::is_constrained_dof + get_primal_constraint_values for Dirichlet rows (empty constraint rows):
 * local row index and value g. */
int feasm_set_constraints(feasm_ctx* ctx, int64_t n_constraints, const uint32_t* rows, const double* values);
/* General constraint rows — hanging nodes, periodic dofs: u_dofs[i] = sum_k coefs[k] u_cols[k] + values[i], k in
 * [row_ptr[i], row_ptr[i+1]) (DofConstraints, include/base/dof_map.h:97-123; an empty row is a Dirichlet value, so this call
 * REPLACES feasm_set_constraints).  Elements that touch a dof with a non-empty row are constrained as
 * DofMap::heterogeneously_constrain_element_matrix_and_vector does it (src/base/dof_map_constraints.C:2519-2614 with
 * build_constraint_matrix_and_vector, :3543-3673, asymmetric_constraint_rows = true): the element's dof list grows by the
 * constraining dofs, K <- C^T K C, F <- C^T (F - K H).  Call it BEFORE feasm_build_pattern (the device-built pattern then
 * holds the couplings DofMap::find_connected_dofs adds, dof_map.C:2915-2975) or pass libMesh's own Graph to
 * feasm_set_pattern.  Rows are expected as DofMap::process_constraints leaves them before any assembly (no constraining
 * dof with a NON-empty row of its own is required — those are substituted, :3595-3673 — but a constraining dof that is itself
 * a Dirichlet dof keeps, in FEASM_CONSTRAIN_HOMOGENEOUS mode, right-hand side 0 here where :2439-2514 would leave (C^T F)_i).
 * One rank per context; feasm_assemble only (not begin / end, not feasm_residual). */
int feasm_set_constraint_rows(feasm_ctx* ctx, int64_t n, const uint32_t* dofs, const uint64_t* row_ptr, const uint32_t* cols,
                              const double* coefs, const double* values);

/* Side (face) loads — the boundary-integral part of the assemble loop, systems_of_equations_ex6.C:240-257:
 *   for the listed (element, side) pairs:  fe_face->reinit(elem, side)  (src/fe/fe_boundary.C:109-240, QGauss(dim-1) of
 *   the volume rule's order)  and  Fe_var[i](a) += JxW_face[qp] * (g[i] * phi_face[a][qp]),
 * added to Fe after the volume terms and before the element constraint in every following feasm_assemble /
 * feasm_element_matrices.  The adaptor fills the list from BoundaryInfo::has_boundary_id(elem, side, id) for sides
 * without a neighbour.  g: n_g = number of variables components (a traction; n_g = 1: a Neumann flux).  HEX8 / HEX20 / HEX27
 * (QUAD4 / QUAD8 / QUAD9 sides) and TET10 (TRI6 sides, QGauss(2, FIFTH) = the 7-point rule of quadrature_gauss_2D.C:234-257).
 * n = 0 clears. */
int feasm_set_side_loads(feasm_ctx* ctx, int64_t n, const uint32_t* elems, const uint8_t* sides, const double* g, int n_g);

/* Dirichlet data by the penalty method — the boundary loop of introduction_ex3.C:386-455 on the listed (element, side)
 * pairs (there: every side with neighbor_ptr(side) == nullptr):
 *   Ke(i,j) += JxW_face[qp]*penalty*phi_face[i][qp]*phi_face[j][qp];   Fe(i) += JxW_face[qp]*penalty*value*phi_face[i][qp];
 * value = the given constant or the exact solution at the side point (introduction_ex3/exact_solution.C:40-47); with
 * several variables every variable's diagonal block gets the term.  Added after the volume terms and the side loads,
 * before the element constraint.  HEX8 / HEX20 / HEX27 / TET10.  n = 0 clears. */
#define FEASM_PENALTY_CONSTANT 0
#define FEASM_PENALTY_EX4_EXACT 1
int feasm_set_side_penalty(feasm_ctx* ctx, int64_t n, const uint32_t* elems, const uint8_t* sides, double penalty, int kind, double value);

/* ---- sparsity pattern ---------------------------------------------------------------------------- */
/* Host graph route = SparseMatrix::update_sparsity_pattern(const SparsityPattern::Graph&)
 * (include/numerics/sparse_matrix.h:175): CSR over the n_rows local rows, columns = global ids ascending. */
int feasm_set_pattern(feasm_ctx* ctx, const uint64_t* row_ptr, const uint32_t* col_idx);
/* Device route = SparsityPattern::Build (src/base/sparsity_pattern.C:96-363) on the GPU from elem_dofs. */
int feasm_build_pattern(feasm_ctx* ctx);
int64_t feasm_nnz(const feasm_ctx* ctx);
int feasm_get_pattern(const feasm_ctx* ctx, uint64_t* row_ptr, uint32_t* col_idx);
/* Builds the element->CSR-slot map and the per-row incidence lists the deterministic scatter uses.  Called
 * implicitly by the first feasm_assemble. */
int feasm_prepare(feasm_ctx* ctx);

/* ---- the hot path ------------------------------------------------------------------------------ */
/* Replaces the whole user assemble loop (the function handed to System::attach_assemble_function,
 * include/systems/system.h:1470): zeroes the matrix and RHS (ImplicitSystem::assemble,
 * src/systems/implicit_system.C:101-105) and assembles every local element.  Asynchronous on the context's
 * stream; feasm_synchronize or any getter waits. */
int feasm_assemble(feasm_ctx* ctx, const feasm_params* params);
int feasm_synchronize(feasm_ctx* ctx);
/* The same assembly in two halves for multi-rank runs ("exchange while the interior assembles"; the role of
 * MatAssemblyBegin/End around the stash, petsc_matrix.C close()): _begin runs the element phase, reduces the rows this
 * rank does NOT own (local rows n_owned..n_rows) and packs them into the send buffer (feasm_pack_ghost's buffer; it
 * must not be packed again); the caller starts its ncclSend/ncclRecv group, then _end reduces the owned rows while
 * the transfer is in flight; feasm_unpack_ghost after both.  Results are identical to feasm_assemble + pack. */
/* "Assemble interface elements first, exchange while the interior assembles" (SURVEY 8e): the caller orders its local
 * elements so that those touching a ghost row come first and names the length of that prefix; feasm_assemble_begin then
 * assembles only the prefix before it reduces and packs the ghost rows, and feasm_assemble_end assembles the rest (under
 * the transfer) before the owned rows.  0 (the default) = every element in _begin.  With LinearPartitioner slabs the
 * prefix is the first element layer of the slab without any reordering.  Results do not depend on it. */
int feasm_set_interface_prefix(feasm_ctx* ctx, int64_t n_elems);
int feasm_assemble_begin(feasm_ctx* ctx, const feasm_params* params);
int feasm_assemble_end(feasm_ctx* ctx);
double* feasm_send_buffer(feasm_ctx* ctx, uint64_t* n_doubles); /* the packed buffer, without packing again */
/* Element-level output for tests: Ke [e1-e0][nd*nd] row-major and Fe [e1-e0][nd], constraints applied iff
 * constrained != 0; host pointers. */
int feasm_element_matrices(feasm_ctx* ctx, const feasm_params* params, int64_t e0, int64_t e1, int constrained,
                           double* Ke, double* Fe);

/* SparseMatrix::zero / NumericVector::zero */
int feasm_zero(feasm_ctx* ctx);
/* SparseMatrix::add_matrix(dm, rows, cols) (sparse_matrix.h:312) and NumericVector::add_vector
 * (numeric_vector.h:519) for a single host element matrix — the slow per-element route that keeps the virtuals
 * of the adaptor classes working.  Rows/cols are local row indices / global column ids. */
int feasm_add_matrix(feasm_ctx* ctx, const double* dm, const uint32_t* rows, int n_rows, const uint32_t* cols,
                     int n_cols);
int feasm_add_vector(feasm_ctx* ctx, const double* v, const uint32_t* rows, int n);
/* SparseMatrix::set(i, j, value) (sparse_matrix.h:295) for a block: A(rows[i], cols[j]) = dm(i,j). */
int feasm_set_matrix(feasm_ctx* ctx, const double* dm, const uint32_t* rows, int n_rows, const uint32_t* cols, int n_cols);
/* A context that holds a matrix only — pattern, values, right-hand side, no mesh: SparseMatrix::init (sparse_matrix.h:189)
 * with a known Graph, the target of clone() / get_transpose().  The per-element virtuals, feasm_zero, the getters,
 * feasm_set_values and feasm_vector_mult work on it; feasm_assemble needs feasm_set_mesh / feasm_set_dofs. */
int feasm_init_matrix(feasm_ctx* ctx, int64_t n_rows, int64_t n_owned, const uint64_t* row_ptr, const uint32_t* col_idx);
/* H2D upload of the CSR values and / or the right-hand side (either may be NULL): the inverse of feasm_get_values. */
int feasm_set_values(feasm_ctx* ctx, const double* csr_values, const double* rhs);
/* feasm_add_matrix without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (tight serving loops, harnesses that forbid it, code meant for CUDA
 * graph capture).  Same arguments, same result as feasm_add_matrix; the staging buffers are a fixed file-scope device
 * array of FEASM_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until the process exits.  Blocks with
 * n_rows or n_cols > FEASM_NOALLOC_MAX_DIM are rejected with FEASM_ERR_ARG (use feasm_add_matrix).  The call performs no
 * cudaMalloc / cudaFree; it blocks until the add has completed.  All calls on one device share the scratch: calls that
 * overlap on the GPU (two contexts of one device driven concurrently) race — serialise them per device. */
#define FEASM_NOALLOC_MAX_DIM 128
#define FEASM_NOALLOC_SCRATCH_BYTES (8 * 128 * 128 + 2 * 4 * 128)
int feasm_add_matrix_noalloc(feasm_ctx* ctx, const double* dm, const uint32_t* rows, int n_rows, const uint32_t* cols,
                             int n_cols);

/* ---- outputs ------------------------------------------------------------------------------------ */
int feasm_get_values(feasm_ctx* ctx, double* csr_values, double* rhs); /* D2H, either may be NULL */
double* feasm_device_values(feasm_ctx* ctx);                           /* device pointers, valid until destroy */
double* feasm_device_rhs(feasm_ctx* ctx);
void* feasm_stream(feasm_ctx* ctx);                                    /* cudaStream_t */

/* ---- the step after the path: SpMV and a Krylov solve on the device matrix ----- */
/* SparseMatrix::vector_mult (sparse_matrix.h:579-587): y = A x, device pointers of n_owned doubles (= n_rows on one rank),
 * on feasm_stream.  On several ranks (feasm_build_halo below) x and y are the owned parts of distributed vectors and the
 * call fetches the entries of x other ranks own before the product. */
int feasm_vector_mult(feasm_ctx* ctx, const double* d_x, double* d_y);
int feasm_vector_mult_host(feasm_ctx* ctx, const double* x, double* y); /* same with host arrays (H2D, product, D2H) */
/* LinearSolver::solve with CG + Jacobi preconditioning on the assembled system A x = rhs.  x0 (host, may be NULL =
 * zero) is the initial guess, x (host) receives the solution; stops at ||r|| <= rel_tol ||b|| or max_its.  Returns
 * FEASM_OK (converged), FEASM_NOT_CONVERGED (max_its reached; outputs valid) or FEASM_ERR_BREAKDOWN.  On several ranks
 * (collective; communicator + feasm_build_halo) x0 / x are the owned parts, every dot product is finished by an all-gather
 * of the ranks' partial sums added in rank order, and all ranks return the same iteration count and residual. */
int feasm_cg_solve(feasm_ctx* ctx, double rel_tol, int max_its, const double* x0, double* x, int* iterations,
                   double* rel_residual);

/* Residual without the global matrix: R = K(u) u - F summed straight from the constrained element
 * matrices, r_i = sum over the (element, local dof) pairs of row i of ( Ke(a,:) . u_e - Fe(a) ) in element order — the
 * work of NonlinearImplicitSystem::ComputeResidual::residual(X, R, S) (include/systems/nonlinear_implicit_system.h:
 * 90-101) / FEMSystem::assembly(get_residual = true, get_jacobian = false) (src/systems/fem_system.C:875) for the
 * linear problems of this library: no CSR values are written or read.  d_u, d_r: device vectors of n_rows entries
 * (feasm_vec_alloc; on several ranks the owned parts, n_owned entries).  Runs the element phase, then one row-parallel
 * pass over the element records.  On several ranks (collective; communicator + feasm_build_halo) the ghost entries of u come
 * from their owners first and the sums of the ghost rows go to their owners, who add them peer by peer in the fixed order
 * of the assembly's exchange plan (DistributedVector::close, distributed_vector.C:399-470). */
int feasm_residual(feasm_ctx* ctx, const feasm_params* params, const double* d_u, double* d_r);

/* ---- device vectors: what a NumericVector back-end needs (numeric_vector.h:119-847; shim/gpu_numeric_vector.h) -- */
/* Blocking calls on feasm_stream.  d_* are device pointers from feasm_vec_alloc (or feasm_device_rhs). */
#define FEASM_COPY_H2D 0
#define FEASM_COPY_D2H 1
#define FEASM_COPY_D2D 2
#define FEASM_VEC_MULT 0        /* pointwise: out = a .* b        (NumericVector::pointwise_mult, operator*=) */
#define FEASM_VEC_DIV 1         /*            out = a ./ b        (pointwise_divide, operator/=) */
#define FEASM_VEC_ABS 2         /*            out = |a|           (abs) */
#define FEASM_VEC_RECIPROCAL 3  /*            out = 1 ./ a        (reciprocal) */
#define FEASM_VEC_ADD_SCALAR 4  /*            out = a + s         (add(T)) */
#define FEASM_VEC_SCALE 5       /*            out = s a           (scale) */
#define FEASM_VEC_SUM 10        /* reductions (fixed tree, deterministic): sum, min, max, l1_norm, linfty_norm, dot */
#define FEASM_VEC_MIN 11
#define FEASM_VEC_MAX 12
#define FEASM_VEC_L1 13
#define FEASM_VEC_LINF 14
#define FEASM_VEC_DOT 15
double* feasm_vec_alloc(feasm_ctx* ctx, int64_t n); /* zero-initialised; NULL on failure */
void feasm_vec_free(feasm_ctx* ctx, double* d_v);
int feasm_vec_copy(feasm_ctx* ctx, int kind, void* dst, const void* src, int64_t n);
int feasm_vec_fill(feasm_ctx* ctx, double* d_y, int64_t n, double value);
int feasm_vec_axpy(feasm_ctx* ctx, double* d_y, double a, const double* d_x, int64_t n); /* y += a x */
int feasm_vec_pointwise(feasm_ctx* ctx, int op, double* d_out, const double* d_a, const double* d_b, double s, int64_t n);
int feasm_vec_reduce(feasm_ctx* ctx, int op, const double* d_x, const double* d_y, int64_t n, double* result);
int feasm_vec_set_entry(feasm_ctx* ctx, double* d_y, int64_t i, double value, int add); /* set(i,v) / add(i,v) */

/* ---- multi-GPU ghost rows (DistributedVector::close / MatAssemblyBegin-End) -------------------- */
/* After feasm_prepare on every rank: describe, per peer, which of this rank's ghost-row slots go to the peer
 * and which owned slots receive the peer's contributions.  send_slots / recv_slots index the CSR value array
 * (values) or, when >= nnz, the RHS (slot - nnz).  The plan is computed by the host (libmesh_b200.partition). */
int feasm_set_exchange_plan(feasm_ctx* ctx, int n_peers, const int32_t* peer_rank, const uint64_t* send_ptr,
                            const uint64_t* send_slots, const uint64_t* recv_ptr, const uint64_t* recv_slots);
/* Packs the ghost contributions into one contiguous device send buffer (peer-major); returns its pointer.
 * The collective itself (ncclSend/ncclRecv group = all-to-all-v) is issued by the host on feasm_stream. */
double* feasm_pack_ghost(feasm_ctx* ctx, uint64_t* n_doubles);
double* feasm_recv_buffer(feasm_ctx* ctx, uint64_t* n_doubles);
/* Adds the received buffer into the owned slots in fixed peer order (deterministic). */
int feasm_unpack_ghost(feasm_ctx* ctx);

/* The exchange itself, below the C ABI (feasm_create(device, ncclComm_t) / feasm_exchange_ghost_rows).  NCCL is
 * bound at run time (libnccl.so.2 of the process, e.g. the one an MPI + NCCL host or torch already loaded).
 *   feasm_set_communicator : hand in an ncclComm_t the host owns (one rank of the communicator per context; NULL = back to
 *                            one rank).  The role of ParallelObject::comm() for the GpuSparseMatrix.
 *   feasm_comm_unique_id / feasm_comm_init : for hosts without an NCCL communicator of their own: rank 0 makes the
 *                            FEASM_COMM_ID_BYTES bootstrap id (ncclGetUniqueId), the host broadcasts the bytes by whatever
 *                            means it has (MPI_Bcast, a file, torch.distributed), every rank calls feasm_comm_init
 *                            (ncclCommInitRank; collective).  The context owns and destroys that communicator.
 *   feasm_exchange_begin   : after feasm_assemble_begin (or feasm_assemble + feasm_pack_ghost): one grouped ncclSend /
 *                            ncclRecv per neighbour of the plan, on the context's high-priority communication stream,
 *                            ordered behind the pack by an event.  Returns at once; the caller goes on to feasm_assemble_end.
 *   feasm_exchange_end     : makes feasm_stream wait for the transfer and adds the received rows (feasm_unpack_ghost).
 *   feasm_exchange_ghost_rows = begin + end (DistributedVector::close, distributed_vector.C:399-470).
 *   feasm_comm_allgather_u64 : setup helper (dof ownership ranges, counts): n values from every rank to every rank. */
#define FEASM_COMM_ID_BYTES 128
int feasm_set_communicator(feasm_ctx* ctx, void* nccl_comm, int n_ranks, int rank);
int feasm_comm_unique_id(void* id_bytes);
int feasm_comm_init(feasm_ctx* ctx, int n_ranks, int rank, const void* id_bytes);
int feasm_exchange_begin(feasm_ctx* ctx);
int feasm_exchange_end(feasm_ctx* ctx);
int feasm_exchange_ghost_rows(feasm_ctx* ctx);
int feasm_comm_allgather_u64(feasm_ctx* ctx, const uint64_t* send, int n, uint64_t* recv);

/* The SETUP of a multi-rank assembly (SparsityPattern::Build::parallel_sync, src/base/sparsity_pattern.C:446-575, plus the
 * fixed routes DistributedVector::close, src/numerics/distributed_vector.C:399-470, walks every assembly).  Local numbering
 * is the host's job (feasm_set_dofs: owned rows first = global id - first_dof[rank], then the ghost rows in ascending
 * global id; the pattern's column ids are GLOBAL dof ids).
 *   feasm_build_exchange : collective over the context's communicator, after feasm_build_pattern.  first_dof[n_ranks + 1] =
 *                          the dof ranges of the ranks (DofMap::first_dof(p), end_dof(last)).  Ships the column lists of the
 *                          ghost rows to their owners, unites them with the owned rows (feasm_set_pattern), fixes the slot
 *                          plan (feasm_set_exchange_plan), builds the scatter maps (feasm_prepare) and picks the interface
 *                          prefix (feasm_set_interface_prefix; environment FEASM_IFACE=0 keeps it off).
 *   feasm_planner_*      : the same logic for hosts that carry the messages themselves (libMesh: TIMPI push_parallel_vector_data
 *                          over MPI; no device and no communicator needed):
 *                            create(rank, n_ranks, first_dof, n_owned, n_rows, global_ids, row_ptr, col_idx) -> NULL on bad input
 *                            message(peer, &words) -> number of uint32 words to send to `peer` (0: none)
 *                            receive(peer, words, n)   the message a peer made for this rank
 *                            finish()                  merged pattern + slot plan
 *                            sizes(out[5])             nnz, peers, send slots, receive slots, pattern grew (0 / 1)
 *                            get(...)                  copies out whatever pointer is non-NULL
 *                          and feasm_apply_plan(ctx, planner) = feasm_set_pattern (if it grew) + feasm_set_exchange_plan +
 *                          feasm_prepare. */
typedef struct feasm_planner feasm_planner;
int feasm_build_exchange(feasm_ctx* ctx, const uint64_t* first_dof);
feasm_planner* feasm_planner_create(int rank, int n_ranks, const uint64_t* first_dof, uint64_t n_owned, uint64_t n_rows,
                                    const uint32_t* global_ids, const uint64_t* row_ptr, const uint32_t* col_idx);
void feasm_planner_destroy(feasm_planner* p);
const char* feasm_planner_error(const feasm_planner* p);
uint64_t feasm_planner_message(feasm_planner* p, int peer, const uint32_t** words);
int feasm_planner_receive(feasm_planner* p, int peer, const uint32_t* words, uint64_t n);
int feasm_planner_finish(feasm_planner* p);
int feasm_planner_sizes(const feasm_planner* p, uint64_t sizes[5]);
int feasm_planner_get(const feasm_planner* p, uint64_t* row_ptr, uint32_t* col_idx, int32_t* peers, uint64_t* send_ptr,
                      uint64_t* send_slots, uint64_t* recv_ptr, uint64_t* recv_slots);
int feasm_apply_plan(feasm_ctx* ctx, const feasm_planner* p);
/* number of peers of the plan the context holds (>= 0); peers[n], send_ptr[n + 1], send_slots, recv_ptr[n + 1], recv_slots
 * are copied out when non-NULL */
int feasm_get_exchange_plan(feasm_ctx* ctx, int32_t* peers, uint64_t* send_ptr, uint64_t* send_slots, uint64_t* recv_ptr,
                            uint64_t* recv_slots);

/* ---- the assembled matrix of several ranks used on the device: the halo of a distributed vector ---------------------------
 * The owned rows hold GLOBAL column ids, so y = A x needs the entries of x that other ranks own: what
 * NumericVector::localize(v_local, send_list) (numeric_vector.h:644-650; DistributedVector::localize,
 * src/numerics/distributed_vector.C:479-521) fetches with the DofMap's send_list.  The library keeps a column-space copy of x
 * of global length (n_global doubles) per context: owned part copied, ghost entries scattered in.
 *   feasm_build_halo     : collective over the context's communicator, after the final pattern (feasm_build_exchange):
 *                          finds the needed columns on the device (columns of the owned rows outside the owned range, plus
 *                          the ids of the ghost rows for feasm_residual), asks their owners once.  Afterwards
 *                          feasm_vector_mult, feasm_cg_solve and feasm_residual work on the distributed system.
 *   feasm_halo_columns   : the needed global ids, ascending (returns the count; gids may be NULL) — for hosts that carry
 *                          the request lists themselves (libMesh: the send_list is exactly this set) ...
 *   feasm_set_halo_plan  : ... and hand the routes in: per peer the owned local rows it needs (send_rows) and the global
 *                          ids this rank receives from it (recv_gids), in the order of the peer's request.
 *   feasm_halo_pack / _recv_buffer / _unpack : the per-product exchange in pieces for such hosts: pack x's requested
 *                          entries (returns the send buffer, peer-major), the host moves the segments into the peers'
 *                          receive buffers, unpack scatters them into the column-space copy; feasm_vector_mult without
 *                          a communicator then uses that copy (and refreshes the owned part itself).
 *   feasm_get_halo_plan  : the plan the context holds (returns the number of peers). */
int feasm_build_halo(feasm_ctx* ctx, const uint64_t* first_dof);
int64_t feasm_halo_columns(feasm_ctx* ctx, uint64_t n_global, uint64_t first, uint32_t* gids);
int feasm_set_halo_plan(feasm_ctx* ctx, uint64_t n_global, uint64_t first, int n_peers, const int32_t* peer_rank,
                        const uint64_t* send_ptr, const uint32_t* send_rows, const uint64_t* recv_ptr, const uint32_t* recv_gids);
double* feasm_halo_pack(feasm_ctx* ctx, const double* d_x, uint64_t* n_doubles);
double* feasm_halo_recv_buffer(feasm_ctx* ctx, uint64_t* n_doubles);
int feasm_halo_unpack(feasm_ctx* ctx);
int feasm_get_halo_plan(feasm_ctx* ctx, int32_t* peers, uint64_t* send_ptr, uint32_t* send_rows, uint64_t* recv_ptr,
                        uint32_t* recv_gids);

/* ---- host-side generators for stand-alone runs (no libMesh present) ---------------------------- */
/* MeshTools::Generation::build_cube (src/mesh/mesh_generation.C:317-1937) for HEX8/HEX20/HEX27/TET10/PRISM18: fills caller arrays.
 * Sizes: feasm_cube_sizes. */
int feasm_cube_sizes(int elem_type, int nx, int ny, int nz, int64_t* n_elem, int64_t* n_nodes);
int feasm_build_cube(int elem_type, int nx, int ny, int nz, const double bounds[6], double* xyz, uint32_t* conn,
                     uint8_t* node_boundary_mask);
/* DofMap::distribute_dofs first-touch numbering (src/base/dof_map.C:1417-1535) for n_parts ranks with the given
 * element owners (NULL = one rank); returns n_dofs, fills node_dof_base[n_nodes] and first_dof[n_parts+1]. */
int64_t feasm_distribute_dofs(int npe, int64_t n_elem, int64_t n_nodes, const uint32_t* conn, int n_vars, int n_parts,
                              const int32_t* elem_pid, uint32_t* node_dof_base, uint32_t* first_dof, int32_t* node_pid);

/* ---- device-side mesh ingest (the step before the path, without the host) --------------- */
/* MeshTools::Generation::build_cube (src/mesh/mesh_generation.C:1285-1340 nodes, :1399-1432 elements, :1779-1785
 * rescale, :1416-1432 boundary ids; TET10: the 24-tet split + all_second_order of :1795-1907) for HEX8 / HEX27 / TET10,
 * generated in the context's device arrays (HEX20, :1559-1592, and PRISM18, :1634-1756: made by feasm_build_cube on the
 * host and uploaded); replaces
 * feasm_build_cube + feasm_set_mesh.  Same node ids, element order and coordinates bit for bit. */
int feasm_generate_cube(feasm_ctx* ctx, int elem_type, int nx, int ny, int nz, const double bounds[6]);
/* DofMap::distribute_dofs (src/base/dof_map.C:1445-1507, one rank, one variable group of n_vars LAGRANGE variables)
 * + DofMap::dof_indices (:2280-2308) for every element, on the device, for whatever mesh the context holds
 * (generated or feasm_set_mesh); replaces feasm_distribute_dofs + feasm_set_dofs.  Returns n_dofs (< 0: error code). */
int64_t feasm_distribute_dofs_device(feasm_ctx* ctx, int n_vars);
/* DirichletBoundary on the boundary ids in id_mask (bit b = id b) for the variables in var_mask, nodal interpolation
 * of the boundary function (DofMap::create_dof_constraints, src/base/dof_map_constraints.C:1736);
 * replaces feasm_set_constraints for generated meshes.  Returns the number of constrained dofs (< 0: error code). */
#define FEASM_DIRICHLET_ZERO 0      /* g = 0 (ex6 clamped sides) */
#define FEASM_DIRICHLET_EX4_EXACT 1 /* g = cos(pi x/2) sin(pi y/2) cos(pi z/2), introduction_ex3/exact_solution.C:40-47 */
int64_t feasm_generate_dirichlet(feasm_ctx* ctx, unsigned id_mask, int kind, unsigned var_mask);
/* Read-backs (D2H; any pointer may be NULL): what the context holds, however it got there. */
int feasm_get_mesh(feasm_ctx* ctx, double* xyz, uint32_t* conn, uint8_t* node_boundary_mask);
int feasm_get_dofs(feasm_ctx* ctx, uint32_t* elem_dofs, uint32_t* node_dof_base);
int feasm_get_constraints(feasm_ctx* ctx, uint8_t* flag, double* value);

/* ---- instrumentation ---------------------------------------------------------------------------- */
/* Milliseconds of the kernels of the last feasm_assemble, measured with CUDA events on the context's stream:
 * [0] element kernel, [1] scatter/row reduction, [2] whole call, [3] element-level constraint pass.  Also the
 * number of kernel launches. */
int feasm_last_timing(feasm_ctx* ctx, float ms[4], int* launches);
/* FP64 pipe peaks of this device, measured now (register-resident DFMA chain and mma.sync.m8n8k4.f64 chain), in
 * TFLOP/s: the FP64 denominators of the roofline bench.py reports (MEASURED_PEAKS.json has no FP64 entry). */
int feasm_measure_fp64_peaks(feasm_ctx* ctx, double* dfma_tflops, double* dmma_tflops);

#ifdef __cplusplus
}
#endif
#endif /* FEASM_H */
