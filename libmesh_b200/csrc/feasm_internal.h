// feasm_internal.h — context and internal interfaces of libfeasm (product code; never includes oracle/).
#ifndef FEASM_INTERNAL_H
#define FEASM_INTERNAL_H

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "feasm.h"

// Reference-element tables for one (element type, quadrature order): what FE::init_shape_functions
// (src/fe/fe.C:440-732) and FEMap::init_reference_to_physical_map (src/fe/fe_map.C:108-444) cache per
// element TYPE in libMesh.  Layout [i][q], q fastest, so lane = q reads are bank-conflict free.
struct FeasmTables {
  int npe = 0, nq = 0;
  std::vector<double> qp;     // nq*3
  std::vector<double> qw;     // nq
  std::vector<double> phi;    // npe*nq
  std::vector<double> dxi, deta, dzeta;
};
bool feasm_make_tables(int elem_type, int qorder, FeasmTables& t, std::string& err);
// The same for the sides of an element type (FE::reinit(elem, side), src/fe/fe_boundary.C:109-240)
struct FeasmSideTables {
  int npe = 0, ns = 0, nqf = 0;    // parent nodes, side nodes, side quadrature points
  std::vector<int> side_nodes;     // [6][ns] parent-local node of side node i
  std::vector<double> qw;          // nqf
  std::vector<double> psi, dxi, deta;  // [ns][nqf] the side's map functions and derivatives
  std::vector<double> phi;         // [6][npe][nqf] parent shapes at the side points
};
bool feasm_make_side_tables(int elem_type, int qorder, FeasmSideTables& t, std::string& err);

// Element records REC[e][a], feasm_recb bytes each: row a of element e = Ke(a, 0..nd-1) followed by Fe(a) at column
// nd, padded to a whole number of 32 B sectors (HEX27: 28 doubles = 224 B), then the position-map row of the (e,a)
// pair (HEX27: 32 B; 256 B per record).  One record is everything one (element, local dof) pair contributes to one CSR
// row — values written by the element kernel every assembly, positions written once by feasm_prepare — so the row
// reduction fetches it as ONE contiguous, sector-aligned run.
#ifdef __CUDACC__
#define FEASM_HD __host__ __device__
#else
#define FEASM_HD
#endif
FEASM_HD constexpr int feasm_ldr(int nd) { return (nd + 1 + 3) & ~3; }
// Position-map row of a record, PLD entries (padded to whole sectors).  Entries 0..nd-1 = column position inside the
// CSR row; entry nd = the image cell that collects Fe (= max_row_len, the cell behind the longest row); entry nd+1 =
// flag of the (e,a) pair (RR_LAST, feasm_rr_dev.cuh).  uint8 when every row has <= 255 entries, else uint16.
FEASM_HD constexpr int feasm_pld(int nd, int pos_bytes) { return (((nd + 2) * pos_bytes + 31) & ~31) / pos_bytes; }
FEASM_HD constexpr int feasm_recb(int nd, int pos_bytes) { return feasm_ldr(nd) * 8 + feasm_pld(nd, pos_bytes) * pos_bytes; }

struct feasm_general;  // feasm_general.cu: general (non-Dirichlet) constraint rows
struct feasm_halo;     // feasm_halo.cu: column-space copy of a distributed vector (product / solve / residual on several ranks)

struct feasm_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  std::string err;
  int sm_count = 148;
  size_t smem_optin = 0;

  // mesh
  int elem_type = 0, npe = 0;
  int64_t n_elem = 0, n_nodes = 0;
  double* d_xyz = nullptr;
  uint32_t* d_conn = nullptr;
  uint8_t* d_node_mask = nullptr;   // [n_nodes] bit b = node lies on boundary id b (feasm_generate_cube only)
  uint32_t* d_node_base = nullptr;  // [n_nodes] first dof of the node (feasm_distribute_dofs_device only)

  // dofs
  int n_vars = 0, nd = 0;
  int64_t n_rows = 0, n_owned = 0;
  uint32_t* d_elem_dofs = nullptr;
  uint32_t* d_gid = nullptr;  // nullptr = identity

  // constraints
  uint8_t* d_con_flag = nullptr;
  double* d_con_val = nullptr;
  int64_t n_con = 0;
  std::vector<uint32_t> h_con_rows;  // the constrained rows as last given (detects a changed constrained set)
  uint32_t* d_con_rows = nullptr;    // [2 * cap]: the current (row) list | the previous one (to clear its flags)
  double* d_con_vals_list = nullptr; // [cap] the values of the current list
  size_t d_con_rows_cap = 0;
  int64_t n_con_dev_prev = 0;        // length of the list the device flags currently reflect
  bool con_flags_untracked = false;  // flags were set on the device without a list (feasm_generate_dirichlet)

  // pattern + system
  uint64_t nnz = 0;
  uint64_t* d_row_ptr = nullptr;
  uint32_t* d_col_idx = nullptr;
  double* d_values = nullptr;
  double* d_rhs = nullptr;
  uint32_t max_row_len = 0;

  // scatter maps (built by feasm_prepare)
  bool prepared = false;
  int pos_bytes = 2;
  uint8_t* d_elem_flag = nullptr;  // [n_elem] element touches a constrained dof (feasm_launch_elem_flags)
  bool elem_flag_valid = false;    // reflects the current constrained set
  uint32_t* d_con_elems = nullptr; // the flagged elements, compacted (any order)
  int64_t n_con_elems = 0;
  uint64_t* d_inc_ptr = nullptr;  // [n_rows+1]
  uint32_t* d_inc = nullptr;      // [n_elem*nd] incidence ids e*nd + a, ascending within a row
  uint32_t max_inc = 0;

  // elements the affine reference-tensor kernel left to the quadrature kernel (feasm_params.affine_tensor)
  uint32_t* d_na_list = nullptr;  // [n_elem] element ids + the counter behind them

  // element buffers
  double* d_kf = nullptr;  // element records [n_elem][nd][feasm_recb bytes]
  double* d_y = nullptr;   // feasm_residual: y(e,a) = Ke(a,:) . u_e - Fe(a), [n_elem][nd]
  size_t d_y_cap = 0;
  int* d_flags = nullptr;  // [0] degenerate count, [1] pattern misses, [2] ticket counter of the element kernel

  // tables on device, keyed by quadrature order
  int tab_qorder = -1;
  int tab_nq = 0;
  double* d_tab = nullptr;  // phi | dxi | deta | dzeta | qw  (4*npe*nq + nq doubles)

  // side loads (feasm_set_side_loads): elements with at least one loaded side, ascending, and their side bits
  int64_t n_side_elems = 0;
  uint32_t* d_side_elems = nullptr;
  uint8_t* d_side_mask = nullptr;
  double side_g[3] = {0., 0., 0.};
  // penalty sides (feasm_set_side_penalty), same list form
  int64_t n_pen_elems = 0;
  uint32_t* d_pen_elems = nullptr;
  uint8_t* d_pen_mask = nullptr;
  double penalty = 0., pen_value = 0.;
  int pen_kind = 0;
  int side_qorder = -1;            // the side tables on the device are for this order
  int side_ns = 0, side_nqf = 0;
  double* d_side_tab = nullptr;    // qw[nqf] | dxi[ns][nqf] | deta[ns][nqf] | phi[6][npe][nqf]
  int* d_side_nodes = nullptr;     // [6][ns]

  // tabulated forcing copy
  double* d_f = nullptr;
  size_t d_f_cap = 0;

  // ghost exchange
  int n_peers = 0;
  std::vector<int32_t> peer_rank;
  std::vector<uint64_t> send_ptr, recv_ptr;
  uint64_t* d_send_slots = nullptr;
  uint64_t* d_recv_slots = nullptr;
  double* d_send_buf = nullptr;
  double* d_recv_buf = nullptr;
  // unpack in ONE launch: the received entries grouped by destination slot, each group in peer order
  uint64_t n_unpack_slots = 0;
  uint64_t* d_unpack_slot = nullptr;  // [n_unpack_slots] destination slot
  uint64_t* d_unpack_seg = nullptr;   // [n_unpack_slots+1] group boundaries in d_unpack_perm
  uint64_t* d_unpack_perm = nullptr;  // [n_recv] index into the receive buffer

  // NCCL exchange below the C ABI (feasm_comm.cu)
  void* comm = nullptr;        // ncclComm_t
  bool comm_owned = false;     // created by feasm_comm_init (destroyed with the context) or handed in by the host
  int comm_rank = 0, comm_size = 1;
  cudaStream_t comm_stream = nullptr;  // high priority
  cudaEvent_t ev_packed = nullptr, ev_exchanged = nullptr;
  bool exchange_open = false;

  // per-element compat route (feasm_add_matrix / feasm_add_vector): staging blocks, kept
  unsigned char* h_stage = nullptr;  // pinned
  unsigned char* d_stage = nullptr;
  size_t stage_cap = 0;
  bool matrix_only = false;  // feasm_init_matrix: pattern + values without a mesh

  feasm_general* gen = nullptr;  // elements widened by hanging-node constraints (feasm_set_constraint_rows)
  feasm_halo* halo = nullptr;    // feasm_build_halo / feasm_set_halo_plan

  float last_ms[4] = {0, 0, 0, 0};
  int last_launches = 0;
  bool assemble_open = false;  // between feasm_assemble_begin and feasm_assemble_end
  int64_t n_iface = 0;         // feasm_set_interface_prefix: elements [0, n_iface) touch ghost rows
  int64_t pending_e0 = -1;     // feasm_assemble_end still has elements [pending_e0, n_elem) to assemble
  feasm_params pending_params{};
  const double* pending_f = nullptr;
};

// Every cudaFree of the library goes through this: a failing free means a stale or foreign pointer — a library bug that
// would otherwise surface later as an unrelated "invalid argument".  It is reported on stderr and the error state cleared.
inline cudaError_t feasm_checked_free(void* p, const char* what, const char* file, int line) {
  if (!p) return cudaSuccess;
  const cudaError_t e = (cudaFree)(p);
  if (e != cudaSuccess) {
    std::fprintf(stderr, "libfeasm: cudaFree(%s = %p) failed at %s:%d: %s\n", what, p, file, line, cudaGetErrorString(e));
    (void)cudaGetLastError();
  }
  return e;
}
#define cudaFree(p) feasm_checked_free((void*)(p), #p, __FILE__, __LINE__)

// -DFEASM_DEBUG_BOUNDS: device-side asserts on every index the kernels take from a precomputed map (position rows of the
// records, ring slots, exchange slots, pattern positions).  A failing assert ends the kernel with cudaErrorAssert and names
// the line; the product build compiles them out.  (make DEBUG_BOUNDS=1 builds libmesh_b200/libfeasm_bounds.so; the GPU suite
// runs against it with FEASM_LIB.)
#ifdef FEASM_DEBUG_BOUNDS
#include <cassert>
#define FEASM_DEV_ASSERT(c) assert(c)
#else
#define FEASM_DEV_ASSERT(c) ((void)0)
#endif

// error helpers
int feasm_fail(feasm_ctx* ctx, int code, const std::string& msg);
inline const char* feasm_basename(const char* path) {
  const char* b = path;
  for (const char* c = path; *c; ++c) if (*c == '/') b = c + 1;
  return b;
}
#define FEASM_CUDA(ctx, call)                                                                         \
  do {                                                                                                \
    cudaError_t e__ = (call);                                                                         \
    if (e__ != cudaSuccess)                                                                           \
      return feasm_fail(ctx, FEASM_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__) +    \
                                                 " (" + feasm_basename(__FILE__) + ":" + std::to_string(__LINE__) + ")"); \
  } while (0)

// feasm_api.cu: bookkeeping shared by the host-array and the device-side ingest
void feasm_reset_mesh(feasm_ctx* ctx, int elem_type, int64_t n_elem, int64_t n_nodes);
int feasm_reset_dofs(feasm_ctx* ctx, int n_vars, int64_t n_rows, int64_t n_owned);
// feasm_pattern.cu
int feasm_build_incidence(feasm_ctx* ctx);
int feasm_device_build_pattern(feasm_ctx* ctx);
int feasm_build_pos(feasm_ctx* ctx);
// feasm_assemble.cu
// d_kf: records of elements [e0, e1), rs doubles per record; constrain_mode = FEASM_CONSTRAIN_* (applied in place by
// the kernel to the elements flagged in d_elem_flag) or < 0 for none
int feasm_launch_element_kernel(feasm_ctx* ctx, const feasm_params& p, const double* d_f, int64_t e0, int64_t e1,
                                double* d_kf, int rs, int constrain_mode);
int feasm_launch_elem_flags(feasm_ctx* ctx);
// d_con_rows[0..n) / d_con_vals_list -> d_con_flag / d_con_val; when set_flags, the n_prev rows of the previous list
// (second half of d_con_rows) are cleared first
int feasm_launch_scatter_constraints(feasm_ctx* ctx, int64_t n, int64_t n_prev, bool set_flags);
// feasm_side.cu: adds the side loads to the Fe column of the records of elements [e0, e1) (no-op without loads)
int feasm_launch_side_loads(feasm_ctx* ctx, int qorder, int64_t e0, int64_t e1, double* d_kf, int rs);
// d_row_list == nullptr: every row; else the n_list rows listed
int feasm_launch_row_reduce(feasm_ctx* ctx, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream, int64_t row0 = 0);
int feasm_launch_row_residual(feasm_ctx* ctx, const double* d_u, double* d_r);
int feasm_launch_pack(feasm_ctx* ctx);
int feasm_launch_unpack(feasm_ctx* ctx);
// feasm_general.cu: general constraint rows (hanging nodes)
void feasm_general_release(feasm_ctx* ctx);
bool feasm_general_active(const feasm_ctx* ctx);
int feasm_general_extend_pattern(feasm_ctx* ctx);                           // after feasm_device_build_pattern
void feasm_general_invalidate_plan(feasm_ctx* ctx);                       // the pattern changed
int feasm_general_prepare(feasm_ctx* ctx);                                  // inside feasm_prepare
int feasm_general_transform(feasm_ctx* ctx, int rs, int constrain_mode);    // after the element kernel
int feasm_general_add(feasm_ctx* ctx);                                      // after the row reduction
// feasm_comm.cu: destroys an owned communicator and the communication stream
void feasm_comm_release(feasm_ctx* ctx);
int feasm_comm_p2p(feasm_ctx* ctx, int n_peers, const int32_t* peer, const uint64_t* send_ptr, const void* d_send,
                   const uint64_t* recv_ptr, void* d_recv, int elem_bytes, cudaStream_t stream);
int feasm_comm_allgather_f64(feasm_ctx* ctx, const double* d_send, double* d_recv, int n, cudaStream_t stream);
// feasm_halo.cu: a matrix whose rows are spread over several ranks (global column ids, owned rows first)
void feasm_halo_release(feasm_ctx* ctx);
bool feasm_halo_ready(const feasm_ctx* ctx);  // a halo plan is set
inline bool feasm_is_distributed(const feasm_ctx* ctx) { return ctx->d_gid || ctx->n_owned != ctx->n_rows || feasm_halo_ready(ctx); }
// the column-space copy of the distributed vector whose owned part is d_x (n_owned doubles), current on ctx->stream when
// the call returns: owned part copied, ghost entries fetched from their owners (NCCL) — or, without a communicator, left as
// the caller's own feasm_halo_pack / transfer / feasm_halo_unpack made them
int feasm_halo_exchange(feasm_ctx* ctx, const double* d_x, const double** d_xfull);
uint32_t feasm_halo_first(const feasm_ctx* ctx);  // global id of local row 0 (0 without a halo plan)
int feasm_residual_distributed(feasm_ctx* ctx, const double* d_u, double* d_r);
// feasm_solve.cu: y[0..n_rows_used) = A x with x indexed by the matrix's column ids
int feasm_spmv_rows(feasm_ctx* ctx, int64_t n_rows_used, const double* d_x, double* d_y);
// mode 0: add, 1: set
int feasm_launch_add_matrix(feasm_ctx* ctx, const double* d_dm, const uint32_t* d_rows, int nr, const uint32_t* d_cols,
                            int nc, int mode = 0);
int feasm_launch_add_vector(feasm_ctx* ctx, const double* d_v, const uint32_t* d_rows, int n);

#endif
