// feasm_api.cu — the C ABI of libfeasm (include/feasm.h): context, uploads, pattern, assemble, outputs.
#include <algorithm>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "feasm_internal.h"

static std::string g_create_err;

int feasm_fail(feasm_ctx* ctx, int code, const std::string& msg) {
  if (ctx) ctx->err = msg; else g_create_err = msg;
  return code;
}

namespace {

template <typename T>
int upload(feasm_ctx* ctx, T** dst, const T* src, size_t n) {
  if (*dst) { cudaFree(*dst); *dst = nullptr; }
  FEASM_CUDA(ctx, cudaMalloc(dst, sizeof(T) * (n ? n : 1)));
  if (n) FEASM_CUDA(ctx, cudaMemcpyAsync(*dst, src, sizeof(T) * n, cudaMemcpyHostToDevice, ctx->stream));
  return FEASM_OK;
}

template <typename T>
void release(T*& p) {
  if (p) cudaFree(p);
  p = nullptr;
}

void drop_maps(feasm_ctx* ctx) {
  ctx->prepared = false;
  release(ctx->d_elem_flag);
  release(ctx->d_na_list);
  release(ctx->d_con_elems);
  ctx->n_con_elems = 0;
  ctx->elem_flag_valid = false;
  release(ctx->d_inc_ptr);
  release(ctx->d_inc);
  release(ctx->d_kf);
  release(ctx->d_y);
  ctx->d_y_cap = 0;
}

int ensure_tables(feasm_ctx* ctx, int qorder) {
  if (ctx->d_tab && ctx->tab_qorder == qorder) return FEASM_OK;
  FeasmTables t;
  std::string err;
  if (!feasm_make_tables(ctx->elem_type, qorder, t, err)) return feasm_fail(ctx, FEASM_ERR_ARG, err);
  std::vector<double> flat;
  flat.insert(flat.end(), t.phi.begin(), t.phi.end());
  flat.insert(flat.end(), t.dxi.begin(), t.dxi.end());
  flat.insert(flat.end(), t.deta.begin(), t.deta.end());
  flat.insert(flat.end(), t.dzeta.begin(), t.dzeta.end());
  flat.insert(flat.end(), t.qw.begin(), t.qw.end());
  int rc = upload(ctx, &ctx->d_tab, flat.data(), flat.size());
  if (rc) return rc;
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->tab_qorder = qorder;
  ctx->tab_nq = t.nq;
  return FEASM_OK;
}

int resolve_qorder(const feasm_ctx* ctx, const feasm_params& p) {
  if (p.quadrature_order) return p.quadrature_order;
  // ex4 hard-codes FIFTH (introduction_ex4.C:352); ex6 uses default_quadrature_order() = 2*order+1
  if (p.problem == FEASM_ELASTICITY) return ctx->elem_type == FEASM_HEX8 ? 3 : 5;
  return 5;
}

int resolve_forcing(feasm_ctx* ctx, const feasm_params& p, int nq, const double** d_f) {
  *d_f = nullptr;
  if (p.problem != FEASM_POISSON || p.forcing != FEASM_FORCING_TABULATED) return FEASM_OK;
  if (p.f_device) { *d_f = p.f_device; return FEASM_OK; }
  if (!p.f_host) return feasm_fail(ctx, FEASM_ERR_ARG, "tabulated forcing requested but no array given");
  const size_t n = (size_t)ctx->n_elem * nq;
  if (ctx->d_f_cap < n) {
    release(ctx->d_f);
    FEASM_CUDA(ctx, cudaMalloc(&ctx->d_f, sizeof(double) * n));
    ctx->d_f_cap = n;
  }
  FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_f, p.f_host, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
  *d_f = ctx->d_f;
  return FEASM_OK;
}

int check_flags(feasm_ctx* ctx) {
  int flags[2] = {0, 0};
  FEASM_CUDA(ctx, cudaMemcpyAsync(flags, ctx->d_flags, sizeof(flags), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (flags[0] || flags[1]) FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_flags, 0, sizeof(flags), ctx->stream));  // reported once
  if (flags[0]) return feasm_fail(ctx, FEASM_ERR_DEGENERATE, "Jacobian at or under tolerance 0 (degenerate map) at " + std::to_string(flags[0]) + " quadrature points");
  if (flags[1]) return feasm_fail(ctx, FEASM_ERR_PATTERN, "add to an entry outside the sparsity pattern");
  return FEASM_OK;
}

}  // namespace

extern "C" {

int feasm_version(void) { return FEASM_VERSION; }

const char* feasm_last_error(const feasm_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int feasm_create(int device, feasm_ctx** out) {
  if (!out) return feasm_fail(nullptr, FEASM_ERR_ARG, "out is NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0)
    return feasm_fail(nullptr, FEASM_ERR_CUDA, std::string("no CUDA device: libfeasm has no CPU fallback (") + cudaGetErrorString(e) + ")");
  if (device < 0 || device >= n) return feasm_fail(nullptr, FEASM_ERR_ARG, "bad device ordinal");
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return feasm_fail(nullptr, FEASM_ERR_CUDA, cudaGetErrorString(e));
  if (prop.major != 10) return feasm_fail(nullptr, FEASM_ERR_CUDA, "libfeasm is built for sm_100a (B200) only; device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor));
  if ((e = cudaSetDevice(device)) != cudaSuccess) return feasm_fail(nullptr, FEASM_ERR_CUDA, cudaGetErrorString(e));
  feasm_ctx* ctx = new feasm_ctx;
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->smem_optin = prop.sharedMemPerBlockOptin;
  if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
    delete ctx;
    return feasm_fail(nullptr, FEASM_ERR_CUDA, cudaGetErrorString(e));
  }
  for (auto& ev : ctx->ev) cudaEventCreate(&ev);
  if (cudaMalloc(&ctx->d_flags, 4 * sizeof(int)) != cudaSuccess) {  // [0..1] flags, [2] the element kernel's ticket counter
    delete ctx;
    return feasm_fail(nullptr, FEASM_ERR_CUDA, "cudaMalloc failed");
  }
  cudaMemset(ctx->d_flags, 0, 4 * sizeof(int));
  *out = ctx;
  return FEASM_OK;
}

void feasm_destroy(feasm_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  feasm_comm_release(ctx);
  feasm_general_release(ctx);
  feasm_halo_release(ctx);
  drop_maps(ctx);
  release(ctx->d_xyz); release(ctx->d_conn); release(ctx->d_elem_dofs); release(ctx->d_gid);
  release(ctx->d_node_mask); release(ctx->d_node_base);
  release(ctx->d_con_flag); release(ctx->d_con_val); release(ctx->d_con_rows); release(ctx->d_con_vals_list);
  release(ctx->d_unpack_perm); release(ctx->d_unpack_seg); release(ctx->d_unpack_slot);
  release(ctx->d_row_ptr); release(ctx->d_col_idx);
  release(ctx->d_values); release(ctx->d_rhs); release(ctx->d_flags); release(ctx->d_tab); release(ctx->d_f);
  release(ctx->d_side_elems); release(ctx->d_side_mask); release(ctx->d_side_tab); release(ctx->d_side_nodes);
  release(ctx->d_pen_elems); release(ctx->d_pen_mask); release(ctx->d_y);
  release(ctx->d_send_slots); release(ctx->d_recv_slots); release(ctx->d_send_buf); release(ctx->d_recv_buf);
  if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
  release(ctx->d_stage);
  for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int feasm_set_mesh(feasm_ctx* ctx, int elem_type, int64_t n_elem, int64_t n_nodes, const double* xyz, const uint32_t* conn) {
  if (!ctx) return FEASM_ERR_ARG;
  if (elem_type != FEASM_HEX8 && elem_type != FEASM_HEX27 && elem_type != FEASM_TET10 && elem_type != FEASM_HEX20 && elem_type != FEASM_PRISM18)
    return feasm_fail(ctx, FEASM_ERR_ARG, "unsupported element type (HEX8, HEX20, HEX27, TET10)");
  if (n_elem < 0 || n_nodes <= 0 || !xyz || !conn) return feasm_fail(ctx, FEASM_ERR_ARG, "bad mesh arguments");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  feasm_reset_mesh(ctx, elem_type, n_elem, n_nodes);
  int rc = upload(ctx, &ctx->d_xyz, xyz, (size_t)n_nodes * 3);
  if (rc) return rc;
  rc = upload(ctx, &ctx->d_conn, conn, (size_t)n_elem * ctx->npe);
  if (rc) return rc;
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

int feasm_update_coordinates(feasm_ctx* ctx, const double* xyz) {
  if (!ctx || !ctx->d_xyz || !xyz) return feasm_fail(ctx, FEASM_ERR_STATE, "set_mesh first");
  FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_xyz, xyz, sizeof(double) * 3 * ctx->n_nodes, cudaMemcpyHostToDevice, ctx->stream));
  return FEASM_OK;
}

int feasm_set_dofs(feasm_ctx* ctx, int n_vars, int64_t n_rows, int64_t n_owned, const uint32_t* elem_dofs, const uint32_t* global_ids) {
  if (!ctx || !ctx->d_conn) return feasm_fail(ctx, FEASM_ERR_STATE, "set_mesh first");
  if (n_vars < 1 || n_rows <= 0 || n_owned < 0 || n_owned > n_rows || !elem_dofs) return feasm_fail(ctx, FEASM_ERR_ARG, "bad dof arguments");
  if (n_rows >= 0xffffffffll) return feasm_fail(ctx, FEASM_ERR_ARG, "dof ids are 32-bit (dof_id_type)");
  const size_t n = (size_t)ctx->n_elem * ctx->npe * n_vars;
  for (size_t i = 0; i < n; ++i)
    if (elem_dofs[i] >= (uint64_t)n_rows) return feasm_fail(ctx, FEASM_ERR_ARG, "elem_dofs entry out of range");
  int rc = upload(ctx, &ctx->d_elem_dofs, elem_dofs, n);
  if (rc) return rc;
  release(ctx->d_gid);
  if (global_ids) { rc = upload(ctx, &ctx->d_gid, global_ids, (size_t)n_rows); if (rc) return rc; }
  release(ctx->d_node_base);  // numbering supplied by the caller: no node -> dof table
  ctx->matrix_only = false;
  return feasm_reset_dofs(ctx, n_vars, n_rows, n_owned);
}

}  // extern "C"

// The bookkeeping of feasm_set_mesh / feasm_set_dofs without the uploads (shared with the device-side ingest,
// feasm_ingest.cu).
void feasm_reset_mesh(feasm_ctx* ctx, int elem_type, int64_t n_elem, int64_t n_nodes) {
  drop_maps(ctx);
  ctx->elem_type = elem_type;
  ctx->npe = elem_type;
  ctx->n_elem = n_elem;
  ctx->n_nodes = n_nodes;
  ctx->n_iface = 0;
  ctx->pending_e0 = -1;
  ctx->tab_qorder = -1;
  ctx->side_qorder = -1;
  release(ctx->d_side_elems); release(ctx->d_side_mask);  // element ids of the previous mesh
  ctx->n_side_elems = 0;
  release(ctx->d_pen_elems); release(ctx->d_pen_mask);
  ctx->n_pen_elems = 0;
  release(ctx->d_node_mask); release(ctx->d_node_base);
}

int feasm_reset_dofs(feasm_ctx* ctx, int n_vars, int64_t n_rows, int64_t n_owned) {
  drop_maps(ctx);
  feasm_general_release(ctx);  // constraint rows name dofs of the previous numbering
  feasm_halo_release(ctx);     // the halo plan belongs to the previous pattern
  // the pattern, the values and the exchange plan belong to the previous dof set: feasm_prepare / feasm_assemble return
  // FEASM_ERR_STATE until feasm_set_pattern / feasm_build_pattern is called again
  release(ctx->d_row_ptr); release(ctx->d_col_idx); release(ctx->d_values);
  ctx->nnz = 0;
  ctx->max_row_len = 0;
  release(ctx->d_send_slots); release(ctx->d_recv_slots); release(ctx->d_send_buf); release(ctx->d_recv_buf);
  release(ctx->d_unpack_perm); release(ctx->d_unpack_seg); release(ctx->d_unpack_slot);
  ctx->n_unpack_slots = 0;
  ctx->n_peers = 0;
  ctx->peer_rank.clear(); ctx->send_ptr.clear(); ctx->recv_ptr.clear();
  ctx->n_vars = n_vars;
  ctx->nd = ctx->npe * n_vars;
  ctx->n_rows = n_rows;
  ctx->n_owned = n_owned;
  release(ctx->d_con_flag); release(ctx->d_con_val);
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_con_flag, n_rows));
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_con_val, sizeof(double) * n_rows));
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_con_flag, 0, n_rows, ctx->stream));
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_con_val, 0, sizeof(double) * n_rows, ctx->stream));
  release(ctx->d_rhs);
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_rhs, sizeof(double) * n_rows));
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_rhs, 0, sizeof(double) * n_rows, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->n_con = 0;
  ctx->h_con_rows.clear();
  release(ctx->d_con_rows); release(ctx->d_con_vals_list);
  ctx->d_con_rows_cap = 0;
  ctx->n_con_dev_prev = 0;
  ctx->con_flags_untracked = false;
  return FEASM_OK;
}

extern "C" {

int feasm_set_constraints(feasm_ctx* ctx, int64_t n, const uint32_t* rows, const double* values) {
  if (!ctx || !ctx->d_con_flag) return feasm_fail(ctx, FEASM_ERR_STATE, "set_dofs first");
  if (n < 0 || (n && (!rows || !values))) return feasm_fail(ctx, FEASM_ERR_ARG, "bad constraint arguments");
  for (int64_t i = 0; i < n; ++i)
    if (rows[i] >= (uint64_t)ctx->n_rows) return feasm_fail(ctx, FEASM_ERR_ARG, "constraint row out of range");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  feasm_general_release(ctx);  // a plain (row, value) list means Dirichlet rows only; feasm_set_constraint_rows rebuilds its own
  // Only the (row, value) LIST crosses the bus (the dense flag / value arrays live on the device and are scattered
  // there).  The constrained SET decides which elements take the element-level constraint pass; when the same rows
  // come again in the same order — every assemble of a steady problem — only the values are refreshed.
  const bool same_set = !ctx->con_flags_untracked && (size_t)n == ctx->h_con_rows.size() &&
                        (n == 0 || std::memcmp(rows, ctx->h_con_rows.data(), sizeof(uint32_t) * n) == 0);
  if ((size_t)n > ctx->d_con_rows_cap || ctx->con_flags_untracked) {
    if ((size_t)n > ctx->d_con_rows_cap) {
      release(ctx->d_con_rows); release(ctx->d_con_vals_list);
      FEASM_CUDA(ctx, cudaMalloc(&ctx->d_con_rows, sizeof(uint32_t) * 2 * n));  // new list | previous list
      FEASM_CUDA(ctx, cudaMalloc(&ctx->d_con_vals_list, sizeof(double) * n));
      ctx->d_con_rows_cap = (size_t)n;
    }
    // the previous list on the device went with the buffer (or never existed): clear flags and values wholesale
    FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_con_flag, 0, ctx->n_rows, ctx->stream));
    FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_con_val, 0, sizeof(double) * ctx->n_rows, ctx->stream));
    ctx->n_con_dev_prev = 0;
    ctx->con_flags_untracked = false;
  }
  if (n) FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_con_vals_list, values, sizeof(double) * n, cudaMemcpyHostToDevice, ctx->stream));
  if (!same_set) {
    // previous rows -> second half, new rows -> first half; the kernel clears the old flags, then sets the new ones
    if (ctx->n_con_dev_prev)
      FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_con_rows + ctx->d_con_rows_cap, ctx->d_con_rows, sizeof(uint32_t) * ctx->n_con_dev_prev,
                                      cudaMemcpyDeviceToDevice, ctx->stream));
    if (n) FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_con_rows, rows, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, ctx->stream));
    ctx->elem_flag_valid = false;
  }
  int rc = feasm_launch_scatter_constraints(ctx, n, same_set ? 0 : ctx->n_con_dev_prev, !same_set);
  if (rc) return rc;
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the caller's arrays may go away
  if (!same_set) ctx->h_con_rows.assign(rows, rows + n);
  ctx->n_con_dev_prev = n;
  ctx->n_con = n;
  return FEASM_OK;
}

static int alloc_values(feasm_ctx* ctx) {
  release(ctx->d_values);
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_values, sizeof(double) * (ctx->nnz ? ctx->nnz : 1)));
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_values, 0, sizeof(double) * ctx->nnz, ctx->stream));
  return FEASM_OK;
}

int feasm_set_pattern(feasm_ctx* ctx, const uint64_t* row_ptr, const uint32_t* col_idx) {
  if (!ctx || (!ctx->d_elem_dofs && !ctx->matrix_only)) return feasm_fail(ctx, FEASM_ERR_STATE, "set_dofs first");
  if (!row_ptr || !col_idx || row_ptr[0] != 0) return feasm_fail(ctx, FEASM_ERR_ARG, "bad pattern");
  for (int64_t r = 0; r < ctx->n_rows; ++r) {
    if (row_ptr[r + 1] < row_ptr[r]) return feasm_fail(ctx, FEASM_ERR_ARG, "row_ptr not monotone");
    for (uint64_t k = row_ptr[r] + 1; k < row_ptr[r + 1]; ++k)
      if (col_idx[k] <= col_idx[k - 1]) return feasm_fail(ctx, FEASM_ERR_ARG, "columns of a row must be strictly ascending");
  }
  ctx->prepared = false;
  feasm_general_invalidate_plan(ctx);
  feasm_halo_release(ctx);
  ctx->nnz = row_ptr[ctx->n_rows];
  int rc = upload(ctx, &ctx->d_row_ptr, row_ptr, (size_t)ctx->n_rows + 1);
  if (rc) return rc;
  rc = upload(ctx, &ctx->d_col_idx, col_idx, (size_t)ctx->nnz);
  if (rc) return rc;
  rc = alloc_values(ctx);
  if (rc) return rc;
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

int feasm_build_pattern(feasm_ctx* ctx) {
  if (!ctx || !ctx->d_elem_dofs) return feasm_fail(ctx, FEASM_ERR_STATE, "set_dofs first");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->prepared = false;
  feasm_general_invalidate_plan(ctx);
  feasm_halo_release(ctx);
  int rc = feasm_device_build_pattern(ctx);
  if (rc) return rc;
  rc = alloc_values(ctx);
  if (rc) return rc;
  return feasm_general_extend_pattern(ctx);  // couplings of elements widened by hanging-node constraints (no-op without them)
}

int64_t feasm_nnz(const feasm_ctx* ctx) { return ctx ? (int64_t)ctx->nnz : -1; }

int feasm_get_pattern(const feasm_ctx* cctx, uint64_t* row_ptr, uint32_t* col_idx) {
  feasm_ctx* ctx = const_cast<feasm_ctx*>(cctx);
  if (!ctx || !ctx->d_row_ptr) return feasm_fail(ctx, FEASM_ERR_STATE, "no pattern");
  if (row_ptr) FEASM_CUDA(ctx, cudaMemcpyAsync(row_ptr, ctx->d_row_ptr, sizeof(uint64_t) * (ctx->n_rows + 1), cudaMemcpyDeviceToHost, ctx->stream));
  if (col_idx) FEASM_CUDA(ctx, cudaMemcpyAsync(col_idx, ctx->d_col_idx, sizeof(uint32_t) * ctx->nnz, cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

int feasm_prepare(feasm_ctx* ctx) {
  if (!ctx || !ctx->d_row_ptr) return feasm_fail(ctx, FEASM_ERR_STATE, "pattern not set: call feasm_set_pattern or feasm_build_pattern");
  if (ctx->prepared) return FEASM_OK;
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = feasm_build_incidence(ctx);
  if (rc) return rc;
  rc = feasm_build_pos(ctx);
  if (rc) return rc;
  rc = feasm_general_prepare(ctx);
  if (rc) return rc;
  ctx->prepared = true;
  return FEASM_OK;
}

// mode 0: the whole assembly.  mode 1 (feasm_assemble_begin): element phase, then the rows this rank does not own
// (local rows n_owned..n_rows) and the pack of the send buffer, so that the exchange can start while
// feasm_assemble_end reduces the owned rows.
static int assemble_impl(feasm_ctx* ctx, const feasm_params* params, int mode) {
  if (!ctx || !params) return FEASM_ERR_ARG;
  if (!ctx->d_conn || !ctx->d_elem_dofs) return feasm_fail(ctx, FEASM_ERR_STATE, "set_mesh / set_dofs first (this context holds a matrix only)");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = feasm_prepare(ctx);
  if (rc) return rc;
  const int qorder = resolve_qorder(ctx, *params);
  rc = ensure_tables(ctx, qorder);
  if (rc) return rc;
  const double* d_f = nullptr;
  rc = resolve_forcing(ctx, *params, ctx->tab_nq, &d_f);
  if (rc) return rc;
  // matrix->zero(); rhs->zero() (implicit_system.C:101-105) is implied: the row reduction overwrites every slot.
  // d_flags is NOT cleared here: a degenerate-map / pattern-miss flag raised by step i must survive the launch of
  // step i+1 until feasm_synchronize / a getter reports (and clears) it.
  FEASM_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
  int launches = 2;
  const int rs = feasm_recb(ctx->nd, ctx->pos_bytes) / 8;  // doubles per element record
  if (!ctx->elem_flag_valid) {
    rc = feasm_launch_elem_flags(ctx);
    if (rc) return rc;
    ++launches;
  }
  // Interface elements first: when the caller named the prefix of elements that touch ghost rows
  // (feasm_set_interface_prefix), feasm_assemble_begin assembles only those, reduces and packs the ghost rows, and the
  // transfer then runs beside the element phase of all the other elements, which feasm_assemble_end launches.
  const bool two_phase = mode == 1 && ctx->n_iface > 0 && ctx->n_iface < ctx->n_elem && ctx->n_owned < ctx->n_rows;
  const int64_t e_first = two_phase ? ctx->n_iface : ctx->n_elem;
  rc = feasm_launch_element_kernel(ctx, *params, d_f, 0, e_first, ctx->d_kf, rs, params->constrain);
  if (rc) return rc;
  if (feasm_general_active(ctx)) {  // elements widened by hanging-node constraints: C^T K C into the side buffer
    if (mode != 0) return feasm_fail(ctx, FEASM_ERR_STATE, "general constraint rows: use feasm_assemble (one rank)");
    if (params->constrain >= 0) {
      rc = feasm_general_transform(ctx, rs, params->constrain);
      if (rc) return rc;
      launches += 2;
    }
  }
  if (ctx->n_con_elems > 0) ++launches;
  if (params->problem == FEASM_POISSON && params->affine_tensor) ++launches;  // the list kernel for non-affine elements
  FEASM_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
  ctx->last_launches = launches;
  ctx->pending_e0 = two_phase ? e_first : -1;
  if (two_phase) { ctx->pending_params = *params; ctx->pending_f = d_f; }
  if (mode == 0) {
    rc = feasm_launch_row_reduce(ctx, nullptr, ctx->n_rows, ctx->stream);
    if (rc) return rc;
    if (feasm_general_active(ctx) && params->constrain >= 0) {
      rc = feasm_general_add(ctx);
      if (rc) return rc;
    }
    FEASM_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
    return FEASM_OK;
  }
  if (ctx->n_owned < ctx->n_rows) {
    rc = feasm_launch_row_reduce(ctx, nullptr, ctx->n_rows, ctx->stream, ctx->n_owned);
    if (rc) return rc;
    ++ctx->last_launches;
  }
  if (ctx->d_send_buf) {
    rc = feasm_launch_pack(ctx);
    if (rc) return rc;
    ++ctx->last_launches;
  }
  ctx->assemble_open = true;
  return FEASM_OK;
}

int feasm_residual(feasm_ctx* ctx, const feasm_params* params, const double* d_u, double* d_r) {
  if (!ctx || !params) return FEASM_ERR_ARG;
  if (ctx && feasm_general_active(ctx)) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_residual: not with general constraint rows (feasm_set_constraint_rows) in this version");
  if (!d_u || !d_r) return feasm_fail(ctx, FEASM_ERR_ARG, "null vector");
  const bool several = feasm_is_distributed(ctx);
  if (several && !feasm_halo_ready(ctx)) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_residual on several ranks: feasm_build_halo first");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = feasm_prepare(ctx);
  if (rc) return rc;
  rc = ensure_tables(ctx, resolve_qorder(ctx, *params));
  if (rc) return rc;
  const double* d_f = nullptr;
  rc = resolve_forcing(ctx, *params, ctx->tab_nq, &d_f);
  if (rc) return rc;
  if (!ctx->elem_flag_valid) {
    rc = feasm_launch_elem_flags(ctx);
    if (rc) return rc;
  }
  const int rs = feasm_recb(ctx->nd, ctx->pos_bytes) / 8;
  rc = feasm_launch_element_kernel(ctx, *params, d_f, 0, ctx->n_elem, ctx->d_kf, rs, params->constrain);
  if (rc) return rc;
  if (several) return feasm_residual_distributed(ctx, d_u, d_r);  // ghost entries of u in, ghost-row sums out
  return feasm_launch_row_residual(ctx, d_u, d_r);
}

int feasm_assemble(feasm_ctx* ctx, const feasm_params* params) { return assemble_impl(ctx, params, 0); }
int feasm_assemble_begin(feasm_ctx* ctx, const feasm_params* params) { return assemble_impl(ctx, params, 1); }

int feasm_assemble_end(feasm_ctx* ctx) {
  if (!ctx) return FEASM_ERR_ARG;
  if (!ctx->assemble_open) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_assemble_end without feasm_assemble_begin");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->assemble_open = false;
  int rc = FEASM_OK;
  if (ctx->pending_e0 >= 0) {  // the elements that touch no ghost row, while the ghost rows travel
    const int rs = feasm_recb(ctx->nd, ctx->pos_bytes) / 8;
    const int64_t e0 = ctx->pending_e0;
    ctx->pending_e0 = -1;
    rc = feasm_launch_element_kernel(ctx, ctx->pending_params, ctx->pending_f, e0, ctx->n_elem,
                                     ctx->d_kf + (size_t)e0 * ctx->nd * rs, rs, ctx->pending_params.constrain);
    if (rc) return rc;
    ++ctx->last_launches;
    FEASM_CUDA(ctx, cudaEventRecord(ctx->ev[1], ctx->stream));
  }
  rc = feasm_launch_row_reduce(ctx, nullptr, ctx->n_owned, ctx->stream);
  if (rc) return rc;
  FEASM_CUDA(ctx, cudaEventRecord(ctx->ev[2], ctx->stream));
  return FEASM_OK;
}

int feasm_set_interface_prefix(feasm_ctx* ctx, int64_t n_elems) {
  if (!ctx) return FEASM_ERR_ARG;
  if (n_elems < 0 || n_elems > ctx->n_elem) return feasm_fail(ctx, FEASM_ERR_ARG, "interface prefix out of range");
  ctx->n_iface = n_elems;
  return FEASM_OK;
}

int feasm_synchronize(feasm_ctx* ctx) {
  if (!ctx) return FEASM_ERR_ARG;
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return check_flags(ctx);
}

int feasm_last_timing(feasm_ctx* ctx, float ms[4], int* launches) {
  if (!ctx) return FEASM_ERR_ARG;
  FEASM_CUDA(ctx, cudaEventSynchronize(ctx->ev[2]));
  cudaEventElapsedTime(&ctx->last_ms[0], ctx->ev[0], ctx->ev[3]);
  cudaEventElapsedTime(&ctx->last_ms[3], ctx->ev[3], ctx->ev[1]);
  cudaEventElapsedTime(&ctx->last_ms[1], ctx->ev[1], ctx->ev[2]);
  cudaEventElapsedTime(&ctx->last_ms[2], ctx->ev[0], ctx->ev[2]);
  if (ms) std::memcpy(ms, ctx->last_ms, sizeof(ctx->last_ms));
  if (launches) *launches = ctx->last_launches;
  return FEASM_OK;
}

int feasm_element_matrices(feasm_ctx* ctx, const feasm_params* params, int64_t e0, int64_t e1, int constrained, double* Ke, double* Fe) {
  if (!ctx || !params) return FEASM_ERR_ARG;
  if (!ctx->d_elem_dofs) return feasm_fail(ctx, FEASM_ERR_STATE, "set_dofs first");
  if (e0 < 0 || e1 > ctx->n_elem || e0 >= e1) return feasm_fail(ctx, FEASM_ERR_ARG, "bad element range");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int qorder = resolve_qorder(ctx, *params);
  int rc = ensure_tables(ctx, qorder);
  if (rc) return rc;
  const double* d_f = nullptr;
  rc = resolve_forcing(ctx, *params, ctx->tab_nq, &d_f);
  if (rc) return rc;
  const size_t ne = (size_t)(e1 - e0), nd = ctx->nd, ldr = feasm_ldr(ctx->nd);
  if (constrained && ctx->n_con > 0 && !ctx->elem_flag_valid) {
    rc = feasm_launch_elem_flags(ctx);
    if (rc) return rc;
  }
  double* d_kf = nullptr;  // scratch records without a position row (record stride = ldr doubles)
  FEASM_CUDA(ctx, cudaMalloc(&d_kf, sizeof(double) * ne * nd * ldr));
  rc = feasm_launch_element_kernel(ctx, *params, d_f, e0, e1, d_kf, (int)ldr, constrained ? params->constrain : -1);
  if (!rc) {
    // scratch records are (Ke row | Fe | pad): strip the padding on the way out
    if (Ke) cudaMemcpy2DAsync(Ke, sizeof(double) * nd, d_kf, sizeof(double) * ldr, sizeof(double) * nd, ne * nd, cudaMemcpyDeviceToHost, ctx->stream);
    if (Fe) cudaMemcpy2DAsync(Fe, sizeof(double), d_kf + nd, sizeof(double) * ldr, sizeof(double), ne * nd, cudaMemcpyDeviceToHost, ctx->stream);
    rc = check_flags(ctx);
  }
  cudaFree(d_kf);
  return rc;
}

int feasm_zero(feasm_ctx* ctx) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_values, 0, sizeof(double) * ctx->nnz, ctx->stream));
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_rhs, 0, sizeof(double) * ctx->n_rows, ctx->stream));
  return FEASM_OK;
}

// Staging area of the per-element compat route: one pinned host block and one device block, grown on demand and kept, so a
// call costs one packed H2D copy and one kernel — no cudaMalloc / cudaFree per call.
static int stage_block(feasm_ctx* ctx, size_t bytes, unsigned char** h, unsigned char** d) {
  if (ctx->stage_cap < bytes) {
    FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    release(ctx->d_stage);
    ctx->h_stage = nullptr; ctx->stage_cap = 0;
    size_t cap = 64 * 1024;
    while (cap < bytes) cap *= 2;
    FEASM_CUDA(ctx, cudaMallocHost(&ctx->h_stage, cap));
    FEASM_CUDA(ctx, cudaMalloc(&ctx->d_stage, cap));
    ctx->stage_cap = cap;
  }
  *h = ctx->h_stage; *d = ctx->d_stage;
  return FEASM_OK;
}

// mode 0: add (SparseMatrix::add_matrix, sparse_matrix.h:312), mode 1: set (SparseMatrix::set, :295)
static int put_block(feasm_ctx* ctx, const double* dm, const uint32_t* rows, int nr, const uint32_t* cols, int nc, int mode) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  if (!dm || !rows || !cols || nr <= 0 || nc <= 0) return feasm_fail(ctx, FEASM_ERR_ARG, "bad add_matrix arguments");
  for (int i = 0; i < nr; ++i)
    if (rows[i] >= (uint64_t)ctx->n_rows) return feasm_fail(ctx, FEASM_ERR_ARG, "row out of range");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t dm_b = sizeof(double) * (size_t)nr * nc, r_b = sizeof(uint32_t) * nr, c_b = sizeof(uint32_t) * nc;
  unsigned char *h = nullptr, *d = nullptr;
  int rc = stage_block(ctx, dm_b + r_b + c_b, &h, &d);
  if (rc) return rc;
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the previous call's copy has left the pinned block
  std::memcpy(h, dm, dm_b); std::memcpy(h + dm_b, rows, r_b); std::memcpy(h + dm_b + r_b, cols, c_b);
  FEASM_CUDA(ctx, cudaMemcpyAsync(d, h, dm_b + r_b + c_b, cudaMemcpyHostToDevice, ctx->stream));
  rc = feasm_launch_add_matrix(ctx, reinterpret_cast<const double*>(d), reinterpret_cast<const uint32_t*>(d + dm_b), nr,
                               reinterpret_cast<const uint32_t*>(d + dm_b + r_b), nc, mode);
  return rc ? rc : check_flags(ctx);  // "add to an entry outside the pattern" is reported by the call that did it
}

int feasm_add_matrix(feasm_ctx* ctx, const double* dm, const uint32_t* rows, int nr, const uint32_t* cols, int nc) {
  return put_block(ctx, dm, rows, nr, cols, nc, 0);
}
int feasm_set_matrix(feasm_ctx* ctx, const double* dm, const uint32_t* rows, int nr, const uint32_t* cols, int nc) {
  return put_block(ctx, dm, rows, nr, cols, nc, 1);
}

int feasm_add_vector(feasm_ctx* ctx, const double* v, const uint32_t* rows, int n) {
  if (!ctx || !ctx->d_rhs) return feasm_fail(ctx, FEASM_ERR_STATE, "no vector");
  if (!v || !rows || n <= 0) return feasm_fail(ctx, FEASM_ERR_ARG, "bad add_vector arguments");
  for (int i = 0; i < n; ++i)
    if (rows[i] >= (uint64_t)ctx->n_rows) return feasm_fail(ctx, FEASM_ERR_ARG, "row out of range");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const size_t v_b = sizeof(double) * (size_t)n, r_b = sizeof(uint32_t) * n;
  unsigned char *h = nullptr, *d = nullptr;
  int rc = stage_block(ctx, v_b + r_b, &h, &d);
  if (rc) return rc;
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  std::memcpy(h, v, v_b); std::memcpy(h + v_b, rows, r_b);
  FEASM_CUDA(ctx, cudaMemcpyAsync(d, h, v_b + r_b, cudaMemcpyHostToDevice, ctx->stream));
  // NumericVector::add_vector (numeric_vector.C:690-697): entries of one call may repeat a row; they are added in order
  return feasm_launch_add_vector(ctx, reinterpret_cast<const double*>(d), reinterpret_cast<const uint32_t*>(d + v_b), n);
}

// A context that holds only a matrix (pattern + values) and a right-hand side: what SparseMatrix::init / clone /
// get_transpose need when no mesh is attached; the per-element virtuals (add_matrix, set, add_vector), zero, the getters and
// vector_mult work on it.
int feasm_init_matrix(feasm_ctx* ctx, int64_t n_rows, int64_t n_owned, const uint64_t* row_ptr, const uint32_t* col_idx) {
  if (!ctx) return FEASM_ERR_ARG;
  if (n_rows <= 0 || n_owned < 0 || n_owned > n_rows || !row_ptr || !col_idx) return feasm_fail(ctx, FEASM_ERR_ARG, "bad matrix arguments");
  if (n_rows >= 0xffffffffll) return feasm_fail(ctx, FEASM_ERR_ARG, "dof ids are 32-bit (dof_id_type)");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  feasm_reset_mesh(ctx, 0, 0, 0);
  release(ctx->d_xyz); release(ctx->d_conn); release(ctx->d_elem_dofs); release(ctx->d_gid);
  int rc = feasm_reset_dofs(ctx, 1, n_rows, n_owned);
  if (rc) return rc;
  ctx->matrix_only = true;
  return feasm_set_pattern(ctx, row_ptr, col_idx);
}

int feasm_set_values(feasm_ctx* ctx, const double* values, const double* rhs) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (values) FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_values, values, sizeof(double) * ctx->nnz, cudaMemcpyHostToDevice, ctx->stream));
  if (rhs) FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_rhs, rhs, sizeof(double) * ctx->n_rows, cudaMemcpyHostToDevice, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

int feasm_get_values(feasm_ctx* ctx, double* values, double* rhs) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  if (values) FEASM_CUDA(ctx, cudaMemcpyAsync(values, ctx->d_values, sizeof(double) * ctx->nnz, cudaMemcpyDeviceToHost, ctx->stream));
  if (rhs) FEASM_CUDA(ctx, cudaMemcpyAsync(rhs, ctx->d_rhs, sizeof(double) * ctx->n_rows, cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return check_flags(ctx);
}

double* feasm_device_values(feasm_ctx* ctx) { return ctx ? ctx->d_values : nullptr; }
double* feasm_device_rhs(feasm_ctx* ctx) { return ctx ? ctx->d_rhs : nullptr; }
void* feasm_stream(feasm_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int feasm_set_exchange_plan(feasm_ctx* ctx, int n_peers, const int32_t* peer_rank, const uint64_t* send_ptr,
                            const uint64_t* send_slots, const uint64_t* recv_ptr, const uint64_t* recv_slots) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  if (n_peers < 0 || (n_peers && (!peer_rank || !send_ptr || !recv_ptr))) return feasm_fail(ctx, FEASM_ERR_ARG, "bad plan");
  feasm_halo_release(ctx);  // its vector routes are derived from this plan
  ctx->n_peers = n_peers;
  ctx->peer_rank.assign(peer_rank, peer_rank + n_peers);
  ctx->send_ptr.assign(send_ptr, send_ptr + n_peers + 1);
  ctx->recv_ptr.assign(recv_ptr, recv_ptr + n_peers + 1);
  const uint64_t ns = n_peers ? send_ptr[n_peers] : 0, nr = n_peers ? recv_ptr[n_peers] : 0;
  const uint64_t lim = ctx->nnz + (uint64_t)ctx->n_rows;
  for (uint64_t i = 0; i < ns; ++i) if (send_slots[i] >= lim) return feasm_fail(ctx, FEASM_ERR_ARG, "send slot out of range");
  for (uint64_t i = 0; i < nr; ++i) if (recv_slots[i] >= lim) return feasm_fail(ctx, FEASM_ERR_ARG, "recv slot out of range");
  int rc = upload(ctx, &ctx->d_send_slots, send_slots, ns);
  if (rc) return rc;
  rc = upload(ctx, &ctx->d_recv_slots, recv_slots, nr);
  if (rc) return rc;
  release(ctx->d_send_buf); release(ctx->d_recv_buf);
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_send_buf, sizeof(double) * (ns ? ns : 1)));
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_recv_buf, sizeof(double) * (nr ? nr : 1)));
  // One unpack launch for all peers: the received entries are grouped by destination slot, each group in receive-buffer
  // (= peer) order, and one thread adds a group to its slot — the same sums, in the same order, as one launch per peer.
  {
    std::vector<uint64_t> perm(nr);
    for (uint64_t i = 0; i < nr; ++i) perm[i] = i;
    std::stable_sort(perm.begin(), perm.end(), [&](uint64_t a, uint64_t b) { return recv_slots[a] < recv_slots[b]; });
    std::vector<uint64_t> uslot, seg;
    for (uint64_t i = 0; i < nr; ++i)
      if (i == 0 || recv_slots[perm[i]] != recv_slots[perm[i - 1]]) { uslot.push_back(recv_slots[perm[i]]); seg.push_back(i); }
    seg.push_back(nr);
    ctx->n_unpack_slots = uslot.size();
    rc = upload(ctx, &ctx->d_unpack_perm, perm.data(), perm.size());
    if (rc) return rc;
    rc = upload(ctx, &ctx->d_unpack_slot, uslot.data(), uslot.size());
    if (rc) return rc;
    rc = upload(ctx, &ctx->d_unpack_seg, seg.data(), seg.size());
    if (rc) return rc;
    FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the vectors go out of scope
  }
  return FEASM_OK;
}

double* feasm_pack_ghost(feasm_ctx* ctx, uint64_t* n_doubles) {
  if (!ctx || !ctx->d_send_buf) return nullptr;
  if (feasm_launch_pack(ctx)) return nullptr;
  if (n_doubles) *n_doubles = ctx->send_ptr.empty() ? 0 : ctx->send_ptr.back();
  return ctx->d_send_buf;
}

double* feasm_send_buffer(feasm_ctx* ctx, uint64_t* n_doubles) {
  if (!ctx || !ctx->d_send_buf) return nullptr;
  if (n_doubles) *n_doubles = ctx->send_ptr.empty() ? 0 : ctx->send_ptr.back();
  return ctx->d_send_buf;
}

double* feasm_recv_buffer(feasm_ctx* ctx, uint64_t* n_doubles) {
  if (!ctx || !ctx->d_recv_buf) return nullptr;
  if (n_doubles) *n_doubles = ctx->recv_ptr.empty() ? 0 : ctx->recv_ptr.back();
  return ctx->d_recv_buf;
}

int feasm_unpack_ghost(feasm_ctx* ctx) {
  if (!ctx || !ctx->d_recv_buf) return feasm_fail(ctx, FEASM_ERR_STATE, "no exchange plan");
  return feasm_launch_unpack(ctx);
}

}  // extern "C"
