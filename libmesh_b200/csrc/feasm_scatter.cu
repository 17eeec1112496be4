// feasm_scatter.cu — the deterministic scatter: row reduction into CSR with the Dirichlet constraints fused in.
//
// Row reduction = SparseMatrix::add_matrix / NumericVector::add_vector (laspack_matrix.C:259-275,
// numeric_vector.C:690-697) turned inside out: a warp sums, for each CSR row it owns, the element rows
// KF[e][a][:] incident to that row in element order through the precomputed position map — no atomics, bitwise
// reproducible, and the same accumulation order as the reference's serial element loop.
// DofMap::heterogeneously_constrain_element_matrix_and_vector / constrain_element_matrix_and_vector
// (src/base/dof_map_constraints.C:2518-2614, :2439-2514) with empty constraint rows act entry-wise on an element
// row, so they are applied here while the row streams through registers:
//   row dof constrained            -> the element contributes K(a,a) = 1 and F(a) = g_a (0 when homogeneous)
//   column dof b constrained only  -> F(a) -= K(a,b) g_b (heterogeneous), K(a,b) = 0
// The "column constrained" bit travels in the top bit of the position map, so the common unconstrained entry costs
// nothing extra.
#include "feasm_rr_dev.cuh"

namespace {
// Stand-alone row reduction: one batch of RB rows per warp.  ND = dofs per element when known at compile time
// (strides and lane loops fold away), 0 = generic (runtime nd).
template <typename PosT, int ND, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
k_row_reduce(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc,
             const unsigned char* __restrict__ pos, const unsigned char* __restrict__ kf, int nd_rt, int RB, int lcap, int cap,
             const uint32_t* __restrict__ row_list, int64_t row0, int64_t n_list, const uint8_t* __restrict__ con_flag,
             const double* __restrict__ con_val, const uint32_t* __restrict__ elem_dofs, int mode,
             double* __restrict__ values, double* __restrict__ rhs) {
  const int nd = ND ? ND : nd_rt;
  const int slotb = feasm_ldr(nd) * 8 + feasm_pld(nd, (int)sizeof(PosT)) * (int)sizeof(PosT);
  const int G = rr_stage(slotb);
  extern __shared__ __align__(16) unsigned char s_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t base = row0 + (blockIdx.x * (int64_t)WARPS + warp) * RB;  // rows [row0, n_list) (positions, with a row list)
  if (base >= n_list) return;
  const int nrows = (int)(n_list - base < RB ? n_list - base : RB);
  const uint32_t warp_bytes = (uint32_t)(RR_NG * G * slotb + cap * 8 + lcap * 4);
  const uint32_t ring_a = (uint32_t)__cvta_generic_to_shared(s_raw) + warp * warp_bytes;
  const uint32_t img_a = ring_a + RR_NG * G * slotb;
  const uint32_t list_a = img_a + cap * 8;
  for (int t = lane; t < cap; t += 32) sts_f64(img_a + 8 * t, 0.);
  __syncwarp();
  rr_process_batch<PosT, ND>(base, nrows, lane, ring_a, img_a, list_a, G, nd_rt, row_ptr, inc_ptr, inc, pos, kf, row_list, con_flag,
                             con_val, elem_dofs, mode, values, rhs);
}

// Fallback for rows longer than the shared-memory budget: accumulate straight into the (zeroed) global row.
template <typename PosT>
__global__ void k_row_reduce_global(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr,
                                    const uint32_t* __restrict__ inc, const PosT* __restrict__ pos,
                                    const double* __restrict__ kf, int nd, int ldr, int pld, int64_t row0, int64_t n_rows,
                                    const uint8_t* __restrict__ con_flag, const double* __restrict__ con_val,
                                    const uint32_t* __restrict__ elem_dofs, int mode, double* __restrict__ values,
                                    double* __restrict__ rhs) {
  const int lane = threadIdx.x & 31;
  const int64_t r = row0 + ((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  if (r >= n_rows) return;
  double* row = values + row_ptr[r];
  const int len = (int)(row_ptr[r + 1] - row_ptr[r]);
  for (int t = lane; t < len; t += 32) row[t] = 0.;
  __syncwarp();
  const int row_con = con_flag[r];
  const double row_g = (row_con && mode == FEASM_CONSTRAIN_HETEROGENEOUS) ? con_val[r] : 0.;
  double b = 0., corr = 0.;
  for (uint64_t k = inc_ptr[r]; k < inc_ptr[r + 1]; ++k) {
    const uint64_t ea = inc[k] & FEASM_INC_MASK;
    const int a = (int)(ea % nd);
    for (int t = lane; t < nd; t += 32) {
      double val = kf[ea * ldr + t];
      const unsigned pp = pos[ea * pld + t];
      if (row_con) val = (t == a) ? 1. : 0.;
      else if (pp & PosBits<PosT>::FLAG) {
        if (mode == FEASM_CONSTRAIN_HETEROGENEOUS) corr += val * con_val[elem_dofs[(ea / nd) * nd + t]];
        val = 0.;
      }
      row[pp & PosBits<PosT>::MASK] += val;
    }
    b += row_con ? row_g : kf[ea * ldr + nd];
    __syncwarp();
  }
  for (int o = 16; o > 0; o >>= 1) corr += __shfl_xor_sync(0xffffffffu, corr, o);
  if (lane == 0) rhs[r] = b - corr;
}

// Element-level constraint application for feasm_element_matrices (tests / debugging): one warp per element,
// in place on KF rows.  dof_map_constraints.C:2518-2614 with C = diag(0 on constrained, 1 else), H = g on constrained.
__global__ void k_constrain_elements(const uint32_t* __restrict__ elem_dofs, const uint8_t* __restrict__ con_flag,
                                     const double* __restrict__ con_val, int nd, int ldr, int mode, int64_t e0, int64_t e1,
                                     double* __restrict__ kf) {
  const int lane = threadIdx.x & 31;
  const int64_t e = e0 + ((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  if (e >= e1) return;
  double* K = kf + (uint64_t)(e - e0) * nd * ldr;
  const uint32_t* dofs = elem_dofs + (uint64_t)e * nd;
  // Fe <- C^T (Fe - K H)
  for (int i = lane; i < nd; i += 32) {
    if (con_flag[dofs[i]]) continue;
    if (mode != FEASM_CONSTRAIN_HETEROGENEOUS) continue;
    double kh = 0.;
    for (int j = 0; j < nd; ++j)
      if (con_flag[dofs[j]]) kh += K[i * ldr + j] * con_val[dofs[j]];
    K[i * ldr + nd] -= kh;
  }
  __syncwarp();
  // K <- C^T K C, then K(i,i) = 1 and Fe(i) = H(i) on constrained rows
  for (int t = lane; t < nd * nd; t += 32) {
    const int i = t / nd, j = t % nd;
    if (con_flag[dofs[i]] || con_flag[dofs[j]]) K[i * ldr + j] = (i == j) ? 1. : 0.;
  }
  for (int i = lane; i < nd; i += 32)
    if (con_flag[dofs[i]]) K[i * ldr + nd] = (mode == FEASM_CONSTRAIN_HETEROGENEOUS) ? con_val[dofs[i]] : 0.;
}

// Refresh the "column dof is constrained" bit of every position-map entry from the current constrained set.
template <typename PosT>
__global__ void k_pos_flags(const uint32_t* __restrict__ elem_dofs, const uint8_t* __restrict__ con_flag, int nd, int pld,
                            uint64_t n_rows_e /* n_elem*nd */, PosT* __restrict__ pos) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  const uint64_t ea = i / pld;
  const int b = (int)(i % pld);
  if (ea >= n_rows_e || b >= nd) return;
  const uint64_t e = ea / nd;
  const unsigned p = pos[i] & PosBits<PosT>::MASK;
  pos[i] = (PosT)(p | (con_flag[elem_dofs[e * nd + b]] ? PosBits<PosT>::FLAG : 0u));
}

// Bit 31 of an incidence id = "the element touches a constrained dof": its rows take the out-of-line path of the row
// reduction.  Refreshed together with the position-map flags.
__global__ void k_inc_flags(const uint32_t* __restrict__ elem_dofs, const uint8_t* __restrict__ con_flag, int nd, uint64_t total,
                            uint32_t* __restrict__ inc) {
  const uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (k >= total) return;
  const uint32_t ea = inc[k] & FEASM_INC_MASK;
  const uint32_t* d = elem_dofs + (uint64_t)(ea / (uint32_t)nd) * nd;
  unsigned any = 0;
  for (int b = 0; b < nd; ++b) any |= con_flag[d[b]];
  inc[k] = ea | (any ? FEASM_INC_SLOW : 0u);
}

__global__ void k_pack(const uint64_t* __restrict__ slots, uint64_t n, uint64_t nnz, const double* __restrict__ values,
                       const double* __restrict__ rhs, double* __restrict__ out) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i < n) {
    const uint64_t s = slots[i];
    out[i] = s < nnz ? values[s] : rhs[s - nnz];
  }
}

__global__ void k_unpack(const uint64_t* __restrict__ slots, uint64_t n, uint64_t nnz, const double* __restrict__ in,
                         double* __restrict__ values, double* __restrict__ rhs) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i < n) {
    const uint64_t s = slots[i];  // slots of one peer are distinct
    if (s < nnz) values[s] += in[i]; else rhs[s - nnz] += in[i];
  }
}

// SparseMatrix::add_matrix for one host-supplied dense block (slow route; one thread per entry, entries of one
// call are distinct slots so there is no race).
__global__ void k_add_matrix(const double* __restrict__ dm, const uint32_t* __restrict__ rows, int nr,
                             const uint32_t* __restrict__ cols, int nc, const uint64_t* __restrict__ row_ptr,
                             const uint32_t* __restrict__ col_idx, double* __restrict__ values, int* flags) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nr * nc) return;
  const int i = t / nc, j = t % nc;
  const double v = dm[t];
  const uint32_t row = rows[i], col = cols[j];
  const uint32_t* c = col_idx + row_ptr[row];
  int lo = 0, hi = (int)(row_ptr[row + 1] - row_ptr[row]);
  const int len = hi;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (c[mid] < col) lo = mid + 1; else hi = mid;
  }
  if (lo < len && c[lo] == col) values[row_ptr[row] + lo] += v;
  else if (v != 0.) atomicAdd(&flags[1], 1);  // zero may be "added" to a missing entry (sparse_matrix.h:300-303)
}

template <typename PosT, int ND>
int launch_row_reduce_t(feasm_ctx* ctx, int mode, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream, int64_t row0) {
  constexpr int WARPS = 4;
  const int cap = ((int)ctx->max_row_len + 1) & ~1;  // even: every warp's ring stays 16 B aligned
  const int ldr = feasm_ldr(ctx->nd), pld = feasm_pld(ctx->nd, (int)sizeof(PosT));
  const int slotb = ldr * 8 + pld * (int)sizeof(PosT);
  const int G = rr_stage(slotb);
  const int lcap = 256;
  int RB = ctx->max_inc ? (int)(lcap / ctx->max_inc) : 32;
  if (RB > 32) RB = 32;
  const size_t warp_bytes = (size_t)RR_NG * G * slotb + (size_t)cap * 8 + (size_t)lcap * 4;
  const size_t smem = WARPS * warp_bytes;
  if (RB >= 1 && smem <= ctx->smem_optin && ctx->nd <= 128) {
    auto kern = k_row_reduce<PosT, ND, WARPS>;
    FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t batches = (n_list - row0 + RB - 1) / RB;
    const unsigned grid = (unsigned)((batches + WARPS - 1) / WARPS);
    if (grid)
      kern<<<grid, WARPS * 32, smem, stream>>>(ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, (const unsigned char*)ctx->d_pos,
                                               (const unsigned char*)ctx->d_kf, ctx->nd, RB, lcap, cap, d_row_list, row0, n_list,
                                               ctx->d_con_flag, ctx->d_con_val, ctx->d_elem_dofs, mode, ctx->d_values, ctx->d_rhs);
  } else {
    if (d_row_list) return feasm_fail(ctx, FEASM_ERR_NOMEM, "rows too long for the shared-memory row reduction");
    const unsigned grid = (unsigned)(((n_list - row0) * 32 + 255) / 256);
    if (grid)
      k_row_reduce_global<PosT><<<grid, 256, 0, stream>>>(ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, (const PosT*)ctx->d_pos, ctx->d_kf,
                                                        ctx->nd, ldr, pld, row0, n_list, ctx->d_con_flag, ctx->d_con_val,
                                                        ctx->d_elem_dofs, mode, ctx->d_values, ctx->d_rhs);
  }
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

template <typename PosT>
int launch_row_reduce(feasm_ctx* ctx, int mode, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream, int64_t row0) {
  switch (ctx->nd) {
    case 8: return launch_row_reduce_t<PosT, 8>(ctx, mode, d_row_list, n_list, stream, row0);
    case 10: return launch_row_reduce_t<PosT, 10>(ctx, mode, d_row_list, n_list, stream, row0);
    case 27: return launch_row_reduce_t<PosT, 27>(ctx, mode, d_row_list, n_list, stream, row0);
    case 81: return launch_row_reduce_t<PosT, 81>(ctx, mode, d_row_list, n_list, stream, row0);
    default: return launch_row_reduce_t<PosT, 0>(ctx, mode, d_row_list, n_list, stream, row0);
  }
}

}  // namespace

int feasm_launch_row_reduce(feasm_ctx* ctx, int mode, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream, int64_t row0) {
  if (ctx->pos_bytes == 1) return launch_row_reduce<uint8_t>(ctx, mode, d_row_list, n_list, stream, row0);
  return launch_row_reduce<uint16_t>(ctx, mode, d_row_list, n_list, stream, row0);
}

int feasm_launch_constrain_elements(feasm_ctx* ctx, int mode, int64_t e0, int64_t e1, double* d_kf) {
  const int64_t n = e1 - e0;
  k_constrain_elements<<<(unsigned)((n * 32 + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_con_flag, ctx->d_con_val,
                                                                                   ctx->nd, feasm_ldr(ctx->nd), mode, e0, e1, d_kf);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_launch_pos_flags(feasm_ctx* ctx) {
  const int pld = feasm_pld(ctx->nd, ctx->pos_bytes);
  const uint64_t rows_e = (uint64_t)ctx->n_elem * ctx->nd, total = rows_e * pld;
  const unsigned grid = (unsigned)((total + 255) / 256);
  if (grid) {
    if (ctx->pos_bytes == 1)
      k_pos_flags<uint8_t><<<grid, 256, 0, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_con_flag, ctx->nd, pld, rows_e, (uint8_t*)ctx->d_pos);
    else
      k_pos_flags<uint16_t><<<grid, 256, 0, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_con_flag, ctx->nd, pld, rows_e, (uint16_t*)ctx->d_pos);
  }
  if (rows_e) k_inc_flags<<<(unsigned)((rows_e + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_con_flag, ctx->nd, rows_e, ctx->d_inc);
  FEASM_CUDA(ctx, cudaGetLastError());
  ctx->pos_flags_valid = true;
  return FEASM_OK;
}

int feasm_launch_pack(feasm_ctx* ctx) {
  const uint64_t n = ctx->send_ptr.empty() ? 0 : ctx->send_ptr.back();
  if (n) k_pack<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_send_slots, n, ctx->nnz, ctx->d_values, ctx->d_rhs, ctx->d_send_buf);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_launch_unpack(feasm_ctx* ctx) {
  // one launch per peer, in peer order: two peers may hit the same owned slot, the order of the adds is fixed
  for (int p = 0; p < ctx->n_peers; ++p) {
    const uint64_t b = ctx->recv_ptr[p], n = ctx->recv_ptr[p + 1] - b;
    if (!n) continue;
    k_unpack<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_recv_slots + b, n, ctx->nnz, ctx->d_recv_buf + b, ctx->d_values, ctx->d_rhs);
  }
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_launch_add_matrix(feasm_ctx* ctx, const double* d_dm, const uint32_t* d_rows, int nr, const uint32_t* d_cols,
                            int nc) {
  const int n = nr * nc;
  k_add_matrix<<<(n + 127) / 128, 128, 0, ctx->stream>>>(d_dm, d_rows, nr, d_cols, nc, ctx->d_row_ptr, ctx->d_col_idx, ctx->d_values, ctx->d_flags);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}
