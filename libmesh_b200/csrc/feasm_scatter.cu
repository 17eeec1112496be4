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
#include "feasm_internal.h"

namespace {

template <typename PosT> struct PosBits;
template <> struct PosBits<uint8_t> { static constexpr unsigned FLAG = 0x80u, MASK = 0x7fu; };
template <> struct PosBits<uint16_t> { static constexpr unsigned FLAG = 0x8000u, MASK = 0x7fffu; };

// A warp owns a batch of 32 consecutive entries of the row list (or 32 consecutive rows).  Lane l holds the
// pointers of row l of the batch; the contributors of the batch are walked in chunks of U as a three-stage software
// pipeline — incidence ids of chunk c+2, element rows + positions of chunk c+1 in flight while chunk c is
// accumulated into the row image in shared memory — so several KB per warp are outstanding at any time.
template <typename PosT, int NDC, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 2)
k_row_reduce(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc,
             const PosT* __restrict__ pos, const double* __restrict__ kf, int nd, int ldr, int pld,
             const uint32_t* __restrict__ row_list, int64_t n_list, int cap, const uint8_t* __restrict__ con_flag,
             const double* __restrict__ con_val, const uint32_t* __restrict__ elem_dofs, int mode,
             double* __restrict__ values, double* __restrict__ rhs) {
  constexpr int U = 8 / NDC;
  constexpr unsigned FULL = 0xffffffffu, INVALID = 0xffffffffu;
  extern __shared__ double s_img[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t base = (blockIdx.x * (int64_t)WARPS + warp) * 32;
  if (base >= n_list) return;
  const int nrows = (int)(n_list - base < 32 ? n_list - base : 32);
  double* img = s_img + (size_t)warp * cap;
  for (int t = lane; t < cap; t += 32) img[t] = 0.;

  // ---- per-lane row descriptors ---------------------------------------------------------------------------
  uint32_t my_r = 0;
  uint32_t my_ip0 = 0;  // incidence offsets fit 32 bits (n_elem*nd < 2^32 is checked at setup)
  uint64_t my_rp0 = 0;
  int my_cnt = 0, my_len = 0, my_cf = 0;
  double my_g = 0.;
  if (lane < nrows) {
    my_r = row_list ? row_list[base + lane] : (uint32_t)(base + lane);
    my_ip0 = (uint32_t)inc_ptr[my_r];
    my_cnt = (int)(inc_ptr[my_r + 1] - inc_ptr[my_r]);
    my_rp0 = row_ptr[my_r];
    my_len = (int)(row_ptr[my_r + 1] - my_rp0);
    my_cf = con_flag[my_r];
    if (my_cf && mode == FEASM_CONSTRAIN_HETEROGENEOUS) my_g = con_val[my_r];
  }
  __syncwarp();

  struct It { int j, k, cnt; uint32_t ip0; };
  auto load_row = [&](It& it) {
    it.cnt = __shfl_sync(FULL, my_cnt, it.j);
    it.ip0 = __shfl_sync(FULL, my_ip0, it.j);
  };
  auto advance = [&](It& it) {
    it.k += U;
    if (it.k >= it.cnt) { ++it.j; it.k = 0; if (it.j < nrows) load_row(it); }
  };
  auto stage1 = [&](const It& it) -> uint32_t {
    return (lane < U && it.k + lane < it.cnt) ? inc[it.ip0 + it.k + lane] : INVALID;
  };
  double vN[U][NDC];
  unsigned pN[U][NDC];
  double fN = 0.;
  auto stage2 = [&](uint32_t ea) {
    fN = (ea != INVALID) ? __ldcs(kf + (uint64_t)ea * ldr + nd) : 0.;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t eu = __shfl_sync(FULL, ea, u);
#pragma unroll
      for (int c = 0; c < NDC; ++c) {
        const int b = lane + 32 * c;
        vN[u][c] = 0.; pN[u][c] = 0;
        if (eu != INVALID && b < nd) {
          vN[u][c] = __ldcs(kf + (uint64_t)eu * ldr + b);
          pN[u][c] = pos[(uint64_t)eu * pld + b];
        }
      }
    }
  };

  It i1{0, 0, 0, 0}, i2, i3;
  load_row(i1);
  i2 = i1; i3 = i1;
  uint32_t ea1 = stage1(i1), ea2, ea3;
  advance(i1);
  ea2 = ea1;
  stage2(ea2);
  ea1 = INVALID;
  if (i1.j < nrows) { ea1 = stage1(i1); advance(i1); }

  double bsum = 0., corr = 0.;
  while (i3.j < nrows) {
    double vC[U][NDC];
    unsigned pC[U][NDC];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int c = 0; c < NDC; ++c) { vC[u][c] = vN[u][c]; pC[u][c] = pN[u][c]; }
    const double fC = fN;
    ea3 = ea2;
    // issue the loads of the next chunk, then the incidence ids of the one after
    advance(i2);
    if (i2.j < nrows) { ea2 = ea1; stage2(ea2); }
    if (i1.j < nrows) { ea1 = stage1(i1); advance(i1); } else ea1 = INVALID;

    // ---- accumulate chunk i3 into the row image ------------------------------------------------------------
    const int row_con = __shfl_sync(FULL, my_cf, i3.j);
    const double row_g = __shfl_sync(FULL, my_g, i3.j);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t eu = __shfl_sync(FULL, ea3, u);
      const double fu = __shfl_sync(FULL, fC, u);
      if (eu != INVALID) {  // warp-uniform
        const int a = (int)(eu % (uint32_t)nd);
#pragma unroll
        for (int c = 0; c < NDC; ++c) {
          const int b = lane + 32 * c;
          if (b < nd) {
            double val = vC[u][c];
            const unsigned pp = pC[u][c];
            if (row_con) {
              val = (b == a) ? 1. : 0.;
            } else if (pp & PosBits<PosT>::FLAG) {
              if (mode == FEASM_CONSTRAIN_HETEROGENEOUS) corr += val * con_val[elem_dofs[(uint64_t)(eu / (uint32_t)nd) * nd + b]];
              val = 0.;
            }
            img[pp & PosBits<PosT>::MASK] += val;  // positions within one element row are distinct
          }
        }
        bsum += row_con ? row_g : fu;
        __syncwarp();
      }
    }
    // ---- last chunk of the row: write the row out -------------------------------------------------------------
    if (i3.k + U >= i3.cnt) {
      const uint64_t rp0 = __shfl_sync(FULL, my_rp0, i3.j);
      const int len = __shfl_sync(FULL, my_len, i3.j);
      const uint32_t r = __shfl_sync(FULL, my_r, i3.j);
      if (__any_sync(FULL, corr != 0.)) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) corr += __shfl_xor_sync(FULL, corr, o);
        bsum -= corr;
      }
      for (int t = lane; t < len; t += 32) {
        __stcs(values + rp0 + t, img[t]);
        img[t] = 0.;
      }
      if (lane == 0) rhs[r] = bsum;
      bsum = 0.; corr = 0.;
      __syncwarp();
    }
    advance(i3);
  }
}

// Fallback for rows longer than the shared-memory budget: accumulate straight into the (zeroed) global row.
template <typename PosT>
__global__ void k_row_reduce_global(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr,
                                    const uint32_t* __restrict__ inc, const PosT* __restrict__ pos,
                                    const double* __restrict__ kf, int nd, int ldr, int pld, int64_t n_rows,
                                    const uint8_t* __restrict__ con_flag, const double* __restrict__ con_val,
                                    const uint32_t* __restrict__ elem_dofs, int mode, double* __restrict__ values,
                                    double* __restrict__ rhs) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rows) return;
  double* row = values + row_ptr[r];
  const int len = (int)(row_ptr[r + 1] - row_ptr[r]);
  for (int t = lane; t < len; t += 32) row[t] = 0.;
  __syncwarp();
  const int row_con = con_flag[r];
  const double row_g = (row_con && mode == FEASM_CONSTRAIN_HETEROGENEOUS) ? con_val[r] : 0.;
  double b = 0., corr = 0.;
  for (uint64_t k = inc_ptr[r]; k < inc_ptr[r + 1]; ++k) {
    const uint64_t ea = inc[k];
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

template <typename PosT, int NDC>
int launch_row_reduce_t(feasm_ctx* ctx, int mode, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream) {
  constexpr int WARPS = 8;
  const int cap = (int)ctx->max_row_len;
  const size_t smem = sizeof(double) * (size_t)WARPS * cap;
  const int ldr = feasm_ldr(ctx->nd), pld = feasm_pld(ctx->nd, (int)sizeof(PosT));
  const PosT* pos = static_cast<const PosT*>(ctx->d_pos);
  if (smem <= ctx->smem_optin) {
    auto kern = k_row_reduce<PosT, NDC, WARPS>;
    FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t batches = (n_list + 31) / 32;
    const unsigned grid = (unsigned)((batches + WARPS - 1) / WARPS);
    if (grid)
      kern<<<grid, WARPS * 32, smem, stream>>>(ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, pos, ctx->d_kf, ctx->nd, ldr, pld,
                                               d_row_list, n_list, cap, ctx->d_con_flag, ctx->d_con_val, ctx->d_elem_dofs,
                                               mode, ctx->d_values, ctx->d_rhs);
  } else {
    if (d_row_list) return feasm_fail(ctx, FEASM_ERR_NOMEM, "rows too long for the shared-memory row reduction");
    const unsigned grid = (unsigned)((ctx->n_rows * 32 + 255) / 256);
    k_row_reduce_global<PosT><<<grid, 256, 0, stream>>>(ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, pos, ctx->d_kf, ctx->nd, ldr,
                                                        pld, ctx->n_rows, ctx->d_con_flag, ctx->d_con_val, ctx->d_elem_dofs,
                                                        mode, ctx->d_values, ctx->d_rhs);
  }
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

template <typename PosT>
int launch_row_reduce(feasm_ctx* ctx, int mode, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream) {
  const int ndc = (ctx->nd + 31) / 32;
  if (ndc == 1) return launch_row_reduce_t<PosT, 1>(ctx, mode, d_row_list, n_list, stream);
  if (ndc == 2) return launch_row_reduce_t<PosT, 2>(ctx, mode, d_row_list, n_list, stream);
  if (ndc == 3) return launch_row_reduce_t<PosT, 3>(ctx, mode, d_row_list, n_list, stream);
  if (ndc == 4) return launch_row_reduce_t<PosT, 4>(ctx, mode, d_row_list, n_list, stream);
  return feasm_fail(ctx, FEASM_ERR_ARG, "more than 128 dofs per element");
}

}  // namespace

int feasm_launch_row_reduce(feasm_ctx* ctx, int mode, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream) {
  if (ctx->pos_bytes == 1) return launch_row_reduce<uint8_t>(ctx, mode, d_row_list, n_list, stream);
  return launch_row_reduce<uint16_t>(ctx, mode, d_row_list, n_list, stream);
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
