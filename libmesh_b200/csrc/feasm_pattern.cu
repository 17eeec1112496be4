// feasm_pattern.cu — setup kernels: row incidence lists, device sparsity build, element->CSR position map.
//
// Replaces SparsityPattern::Build (src/base/sparsity_pattern.C:96-363: per element sorted_connected_dofs, merged
// row by row into a sorted-unique Graph) with a row-parallel formulation: row r = sorted unique union of the
// dof lists of the elements incident to r.  The result is bit-identical to the host Graph.  The position map
// replaces the per-entry std::equal_range of LaspackMatrix::add_matrix (src/numerics/laspack_matrix.C:713-732)
// with one precomputed uint16 per (element, i, j).
#include <algorithm>
#include <cub/cub.cuh>

#include "feasm_internal.h"

namespace {

__global__ void k_count_incidence(const uint32_t* __restrict__ elem_dofs, uint64_t total, unsigned long long* cnt) {
  uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i < total) atomicAdd(&cnt[elem_dofs[i]], 1ull);  // counts are order independent
}

__global__ void k_fill_incidence(const uint32_t* __restrict__ elem_dofs, uint64_t total,
                                 const uint64_t* __restrict__ inc_ptr, unsigned long long* cursor, uint32_t* inc) {
  uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i < total) {
    const uint32_t r = elem_dofs[i];
    const unsigned long long k = atomicAdd(&cursor[r], 1ull);
    inc[inc_ptr[r] + k] = (uint32_t)i;  // i = e*nd + a
  }
}

// Lists are short (<= a few dozen): per-row insertion sort makes the order deterministic (ascending e*nd+a =
// element iteration order, the order the reference accumulates in).
__global__ void k_sort_incidence(const uint64_t* __restrict__ inc_ptr, uint32_t* inc, int64_t n_rows) {
  int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= n_rows) return;
  uint32_t* a = inc + inc_ptr[r];
  const int n = (int)(inc_ptr[r + 1] - inc_ptr[r]);
  for (int i = 1; i < n; ++i) {
    const uint32_t v = a[i];
    int j = i - 1;
    while (j >= 0 && a[j] > v) { a[j + 1] = a[j]; --j; }
    a[j + 1] = v;
  }
}

__global__ void k_max_u64_diff(const uint64_t* __restrict__ ptr, int64_t n, unsigned int* out) {
  int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r < n) atomicMax(out, (unsigned int)(ptr[r + 1] - ptr[r]));
}

// One warp per row.  Candidates (global column ids of all dofs of all incident elements) are gathered into
// shared memory, bitonic-sorted, and the unique ones counted (pass 0) or written (pass 1).
template <int PASS>
__global__ void k_row_pattern(const uint32_t* __restrict__ elem_dofs, const uint32_t* __restrict__ gid, int nd,
                              const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc, int64_t n_rows,
                              int cap /*pow2*/, uint32_t* __restrict__ counts, const uint64_t* __restrict__ row_ptr,
                              uint32_t* __restrict__ col_idx) {
  extern __shared__ uint32_t s_all[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r = blockIdx.x * (int64_t)(blockDim.x >> 5) + warp;
  if (r >= n_rows) return;
  uint32_t* s = s_all + (size_t)warp * cap;
  const uint64_t k0 = inc_ptr[r], k1 = inc_ptr[r + 1];
  const int n = (int)(k1 - k0) * nd;
  int p2 = 32;
  while (p2 < n) p2 <<= 1;
  for (int t = lane; t < p2; t += 32) {
    uint32_t v = 0xffffffffu;
    if (t < n) {
      const uint32_t ea = inc[k0 + t / nd];
      const uint32_t dof = elem_dofs[(uint64_t)(ea / nd) * nd + (t % nd)];
      v = gid ? gid[dof] : dof;
    }
    s[t] = v;
  }
  __syncwarp();
  for (int k = 2; k <= p2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < p2; t += 32) {
        const int u = t ^ j;
        if (u > t) {
          const uint32_t a = s[t], b = s[u];
          const bool up = ((t & k) == 0);
          if ((a > b) == up) { s[t] = b; s[u] = a; }
        }
      }
      __syncwarp();
    }
  // unique
  uint32_t base = 0;
  for (int t0 = 0; t0 < p2; t0 += 32) {
    const int t = t0 + lane;
    const uint32_t v = s[t];
    const bool keep = (v != 0xffffffffu) && (t == 0 || s[t - 1] != v);
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (PASS == 1 && keep) col_idx[row_ptr[r] + base + __popc(m & ((1u << lane) - 1))] = v;
    base += __popc(m);
  }
  if (PASS == 0 && lane == 0) counts[r] = base;
}

__global__ void k_counts_to_u64(const uint32_t* __restrict__ c, uint64_t* out, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = c[i];
}

// Position-map row of record (e,a): entry b < nd = index of column gid(dof_b) inside CSR row dof_a; entry nd = the Fe
// cell; entry nd+1 = RR_LAST when (e,a) is the last contributor of its row (feasm_internal.h).
template <typename PosT>
__global__ void k_build_pos(const uint32_t* __restrict__ elem_dofs, const uint32_t* __restrict__ gid, int nd, int pld,
                            int64_t n_elem, const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col_idx,
                            const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc, unsigned fe_cell,
                            unsigned char* __restrict__ rec, int recb, int pos_off, int* flags) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  const uint64_t total = (uint64_t)n_elem * nd * pld;
  if (i >= total) return;
  const uint64_t ea = i / pld;
  const int b = (int)(i % pld);
  PosT* pos = reinterpret_cast<PosT*>(rec + ea * recb + pos_off) + b;
  const uint32_t row = elem_dofs[ea];
  if (b >= nd) {
    PosT v = 0;
    if (b == nd) v = (PosT)fe_cell;
    else if (b == nd + 1) v = (PosT)(inc[inc_ptr[row + 1] - 1] == (uint32_t)ea ? 1u : 0u);
    *pos = v;
    return;
  }
  const uint64_t e = ea / nd;
  const uint32_t dcol = elem_dofs[e * nd + b];
  const uint32_t col = gid ? gid[dcol] : dcol;
  const uint32_t* c = col_idx + row_ptr[row];
  int lo = 0, hi = (int)(row_ptr[row + 1] - row_ptr[row]);
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (c[mid] < col) lo = mid + 1; else hi = mid;
  }
  if (lo >= (int)(row_ptr[row + 1] - row_ptr[row]) || c[lo] != col) {
    atomicAdd(&flags[1], 1);
    lo = 0;
  }
  *pos = (PosT)lo;
}

inline unsigned grid_for(uint64_t n, int block) { return (unsigned)((n + block - 1) / block); }

}  // namespace

int feasm_build_incidence(feasm_ctx* ctx) {
  if (ctx->d_inc_ptr) return FEASM_OK;
  const uint64_t total = (uint64_t)ctx->n_elem * ctx->nd;
  if (total >= 0x80000000ull) return feasm_fail(ctx, FEASM_ERR_ARG, "n_elem*nd exceeds 2^31 incidence ids");
  const int64_t nr = ctx->n_rows;
  unsigned long long* d_cnt = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&d_cnt, sizeof(unsigned long long) * (nr + 1)));
  FEASM_CUDA(ctx, cudaMemsetAsync(d_cnt, 0, sizeof(unsigned long long) * (nr + 1), ctx->stream));
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_inc_ptr, sizeof(uint64_t) * (nr + 1)));
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_inc, sizeof(uint32_t) * total));
  k_count_incidence<<<grid_for(total, 256), 256, 0, ctx->stream>>>(ctx->d_elem_dofs, total, d_cnt);
  void* tmp = nullptr;
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, (uint64_t*)d_cnt, ctx->d_inc_ptr, nr + 1, ctx->stream);
  FEASM_CUDA(ctx, cudaMalloc(&tmp, tmp_bytes));
  cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, (uint64_t*)d_cnt, ctx->d_inc_ptr, nr + 1, ctx->stream);
  FEASM_CUDA(ctx, cudaMemsetAsync(d_cnt, 0, sizeof(unsigned long long) * (nr + 1), ctx->stream));
  k_fill_incidence<<<grid_for(total, 256), 256, 0, ctx->stream>>>(ctx->d_elem_dofs, total, ctx->d_inc_ptr, d_cnt, ctx->d_inc);
  k_sort_incidence<<<grid_for(nr, 128), 128, 0, ctx->stream>>>(ctx->d_inc_ptr, ctx->d_inc, nr);
  unsigned int* d_max = (unsigned int*)d_cnt;
  FEASM_CUDA(ctx, cudaMemsetAsync(d_max, 0, sizeof(unsigned int), ctx->stream));
  k_max_u64_diff<<<grid_for(nr, 256), 256, 0, ctx->stream>>>(ctx->d_inc_ptr, nr, d_max);
  FEASM_CUDA(ctx, cudaMemcpyAsync(&ctx->max_inc, d_max, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(tmp);
  cudaFree(d_cnt);
  return FEASM_OK;
}

int feasm_device_build_pattern(feasm_ctx* ctx) {
  int rc = feasm_build_incidence(ctx);
  if (rc) return rc;
  const int64_t nr = ctx->n_rows;
  int cap = 32;
  while (cap < (int)ctx->max_inc * ctx->nd) cap <<= 1;
  int warps = 8;
  while (warps > 1 && (size_t)warps * cap * 4 > 96 * 1024) warps >>= 1;
  const size_t smem = (size_t)warps * cap * 4;
  if (smem > ctx->smem_optin) return feasm_fail(ctx, FEASM_ERR_NOMEM, "row candidate list exceeds shared memory");
  FEASM_CUDA(ctx, cudaFuncSetAttribute(k_row_pattern<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  FEASM_CUDA(ctx, cudaFuncSetAttribute(k_row_pattern<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  uint32_t* d_counts = nullptr;
  uint64_t* d_counts64 = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&d_counts, sizeof(uint32_t) * (nr + 1)));
  FEASM_CUDA(ctx, cudaMalloc(&d_counts64, sizeof(uint64_t) * (nr + 1)));
  FEASM_CUDA(ctx, cudaMemsetAsync(d_counts, 0, sizeof(uint32_t) * (nr + 1), ctx->stream));
  if (ctx->d_row_ptr) cudaFree(ctx->d_row_ptr);
  if (ctx->d_col_idx) cudaFree(ctx->d_col_idx);
  ctx->d_col_idx = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_row_ptr, sizeof(uint64_t) * (nr + 1)));
  const unsigned grid = grid_for(nr, warps);
  k_row_pattern<0><<<grid, warps * 32, smem, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_gid, ctx->nd, ctx->d_inc_ptr,
                                                             ctx->d_inc, nr, cap, d_counts, nullptr, nullptr);
  k_counts_to_u64<<<grid_for(nr + 1, 256), 256, 0, ctx->stream>>>(d_counts, d_counts64, nr + 1);
  void* tmp = nullptr;
  size_t tmp_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_counts64, ctx->d_row_ptr, nr + 1, ctx->stream);
  FEASM_CUDA(ctx, cudaMalloc(&tmp, tmp_bytes));
  cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, d_counts64, ctx->d_row_ptr, nr + 1, ctx->stream);
  uint64_t nnz = 0;
  FEASM_CUDA(ctx, cudaMemcpyAsync(&nnz, ctx->d_row_ptr + nr, sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->nnz = nnz;
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_col_idx, sizeof(uint32_t) * (nnz ? nnz : 1)));
  k_row_pattern<1><<<grid, warps * 32, smem, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_gid, ctx->nd, ctx->d_inc_ptr,
                                                             ctx->d_inc, nr, cap, d_counts, ctx->d_row_ptr, ctx->d_col_idx);
  FEASM_CUDA(ctx, cudaGetLastError());
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(tmp);
  cudaFree(d_counts);
  cudaFree(d_counts64);
  return FEASM_OK;
}

int feasm_build_pos(feasm_ctx* ctx) {
  unsigned int* d_max = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&d_max, sizeof(unsigned int)));
  FEASM_CUDA(ctx, cudaMemsetAsync(d_max, 0, sizeof(unsigned int), ctx->stream));
  k_max_u64_diff<<<grid_for(ctx->n_rows, 256), 256, 0, ctx->stream>>>(ctx->d_row_ptr, ctx->n_rows, d_max);
  FEASM_CUDA(ctx, cudaMemcpyAsync(&ctx->max_row_len, d_max, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(d_max);
  if (ctx->max_row_len > 32767u) return feasm_fail(ctx, FEASM_ERR_ARG, "row longer than 32767 entries: position map overflows");
  ctx->pos_bytes = ctx->max_row_len <= 255u ? 1 : 2;  // positions and the Fe cell (= max_row_len) fit the entry
  const int pld = feasm_pld(ctx->nd, ctx->pos_bytes), recb = feasm_recb(ctx->nd, ctx->pos_bytes);
  const uint64_t total = (uint64_t)ctx->n_elem * ctx->nd * pld;
  // the element records: values part rewritten by every assembly, position rows written here once
  if (ctx->d_kf) cudaFree(ctx->d_kf);
  ctx->d_kf = nullptr;
  const size_t rec_bytes = (size_t)ctx->n_elem * ctx->nd * recb + 256;  // slack: the row reduction prefetches one record ahead
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_kf, rec_bytes));
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_kf, 0, rec_bytes, ctx->stream));
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_flags, 0, 2 * sizeof(int), ctx->stream));
  if (total) {
    unsigned char* rec = reinterpret_cast<unsigned char*>(ctx->d_kf);
    const int pos_off = feasm_ldr(ctx->nd) * 8;
    if (ctx->pos_bytes == 1)
      k_build_pos<uint8_t><<<grid_for(total, 256), 256, 0, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_gid, ctx->nd, pld, ctx->n_elem,
                                                                         ctx->d_row_ptr, ctx->d_col_idx, ctx->d_inc_ptr, ctx->d_inc,
                                                                         ctx->max_row_len, rec, recb, pos_off, ctx->d_flags);
    else
      k_build_pos<uint16_t><<<grid_for(total, 256), 256, 0, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_gid, ctx->nd, pld, ctx->n_elem,
                                                                          ctx->d_row_ptr, ctx->d_col_idx, ctx->d_inc_ptr, ctx->d_inc,
                                                                          ctx->max_row_len, rec, recb, pos_off, ctx->d_flags);
  }
  int flags[2] = {0, 0};
  FEASM_CUDA(ctx, cudaMemcpyAsync(flags, ctx->d_flags, sizeof(flags), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (flags[1]) return feasm_fail(ctx, FEASM_ERR_PATTERN, "element couples dofs that are not in the sparsity pattern");
  return FEASM_OK;
}
