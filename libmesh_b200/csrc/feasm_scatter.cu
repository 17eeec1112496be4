// feasm_scatter.cu — the deterministic scatter: row reduction of the element records into CSR.
//
// Row reduction = SparseMatrix::add_matrix / NumericVector::add_vector (laspack_matrix.C:259-275,
// numeric_vector.C:690-697) turned inside out: a warp sums, for each CSR row it owns, the element records
// REC[e][a] incident to that row in element order through the position row each record carries — no atomics, bitwise
// reproducible, and the same accumulation order as the reference's serial element loop.  The Dirichlet constraints
// (DofMap::heterogeneously_constrain_element_matrix_and_vector, src/base/dof_map_constraints.C:2518-2614) were applied
// to the records by the element kernel, as the reference applies them to Ke/Fe before add_matrix.
#include <cstdlib>

#include "feasm_rr_dev.cuh"

namespace {
// Stand-alone row reduction: one batch of RB rows per warp.  ND = dofs per element when known at compile time
// (strides and lane loops fold away), 0 = generic (runtime nd).
template <typename PosT, int ND, int WARPS, bool BULK>
__global__ void __launch_bounds__(WARPS * 32)
k_row_reduce(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc,
             const unsigned char* __restrict__ rec, int nd_rt, int RB, int lcap, int cap, int fe_cell, int G,
             const uint32_t* __restrict__ row_list, int64_t row0, int64_t n_list, double* __restrict__ values,
             double* __restrict__ rhs) {
  const int nd = ND ? ND : nd_rt;
  const int recb = feasm_recb(nd, (int)sizeof(PosT));
  extern __shared__ __align__(16) unsigned char s_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t base = row0 + (blockIdx.x * (int64_t)WARPS + warp) * RB;  // rows [row0, n_list) (positions, with a row list)
  if (base >= n_list) return;
  const int nrows = (int)(n_list - base < RB ? n_list - base : RB);
  const uint32_t warp_bytes = (uint32_t)(RR_NG * G * recb + cap * 8 + lcap * 4 + 8 * RR_NG);
  const uint32_t ring_a = (uint32_t)__cvta_generic_to_shared(s_raw) + warp * warp_bytes;
  const uint32_t img_a = ring_a + RR_NG * G * recb;
  const uint32_t list_a = img_a + cap * 8;
  const uint32_t bar_a = list_a + lcap * 4;
  for (int t = lane; t < cap; t += 32) sts_f64(img_a + 8 * t, 0.);
  if (BULK) {
    if (lane < RR_NG) mbar_init(bar_a + 8 * lane, 1);
    mbar_init_fence();
  }
  __syncwarp();
  uint32_t phase = 0;
  rr_process_batch<PosT, ND, BULK>(base, nrows, lane, ring_a, img_a, list_a, G, nd_rt, fe_cell, row_ptr, inc_ptr, inc, rec, row_list,
                                   values, rhs, bar_a, phase);
}

// Fallback for rows longer than the shared-memory budget: one warp per row accumulates straight into the (zeroed)
// global row, contributors in element order.
// ---- 8 or 10 dofs per element (HEX8 / TET10, one variable): four or two rows per warp instruction ------------------------
// With lane = local dof only 9 of 32 lanes carry a contributor and the ring kernel above is instruction-bound for HEX8
// (ncu: issue slots 87 % busy at half the HBM peak).  Here a group of 8 (16 for TET10) lanes owns a row: lane = local dof b of the
// contributor; the incidence ids of up to 8 contributors are fetched by the group's lanes and broadcast with a
// width-8 shuffle, all their record entries (value, position byte; lane 0 also Fe) are loaded straight from global
// memory — 16 independent loads per lane in flight, no staging ring — and then added into the group's row image in
// shared memory strictly in element order (bitwise the same sums as k_row_reduce).  Rows are dealt to consecutive
// groups, so a warp writes four consecutive CSR rows.
template <typename PosT, int ND, int GROUP>
__global__ void __launch_bounds__(256)
k_row_reduce_group(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc,
                 const unsigned char* __restrict__ rec, int img_cells, const uint32_t* __restrict__ row_list, int64_t row0,
                 int64_t n_list, double* __restrict__ values, double* __restrict__ rhs) {
  static_assert((GROUP == 8 || GROUP == 16) && ND <= GROUP, "a group of 8 or 16 lanes, one lane per local dof");
  constexpr int ldrb = feasm_ldr(ND) * 8, recb = feasm_recb(ND, (int)sizeof(PosT));
  constexpr int GPB = 256 / GROUP;  // groups per block
  extern __shared__ double s_img[];  // [groups of the block][img_cells]
  const int lane = threadIdx.x & 31, gl = lane & (GROUP - 1);
  const unsigned gmask = (GROUP == 8 ? 0xffu : 0xffffu) << (lane & ~(GROUP - 1));
  double* img = s_img + (threadIdx.x / GROUP) * img_cells;
  for (int t = gl; t < img_cells; t += GROUP) img[t] = 0.;
  __syncwarp(gmask);
  const int64_t n_groups = (int64_t)gridDim.x * GPB;
  for (int64_t i = row0 + blockIdx.x * (int64_t)GPB + (threadIdx.x / GROUP); i < n_list; i += n_groups) {
    const uint32_t r = row_list ? row_list[i] : (uint32_t)i;
    const uint64_t i0 = inc_ptr[r], p0 = row_ptr[r];
    const int cnt = (int)(inc_ptr[r + 1] - i0), len = (int)(row_ptr[r + 1] - p0);
    double fe = 0.;
    for (int c0 = 0; c0 < cnt; c0 += 8) {
      const int n = cnt - c0 < 8 ? cnt - c0 : 8;
      const uint32_t my_id = gl < n ? inc[i0 + c0 + gl] : 0u;
      double v[8], f[8];
      unsigned pp[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint32_t id = __shfl_sync(gmask, my_id, c, GROUP);
        v[c] = 0.; f[c] = 0.; pp[c] = 0;
        if (c < n && gl < ND) {
          const unsigned char* R = rec + (uint64_t)id * (uint32_t)recb;
          v[c] = reinterpret_cast<const double*>(R)[gl];
          pp[c] = reinterpret_cast<const PosT*>(R + ldrb)[gl];
          if (gl == 0) f[c] = reinterpret_cast<const double*>(R)[ND];
        }
      }
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        if (c < n && gl < ND) {  // positions within one record are distinct; contributors follow one another in element order
          img[pp[c]] += v[c];
          fe += f[c];
        }
        __syncwarp(gmask);
      }
    }
    for (int t = gl; t < len; t += GROUP) {
      __stcs(values + p0 + t, img[t]);
      img[t] = 0.;
    }
    if (gl == 0) rhs[r] = fe;
    __syncwarp(gmask);
  }
}

// Residual without the global matrix: r_i = sum over the (e, a) pairs of row i of ( Ke(a,:) . u_e - Fe(a) ), from the
// element records (the rows of the constrained element matrices), in element order — what a
// NonlinearImplicitSystem::ComputeResidual / FEMSystem::assembly(get_residual = true, get_jacobian = false) computes
// for this linear problem, R = K(u) u - F.  Two passes: k_elem_apply, one warp per element, gathers u_e ONCE per element
// (summing per row instead would gather it once per record: 729 instead of 27 loads of u per HEX27 element — measured
// 5.2 ms, L2-bound), reads the element's records row by row, coalesced, and leaves y(e,a) = Ke(a,:) . u_e - Fe(a);
// k_row_sum adds the y of a row's incidences in element order.  Fixed shuffle tree and fixed order: deterministic.
__global__ void k_elem_apply(const unsigned char* __restrict__ rec, int recb, const uint32_t* __restrict__ elem_dofs, int nd,
                             const double* __restrict__ u, int64_t n_elem, double* __restrict__ y) {
  extern __shared__ double s_ue[];  // [warps][nd]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  double* ue = s_ue + (size_t)warp * nd;
  for (int64_t e = blockIdx.x * (int64_t)warps + warp; e < n_elem; e += (int64_t)gridDim.x * warps) {
    for (int b = lane; b < nd; b += 32) ue[b] = u[elem_dofs[e * nd + b]];
    __syncwarp();
    const unsigned char* R0 = rec + (uint64_t)e * nd * (uint32_t)recb;
    constexpr int RIF = 8;  // rows in flight
    for (int a0 = 0; a0 < nd; a0 += RIF) {
      double s[RIF];
#pragma unroll
      for (int j = 0; j < RIF; ++j) s[j] = 0.;
      for (int b = lane; b <= nd; b += 32) {  // b outermost: the RIF row loads of one pass are independent and go out together
        const double ub = b < nd ? ue[b] : -1.;  // entry nd is Fe: y = Ke(a,:) . u_e - Fe(a)
        double v[RIF];
#pragma unroll
        for (int j = 0; j < RIF; ++j)
          v[j] = a0 + j < nd ? reinterpret_cast<const double*>(R0 + (size_t)(a0 + j) * recb)[b] : 0.;
#pragma unroll
        for (int j = 0; j < RIF; ++j) s[j] += v[j] * ub;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1)
#pragma unroll
        for (int j = 0; j < RIF; ++j) s[j] += __shfl_xor_sync(0xffffffffu, s[j], o);
      double mine = s[0];
#pragma unroll
      for (int j = 1; j < RIF; ++j) mine = lane == j ? s[j] : mine;
      if (lane < RIF && a0 + lane < nd) y[e * nd + a0 + lane] = mine;
    }
    __syncwarp();
  }
}

__global__ void k_row_sum(const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc, const double* __restrict__ y,
                          int64_t n_rows, double* __restrict__ r) {
  const int64_t row = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (row >= n_rows) return;
  double acc = 0.;
  for (uint64_t k = inc_ptr[row]; k < inc_ptr[row + 1]; ++k) acc += y[inc[k]];
  r[row] = acc;
}

template <typename PosT>
__global__ void k_row_reduce_global(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr,
                                    const uint32_t* __restrict__ inc, const unsigned char* __restrict__ rec, int nd,
                                    int64_t row0, int64_t n_rows, double* __restrict__ values, double* __restrict__ rhs) {
  const int lane = threadIdx.x & 31;
  const int64_t r = row0 + ((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  if (r >= n_rows) return;
  const int ldrb = feasm_ldr(nd) * 8, recb = feasm_recb(nd, (int)sizeof(PosT));
  double* row = values + row_ptr[r];
  const int len = (int)(row_ptr[r + 1] - row_ptr[r]);
  for (int t = lane; t < len; t += 32) row[t] = 0.;
  __syncwarp();
  double b = 0.;
  for (uint64_t k = inc_ptr[r]; k < inc_ptr[r + 1]; ++k) {
    const unsigned char* rc = rec + (uint64_t)inc[k] * recb;
    const double* val = reinterpret_cast<const double*>(rc);
    const PosT* pp = reinterpret_cast<const PosT*>(rc + ldrb);
    for (int t = lane; t < nd; t += 32) row[pp[t]] += val[t];
    b += val[nd];
    __syncwarp();
  }
  if (lane == 0) rhs[r] = b;
}

// elem_flag[e] = element e touches a constrained dof: its records get the in-place constraint pass of the element
// kernel (constrain_element, feasm_elem_dev.cuh).  Refreshed when the constrained SET changes.
__global__ void k_elem_flags(const uint32_t* __restrict__ elem_dofs, const uint8_t* __restrict__ con_flag, int nd, int64_t n_elem,
                             uint8_t* __restrict__ elem_flag, uint32_t* __restrict__ list, unsigned* __restrict__ count) {
  const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e >= n_elem) return;
  const uint32_t* d = elem_dofs + e * nd;
  unsigned any = 0;
  for (int b = 0; b < nd; ++b) any |= con_flag[d[b]];
  // 1: Dirichlet dofs only (constrained in place); 2: touches a dof with a non-empty constraint row — the element is widened
  // by feasm_general.cu and must reach its records unconstrained
  const unsigned f = (any & 2u) ? 2u : (any ? 1u : 0u);
  elem_flag[e] = (uint8_t)f;
  if (f == 1u) list[atomicAdd(count, 1u)] = (uint32_t)e;  // order is irrelevant: elements are constrained independently
}

__global__ void k_pack(const uint64_t* __restrict__ slots, uint64_t n, uint64_t nnz, uint64_t n_rows, const double* __restrict__ values,
                       const double* __restrict__ rhs, double* __restrict__ out) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i < n) {
    const uint64_t s = slots[i];
    FEASM_DEV_ASSERT(s < nnz + n_rows);
    out[i] = s < nnz ? values[s] : rhs[s - nnz];
  }
}

// One thread per destination slot: the entries the peers sent for it are added in receive-buffer (= fixed peer) order.
__global__ void k_unpack(const uint64_t* __restrict__ uslot, const uint64_t* __restrict__ seg, const uint64_t* __restrict__ perm,
                         uint64_t n_slots, uint64_t nnz, uint64_t n_rows, uint64_t n_recv, const double* __restrict__ in,
                         double* __restrict__ values, double* __restrict__ rhs) {
  const uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (k >= n_slots) return;
  const uint64_t s = uslot[k];
  FEASM_DEV_ASSERT(s < nnz + n_rows);
  double* dst = s < nnz ? values + s : rhs + (s - nnz);
  double acc = *dst;
  for (uint64_t i = seg[k]; i < seg[k + 1]; ++i) { FEASM_DEV_ASSERT(perm[i] < n_recv); acc += in[perm[i]]; }
  *dst = acc;
}

// feasm_set_constraints on the device: clear the flags of the previous list, then scatter the new (row, value) list
__global__ void k_clear_constraints(const uint32_t* __restrict__ rows, int64_t n, uint8_t* __restrict__ flag, double* __restrict__ val) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) { flag[rows[i]] = 0; val[rows[i]] = 0.; }
}
__global__ void k_scatter_constraints(const uint32_t* __restrict__ rows, const double* __restrict__ vals, int64_t n, bool set_flags,
                                      uint8_t* __restrict__ flag, double* __restrict__ val) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) { if (set_flags) flag[rows[i]] = 1; val[rows[i]] = vals[i]; }
}

// SparseMatrix::add_matrix for one host-supplied dense block (slow route; one thread per entry, entries of one
// call are distinct slots so there is no race).
__global__ void k_add_matrix(const double* __restrict__ dm, const uint32_t* __restrict__ rows, int nr,
                             const uint32_t* __restrict__ cols, int nc, const uint64_t* __restrict__ row_ptr,
                             const uint32_t* __restrict__ col_idx, double* __restrict__ values, int* flags, int mode) {
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
  // add: atomic, so a block that names a dof twice (repeated ids in rows / cols) loses nothing
  if (lo < len && c[lo] == col) { if (mode) values[row_ptr[row] + lo] = v; else atomicAdd(&values[row_ptr[row] + lo], v); }
  else if (v != 0.) atomicAdd(&flags[1], 1);  // zero may be "added" to a missing entry (sparse_matrix.h:300-303)
}

// NumericVector::add_vector for one host-supplied list: one thread walks it, so repeated rows add in order
__global__ void k_add_vector(const double* __restrict__ v, const uint32_t* __restrict__ rows, int n, double* __restrict__ rhs) {
  for (int i = 0; i < n; ++i) rhs[rows[i]] += v[i];
}

template <typename PosT, int ND>
int launch_row_reduce_t(feasm_ctx* ctx, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream, int64_t row0) {
  // nd = 10 (TET10, 16-lane groups) was tried and is slower than the ring kernel (0.75 vs 0.52 ms at 786 k tets): rows have
  // 2..48 contributors there, the two groups of a warp wait for each other
  if (ND == 8) {
    constexpr int GROUP = ND == 8 ? 8 : 16, GPB = 256 / GROUP;
    static const bool ring = getenv("FEASM_RR_GROUP") && atoi(getenv("FEASM_RR_GROUP")) == 0;  // A/B knob: 0 = the ring kernel
    const int img_cells = (int)ctx->max_row_len + 1;
    const size_t smem = sizeof(double) * GPB * (size_t)img_cells;
    if (!ring && smem <= 48 * 1024) {
      const int64_t rows = n_list - row0;
      int64_t grid = (rows + GPB - 1) / GPB;
      const int64_t cap = (int64_t)ctx->sm_count * 8 * 4;  // a few rows per group
      if (grid > cap) grid = cap;
      if (grid > 0)
        k_row_reduce_group<PosT, (ND == 8 || ND == 10) ? ND : 8, GROUP><<<(unsigned)grid, 256, smem, stream>>>(
            ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, reinterpret_cast<const unsigned char*>(ctx->d_kf), img_cells, d_row_list, row0, n_list,
            ctx->d_values, ctx->d_rhs);
      FEASM_CUDA(ctx, cudaGetLastError());
      return FEASM_OK;
    }
  }
  constexpr int WARPS = 4;
  const int cap = ((int)ctx->max_row_len + 2) & ~1;  // row cells + the Fe cell; even: every warp's ring stays 16 B aligned
  const int recb = feasm_recb(ctx->nd, (int)sizeof(PosT));
  static const int g_env = getenv("FEASM_RR_G") ? atoi(getenv("FEASM_RR_G")) : 0;  // tuning knob
  const int G = g_env > 0 ? g_env : rr_stage(recb);
  const int lcap = 256;
  int RB = ctx->max_inc ? (int)(lcap / ctx->max_inc) : 32;
  if (RB > 32) RB = 32;
  const size_t warp_bytes = (size_t)RR_NG * G * recb + (size_t)cap * 8 + (size_t)lcap * 4 + 8 * RR_NG;
  const size_t smem = WARPS * warp_bytes;
  const unsigned char* rec = reinterpret_cast<const unsigned char*>(ctx->d_kf);
  if (RB >= 1 && smem <= ctx->smem_optin && ctx->nd <= 128) {
    static const bool bulk = !(getenv("FEASM_RR_BULK") && atoi(getenv("FEASM_RR_BULK")) == 0);  // 0: cp.async ring fill
    auto kern = bulk && G <= 32 ? k_row_reduce<PosT, ND, WARPS, true> : k_row_reduce<PosT, ND, WARPS, false>;
    FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t batches = (n_list - row0 + RB - 1) / RB;
    const unsigned grid = (unsigned)((batches + WARPS - 1) / WARPS);
    if (grid)
      kern<<<grid, WARPS * 32, smem, stream>>>(ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, rec, ctx->nd, RB, lcap, cap,
                                               (int)ctx->max_row_len, G, d_row_list, row0, n_list, ctx->d_values, ctx->d_rhs);
  } else {
    if (d_row_list) return feasm_fail(ctx, FEASM_ERR_NOMEM, "rows too long for the shared-memory row reduction");
    const unsigned grid = (unsigned)(((n_list - row0) * 32 + 255) / 256);
    if (grid)
      k_row_reduce_global<PosT><<<grid, 256, 0, stream>>>(ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, rec, ctx->nd, row0, n_list,
                                                          ctx->d_values, ctx->d_rhs);
  }
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

template <typename PosT>
int launch_row_reduce(feasm_ctx* ctx, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream, int64_t row0) {
  switch (ctx->nd) {
    case 8: return launch_row_reduce_t<PosT, 8>(ctx, d_row_list, n_list, stream, row0);
    case 10: return launch_row_reduce_t<PosT, 10>(ctx, d_row_list, n_list, stream, row0);
    case 18: return launch_row_reduce_t<PosT, 18>(ctx, d_row_list, n_list, stream, row0);
    case 20: return launch_row_reduce_t<PosT, 20>(ctx, d_row_list, n_list, stream, row0);
    case 27: return launch_row_reduce_t<PosT, 27>(ctx, d_row_list, n_list, stream, row0);
    case 81: return launch_row_reduce_t<PosT, 81>(ctx, d_row_list, n_list, stream, row0);
    default: return launch_row_reduce_t<PosT, 0>(ctx, d_row_list, n_list, stream, row0);
  }
}

}  // namespace

int feasm_launch_row_reduce(feasm_ctx* ctx, const uint32_t* d_row_list, int64_t n_list, cudaStream_t stream, int64_t row0) {
  if (ctx->pos_bytes == 1) return launch_row_reduce<uint8_t>(ctx, d_row_list, n_list, stream, row0);
  return launch_row_reduce<uint16_t>(ctx, d_row_list, n_list, stream, row0);
}

int feasm_launch_row_residual(feasm_ctx* ctx, const double* d_u, double* d_r) {
  const int recb = feasm_recb(ctx->nd, ctx->pos_bytes);
  const size_t ny = (size_t)ctx->n_elem * ctx->nd;
  if (ctx->d_y_cap < ny) {
    if (ctx->d_y) cudaFree(ctx->d_y);
    ctx->d_y = nullptr; ctx->d_y_cap = 0;
    FEASM_CUDA(ctx, cudaMalloc(&ctx->d_y, sizeof(double) * (ny ? ny : 1)));
    ctx->d_y_cap = ny;
  }
  if (ctx->n_elem) {
    const int warps = 8;
    const int64_t grid = (ctx->n_elem + warps - 1) / warps;  // one element per warp: the block scheduler evens out the tail
    k_elem_apply<<<(unsigned)grid, warps * 32, sizeof(double) * warps * ctx->nd, ctx->stream>>>(
        reinterpret_cast<const unsigned char*>(ctx->d_kf), recb, ctx->d_elem_dofs, ctx->nd, d_u, ctx->n_elem, ctx->d_y);
  }
  if (ctx->n_rows)
    k_row_sum<<<(unsigned)((ctx->n_rows + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_inc_ptr, ctx->d_inc, ctx->d_y, ctx->n_rows, d_r);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_launch_elem_flags(feasm_ctx* ctx) {
  const size_t ne = (size_t)(ctx->n_elem ? ctx->n_elem : 1);
  if (!ctx->d_elem_flag) FEASM_CUDA(ctx, cudaMalloc(&ctx->d_elem_flag, ne));
  if (!ctx->d_con_elems) FEASM_CUDA(ctx, cudaMalloc(&ctx->d_con_elems, sizeof(uint32_t) * (ne + 1)));
  unsigned* d_count = ctx->d_con_elems + ne;  // the counter lives behind the list
  FEASM_CUDA(ctx, cudaMemsetAsync(d_count, 0, sizeof(unsigned), ctx->stream));
  if (ctx->n_elem)
    k_elem_flags<<<(unsigned)((ctx->n_elem + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_con_flag, ctx->nd, ctx->n_elem,
                                                                                 ctx->d_elem_flag, ctx->d_con_elems, d_count);
  FEASM_CUDA(ctx, cudaGetLastError());
  unsigned count = 0;
  FEASM_CUDA(ctx, cudaMemcpyAsync(&count, d_count, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->n_con_elems = count;
  ctx->elem_flag_valid = true;
  return FEASM_OK;
}

int feasm_launch_pack(feasm_ctx* ctx) {
  const uint64_t n = ctx->send_ptr.empty() ? 0 : ctx->send_ptr.back();
  if (n) k_pack<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_send_slots, n, ctx->nnz, (uint64_t)ctx->n_rows, ctx->d_values, ctx->d_rhs,
                                                                             ctx->d_send_buf);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_launch_unpack(feasm_ctx* ctx) {
  const uint64_t n = ctx->n_unpack_slots;
  if (n)
    k_unpack<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_unpack_slot, ctx->d_unpack_seg, ctx->d_unpack_perm, n, ctx->nnz,
                                                                    (uint64_t)ctx->n_rows, ctx->n_peers ? ctx->recv_ptr[ctx->n_peers] : 0,
                                                                    ctx->d_recv_buf, ctx->d_values, ctx->d_rhs);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_launch_scatter_constraints(feasm_ctx* ctx, int64_t n, int64_t n_prev, bool set_flags) {
  {
    const cudaError_t pre = cudaGetLastError();  // an error an earlier, unchecked call left behind is not this launch's
    if (pre != cudaSuccess) return feasm_fail(ctx, FEASM_ERR_CUDA, std::string("error pending before feasm_set_constraints: ") + cudaGetErrorString(pre));
  }
  if (set_flags && n_prev)
    k_clear_constraints<<<(unsigned)((n_prev + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_con_rows + ctx->d_con_rows_cap, n_prev, ctx->d_con_flag,
                                                                                    ctx->d_con_val);
  if (n)
    k_scatter_constraints<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_con_rows, ctx->d_con_vals_list, n, set_flags,
                                                                                 ctx->d_con_flag, ctx->d_con_val);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_launch_add_matrix(feasm_ctx* ctx, const double* d_dm, const uint32_t* d_rows, int nr, const uint32_t* d_cols,
                            int nc, int mode) {
  const int n = nr * nc;
  k_add_matrix<<<(n + 127) / 128, 128, 0, ctx->stream>>>(d_dm, d_rows, nr, d_cols, nc, ctx->d_row_ptr, ctx->d_col_idx, ctx->d_values, ctx->d_flags,
                                                         mode);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_launch_add_vector(feasm_ctx* ctx, const double* d_v, const uint32_t* d_rows, int n) {
  k_add_vector<<<1, 1, 0, ctx->stream>>>(d_v, d_rows, n, ctx->d_rhs);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}
