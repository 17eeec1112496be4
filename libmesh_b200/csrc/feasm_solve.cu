// feasm_solve.cu — the step right after the assembly path: y = A x on the assembled CSR matrix and
// a Jacobi-preconditioned conjugate-gradient solve, so a LinearImplicitSystem::solve can finish without copying the
// matrix back to the host.
//
//   feasm_vector_mult  = SparseMatrix::vector_mult (include/numerics/sparse_matrix.h:579-587)
//   feasm_cg_solve     = LinearSolver<T>::solve with CG + JACOBI_PRECOND (include/solvers/linear_solver.h,
//                        src/solvers/: the default Krylov route of the library back-ends)
//
// Everything is deterministic: the SpMV sums a row in column order by one warp with a fixed shuffle tree, the dot
// products use a fixed two-level reduction (no atomics).  On several ranks (rows spread over the ranks, global column ids;
// feasm_halo.cu) the product first fetches the ghost entries of x from their owners, and every dot product is finished by
// an all-gather of the ranks' partial sums that every rank adds in rank order — so all ranks hold bitwise the same scalars
// and the iterates do not depend on how NCCL would have ordered a reduction.
#include <cmath>
#include <cstdlib>
#include <vector>

#include "feasm_internal.h"

namespace {

// One warp per row: lanes stride over the row's entries, fixed shuffle tree.  HBM-bound: 12 B per non-zero.
__global__ void k_spmv(const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col_idx, const double* __restrict__ values,
                       const double* __restrict__ x, int64_t n_rows, double* __restrict__ y) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rows) return;
  const uint64_t p0 = row_ptr[r], p1 = row_ptr[r + 1];
  double s = 0.;
  for (uint64_t k0 = p0; k0 < p1; k0 += 128) {  // four independent (value, column, x) load chains per lane
    double v[4];
    uint32_t c[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint64_t k = k0 + lane + 32 * u;
      const bool ok = k < p1;
      v[u] = ok ? __ldcs(values + k) : 0.;
      c[u] = ok ? __ldcs(col_idx + k) : 0xffffffffu;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (c[u] != 0xffffffffu) s += v[u] * x[c[u]];  // padding lanes add nothing (0 * x[0] would spread a non-finite x[0])
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) y[r] = s;
}

// The same product, software-pipelined over the rows a warp owns (persistent grid, rows r, r + W, r + 2W, ...): a row is
// three dependent memory round trips (row_ptr -> values / columns -> x) and little else, so one-row-at-a-time leaves the
// memory system idle two thirds of the time (3.6 TB/s = 55 % of the HBM peak at 96^3).  Here the row pointers of row i+2,
// the first 128 values / columns of row i+1 and the x gather of row i are in flight together.  Same order of the adds
// per lane and the same shuffle tree: bitwise the sums of k_spmv.
__global__ void __launch_bounds__(256)
k_spmv_pipe(const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col_idx, const double* __restrict__ values,
            const double* __restrict__ x, int64_t n_rows, double* __restrict__ y) {
  const int lane = threadIdx.x & 31;
  const int64_t W = ((int64_t)gridDim.x * blockDim.x) >> 5;
  int64_t r = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rows) return;
  auto load_chunk = [&](uint64_t p0, uint64_t p1, double (&v)[4], uint32_t (&c)[4]) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint64_t k = p0 + lane + 32 * u;
      const bool ok = k < p1;
      v[u] = ok ? __ldcs(values + k) : 0.;
      c[u] = ok ? __ldcs(col_idx + k) : 0xffffffffu;
    }
  };
  // prologue: row r's pointers and first chunk, row r+W's pointers
  uint64_t p0 = row_ptr[r], p1 = row_ptr[r + 1];
  uint64_t q0 = 0, q1 = 0;
  if (r + W < n_rows) { q0 = row_ptr[r + W]; q1 = row_ptr[r + W + 1]; }
  double v[4];
  uint32_t c[4];
  load_chunk(p0, p1, v, c);
  for (; r < n_rows; r += W) {
    // stage A: pointers of the row after next; stage B: first chunk of the next row
    uint64_t s0 = 0, s1 = 0;
    if (r + 2 * W < n_rows) { s0 = row_ptr[r + 2 * W]; s1 = row_ptr[r + 2 * W + 1]; }
    double vn[4] = {0., 0., 0., 0.};
    uint32_t cn[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
    if (r + W < n_rows) load_chunk(q0, q1, vn, cn);
    // stage C: this row
    double s = 0.;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (c[u] != 0xffffffffu) s += v[u] * x[c[u]];
    for (uint64_t k0 = p0 + 128; k0 < p1; k0 += 128) {  // rows longer than 128 entries (several variables per node)
      double v2[4];
      uint32_t c2[4];
      load_chunk(k0, p1, v2, c2);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (c2[u] != 0xffffffffu) s += v2[u] * x[c2[u]];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) y[r] = s;
    p0 = q0; p1 = q1; q0 = s0; q1 = s1;
#pragma unroll
    for (int u = 0; u < 4; ++u) { v[u] = vn[u]; c[u] = cn[u]; }
  }
}

// The same product as a STREAM over the non-zeros: a warp owns 32 consecutive rows = one contiguous run of the value and
// column arrays (2000 entries on average at HEX27), which all 32 lanes read fully coalesced whatever the row lengths are
// (one-warp-per-row leaves 3/4 of its load slots empty on a 27-entry row), multiply with the gathered x and park in shared
// memory, CH entries at a time; then lane i adds up the entries of row r0 + i left to right.  Rows of FE matrices have odd
// lengths (27, 45, 75, 125), so the lanes' walks through the buffer fall into different banks.  The sum of a row is the
// plain left-to-right sum of its entries in column order (deterministic; not the order of k_spmv's strided partial sums).
template <int WARPS, int CH>
__global__ void __launch_bounds__(WARPS * 32)
k_spmv_stream(const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col_idx, const double* __restrict__ values,
              const double* __restrict__ x, int64_t n_rows, double* __restrict__ y) {
  static_assert(CH % 128 == 0, "a 128-entry block never straddles two chunks");
  __shared__ __align__(16) double s_prod[WARPS][CH];  // (static: the same kernel with dynamic shared memory compiled to 48 instead of 62
                                        // registers and ran 4.02 instead of 4.40 TB/s)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* prod = s_prod[warp];
  const int64_t n_batches = (n_rows + 31) >> 5, stride = (int64_t)gridDim.x * WARPS;
  int64_t b = (int64_t)blockIdx.x * WARPS + warp;
  if (b >= n_batches) return;
  auto row_bounds = [&](int64_t bb, uint64_t& rs, uint64_t& re) {
    const int64_t r = bb * 32 + lane;
    rs = row_ptr[r < n_rows ? r : n_rows];
    re = row_ptr[r + 1 < n_rows ? r + 1 : n_rows];
  };
  uint64_t rs, re, rs_n = 0, re_n = 0;
  row_bounds(b, rs, re);
  for (; b < n_batches; b += stride) {
    if (b + stride < n_batches) row_bounds(b + stride, rs_n, re_n);  // the next batch's row pointers travel under this batch
    const uint64_t p0 = __shfl_sync(0xffffffffu, rs, 0), p1 = __shfl_sync(0xffffffffu, re, 31);
    auto load = [&](uint64_t k0, double (&v)[4], uint32_t (&c)[4]) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint64_t k = k0 + lane + 32 * u;
        const bool ok = k < p1;
        v[u] = ok ? __ldcs(values + k) : 0.;
        c[u] = ok ? __ldcs(col_idx + k) : 0u;
      }
    };
    // two blocks of 128 entries in flight per warp: the (value, column) loads of block j+1 and the x gather of block j;
    // the first block of the next chunk is loaded before this chunk's rows are added up
    double v[4];
    uint32_t c[4];
    load(p0, v, c);
    double acc = 0.;
    for (uint64_t c0 = p0; c0 < p1; c0 += CH) {
      const uint64_t c1 = c0 + CH < p1 ? c0 + CH : p1;
      for (uint64_t k0 = c0; k0 < c1; k0 += 128) {
        double vn[4];
        uint32_t cn[4];
        load(k0 + 128, vn, cn);
        double xv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) xv[u] = x[c[u]];  // (lanes past the end read x[0]; their product is not stored)
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint64_t k = k0 + lane + 32 * u;
          if (k < c1) { FEASM_DEV_ASSERT(k - c0 < (uint64_t)CH); prod[k - c0] = v[u] * xv[u]; }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) { v[u] = vn[u]; c[u] = cn[u]; }
      }
      __syncwarp();
      const uint64_t lo = rs > c0 ? rs : c0, hi = re < c1 ? re : c1;
      if (lo < hi) {  // left to right, two entries per shared-memory load where the pair is 16 B aligned: the lanes' walks are
                      // scattered, so a load instruction costs its wavefronts whatever it returns (ncu: LSU data pipe 92 % busy,
                      // 58 % of its wavefronts from shared memory).  (Four accumulators per row measured slower: 4.40 -> 3.88 TB/s.)
        uint32_t i = (uint32_t)(lo - c0);
        const uint32_t e = (uint32_t)(hi - c0);
        FEASM_DEV_ASSERT(e <= (uint32_t)CH);
        if (i & 1) { acc += prod[i]; ++i; }
        for (; i + 2 <= e; i += 2) {
          const double2 t = *reinterpret_cast<const double2*>(prod + i);
          acc += t.x;
          acc += t.y;
        }
        if (i < e) acc += prod[i];
      }
      __syncwarp();
    }
    const int64_t r = b * 32 + lane;
    if (r < n_rows) y[r] = acc;
    rs = rs_n; re = re_n;
  }
}

// 1 / A(r,r) (binary search of the diagonal in the sorted row); rows without a diagonal entry get 1.
__global__ void k_inv_diag(const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col_idx, const double* __restrict__ values,
                           const uint32_t* __restrict__ gid, uint32_t first, int64_t n_rows, double* __restrict__ dinv) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= n_rows) return;
  const uint32_t col = gid ? gid[r] : first + (uint32_t)r;
  const uint32_t* c = col_idx + row_ptr[r];
  int lo = 0, hi = (int)(row_ptr[r + 1] - row_ptr[r]);
  const int len = hi;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (c[mid] < col) lo = mid + 1; else hi = mid;
  }
  const double d = (lo < len && c[lo] == col) ? values[row_ptr[r] + lo] : 1.;
  dinv[r] = d != 0. ? 1. / d : 1.;
}

constexpr int DOT_BLOCK = 256, DOT_GRID = 592;  // 4 CTAs per SM; partials reduced by one CTA in a fixed order

__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.;
  if (threadIdx.x < 32) {
    t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  __syncthreads();
  return t;  // valid on thread 0
}

// partial[b] = sum over this CTA's grid-stride slice of a[i]*b[i] (b == nullptr: a[i]*a[i]*w[i] when w given)
__global__ void k_dot_partial(const double* __restrict__ a, const double* __restrict__ b, int64_t n, double* __restrict__ partial) {
  __shared__ double sh[32];
  double s = 0.;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s += a[i] * b[i];
  s = block_sum(s, sh);
  if (threadIdx.x == 0) partial[blockIdx.x] = s;
}
__global__ void k_dot_final(const double* __restrict__ partial, int n, double* __restrict__ out) {
  __shared__ double sh[32];
  double s = 0.;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
  s = block_sum(s, sh);
  if (threadIdx.x == 0) *out = s;
}

// s[j] = sum over the ranks q = 0, 1, ... of gathered[q * cnt + j], in rank order (every rank computes the same bits)
__global__ void k_rank_sum(const double* __restrict__ gathered, int world, int cnt, double* __restrict__ s) {
  const int j = threadIdx.x;
  if (j >= cnt) return;
  double t = 0.;
  for (int q = 0; q < world; ++q) t += gathered[q * cnt + j];
  s[j] = t;
}

// x += alpha p; r -= alpha q; z = dinv .* r     (alpha = rz / pq read from device scalars)
__global__ void k_cg_update(int64_t n, const double* __restrict__ rz, const double* __restrict__ pq, const double* __restrict__ p,
                            const double* __restrict__ q, const double* __restrict__ dinv, double* __restrict__ x, double* __restrict__ r,
                            double* __restrict__ z) {
  const double pqv = *pq;
  if (!(pqv > 0.)) return;  // breakdown (p.Ap <= 0): leave x and r as they are; the host reads pq and reports it
  const double alpha = *rz / pqv;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    x[i] += alpha * p[i];
    const double ri = r[i] - alpha * q[i];
    r[i] = ri;
    z[i] = dinv[i] * ri;
  }
}
// p = z + (rz_new / rz_old) p
__global__ void k_cg_direction(int64_t n, const double* __restrict__ rz_new, const double* __restrict__ rz_old, const double* __restrict__ z,
                               double* __restrict__ p) {
  const double beta = *rz_new / *rz_old;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = z[i] + beta * p[i];
}
// r = b - q; z = dinv .* r; p = z
__global__ void k_cg_start(int64_t n, const double* __restrict__ b, const double* __restrict__ q, const double* __restrict__ dinv,
                           double* __restrict__ r, double* __restrict__ z, double* __restrict__ p) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const double ri = b[i] - q[i];
    r[i] = ri;
    const double zi = dinv[i] * ri;
    z[i] = zi;
    p[i] = zi;
  }
}

// ---- device vector algebra behind NumericVector (include/numerics/numeric_vector.h) ------------------------------
__global__ void k_vec_fill(double* __restrict__ y, int64_t n, double v) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] = v;
}
__global__ void k_vec_axpy(double* __restrict__ y, double a, const double* __restrict__ x, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] += a * x[i];
}
// out = a (op) b elementwise; ops that take one operand read a only
__global__ void k_vec_pointwise(int op, double* __restrict__ out, const double* __restrict__ a, const double* __restrict__ b, double s, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double v = a[i];
    switch (op) {
      case FEASM_VEC_MULT: v = v * b[i]; break;
      case FEASM_VEC_DIV: v = v / b[i]; break;
      case FEASM_VEC_ABS: v = fabs(v); break;
      case FEASM_VEC_RECIPROCAL: v = 1. / v; break;
      case FEASM_VEC_ADD_SCALAR: v = v + s; break;
      case FEASM_VEC_SCALE: v = v * s; break;
    }
    out[i] = v;
  }
}
__device__ __forceinline__ double red_combine(int op, double a, double b) {
  switch (op) {
    case FEASM_VEC_MIN: return a < b ? a : b;
    case FEASM_VEC_MAX: case FEASM_VEC_LINF: return a > b ? a : b;
    default: return a + b;
  }
}
__device__ __forceinline__ double red_identity(int op) {
  return op == FEASM_VEC_MIN ? 1.7976931348623157e308 : (op == FEASM_VEC_MAX ? -1.7976931348623157e308 : 0.);
}
__device__ __forceinline__ double red_map(int op, double v) { return (op == FEASM_VEC_L1 || op == FEASM_VEC_LINF) ? fabs(v) : v; }
// fixed two-level tree (deterministic): grid-stride partials, then one CTA over the partials (x == nullptr: partials in)
__global__ void k_vec_reduce(int op, const double* __restrict__ x, int64_t n, double* __restrict__ out, bool final_pass) {
  __shared__ double sh[32];
  double v = red_identity(op);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    v = red_combine(op, v, final_pass ? x[i] : red_map(op, x[i]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = red_combine(op, v, __shfl_xor_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    double t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : red_identity(op);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t = red_combine(op, t, __shfl_xor_sync(0xffffffffu, t, o));
    if (threadIdx.x == 0) out[blockIdx.x] = t;
  }
}
__global__ void k_vec_set_one(double* __restrict__ y, int64_t i, double v, int add) { y[i] = add ? y[i] + v : v; }

int spmv(feasm_ctx* ctx, int64_t n_rows, const double* d_x, double* d_y) {
  const bool pipe = !(getenv("FEASM_SPMV_PIPE") && atoi(getenv("FEASM_SPMV_PIPE")) == 0);  // A/B knob: 0 = one row at a time
  // A/B knob: 0 = never the stream kernel, 1 = always; default: matrices whose rows average at most 100 entries (one
  // variable per node: 27 ... 125 entries), where it beats one-warp-per-row by 1.1x (HEX27) to 2.6x (HEX8)
  const int stream_mode = getenv("FEASM_SPMV_STREAM") ? atoi(getenv("FEASM_SPMV_STREAM")) : -1;
  if (n_rows && (stream_mode == 1 || (stream_mode < 0 && ctx->nnz <= (uint64_t)ctx->n_rows * 100))) {
    constexpr int W = 8, CH = 768;
    static int per_sm = 0;  // resident CTAs per SM: the persistent grid is one wave
    if (!per_sm && (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_spmv_stream<W, CH>, W * 32, 0) != cudaSuccess || per_sm < 1)) per_sm = 2;
    const int64_t want_b = ((n_rows + 31) / 32 + W - 1) / W, cap = (int64_t)ctx->sm_count * per_sm;
    k_spmv_stream<W, CH><<<(unsigned)(want_b < cap ? want_b : cap), W * 32, 0, ctx->stream>>>(ctx->d_row_ptr, ctx->d_col_idx, ctx->d_values, d_x, n_rows, d_y);
    FEASM_CUDA(ctx, cudaGetLastError());
    return FEASM_OK;
  }
  const int64_t want = (n_rows * 32 + 255) / 256;
  // long rows (several variables per node: 375 entries at 3 x HEX27) already keep enough bytes in flight per warp and lose
  // to the pipeline's extra registers (measured: 5.6 vs 4.6 TB/s at 3 x 48^3); short rows gain 3-8 %
  if (pipe && ctx->nnz <= (uint64_t)ctx->n_rows * 160) {  // (the average over owned and ghost rows: the same element type)
    static int per_sm = 0;  // resident CTAs per SM (registers decide: 4 at 54 registers): the persistent grid is one wave
    if (!per_sm && (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_spmv_pipe, 256, 0) != cudaSuccess || per_sm < 1)) per_sm = 4;
    const int64_t cap = (int64_t)ctx->sm_count * per_sm;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    if (grid) k_spmv_pipe<<<grid, 256, 0, ctx->stream>>>(ctx->d_row_ptr, ctx->d_col_idx, ctx->d_values, d_x, n_rows, d_y);
  } else if (want) {
    k_spmv<<<(unsigned)want, 256, 0, ctx->stream>>>(ctx->d_row_ptr, ctx->d_col_idx, ctx->d_values, d_x, n_rows, d_y);
  }
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int dot(feasm_ctx* ctx, const double* a, const double* b, int64_t n, double* d_partial, double* d_out) {
  k_dot_partial<<<DOT_GRID, DOT_BLOCK, 0, ctx->stream>>>(a, b, n, d_partial);
  k_dot_final<<<1, DOT_BLOCK, 0, ctx->stream>>>(d_partial, DOT_GRID, d_out);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

// y = A x on the owned rows; on several ranks x's ghost entries come from their owners first (feasm_halo_exchange)
int product(feasm_ctx* ctx, const double* d_x, double* d_y) {
  const double* xc = nullptr;
  const int rc = feasm_halo_exchange(ctx, d_x, &xc);
  return rc ? rc : spmv(ctx, ctx->n_owned, xc, d_y);
}

// finishes `cnt` dot products whose local sums sit in d_s[0..cnt): all-gather, then the same rank-ordered sum everywhere
int rank_sum(feasm_ctx* ctx, double* d_s, int cnt, double* d_gather) {
  if (ctx->comm_size <= 1 || !ctx->comm) return FEASM_OK;
  const int rc = feasm_comm_allgather_f64(ctx, d_s, d_gather, cnt, ctx->stream);
  if (rc) return rc;
  k_rank_sum<<<1, 32, 0, ctx->stream>>>(d_gather, ctx->comm_size, cnt, d_s);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

}  // namespace

int feasm_spmv_rows(feasm_ctx* ctx, int64_t n_rows_used, const double* d_x, double* d_y) { return spmv(ctx, n_rows_used, d_x, d_y); }

int feasm_vector_mult(feasm_ctx* ctx, const double* d_x, double* d_y) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  if (!d_x || !d_y) return feasm_fail(ctx, FEASM_ERR_ARG, "null vector");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  return product(ctx, d_x, d_y);
}

int feasm_vector_mult_host(feasm_ctx* ctx, const double* x, double* y) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  if (!x || !y) return feasm_fail(ctx, FEASM_ERR_ARG, "null vector");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int64_t n = ctx->n_owned;  // the owned part of x and y (all rows on one rank)
  const size_t nb = sizeof(double) * (size_t)(n ? n : 1);
  double* w = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&w, 2 * nb));
  int rc = FEASM_OK;
  if (cudaMemcpyAsync(w, x, nb, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) rc = feasm_fail(ctx, FEASM_ERR_CUDA, "H2D x");
  if (!rc) rc = feasm_vector_mult(ctx, w, w + (n ? n : 1));
  if (!rc && (cudaMemcpyAsync(y, w + (n ? n : 1), nb, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
              cudaStreamSynchronize(ctx->stream) != cudaSuccess))
    rc = feasm_fail(ctx, FEASM_ERR_CUDA, "D2H y");
  cudaFree(w);
  return rc;
}

int feasm_cg_solve(feasm_ctx* ctx, double rel_tol, int max_its, const double* x0, double* x, int* iterations, double* rel_residual) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  if (!x || max_its < 0 || !(rel_tol >= 0.)) return feasm_fail(ctx, FEASM_ERR_ARG, "bad solver arguments");
  const bool several = feasm_is_distributed(ctx);
  if (several && !ctx->comm) return feasm_fail(ctx, FEASM_ERR_STATE, "cg_solve on several ranks needs the context's communicator (feasm_comm_init / feasm_set_communicator) and feasm_build_halo");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int64_t n = ctx->n_owned;  // every vector is the owned part; x0 / x are the owned parts on the host
  uint32_t first = 0;
  if (several && !feasm_halo_ready(ctx)) return feasm_fail(ctx, FEASM_ERR_STATE, "cg_solve on several ranks: feasm_build_halo first");
  if (n == 0 && !several) {
    if (iterations) *iterations = 0;
    if (rel_residual) *rel_residual = 0.;
    return FEASM_OK;
  }
  const size_t nb = sizeof(double) * (size_t)(n ? n : 1);
  double* w = nullptr;  // x | r | z | p | q | dinv | partials | scalars(4)
  const int world = several ? ctx->comm_size : 1;
  const size_t n1 = (size_t)(n ? n : 1);
  FEASM_CUDA(ctx, cudaMalloc(&w, 6 * nb + sizeof(double) * (DOT_GRID + 4 + 2 * world)));
  double *d_x = w, *d_r = w + n1, *d_z = w + 2 * n1, *d_p = w + 3 * n1, *d_q = w + 4 * n1, *d_dinv = w + 5 * n1;
  double* d_part = w + 6 * n1;
  double* d_s = d_part + DOT_GRID;  // [0] rz, [1] pq, [2] rz_new, [3] rr
  double* d_gather = d_s + 4;       // [world][<= 2] partial sums of the ranks
  int rc = FEASM_OK, its = 0;
  double rel = 0.;
  bool breakdown = false;
  auto fail = [&](int code) { cudaFree(w); return code; };
  const unsigned vg = DOT_GRID;
  if (x0) { if (cudaMemcpyAsync(d_x, x0, nb, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) return fail(feasm_fail(ctx, FEASM_ERR_CUDA, "H2D x0")); }
  else if (cudaMemsetAsync(d_x, 0, nb, ctx->stream) != cudaSuccess) return fail(feasm_fail(ctx, FEASM_ERR_CUDA, "memset x"));
  first = several ? feasm_halo_first(ctx) : 0;
  if (n) k_inv_diag<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_row_ptr, ctx->d_col_idx, ctx->d_values, ctx->d_gid, first, n, d_dinv);
  if ((rc = product(ctx, d_x, d_q))) return fail(rc);
  k_cg_start<<<vg, 256, 0, ctx->stream>>>(n, ctx->d_rhs, d_q, d_dinv, d_r, d_z, d_p);
  double h[4] = {0, 0, 0, 0};
  // ||b|| for the relative test, r.z for the first direction
  if ((rc = dot(ctx, ctx->d_rhs, ctx->d_rhs, n, d_part, d_s + 3))) return fail(rc);
  if ((rc = rank_sum(ctx, d_s + 3, 1, d_gather))) return fail(rc);
  if ((rc = dot(ctx, d_r, d_z, n, d_part, d_s + 0))) return fail(rc);
  if ((rc = rank_sum(ctx, d_s + 0, 1, d_gather))) return fail(rc);
  if (cudaMemcpyAsync(h, d_s, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess)
    return fail(feasm_fail(ctx, FEASM_ERR_CUDA, "cg: scalar read"));
  const double bnorm = std::sqrt(h[3]) > 0. ? std::sqrt(h[3]) : 1.;
  if ((rc = dot(ctx, d_r, d_r, n, d_part, d_s + 3))) return fail(rc);
  if ((rc = rank_sum(ctx, d_s + 3, 1, d_gather))) return fail(rc);
  if (cudaMemcpyAsync(h, d_s, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess)
    return fail(feasm_fail(ctx, FEASM_ERR_CUDA, "cg: scalar read"));
  rel = std::sqrt(h[3]) / bnorm;
  while (its < max_its && rel > rel_tol) {
    if ((rc = product(ctx, d_p, d_q))) return fail(rc);
    if ((rc = dot(ctx, d_p, d_q, n, d_part, d_s + 1))) return fail(rc);
    if ((rc = rank_sum(ctx, d_s + 1, 1, d_gather))) return fail(rc);
    k_cg_update<<<vg, 256, 0, ctx->stream>>>(n, d_s + 0, d_s + 1, d_p, d_q, d_dinv, d_x, d_r, d_z);
    if ((rc = dot(ctx, d_r, d_z, n, d_part, d_s + 2))) return fail(rc);
    if ((rc = dot(ctx, d_r, d_r, n, d_part, d_s + 3))) return fail(rc);
    if ((rc = rank_sum(ctx, d_s + 2, 2, d_gather))) return fail(rc);  // r.z and r.r in one all-gather
    k_cg_direction<<<vg, 256, 0, ctx->stream>>>(n, d_s + 2, d_s + 0, d_z, d_p);
    if (cudaMemcpyAsync(d_s + 0, d_s + 2, sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream) != cudaSuccess ||
        cudaMemcpyAsync(h, d_s, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess)
      return fail(feasm_fail(ctx, FEASM_ERR_CUDA, "cg: scalar read"));
    if (!(h[1] > 0.)) { breakdown = true; break; }  // p.Ap <= 0: not positive definite on this direction; x, r untouched by this step
    ++its;
    rel = std::sqrt(h[3]) / bnorm;
  }
  if (cudaMemcpyAsync(x, d_x, nb, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess)
    return fail(feasm_fail(ctx, FEASM_ERR_CUDA, "cg: D2H x"));
  cudaFree(w);
  if (iterations) *iterations = its;
  if (rel_residual) *rel_residual = rel;
  if (breakdown) return feasm_fail(ctx, FEASM_ERR_BREAKDOWN, "cg: p.Ap <= 0 (matrix not positive definite on the search direction); x holds the last good iterate");
  return rel > rel_tol ? FEASM_NOT_CONVERGED : FEASM_OK;
}


// ---- device vectors for a NumericVector back-end (shim/gpu_numeric_vector.h).  Blocking calls: each returns after its
// work on feasm_stream has completed, like the host-library back-ends. ------------------------------------------------
double* feasm_vec_alloc(feasm_ctx* ctx, int64_t n) {
  if (!ctx || n < 0) return nullptr;
  if (cudaSetDevice(ctx->device) != cudaSuccess) return nullptr;
  double* p = nullptr;
  if (cudaMalloc(&p, sizeof(double) * (size_t)(n ? n : 1)) != cudaSuccess) { feasm_fail(ctx, FEASM_ERR_NOMEM, "feasm_vec_alloc"); return nullptr; }
  cudaMemsetAsync(p, 0, sizeof(double) * (size_t)(n ? n : 1), ctx->stream);
  cudaStreamSynchronize(ctx->stream);
  return p;
}
void feasm_vec_free(feasm_ctx* ctx, double* d_v) {
  if (!ctx || !d_v) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  cudaFree(d_v);
}
int feasm_vec_copy(feasm_ctx* ctx, int kind, void* dst, const void* src, int64_t n) {
  if (!ctx || !dst || !src || n < 0) return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_copy");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const cudaMemcpyKind k = kind == FEASM_COPY_H2D ? cudaMemcpyHostToDevice : (kind == FEASM_COPY_D2H ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice);
  FEASM_CUDA(ctx, cudaMemcpyAsync(dst, src, sizeof(double) * (size_t)n, k, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}
int feasm_vec_fill(feasm_ctx* ctx, double* d_y, int64_t n, double value) {
  if (!ctx || !d_y) return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_fill");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n) k_vec_fill<<<DOT_GRID, 256, 0, ctx->stream>>>(d_y, n, value);
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}
int feasm_vec_axpy(feasm_ctx* ctx, double* d_y, double a, const double* d_x, int64_t n) {
  if (!ctx || !d_y || !d_x) return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_axpy");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n) k_vec_axpy<<<DOT_GRID, 256, 0, ctx->stream>>>(d_y, a, d_x, n);
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}
int feasm_vec_pointwise(feasm_ctx* ctx, int op, double* d_out, const double* d_a, const double* d_b, double s, int64_t n) {
  if (!ctx || !d_out || !d_a || op < FEASM_VEC_MULT || op > FEASM_VEC_SCALE) return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_pointwise");
  if ((op == FEASM_VEC_MULT || op == FEASM_VEC_DIV) && !d_b) return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_pointwise: second operand");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n) k_vec_pointwise<<<DOT_GRID, 256, 0, ctx->stream>>>(op, d_out, d_a, d_b, s, n);
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}
int feasm_vec_reduce(feasm_ctx* ctx, int op, const double* d_x, const double* d_y, int64_t n, double* result) {
  if (!ctx || !d_x || !result) return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_reduce");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  double* w = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&w, sizeof(double) * (DOT_GRID + 1)));
  int rc = FEASM_OK;
  if (op == FEASM_VEC_DOT) {
    if (!d_y) { cudaFree(w); return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_reduce: dot needs two vectors"); }
    rc = dot(ctx, d_x, d_y, n, w, w + DOT_GRID);
  } else if (op >= FEASM_VEC_SUM && op <= FEASM_VEC_LINF) {
    k_vec_reduce<<<DOT_GRID, DOT_BLOCK, 0, ctx->stream>>>(op, d_x, n, w, false);
    k_vec_reduce<<<1, DOT_BLOCK, 0, ctx->stream>>>(op, w, DOT_GRID, w + DOT_GRID, true);
  } else {
    rc = feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_reduce: unknown op");
  }
  if (!rc && (cudaMemcpyAsync(result, w + DOT_GRID, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
              cudaStreamSynchronize(ctx->stream) != cudaSuccess))
    rc = feasm_fail(ctx, FEASM_ERR_CUDA, "feasm_vec_reduce");
  cudaFree(w);
  return rc;
}
int feasm_vec_set_entry(feasm_ctx* ctx, double* d_y, int64_t i, double value, int add) {
  if (!ctx || !d_y || i < 0) return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_vec_set_entry");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  k_vec_set_one<<<1, 1, 0, ctx->stream>>>(d_y, i, value, add);
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}
