// feasm_assemble.cu — the hot path: element kernel (geometry + shapes + Ke/Fe + Dirichlet) and the
// deterministic row reduction into CSR.  sm_100a only.
//
// Element kernel = FEMap::compute_map (src/fe/fe_map.C:958-1077, :1323-1372) + H1FETransformation::map_dphi
// (src/fe/h1_fe_transformation.C:150-187) + the user contraction (introduction_ex4.C:464-537) +
// DofMap::heterogeneously_constrain_element_matrix_and_vector (src/base/dof_map_constraints.C:2518-2614),
// one warp per element:
//   A. lane = quadrature point: J = sum_i X_i (x) dphi_i, det, JxW, inverse-Jacobian rows, x_q, forcing
//   B. lane = quadrature point: physical gradients G[i][d*NQP+q] into shared memory (never to HBM)
//   C. Ke = G W G^T on the FP64 tensor pipe: mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4), upper-triangular 8x8
//      blocks only (Ke is symmetric), accumulators in registers
//   D. stage Ke in shared memory, Fe, Dirichlet epilogue, coalesced store
// Row reduction = SparseMatrix::add_matrix / NumericVector::add_vector (laspack_matrix.C:259-275,
// numeric_vector.C:690-697) turned inside out: one warp per CSR row sums the element rows incident to it in
// element order through the precomputed position map — no atomics, bitwise reproducible, and the same
// accumulation order as the reference's serial loop.
#include "feasm_internal.h"

namespace {

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

template <int NPE, int NQ>
struct ElemCfg {
  static constexpr int NB = (NPE + 7) / 8;          // 8-row blocks of Ke
  static constexpr int IP = NB * 8;                 // padded rows
  static constexpr int NQP = (NQ + 3) & ~3;         // quadrature points padded to the MMA k step
  static constexpr int K = 3 * NQP;                 // contraction length: (d, q)
  static constexpr int LDK = ((K + 11) / 16) * 16 + 4;  // row stride == 4 (mod 16): conflict-free fragment loads
  static constexpr int NBLK = NB * (NB + 1) / 2;
  static constexpr int TAB = 4 * NPE * NQ + NQ;     // phi | dxi | deta | dzeta | qw
  // per-warp shared doubles: G | X | Wk | fw | cg ; then uint32 dofs | cf
  static constexpr int G_SZ = IP * LDK;
  static constexpr int X_SZ = ((NPE * 3 + 1) / 2) * 2;
  static constexpr int W_SZ = K;
  static constexpr int F_SZ = 32;
  static constexpr int CG_SZ = ((NPE + 1) / 2) * 2;
  static constexpr int WARP_DBL = G_SZ + X_SZ + W_SZ + F_SZ + CG_SZ + NPE /* 2*NPE uint32 */ + 2;
  static_assert(NQP <= 32, "lane = quadrature point needs nq <= 32");
  static_assert(NPE <= 32, "lane = shape function needs npe <= 32");
  static_assert(NPE * NPE <= G_SZ, "Ke staging aliases G");
};

// introduction_ex3/exact_solution.C:40-47
__device__ __forceinline__ double exact_solution_dev(double x, double y, double z) {
  const double pi = 3.141592653589793;  // acos(-1.)
  return cos(.5 * pi * x) * sin(.5 * pi * y) * cos(.5 * pi * z);
}
// introduction_ex4.C:500-537
__device__ __forceinline__ double ex4_forcing_dev(double x, double y, double z) {
  const double eps = 1.e-3;
  const double u0 = exact_solution_dev(x, y, z);
  const double uxx = (exact_solution_dev(x - eps, y, z) + exact_solution_dev(x + eps, y, z) + -2. * u0) / eps / eps;
  const double uyy = (exact_solution_dev(x, y - eps, z) + exact_solution_dev(x, y + eps, z) + -2. * u0) / eps / eps;
  const double uzz = (exact_solution_dev(x, y, z - eps) + exact_solution_dev(x, y, z + eps) + -2. * u0) / eps / eps;
  return -(uxx + uyy + uzz);
}

template <int NPE, int NQ, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
k_elem_poisson(const double* __restrict__ xyz, const uint32_t* __restrict__ conn, const uint32_t* __restrict__ elem_dofs,
               const uint8_t* __restrict__ con_flag, const double* __restrict__ con_val, const double* __restrict__ tab,
               int forcing, const double* __restrict__ f_tab, int constrain_mode, int apply_constraints, int64_t e0,
               int64_t e1, double* __restrict__ ke, double* __restrict__ fe, int* flags) {
  using C = ElemCfg<NPE, NQ>;
  extern __shared__ double smem[];
  double* s_tab = smem;  // TAB doubles
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = threadIdx.x; t < C::TAB; t += WARPS * 32) s_tab[t] = tab[t];
  const double* s_phi = s_tab;
  const double* s_dxi = s_tab + NPE * NQ;
  const double* s_deta = s_tab + 2 * NPE * NQ;
  const double* s_dzeta = s_tab + 3 * NPE * NQ;
  const double* s_qw = s_tab + 4 * NPE * NQ;
  double* G = smem + ((C::TAB + 1) & ~1) + (size_t)warp * C::WARP_DBL;
  double* X = G + C::G_SZ;
  double* Wk = X + C::X_SZ;
  double* fw = Wk + C::W_SZ;
  double* cg = fw + C::F_SZ;
  uint32_t* dofs = reinterpret_cast<uint32_t*>(cg + C::CG_SZ);
  uint32_t* cf = dofs + NPE;
  __syncthreads();

  const int64_t stride = (int64_t)gridDim.x * WARPS;
  for (int64_t e = e0 + (int64_t)blockIdx.x * WARPS + warp; e < e1; e += stride) {
    // ---- gather: coordinates (coalesced over the 3 doubles of a node) and dof ids ------------------------
    for (int t = lane; t < NPE * 3; t += 32) X[t] = xyz[(uint64_t)conn[e * NPE + t / 3] * 3 + t % 3];
    if (lane < NPE) dofs[lane] = elem_dofs[e * NPE + lane];
    __syncwarp();

    // ---- A: reference -> physical map at quadrature point q = lane (fe_map.C:994-1068) -------------------
    double J[9];  // dxidx dxidy dxidz | detadx .. | dzetadx ..
    double JxW = 0.;
    if (lane < NQ) {
      double x = 0, y = 0, z = 0;
      double xa = 0, ya = 0, za = 0, xb = 0, yb = 0, zb = 0, xc = 0, yc = 0, zc = 0;
#pragma unroll 3
      for (int i = 0; i < NPE; ++i) {
        const double px = X[3 * i], py = X[3 * i + 1], pz = X[3 * i + 2];
        const double p = s_phi[i * NQ + lane], a = s_dxi[i * NQ + lane], b = s_deta[i * NQ + lane], c = s_dzeta[i * NQ + lane];
        x += p * px; y += p * py; z += p * pz;
        xa += a * px; ya += a * py; za += a * pz;
        xb += b * px; yb += b * py; zb += b * pz;
        xc += c * px; yc += c * py; zc += c * pz;
      }
      const double jac = (xa * (yb * zc - zb * yc) + ya * (zb * xc - xb * zc) + za * (xb * yc - yb * xc));
      if (!(jac > 0.)) atomicAdd(&flags[0], 1);  // check_for_degenerate_map, fe_map.C:467-506
      JxW = jac * s_qw[lane];
      const double inv = 1. / jac;
      J[0] = (yb * zc - zb * yc) * inv; J[1] = (zb * xc - xb * zc) * inv; J[2] = (xb * yc - yb * xc) * inv;
      J[3] = (za * yc - ya * zc) * inv; J[4] = (xa * zc - za * xc) * inv; J[5] = (ya * xc - xa * yc) * inv;
      J[6] = (ya * zb - za * yb) * inv; J[7] = (za * xb - xa * zb) * inv; J[8] = (xa * yb - ya * xb) * inv;
      double f = 0.;
      if (forcing == FEASM_FORCING_EX4_FD) f = ex4_forcing_dev(x, y, z);
      else if (forcing == FEASM_FORCING_TABULATED) f = f_tab[e * NQ + lane];
      fw[lane] = JxW * f;
    } else {
#pragma unroll
      for (int d = 0; d < 9; ++d) J[d] = 0.;
      fw[lane] = 0.;
    }
    // ---- B: physical gradients (h1_fe_transformation.C:150-187), G[i][d*NQP + q] -------------------------
    if (lane < C::NQP) {
      const int q = lane < NQ ? lane : 0;
      const double live = lane < NQ ? 1. : 0.;
#pragma unroll
      for (int d = 0; d < 3; ++d) Wk[d * C::NQP + lane] = JxW;
#pragma unroll 3
      for (int i = 0; i < NPE; ++i) {
        const double a = s_dxi[i * NQ + q] * live, b = s_deta[i * NQ + q] * live, c = s_dzeta[i * NQ + q] * live;
        G[i * C::LDK + lane] = a * J[0] + b * J[3] + c * J[6];
        G[i * C::LDK + C::NQP + lane] = a * J[1] + b * J[4] + c * J[7];
        G[i * C::LDK + 2 * C::NQP + lane] = a * J[2] + b * J[5] + c * J[8];
      }
    }
    for (int t = NPE * C::LDK + lane; t < C::IP * C::LDK; t += 32) G[t] = 0.;  // pad rows (aliased by staging)
    __syncwarp();

    // ---- C: Ke = G W G^T, upper-triangular 8x8 blocks, FP64 MMA ----------------------------------------
    double acc[C::NBLK][2];
#pragma unroll
    for (int b = 0; b < C::NBLK; ++b) { acc[b][0] = 0.; acc[b][1] = 0.; }
    {
      const double* g = G + (lane >> 2) * C::LDK + (lane & 3);
#pragma unroll 3
      for (int k0 = 0; k0 < C::K; k0 += 4) {
        const double w = Wk[k0 + (lane & 3)];
        double a[C::NB];
#pragma unroll
        for (int ib = 0; ib < C::NB; ++ib) a[ib] = g[ib * 8 * C::LDK + k0];
        int blk = 0;
#pragma unroll
        for (int jb = 0; jb < C::NB; ++jb) {
          const double b = a[jb] * w;
#pragma unroll
          for (int ib = 0; ib <= jb; ++ib) { dmma884(acc[blk][0], acc[blk][1], a[ib], b); ++blk; }
        }
      }
    }
    __syncwarp();

    // ---- D: stage Ke (aliases G), Fe, Dirichlet, store ---------------------------------------------------
    double* S = G;
    {
      int blk = 0;
#pragma unroll
      for (int jb = 0; jb < C::NB; ++jb)
#pragma unroll
        for (int ib = 0; ib <= jb; ++ib) {
          const int i = ib * 8 + (lane >> 2);
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const int j = jb * 8 + 2 * (lane & 3) + t;
            if (i < NPE && j < NPE) {
              S[i * NPE + j] = acc[blk][t];
              if (ib != jb) S[j * NPE + i] = acc[blk][t];
            }
          }
          ++blk;
        }
    }
    double Fe = 0.;
    unsigned my_con = 0;
    if (lane < NPE) {
#pragma unroll 3
      for (int q = 0; q < NQ; ++q) Fe += fw[q] * s_phi[lane * NQ + q];  // introduction_ex4.C:536-537
      if (apply_constraints) {
        my_con = con_flag[dofs[lane]];
        cf[lane] = my_con;
        cg[lane] = my_con ? con_val[dofs[lane]] : 0.;
      }
    }
    const unsigned any = __ballot_sync(0xffffffffu, my_con != 0);
    __syncwarp();
    if (any) {
      if (lane < NPE) {
        if (constrain_mode == FEASM_CONSTRAIN_HETEROGENEOUS) {
          double KH = 0.;  // K*H, dof_map_constraints.C:2557-2563
          for (int j = 0; j < NPE; ++j)
            if (cf[j]) KH += S[lane * NPE + j] * cg[j];
          Fe = my_con ? cg[lane] : (Fe - KH);
        } else {
          Fe = my_con ? 0. : Fe;
        }
      }
      __syncwarp();
      for (int t = lane; t < NPE * NPE; t += 32) {
        const int i = t / NPE, j = t - i * NPE;
        if (cf[i] | cf[j]) S[t] = (i == j) ? 1. : 0.;
      }
      __syncwarp();
    }
    double* out = ke + (uint64_t)(e - e0) * (NPE * NPE);
    for (int t = lane; t < NPE * NPE; t += 32) out[t] = S[t];
    if (lane < NPE) fe[(uint64_t)(e - e0) * NPE + lane] = Fe;
    __syncwarp();
  }
}

// One warp per CSR row: values[row] = sum over incident (element, local row) pairs, in element order.
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
k_row_reduce(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc,
             const uint16_t* __restrict__ pos, const double* __restrict__ ke, const double* __restrict__ fe, int nd,
             int64_t n_rows, int cap, double* __restrict__ values, double* __restrict__ rhs) {
  extern __shared__ double s_rows[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t r = blockIdx.x * (int64_t)WARPS + warp;
  if (r >= n_rows) return;
  double* buf = s_rows + (size_t)warp * cap;
  const uint64_t p0 = row_ptr[r];
  const int len = (int)(row_ptr[r + 1] - p0);
  for (int t = lane; t < len; t += 32) buf[t] = 0.;
  __syncwarp();
  double b = 0.;
  const uint64_t k1 = inc_ptr[r + 1];
  for (uint64_t k = inc_ptr[r]; k < k1; ++k) {
    const uint64_t ea = inc[k];
    const double* src = ke + ea * nd;
    const uint16_t* pp = pos + ea * nd;
    for (int t = lane; t < nd; t += 32) buf[pp[t]] += src[t];  // positions within one element row are distinct
    if (lane == 0) b += fe[ea];
    __syncwarp();
  }
  for (int t = lane; t < len; t += 32) values[p0 + t] = buf[t];
  if (lane == 0) rhs[r] = b;
}

// Fallback for rows longer than the shared-memory budget: accumulate straight into the (zeroed) global row.
__global__ void k_row_reduce_global(const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr,
                                    const uint32_t* __restrict__ inc, const uint16_t* __restrict__ pos,
                                    const double* __restrict__ ke, const double* __restrict__ fe, int nd, int64_t n_rows,
                                    double* __restrict__ values, double* __restrict__ rhs) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rows) return;
  double* row = values + row_ptr[r];
  const int len = (int)(row_ptr[r + 1] - row_ptr[r]);
  for (int t = lane; t < len; t += 32) row[t] = 0.;
  __syncwarp();
  double b = 0.;
  for (uint64_t k = inc_ptr[r]; k < inc_ptr[r + 1]; ++k) {
    const uint64_t ea = inc[k];
    for (int t = lane; t < nd; t += 32) row[pos[ea * nd + t]] += ke[ea * nd + t];
    if (lane == 0) b += fe[ea];
    __syncwarp();
  }
  if (lane == 0) rhs[r] = b;
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

template <int NPE, int NQ, int WARPS>
int launch_poisson(feasm_ctx* ctx, const feasm_params& p, const double* d_f, int64_t e0, int64_t e1, int constrained,
                   double* d_ke, double* d_fe) {
  using C = ElemCfg<NPE, NQ>;
  const size_t smem = sizeof(double) * (((C::TAB + 1) & ~1) + (size_t)WARPS * C::WARP_DBL);
  auto kern = k_elem_poisson<NPE, NQ, WARPS>;
  FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t blocks = (e1 - e0 + WARPS - 1) / WARPS;
  if (blocks > ctx->sm_count) blocks = ctx->sm_count;
  if (blocks < 1) blocks = 1;
  kern<<<(unsigned)blocks, WARPS * 32, smem, ctx->stream>>>(ctx->d_xyz, ctx->d_conn, ctx->d_elem_dofs, ctx->d_con_flag,
                                                           ctx->d_con_val, ctx->d_tab, p.forcing, d_f, p.constrain,
                                                           constrained, e0, e1, d_ke, d_fe, ctx->d_flags);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

}  // namespace

int feasm_launch_element_kernel(feasm_ctx* ctx, const feasm_params& p, int qorder, const double* d_f, int64_t e0,
                                int64_t e1, int constrained, double* d_ke, double* d_fe) {
  if (p.problem == FEASM_POISSON) {
    if (ctx->elem_type == FEASM_HEX27 && ctx->tab_nq == 27) return launch_poisson<27, 27, 8>(ctx, p, d_f, e0, e1, constrained, d_ke, d_fe);
    if (ctx->elem_type == FEASM_HEX8 && ctx->tab_nq == 8) return launch_poisson<8, 8, 8>(ctx, p, d_f, e0, e1, constrained, d_ke, d_fe);
    if (ctx->elem_type == FEASM_HEX8 && ctx->tab_nq == 27) return launch_poisson<8, 27, 8>(ctx, p, d_f, e0, e1, constrained, d_ke, d_fe);
    if (ctx->elem_type == FEASM_TET10 && ctx->tab_nq == 14) return launch_poisson<10, 14, 8>(ctx, p, d_f, e0, e1, constrained, d_ke, d_fe);
    return feasm_fail(ctx, FEASM_ERR_ARG, "no Poisson kernel for this element type / quadrature order");
  }
  return feasm_fail(ctx, FEASM_ERR_ARG, "problem not implemented");
}

int feasm_launch_row_reduce(feasm_ctx* ctx) {
  constexpr int WARPS = 8;
  const int cap = (int)ctx->max_row_len;
  const size_t smem = sizeof(double) * (size_t)WARPS * cap;
  if (smem <= ctx->smem_optin) {
    auto kern = k_row_reduce<WARPS>;
    FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = (unsigned)((ctx->n_rows + WARPS - 1) / WARPS);
    kern<<<grid, WARPS * 32, smem, ctx->stream>>>(ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, ctx->d_pos, ctx->d_ke,
                                                 ctx->d_fe, ctx->nd, ctx->n_rows, cap, ctx->d_values, ctx->d_rhs);
  } else {
    const unsigned grid = (unsigned)((ctx->n_rows * 32 + 255) / 256);
    k_row_reduce_global<<<grid, 256, 0, ctx->stream>>>(ctx->d_row_ptr, ctx->d_inc_ptr, ctx->d_inc, ctx->d_pos, ctx->d_ke,
                                                       ctx->d_fe, ctx->nd, ctx->n_rows, ctx->d_values, ctx->d_rhs);
  }
  FEASM_CUDA(ctx, cudaGetLastError());
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
