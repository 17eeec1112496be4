// feasm_element.cu — the element kernel: geometry + shape functions + Ke/Fe contraction.  sm_100a only.
//
// Element kernel = FEMap::compute_map (src/fe/fe_map.C:958-1077, :1167-1244, :1323-1372) +
// H1FETransformation::map_dphi (src/fe/h1_fe_transformation.C:150-187) + the user contraction
// (introduction_ex4.C:464-537).  One warp per element, persistent grid (one CTA per SM):
//   0. gather: the next element's node coordinates are prefetched into registers one element ahead
//   A. lane = quadrature point: affine test (Elem::has_affine_map), then either the affine branch (Jacobian at qp 0
//      computed once by 9 lanes, JxW[p] = JxW[0]/qw[0]*qw[p]) or the general branch (J = sum_i X_i (x) dphi_i per qp);
//      det, inverse-Jacobian rows scaled by sqrt(det), x_q, forcing
//   B+C. Ke = G G^T with G[i][(d,q)] = sqrt(JxW_q) dphi_i/dx_d(q).  G never exists in memory: each lane builds the
//      MMA fragments it owns straight from the reference-gradient table in shared memory (3 LDS + 9 DFMA per
//      fragment triple) and feeds mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4).  Ke is symmetric, so only the upper-
//      triangular 8x8 blocks of the first 8*floor(n/8) rows go through DMMA (A and B fragments are the same
//      register); the n mod 8 leftover rows (3 for HEX27, 2 for TET10) would waste 5/8 of a DMMA block each, so
//      they are contracted with plain DFMA against fragments broadcast by shuffle.
//   D. Fe, then fragments stored straight into the values part of the element's records REC[e][i] (Ke row | Fe,
//      feasm_internal.h), mirrored for the lower triangle.
//   E. Dirichlet constraints: a second, small kernel (k_constrain_list) over the elements that touch a constrained dof.
#include <cstdlib>
#include <type_traits>

#include "feasm_elem_dev.cuh"

#ifndef FEASM_ELAST_WARPS
#define FEASM_ELAST_WARPS 8
#endif
#ifndef FEASM_ELAST_UNROLL
#define FEASM_ELAST_UNROLL 1
#endif

namespace {
constexpr int kElastUnroll = FEASM_ELAST_UNROLL;  // k groups unrolled in the elasticity products (tuning knob)
template <int NPE, int NQ, int WARPS, int RS>
__global__ void __launch_bounds__(WARPS * 32, 1)
k_elem_poisson(const double* __restrict__ xyz, const uint32_t* __restrict__ conn, const double* __restrict__ tab,
               int forcing, const double* __restrict__ f_tab, int64_t e0, int64_t e1, double* __restrict__ kf, int rs_rt,
               int* flags, const uint8_t* __restrict__ elem_flag, ElemConArgs con, unsigned* ticket) {
  using C = ElemCfg<NPE, NQ>;
  const int rs = RS ? RS : rs_rt;  // record stride in doubles: a compile-time constant for the common layout
  extern __shared__ double smem[];
  double* s_tab = smem;               // phi | dxi | deta | dzeta | qw, [i][q]: lane = q reads are conflict free
  double* s_tw = smem + C::TAB_SZ;    // [c][row][TS]: sqrt(w_q) * d phi_row / d xi_c (q), zero padded
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  elem_load_tables<NPE, NQ>(s_tab, s_tw, tab, threadIdx.x, WARPS * 32);
  double* X = s_tw + C::TW_SZ + (size_t)warp * C::WARP_DBL;  // node coordinates of the current element
  double* fw = X + C::X_SZ;                                  // JxW_q * f(x_q)
  // element-level Dirichlet constraint, fused (elem_flag != nullptr): the few elements that touch a constrained dof put
  // their Ke | Fe into this per-warp tile, are constrained there, and leave as whole rows
  double* tile = s_tw + C::TW_SZ + (size_t)WARPS * C::WARP_DBL + (size_t)warp * (NPE * C::LDR);
  const double* twp = s_tw + (lane >> 2) * C::TS + (lane & 3);  // + c*NB*8*TS + ib*8*TS + 4*qg
  __syncthreads();

  // Elements are dealt out by a ticket counter, one element per ticket (dynamic: a static round-robin puts every
  // boundary element of a structured mesh on the same few warps whenever the stride and a mesh dimension share a factor,
  // and boundary elements cost more — the constraint below).  Three tickets are in flight per warp: the element being
  // computed (coordinates in registers), the next one (node ids in registers), and one whose ticket is being fetched,
  // so neither the atomic nor the dependent conn -> xyz load chain is waited for.
  const int64_t n_todo = e1 - e0;
#ifdef FEASM_DEV_STATIC  // development A/B: static round-robin
  int64_t next_static = (int64_t)blockIdx.x * WARPS + warp;
  auto take = [&]() -> int64_t { const int64_t t = next_static; next_static += (int64_t)gridDim.x * WARPS; return t; };
  auto take_raw = [&]() -> unsigned { return (unsigned)take(); };
  auto bcast = [&](unsigned t) -> int64_t { return (int64_t)t; };
#else
  auto take_raw = [&]() -> unsigned { return lane == 0 ? atomicAdd(ticket, 1u) : 0u; };  // lane 0 holds the ticket ...
  auto bcast = [&](unsigned t) -> int64_t { return (int64_t)__shfl_sync(0xffffffffu, t, 0); };  // ... until it is needed
  auto take = [&]() -> int64_t { return bcast(take_raw()); };
#endif
  int64_t ta = take(), tb = take(), tc = take();
  double nx = 0., ny = 0., nz = 0.;
  uint32_t nnode = 0;
  unsigned nflag = 0;  // constraint flag of the element whose coordinates are in nx, ny, nz
  if (ta < n_todo) {
    if (lane < NPE) {
      const uint64_t node = conn[(e0 + ta) * NPE + lane];
      nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
      if (tb < n_todo) nnode = conn[(e0 + tb) * NPE + lane];
    }
    if (elem_flag) nflag = elem_flag[e0 + ta];
  }
  while (ta < n_todo) {
    const int64_t e = e0 + ta;
    const bool flagged = nflag == 1;  // 2: widened by hanging-node constraints, stored unconstrained (feasm_general.cu)
    if (lane < NPE) {
      X[3 * lane] = nx; X[3 * lane + 1] = ny; X[3 * lane + 2] = nz;
      if (tb < n_todo) {
        const uint64_t node = nnode;
        nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
        if (tc < n_todo) nnode = conn[(e0 + tc) * NPE + lane];
      }
    }
    if (elem_flag && tb < n_todo) nflag = elem_flag[e0 + tb];
    const unsigned td_raw = take_raw();  // the atomic is in flight during the element; its value is broadcast at the end
    // a flagged element's constraint data (dof -> constrained? -> value), fetched now, used after the integrals
    unsigned cmask = 0;
    double g_lane = 0.;
    if (flagged) {
      int cj = 0;
      if (lane < NPE) {
        const uint32_t dj = con.elem_dofs[(uint64_t)e * NPE + lane];
        cj = con.con_flag[dj];
        if (cj && con.mode == FEASM_CONSTRAIN_HETEROGENEOUS) g_lane = con.con_val[dj];
      }
      cmask = __ballot_sync(0xffffffffu, cj);
    }

    double* out = kf + (uint64_t)ta * (NPE * rs);
    ElemRegs<NPE, NQ> R;
    __syncwarp();
    elem_compute<NPE, NQ>(X, fw, s_tab, twp, lane, forcing, f_tab ? f_tab + e * NQ : nullptr, flags, R);
    if (!flagged) {
      elem_store<NPE, NQ>(R, out, rs, lane);
    } else {
      // DofMap::heterogeneously_constrain_element_matrix_and_vector (dof_map_constraints.C:2518-2614) on Ke / Fe in the
      // tile, where the reference applies it: after the element integrals, before add_matrix
      elem_store<NPE, NQ>(R, tile, C::LDR, lane);
      constrain_tile<NPE, C::LDR>(tile, cmask, g_lane, lane);
      constexpr int H = C::LDR / 2;
      for (int t = lane; t < NPE * H; t += 32) {
        const int row = t / H, c = t % H;
        *reinterpret_cast<double2*>(out + (size_t)row * rs + 2 * c) = *reinterpret_cast<const double2*>(tile + row * C::LDR + 2 * c);
      }
    }
    __syncwarp();
    ta = tb; tb = tc; tc = bcast(td_raw);
  }
}

template <int NPE, int NQ, int WARPS, int RS = 0>
int launch_poisson(feasm_ctx* ctx, const feasm_params& p, const double* d_f, int64_t e0, int64_t e1, double* d_kf, int rs, int constrain_mode,
                   bool* fused) {
  using C = ElemCfg<NPE, NQ>;
  const size_t smem = sizeof(double) * (size_t)(C::SMEM_DBL(WARPS) + WARPS * NPE * C::LDR);
  auto kern = k_elem_poisson<NPE, NQ, WARPS, RS>;
  FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t blocks = (e1 - e0 + WARPS - 1) / WARPS;
  if (blocks > ctx->sm_count) blocks = ctx->sm_count;
  if (blocks < 1) blocks = 1;
  // the element-level constraint rides in this kernel when there is one to apply (one variable: nd == NPE)
  // ... and no side term has to be added between the volume integrals and the constraint (feasm_launch_side_loads)
  const bool fuse_ok = constrain_mode >= 0 && ctx->n_con_elems > 0 && ctx->nd == NPE && ctx->d_elem_flag && ctx->n_side_elems == 0 &&
                       ctx->n_pen_elems == 0;
#ifdef FEASM_DEV_NOFUSE
  const bool fuse = false;
  (void)fuse_ok;
#else
  const bool fuse = fuse_ok;
#endif
  const ElemConArgs con{ctx->d_elem_dofs, ctx->d_con_flag, ctx->d_con_val, constrain_mode};
  if (e1 - e0 + 3ll * blocks * WARPS >= 0xffffffffll) return feasm_fail(ctx, FEASM_ERR_ARG, "element range exceeds the 32-bit ticket counter");
  unsigned* ticket = reinterpret_cast<unsigned*>(ctx->d_flags + 2);
  FEASM_CUDA(ctx, cudaMemsetAsync(ticket, 0, sizeof(unsigned), ctx->stream));
  kern<<<(unsigned)blocks, WARPS * 32, smem, ctx->stream>>>(ctx->d_xyz, ctx->d_conn, ctx->d_tab, p.forcing, d_f, e0, e1, d_kf, rs,
                                                           ctx->d_flags, fuse ? ctx->d_elem_flag : nullptr, con, ticket);
  FEASM_CUDA(ctx, cudaGetLastError());
  *fused = fuse;
  return FEASM_OK;
}


// ---- HEX8 with the 2x2x2 rule: four elements per warp ---------------------------------------------------------------
// With "lane = quadrature point" an element of 8 points uses a quarter of the warp in phase A, and phase A (map + the
// nine sin/cos of the FD forcing per point) is where a HEX8 element spends its time: k_elem_poisson<8,8> ran 128^3 HEX8
// at the per-element cost of HEX27's phase A.  Here lane = (element slot s = lane>>3, quadrature point / node q = lane&7):
// one pass of phase A serves four consecutive elements at full lane use; the 8x8 contractions (6 DMMA k steps each)
// follow one element after the other, each taking the S of its points from the slot's lanes by shuffle; Fe and the
// stores are again slot-parallel.  Same reference semantics as elem_process (affine test, compute_affine_map's
// "Jacobian of point 0", JxW[p] = JxW[0]/qw[0]*qw[p]); sums run in node order.
template <int WARPS, int RS>
__global__ void __launch_bounds__(WARPS * 32, 1)
k_elem_poisson_hex8x4(const double* __restrict__ xyz, const uint32_t* __restrict__ conn, const double* __restrict__ tab, int forcing,
                      const double* __restrict__ f_tab, int64_t e0, int64_t e1, double* __restrict__ kf, int rs_rt, int* flags) {
  constexpr int NPE = 8, NQ = 8;
  using C = ElemCfg<NPE, NQ>;
  constexpr unsigned FULL = 0xffffffffu;
  const int rs = RS ? RS : rs_rt;
  extern __shared__ double smem[];
  double* s_tab = smem;
  double* s_tw = smem + C::TAB_SZ;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  elem_load_tables<NPE, NQ>(s_tab, s_tw, tab, threadIdx.x, WARPS * 32);
  double* Xw = s_tw + C::TW_SZ + (size_t)warp * (4 * 24 + 32);  // [slot][node][3] | fw[slot][q]
  double* fw = Xw + 4 * 24;
  const int slot = lane >> 3, q = lane & 7, g = lane >> 2, t4 = lane & 3;
  const double* X = Xw + slot * 24;
  const double* twp = s_tw + g * C::TS + t4;
  const double* s_phi = s_tab;
  const double* s_dxi = s_tab + NPE * NQ;
  const double* s_deta = s_dxi + NPE * NQ;
  const double* s_dzeta = s_deta + NPE * NQ;
  const double* s_qw = s_tab + 4 * NPE * NQ;
  __syncthreads();

  const int64_t stride = (int64_t)gridDim.x * WARPS * 4;
  int64_t eb = e0 + ((int64_t)blockIdx.x * WARPS + warp) * 4;  // first element of this warp's group
  // one group ahead: coordinates of node q of element slot; two ahead: its node id (as in k_elem_poisson)
  auto elem_of = [&](int64_t base) { const int64_t e = base + slot; return e < e1 ? e : e1 - 1; };
  double nx = 0., ny = 0., nz = 0.;
  uint32_t nnode = 0;
  if (eb < e1) {
    const uint64_t node = conn[elem_of(eb) * NPE + q];
    nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
    if (eb + stride < e1) nnode = conn[elem_of(eb + stride) * NPE + q];
  }
  for (; eb < e1; eb += stride) {
    const int64_t e = eb + slot;
    const bool valid = e < e1;  // a short last group recomputes element e1-1 in the spare slots and stores nothing
    Xw[slot * 24 + 3 * q] = nx; Xw[slot * 24 + 3 * q + 1] = ny; Xw[slot * 24 + 3 * q + 2] = nz;
    if (eb + stride < e1) {
      const uint64_t node = nnode;
      nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
      if (eb + 2 * stride < e1) nnode = conn[elem_of(eb + 2 * stride) * NPE + q];
    }
    __syncwarp();

    // ---- A: lane = (slot, point) -----------------------------------------------------------------------------------
    const unsigned okmask = __ballot_sync(FULL, affine_check_lane<NPE>(X, q));
    const bool affine = ((okmask >> (8 * slot)) & 0xffu) == 0xffu;
    const int qj = affine ? 0 : q;  // compute_affine_map: the Jacobian of point 0 serves every point (fe_map.C:1167-1244)
    double x = 0., y = 0., z = 0., xa = 0., ya = 0., za = 0., xb = 0., yb = 0., zb = 0., xc = 0., yc = 0., zc = 0.;
#pragma unroll
    for (int i = 0; i < NPE; ++i) {
      const double px = X[3 * i], py = X[3 * i + 1], pz = X[3 * i + 2];
      const double p = s_phi[i * NQ + q], a = s_dxi[i * NQ + qj], b = s_deta[i * NQ + qj], c = s_dzeta[i * NQ + qj];
      x += p * px; y += p * py; z += p * pz;
      // Jacobian sums in the reference's order and roundings (fe_map.C:994-1008; see geom_phase)
      xa = __dadd_rn(xa, __dmul_rn(a, px)); ya = __dadd_rn(ya, __dmul_rn(a, py)); za = __dadd_rn(za, __dmul_rn(a, pz));
      xb = __dadd_rn(xb, __dmul_rn(b, px)); yb = __dadd_rn(yb, __dmul_rn(b, py)); zb = __dadd_rn(zb, __dmul_rn(b, pz));
      xc = __dadd_rn(xc, __dmul_rn(c, px)); yc = __dadd_rn(yc, __dmul_rn(c, py)); zc = __dadd_rn(zc, __dmul_rn(c, pz));
    }
    const double jac = (xa * (yb * zc - zb * yc) + ya * (zb * xc - xb * zc) + za * (xb * yc - yb * xc));  // fe_map.C:1046-1048
    const bool bad = !(jac > 0.);
    if (bad && valid) atomicAdd(&flags[0], 1);
    double JxW = jac * s_qw[q];
    if (affine && q != 0) {  // fe_map.C:1242: JxW[0] / qw[0] * qw[p], the quotient correctly rounded (Markstein, see geom_phase)
      const double w0 = s_qw[0], rc = 1. / w0, j0 = __dmul_rn(jac, w0), q0 = __dmul_rn(j0, rc);
      JxW = __dmul_rn(__fma_rn(__fma_rn(-q0, w0, j0), rc, q0), s_qw[q]);
    }
    const double sc = bad ? 0. : rsqrt(jac);
    double S[9];
    S[0] = (yb * zc - zb * yc) * sc; S[1] = (zb * xc - xb * zc) * sc; S[2] = (xb * yc - yb * xc) * sc;
    S[3] = (za * yc - ya * zc) * sc; S[4] = (xa * zc - za * xc) * sc; S[5] = (ya * xc - xa * yc) * sc;
    S[6] = (ya * zb - za * yb) * sc; S[7] = (za * xb - xa * zb) * sc; S[8] = (xa * yb - ya * xb) * sc;
    {
      double f = 0.;
      if (forcing == FEASM_FORCING_EX4_FD) {
        const bool big = !(fabs(x) + fabs(y) + fabs(z) < 2.e4);
        f = __any_sync(FULL, big) ? ex4_forcing_lib(x, y, z) : ex4_forcing_t<true>(x, y, z);
      } else if (forcing == FEASM_FORCING_TABULATED) {
        f = f_tab[(valid ? e : e1 - 1) * NQ + q];
      }
      fw[lane] = JxW * f;
    }

    // ---- B+C: Ke_s = G_s G_s^T, one slot after the other (h1_fe_transformation.C:150-187, introduction_ex4.C:464-467) ----
    double ta[C::QG], tb[C::QG], tc[C::QG];  // the lane's entries of the scaled reference-gradient table: the same for all slots
#pragma unroll
    for (int qg = 0; qg < C::QG; ++qg) {
      ta[qg] = twp[4 * qg]; tb[qg] = twp[8 * C::TS + 4 * qg]; tc[qg] = twp[16 * C::TS + 4 * qg];
    }
    double acc[4][2];
#pragma unroll
    for (int s4 = 0; s4 < 4; ++s4) {
      acc[s4][0] = 0.; acc[s4][1] = 0.;
#pragma unroll
      for (int qg = 0; qg < C::QG; ++qg) {
        const int src = 8 * s4 + 4 * qg + t4;
        double Sq[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) Sq[k] = __shfl_sync(FULL, S[k], src);
#pragma unroll
        for (int d = 0; d < 3; ++d) {
          const double G = ta[qg] * Sq[d] + tb[qg] * Sq[3 + d] + tc[qg] * Sq[6 + d];
          dmma884(acc[s4][0], acc[s4][1], G, G);
        }
      }
    }
    __syncwarp();  // fw

    // ---- D: Fe (lane = (slot, shape function)) and the stores ------------------------------------------------------------
    double* out = kf + (uint64_t)(eb - e0) * (NPE * rs);
    if (valid) {
      double Fe = 0.;
      if (forcing != FEASM_FORCING_NONE) {
#pragma unroll
        for (int k = 0; k < NQ; ++k) Fe += fw[8 * slot + k] * s_phi[q * NQ + k];
      }
      out[(uint64_t)slot * (NPE * rs) + q * rs + NPE] = Fe;
    }
#pragma unroll
    for (int s4 = 0; s4 < 4; ++s4)
      if (eb + s4 < e1)
        *reinterpret_cast<double2*>(out + (uint64_t)s4 * (NPE * rs) + g * rs + 2 * t4) = make_double2(acc[s4][0], acc[s4][1]);
    __syncwarp();
  }
}

template <int WARPS, int RS>
int launch_poisson_hex8x4(feasm_ctx* ctx, const feasm_params& p, const double* d_f, int64_t e0, int64_t e1, double* d_kf, int rs) {
  using C = ElemCfg<8, 8>;
  const size_t smem = sizeof(double) * (size_t)(C::TAB_SZ + C::TW_SZ + WARPS * (4 * 24 + 32));
  auto kern = k_elem_poisson_hex8x4<WARPS, RS>;
  FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t blocks = (e1 - e0 + 4 * WARPS - 1) / (4 * WARPS);
  if (blocks > ctx->sm_count) blocks = ctx->sm_count;
  if (blocks < 1) blocks = 1;
  kern<<<(unsigned)blocks, WARPS * 32, smem, ctx->stream>>>(ctx->d_xyz, ctx->d_conn, ctx->d_tab, p.forcing, d_f, e0, e1, d_kf, rs, ctx->d_flags);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}


// ---- TET10 with the 14-point rule: two elements per warp -----------------------------------------------------------------
// The same idea as k_elem_poisson_hex8x4 for 14 quadrature points / 10 nodes: lane = (slot = lane>>4, point or node
// lane&15).  Phase A serves two consecutive tets per pass; the contraction (one 8x8 DMMA block + two leftover rows by
// DFMA, as in elem_process) runs slot after slot inside each k group so the table entries are loaded once.
template <int WARPS, int RS>
__global__ void __launch_bounds__(WARPS * 32, 1)
k_elem_poisson_tet10x2(const double* __restrict__ xyz, const uint32_t* __restrict__ conn, const double* __restrict__ tab, int forcing,
                       const double* __restrict__ f_tab, int64_t e0, int64_t e1, double* __restrict__ kf, int rs_rt, int* flags) {
  constexpr int NPE = 10, NQ = 14;
  using C = ElemCfg<NPE, NQ>;
  static_assert(C::NBF == 1 && C::REM == 2 && C::NB == 2 && C::QG == 4, "layout assumed below");
  constexpr unsigned FULL = 0xffffffffu;
  const int rs = RS ? RS : rs_rt;
  extern __shared__ double smem[];
  double* s_tab = smem;
  double* s_tw = smem + C::TAB_SZ;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  elem_load_tables<NPE, NQ>(s_tab, s_tw, tab, threadIdx.x, WARPS * 32);
  double* Xw = s_tw + C::TW_SZ + (size_t)warp * (2 * 30 + 32);  // [slot][node][3] | fw[slot][16]
  double* fw = Xw + 2 * 30;
  const int slot = lane >> 4, l16 = lane & 15, g = lane >> 2, t4 = lane & 3;
  const int q = l16 < NQ ? l16 : 0;  // pad lanes repeat point 0; their table columns are zero
  const double* X = Xw + slot * 30;
  const double* twp = s_tw + g * C::TS + t4;
  const double* s_phi = s_tab;
  const double* s_dxi = s_tab + NPE * NQ;
  const double* s_deta = s_dxi + NPE * NQ;
  const double* s_dzeta = s_deta + NPE * NQ;
  const double* s_qw = s_tab + 4 * NPE * NQ;
  __syncthreads();

  const int64_t stride = (int64_t)gridDim.x * WARPS * 2;
  int64_t eb = e0 + ((int64_t)blockIdx.x * WARPS + warp) * 2;
  auto elem_of = [&](int64_t base) { const int64_t e = base + slot; return e < e1 ? e : e1 - 1; };
  double nx = 0., ny = 0., nz = 0.;
  uint32_t nnode = 0;
  if (eb < e1 && l16 < NPE) {
    const uint64_t node = conn[elem_of(eb) * NPE + l16];
    nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
    if (eb + stride < e1) nnode = conn[elem_of(eb + stride) * NPE + l16];
  }
  for (; eb < e1; eb += stride) {
    const int64_t e = eb + slot;
    const bool valid = e < e1;
    if (l16 < NPE) {
      Xw[slot * 30 + 3 * l16] = nx; Xw[slot * 30 + 3 * l16 + 1] = ny; Xw[slot * 30 + 3 * l16 + 2] = nz;
      if (eb + stride < e1) {
        const uint64_t node = nnode;
        nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
        if (eb + 2 * stride < e1) nnode = conn[elem_of(eb + 2 * stride) * NPE + l16];
      }
    }
    __syncwarp();

    // ---- A: lane = (slot, point) -----------------------------------------------------------------------------------
    const unsigned okmask = __ballot_sync(FULL, affine_check_lane<NPE>(X, l16));
    const bool affine = ((okmask >> (16 * slot)) & 0xffffu) == 0xffffu;
    const int qj = affine ? 0 : q;
    double x = 0., y = 0., z = 0., xa = 0., ya = 0., za = 0., xb = 0., yb = 0., zb = 0., xc = 0., yc = 0., zc = 0.;
#pragma unroll
    for (int i = 0; i < NPE; ++i) {
      const double px = X[3 * i], py = X[3 * i + 1], pz = X[3 * i + 2];
      const double p = s_phi[i * NQ + q], a = s_dxi[i * NQ + qj], b = s_deta[i * NQ + qj], c = s_dzeta[i * NQ + qj];
      x += p * px; y += p * py; z += p * pz;
      // Jacobian sums in the reference's order and roundings (fe_map.C:994-1008; see geom_phase)
      xa = __dadd_rn(xa, __dmul_rn(a, px)); ya = __dadd_rn(ya, __dmul_rn(a, py)); za = __dadd_rn(za, __dmul_rn(a, pz));
      xb = __dadd_rn(xb, __dmul_rn(b, px)); yb = __dadd_rn(yb, __dmul_rn(b, py)); zb = __dadd_rn(zb, __dmul_rn(b, pz));
      xc = __dadd_rn(xc, __dmul_rn(c, px)); yc = __dadd_rn(yc, __dmul_rn(c, py)); zc = __dadd_rn(zc, __dmul_rn(c, pz));
    }
    const double jac = (xa * (yb * zc - zb * yc) + ya * (zb * xc - xb * zc) + za * (xb * yc - yb * xc));
    const bool bad = !(jac > 0.);
    if (bad && valid && l16 < NQ) atomicAdd(&flags[0], 1);
    double JxW = jac * s_qw[q];
    if (affine && q != 0) {  // fe_map.C:1242: JxW[0] / qw[0] * qw[p], the quotient correctly rounded (Markstein, see geom_phase)
      const double w0 = s_qw[0], rc = 1. / w0, j0 = __dmul_rn(jac, w0), q0 = __dmul_rn(j0, rc);
      JxW = __dmul_rn(__fma_rn(__fma_rn(-q0, w0, j0), rc, q0), s_qw[q]);
    }
    const double sc = bad ? 0. : rsqrt(jac);
    double S[9];
    S[0] = (yb * zc - zb * yc) * sc; S[1] = (zb * xc - xb * zc) * sc; S[2] = (xb * yc - yb * xc) * sc;
    S[3] = (za * yc - ya * zc) * sc; S[4] = (xa * zc - za * xc) * sc; S[5] = (ya * xc - xa * yc) * sc;
    S[6] = (ya * zb - za * yb) * sc; S[7] = (za * xb - xa * zb) * sc; S[8] = (xa * yb - ya * xb) * sc;
    {
      double f = 0.;
      if (forcing == FEASM_FORCING_EX4_FD) {
        const bool big = !(fabs(x) + fabs(y) + fabs(z) < 2.e4);
        f = __any_sync(FULL, big) ? ex4_forcing_lib(x, y, z) : ex4_forcing_t<true>(x, y, z);
      } else if (forcing == FEASM_FORCING_TABULATED && l16 < NQ) {
        f = f_tab[(valid ? e : e1 - 1) * NQ + l16];
      }
      fw[lane] = l16 < NQ ? JxW * f : 0.;
    }

    // ---- B+C: Ke_s = G_s G_s^T; rows 0-7 by DMMA, rows 8-9 by DFMA against shuffled fragments -----------------------------
    double acc[2][2] = {{0., 0.}, {0., 0.}};
    double pr[2][2][2] = {{{0., 0.}, {0., 0.}}, {{0., 0.}, {0., 0.}}};  // [slot][row block][leftover row]
#pragma unroll
    for (int qg = 0; qg < C::QG; ++qg) {
      double ta[2], tb[2], tc[2];
#pragma unroll
      for (int ib = 0; ib < 2; ++ib) {
        ta[ib] = twp[ib * 8 * C::TS + 4 * qg];
        tb[ib] = twp[2 * 8 * C::TS + ib * 8 * C::TS + 4 * qg];
        tc[ib] = twp[4 * 8 * C::TS + ib * 8 * C::TS + 4 * qg];
      }
#pragma unroll
      for (int s2 = 0; s2 < 2; ++s2) {
        const int src = 16 * s2 + 4 * qg + t4;
        double Sq[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) Sq[k] = __shfl_sync(FULL, S[k], src);
#pragma unroll
        for (int d = 0; d < 3; ++d) {
          const double G0 = ta[0] * Sq[d] + tb[0] * Sq[3 + d] + tc[0] * Sq[6 + d];
          const double G1 = ta[1] * Sq[d] + tb[1] * Sq[3 + d] + tc[1] * Sq[6 + d];
          dmma884(acc[s2][0], acc[s2][1], G0, G0);
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const double gr = __shfl_sync(FULL, G1, 4 * r + t4);
            pr[s2][0][r] += G0 * gr;
            pr[s2][1][r] += G1 * gr;
          }
        }
      }
    }
#pragma unroll
    for (int s2 = 0; s2 < 2; ++s2)
#pragma unroll
      for (int ib = 0; ib < 2; ++ib)
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          pr[s2][ib][r] += __shfl_xor_sync(FULL, pr[s2][ib][r], 1);
          pr[s2][ib][r] += __shfl_xor_sync(FULL, pr[s2][ib][r], 2);
        }
    __syncwarp();  // fw

    // ---- D: Fe (lane = (slot, shape function)) and the stores ------------------------------------------------------------
    double* out = kf + (uint64_t)(eb - e0) * (NPE * rs);
    if (valid && l16 < NPE) {
      double Fe = 0.;
      if (forcing != FEASM_FORCING_NONE) {
#pragma unroll
        for (int k = 0; k < NQ; ++k) Fe += fw[16 * slot + k] * s_phi[l16 * NQ + k];
      }
      out[(uint64_t)slot * (NPE * rs) + l16 * rs + NPE] = Fe;
    }
#pragma unroll
    for (int s2 = 0; s2 < 2; ++s2) {
      if (eb + s2 >= e1) continue;
      double* o = out + (uint64_t)s2 * (NPE * rs);
      *reinterpret_cast<double2*>(o + g * rs + 2 * t4) = make_double2(acc[s2][0], acc[s2][1]);
#pragma unroll
      for (int r = 0; r < 2; ++r)
        if (t4 == r) {
          o[g * rs + 8 + r] = pr[s2][0][r];
          o[(8 + r) * rs + g] = pr[s2][0][r];
          if (g < 2) o[(8 + g) * rs + 8 + r] = pr[s2][1][r];
        }
    }
    __syncwarp();
  }
}

template <int WARPS, int RS>
int launch_poisson_tet10x2(feasm_ctx* ctx, const feasm_params& p, const double* d_f, int64_t e0, int64_t e1, double* d_kf, int rs) {
  using C = ElemCfg<10, 14>;
  const size_t smem = sizeof(double) * (size_t)(C::TAB_SZ + C::TW_SZ + WARPS * (2 * 30 + 32));
  auto kern = k_elem_poisson_tet10x2<WARPS, RS>;
  FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int64_t blocks = (e1 - e0 + 2 * WARPS - 1) / (2 * WARPS);
  if (blocks > ctx->sm_count) blocks = ctx->sm_count;
  if (blocks < 1) blocks = 1;
  kern<<<(unsigned)blocks, WARPS * 32, smem, ctx->stream>>>(ctx->d_xyz, ctx->d_conn, ctx->d_tab, p.forcing, d_f, e0, e1, d_kf, rs, ctx->d_flags);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}


// ---- affine elements: reference-tensor contraction (feasm_params.affine_tensor) ----------------------------------
// On an element with an affine map (Elem::has_affine_map; FEMap::compute_affine_map, fe_map.C:1167-1244) the inverse
// Jacobian is one matrix for all quadrature points and JxW_q = |J| w_q, so the quadrature loop of
// introduction_ex4.C:464-467 factors:
//   Ke(a,b) = sum_q JxW_q grad phi_a . grad phi_b = sum_{c,c'} G_cc' M_cc'(a,b),
//   G_cc' = |J| sum_d dxi_c/dx_d dxi_c'/dx_d (3x3, symmetric),  M_cc'(a,b) = sum_q w_q d_c phi^_a(q) d_c' phi^_b(q).
// The six tables N_k = M_cc (k = cc) and M_cc' + M_c'c (c < c') depend on the element TYPE only and live in shared
// memory: an element costs 6 FMA per Ke entry instead of a 3*nq-long contraction (HEX27: 4.4 k instead of 59 k FMA), and
// the kernel becomes a streaming write of the records.  Fe still needs the forcing at the quadrature points (same code
// as the quadrature kernel).  Elements whose map is not affine are appended to a list and go through the quadrature
// loop (k_elem_poisson_list).  Same Ke up to rounding: the parity tests hold it to the same 1e-12.
template <int NPE, int NQ, int WARPS, int EB>
__global__ void __launch_bounds__(WARPS * 32, 1)
k_elem_poisson_affine(const double* __restrict__ xyz, const uint32_t* __restrict__ conn, const double* __restrict__ tab, int forcing,
                      const double* __restrict__ f_tab, int64_t e0, int64_t e1, double* __restrict__ kf, int rs, int* flags,
                      uint32_t* __restrict__ na_list, unsigned* na_count) {
  using C = ElemCfg<NPE, NQ>;
  constexpr int NLD = NPE + 1;  // row stride of the N tables
  extern __shared__ double smem[];
  double* s_tab = smem;
  double* s_N = smem + C::TAB_SZ;  // [6][NPE][NLD]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = threadIdx.x; t < C::TAB; t += WARPS * 32) s_tab[t] = tab[t];
  __syncthreads();
  {
    const double* s_qw = s_tab + 4 * NPE * NQ;
    for (int t = threadIdx.x; t < 6 * NPE * NPE; t += WARPS * 32) {
      const int k = t / (NPE * NPE), a = (t / NPE) % NPE, b = t % NPE;
      const int c = k < 3 ? k : (k == 5 ? 1 : 0), cp = k < 3 ? k : (k == 3 ? 1 : 2);
      const double* da = s_tab + (1 + c) * NPE * NQ;
      const double* db = s_tab + (1 + cp) * NPE * NQ;
      double m = 0.;
      for (int q = 0; q < NQ; ++q) {
        double term = da[a * NQ + q] * db[b * NQ + q];
        if (c != cp) term += db[a * NQ + q] * da[b * NQ + q];
        m += s_qw[q] * term;
      }
      s_N[(k * NPE + a) * NLD + b] = m;
    }
  }
  double* X = s_N + 6 * NPE * NLD + (size_t)warp * C::WARP_DBL;
  double* fw = X + C::X_SZ;
  const double* s_phi = s_tab;
  __syncthreads();

  const int64_t stride = (int64_t)gridDim.x * WARPS;
  int64_t e = e0 + (int64_t)blockIdx.x * WARPS + warp;
  double nx = 0., ny = 0., nz = 0.;  // two-deep prefetch as in k_elem_poisson
  uint32_t nnode = 0;
  if (e < e1 && lane < NPE) {
    const uint64_t node = conn[e * NPE + lane];
    nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
    if (e + stride < e1) nnode = conn[(e + stride) * NPE + lane];
  }
  // EB elements per pass: geometry and Fe one after the other, then ONE sweep over the N tables feeds all of them (the
  // tables are 6*NPE*NPE*8 B of shared-memory reads per sweep; the LSU data pipe, not the FP64 pipe, bounds this kernel)
  while (e < e1) {
    double g[EB][6];
    double* out[EB];
#pragma unroll
    for (int j = 0; j < EB; ++j) {
      out[j] = nullptr;
#pragma unroll
      for (int k = 0; k < 6; ++k) g[j][k] = 0.;
      if (e >= e1) continue;  // warp-uniform
      if (lane < NPE) {
        X[3 * lane] = nx; X[3 * lane + 1] = ny; X[3 * lane + 2] = nz;
        if (e + stride < e1) {
          const uint64_t node = nnode;
          nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
          if (e + 2 * stride < e1) nnode = conn[(e + 2 * stride) * NPE + lane];
        }
      }
      __syncwarp();
      double S[9], JxW, x, y, z;
      const bool affine = geom_phase<NPE, NQ>(X, s_tab, lane, forcing == FEASM_FORCING_EX4_FD, flags, S, JxW, x, y, z);
      if (!affine) {  // leave it to the quadrature loop
        if (lane == 0) na_list[atomicAdd(na_count, 1u)] = (uint32_t)e;
      } else {
        double f = 0.;
        if (forcing == FEASM_FORCING_EX4_FD) {
          const bool big = !(fabs(x) + fabs(y) + fabs(z) < 2.e4);
          f = __any_sync(0xffffffffu, big) ? ex4_forcing_lib(x, y, z) : ex4_forcing_t<true>(x, y, z);
        } else if (forcing == FEASM_FORCING_TABULATED && lane < NQ) {
          f = f_tab[e * NQ + lane];
        }
        fw[lane] = lane < NQ ? JxW * f : 0.;
        __syncwarp();
        // S = sqrt|J| * inverse-Jacobian rows (the same on every lane): G_cc' = sum_d S[c][d] S[c'][d]
        g[j][0] = S[0] * S[0] + S[1] * S[1] + S[2] * S[2]; g[j][1] = S[3] * S[3] + S[4] * S[4] + S[5] * S[5];
        g[j][2] = S[6] * S[6] + S[7] * S[7] + S[8] * S[8]; g[j][3] = S[0] * S[3] + S[1] * S[4] + S[2] * S[5];
        g[j][4] = S[0] * S[6] + S[1] * S[7] + S[2] * S[8]; g[j][5] = S[3] * S[6] + S[4] * S[7] + S[5] * S[8];
        out[j] = kf + (uint64_t)(e - e0) * (NPE * rs);
        if (lane < NPE) {
          double Fe = 0., f1 = 0., f2 = 0.;
          if (forcing != FEASM_FORCING_NONE) {
#pragma unroll
            for (int q = 0; q + 2 < NQ; q += 3) {
              Fe += fw[q] * s_phi[lane * NQ + q];
              f1 += fw[q + 1] * s_phi[lane * NQ + q + 1];
              f2 += fw[q + 2] * s_phi[lane * NQ + q + 2];
            }
#pragma unroll
            for (int q = NQ - NQ % 3; q < NQ; ++q) Fe += fw[q] * s_phi[lane * NQ + q];
            Fe = (Fe + f1) + f2;
          }
          out[j][lane * rs + NPE] = Fe;
        }
      }
      __syncwarp();
      e += stride;
    }
    if (lane < NPE) {
      const double* Nl = s_N + lane;
#pragma unroll 3
      for (int a = 0; a < NPE; ++a) {
        const double* r = Nl + a * NLD;
        const double n0 = r[0], n1 = r[NPE * NLD], n2 = r[2 * NPE * NLD], n3 = r[3 * NPE * NLD], n4 = r[4 * NPE * NLD],
                     n5 = r[5 * NPE * NLD];
#pragma unroll
        for (int j = 0; j < EB; ++j)
          if (out[j])  // warp-uniform
            out[j][a * rs + lane] = fma(g[j][5], n5, fma(g[j][4], n4, fma(g[j][3], n3, fma(g[j][2], n2, fma(g[j][1], n1, g[j][0] * n0)))));
      }
    }
    __syncwarp();
  }
}

// The quadrature kernel over a device-side element list (the non-affine leftovers of the kernel above).
template <int NPE, int NQ, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
k_elem_poisson_list(const double* __restrict__ xyz, const uint32_t* __restrict__ conn, const double* __restrict__ tab, int forcing,
                    const double* __restrict__ f_tab, int64_t e0, double* __restrict__ kf, int rs, int* flags,
                    const uint32_t* __restrict__ list, const unsigned* __restrict__ list_n) {
  using C = ElemCfg<NPE, NQ>;
  const unsigned n = *list_n;
  if (n == 0) return;
  extern __shared__ double smem[];
  double* s_tab = smem;
  double* s_tw = smem + C::TAB_SZ;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  elem_load_tables<NPE, NQ>(s_tab, s_tw, tab, threadIdx.x, WARPS * 32);
  double* X = s_tw + C::TW_SZ + (size_t)warp * C::WARP_DBL;
  double* fw = X + C::X_SZ;
  const double* twp = s_tw + (lane >> 2) * C::TS + (lane & 3);
  __syncthreads();
  for (unsigned i = blockIdx.x * WARPS + warp; i < n; i += gridDim.x * WARPS) {
    const int64_t e = list[i];
    if (lane < NPE) {
      const uint64_t node = conn[e * NPE + lane];
      X[3 * lane] = xyz[node * 3]; X[3 * lane + 1] = xyz[node * 3 + 1]; X[3 * lane + 2] = xyz[node * 3 + 2];
    }
    __syncwarp();
    elem_process<NPE, NQ>(X, fw, s_tab, twp, lane, forcing, f_tab ? f_tab + e * NQ : nullptr, flags,
                          kf + (uint64_t)(e - e0) * (NPE * rs), rs);
    __syncwarp();
  }
}

template <int NPE, int NQ, int WARPS, int LWARPS, int EB = 4>
int launch_poisson_affine(feasm_ctx* ctx, const feasm_params& p, const double* d_f, int64_t e0, int64_t e1, double* d_kf, int rs) {
  using C = ElemCfg<NPE, NQ>;
  if (!ctx->d_na_list) FEASM_CUDA(ctx, cudaMalloc(&ctx->d_na_list, sizeof(uint32_t) * ((size_t)ctx->n_elem + 1)));
  unsigned* d_count = ctx->d_na_list + ctx->n_elem;
  FEASM_CUDA(ctx, cudaMemsetAsync(d_count, 0, sizeof(unsigned), ctx->stream));
  {
    const size_t smem = sizeof(double) * (size_t)(C::TAB_SZ + 6 * NPE * (NPE + 1) + WARPS * C::WARP_DBL);
    auto kern = k_elem_poisson_affine<NPE, NQ, WARPS, EB>;
    FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = (e1 - e0 + WARPS - 1) / WARPS;
    if (blocks > ctx->sm_count) blocks = ctx->sm_count;
    if (blocks < 1) blocks = 1;
    kern<<<(unsigned)blocks, WARPS * 32, smem, ctx->stream>>>(ctx->d_xyz, ctx->d_conn, ctx->d_tab, p.forcing, d_f, e0, e1, d_kf, rs,
                                                             ctx->d_flags, ctx->d_na_list, d_count);
    FEASM_CUDA(ctx, cudaGetLastError());
  }
  {
    const size_t smem = sizeof(double) * (size_t)C::SMEM_DBL(LWARPS);
    auto kern = k_elem_poisson_list<NPE, NQ, LWARPS>;
    FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)ctx->sm_count, LWARPS * 32, smem, ctx->stream>>>(ctx->d_xyz, ctx->d_conn, ctx->d_tab, p.forcing, d_f, e0, d_kf, rs,
                                                                     ctx->d_flags, ctx->d_na_list, d_count);
    FEASM_CUDA(ctx, cudaGetLastError());
  }
  return FEASM_OK;
}


// ---- 3-variable isotropic linear elasticity (systems_of_equations_ex6.C:220-237) --------------------------------
// Ke[(i,a),(k,b)] = sum_q JxW sum_jl C_ijkl dphi_b,l dphi_a,j with C_ijkl = l1 d_ij d_kl + l2 (d_ik d_jl + d_il d_jk)
//                 = l1 P_ik[a][b] + l2 P_ki[a][b] + l2 d_ik (P_00 + P_11 + P_22)[a][b],
// P_dd'[a][b] = sum_q G_d[a][q] G_d'[b][q], G as in the Poisson kernel.  So the 81x81 element matrix needs only the
// six distinct 27x27 products P_dd' (d <= d'; P_d'd = P_dd'^T): 7 k-steps of DMMA each instead of a 168-long
// contraction over 81 rows — the isotropic structure cuts the tensor work ninefold (SURVEY a7 allows either count;
// bench/DESIGN quote the flops of THIS algorithm).  One warp per element; each P goes through a shared-memory tile
// (odd row stride: row and column reads are both conflict free) from which the nine variable blocks are combined
// and stored.  Fe_(i,a) = f_i sum_q JxW phi_a(q) (body force; the traction term is added by k_side_loads, feasm_side.cu).
// Shared memory per warp: the three symmetric products P_00, P_11, P_22 packed (upper triangle, row-major) and TWO full
// tiles for the other products, so that every record row of the element (81 entries + Fe) is formed and written ONCE,
// whole, rows in ascending order: the element's 70 KB of records leave the SM as one sequential stream.  Row block 0 needs
// P_01 and P_02, block 1 P_01 and P_12, block 2 P_02 and P_12: with two tiles P_02 is computed twice (+112 of 546 DMMA) —
// the price of 8 warps per SM instead of the 6 that three tiles allow (measured: 6 warps with all six products resident
// ran 6.48 ms, the DRAM-side store cost gone but the DMMA chains of 1-2 warps per sub-partition exposed).  (The first form wrote each 27-wide column block of a row from a different product, microseconds apart: 216 B
// pieces at an 864 B stride, every row touched three times with partial sectors at both ends; the stores cost 3.4 of the
// kernel's 6.6 ms, 2.4 ms of it on the DRAM side — profiles/r2/ncu_full_elast_64.md.)
template <int NPE>
struct ElastTiles {
  static constexpr int TLD = NPE | 1;                         // odd stride: row and column reads are both conflict free
  static constexpr int TILE = ((NPE * TLD + 1) & ~1);
  static constexpr bool PACKED = NPE > 10;                    // small elements have the room for full symmetric tiles
  static constexpr int PK = PACKED ? ((NPE * (NPE + 1) / 2 + 1) & ~1) : TILE;  // symmetric tile: packed upper triangle, or full
  static constexpr int PER_WARP = 3 * PK + 2 * TILE + 32 + ((NPE * 3 + 1) & ~1) + 32;  // + m[a] = sum_q JxW phi_a, + X | fw
  __host__ __device__ static constexpr int pk(int r, int c) { return r * NPE - r * (r - 1) / 2 + (c - r); }  // r <= c
};

template <int NPE, int NQ, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
k_elem_elasticity(const double* __restrict__ xyz, const uint32_t* __restrict__ conn, const double* __restrict__ tab, double lam1,
                  double lam2, double f0, double f1, double f2, int64_t e0, int64_t e1, double* __restrict__ kf, int rs,
                  int* flags) {
  using C = ElemCfg<NPE, NQ>;
  using T = ElastTiles<NPE>;
  constexpr int NBP = (NPE + 7) / 8;   // 8-row blocks, padded (all go through DMMA here)
  constexpr int ND = 3 * NPE;
  const int LDR = rs;  // record stride in doubles
  constexpr int TLD = T::TLD;
  extern __shared__ double smem[];
  double* s_tab = smem;
  double* s_tw = smem + C::TAB_SZ;      // [c][row][TS], rows padded to NBP*8
  constexpr int TWE = 3 * NBP * 8 * C::TS;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = threadIdx.x; t < C::TAB; t += WARPS * 32) s_tab[t] = tab[t];
  for (int t = threadIdx.x; t < TWE; t += WARPS * 32) {
    const int c = t / (NBP * 8 * C::TS), row = (t / C::TS) % (NBP * 8), q = t % C::TS;
    s_tw[t] = (row < NPE && q < NQ) ? sqrt(tab[4 * NPE * NQ + q]) * tab[(1 + c) * NPE * NQ + row * NQ + q] : 0.;
  }
  const double* s_phi = s_tab;
  double* wbase = s_tw + TWE + (size_t)warp * T::PER_WARP;
  double* diag = wbase;                  // P_00 | P_11 | P_22, packed
  double* off = wbase + 3 * T::PK;       // two full tiles
  double* mfe = off + 2 * T::TILE;       // m[a]
  double* X = mfe + 32;                  // node coordinates of the element
  double* fw = X + C::X_SZ;              // JxW per quadrature point
  const int g = lane >> 2, t4 = lane & 3;
  const double* twp = s_tw + g * C::TS + t4;
  __syncthreads();

  const int64_t stride = (int64_t)gridDim.x * WARPS;
  // Two-deep prefetch of the element's inputs: the node coordinates of the element after this one and the node ids of the
  // one after that are in flight while this one is computed.
  const int64_t e_first = e0 + (int64_t)blockIdx.x * WARPS + warp;
  const bool nl = lane < NPE;
  uint64_t node1 = 0;
  double cx = 0., cy = 0., cz = 0.;
  if (nl && e_first < e1) {
    const uint64_t node = conn[e_first * NPE + lane];
    cx = xyz[node * 3]; cy = xyz[node * 3 + 1]; cz = xyz[node * 3 + 2];
  }
  if (nl && e_first + stride < e1) node1 = conn[(e_first + stride) * NPE + lane];
  for (int64_t e = e_first; e < e1; e += stride) {
    if (nl) { X[3 * lane] = cx; X[3 * lane + 1] = cy; X[3 * lane + 2] = cz; }
    __syncwarp();
    if (nl && e + stride < e1) { cx = xyz[node1 * 3]; cy = xyz[node1 * 3 + 1]; cz = xyz[node1 * 3 + 2]; }
    if (nl && e + 2 * stride < e1) node1 = conn[(e + 2 * stride) * NPE + lane];
    double S[9], JxW, x, y, z;
    const bool affine = geom_phase<NPE, NQ>(X, s_tab, lane, false, flags, S, JxW, x, y, z);
    fw[lane] = lane < NQ ? JxW : 0.;
    __syncwarp();
    if (nl) {  // Fe (systems_of_equations_ex6.C:233-236): body force f_i, the same sum for every variable up to the factor
      double m = 0.;
      for (int q = 0; q < NQ; ++q) m += fw[q] * s_phi[lane * NQ + q];
      mfe[lane] = m;
    }
    __syncwarp();
    double* out = kf + (uint64_t)(e - e0) * (ND * LDR);

    // P_dd' for one (d, d') pair.  SYM (d == d'): only the blocks on and above the diagonal go through DMMA, and only
    // the upper triangle is stored (packed); else the full tile Tl[a*TLD + b].
    auto product = [&](int d, int dp, double* Tl, auto sym) {
      constexpr bool SYM = decltype(sym)::value;
      double acc[NBP][NBP][2];
#pragma unroll
      for (int ib = 0; ib < NBP; ++ib)
#pragma unroll
        for (int jb = 0; jb < NBP; ++jb) { acc[ib][jb][0] = 0.; acc[ib][jb][1] = 0.; }
#pragma unroll kElastUnroll
      for (int qg = 0; qg < C::QG; ++qg) {
        double s0 = S[d], s1 = S[3 + d], s2 = S[6 + d], r0 = S[dp], r1 = S[3 + dp], r2 = S[6 + dp];
        if (!affine) {
          const int src = 4 * qg + t4;
          s0 = __shfl_sync(0xffffffffu, s0, src); s1 = __shfl_sync(0xffffffffu, s1, src); s2 = __shfl_sync(0xffffffffu, s2, src);
          r0 = __shfl_sync(0xffffffffu, r0, src); r1 = __shfl_sync(0xffffffffu, r1, src); r2 = __shfl_sync(0xffffffffu, r2, src);
        }
        double A[NBP], B[NBP];
#pragma unroll
        for (int ib = 0; ib < NBP; ++ib) {
          const double a = twp[ib * 8 * C::TS + 4 * qg];
          const double b = twp[NBP * 8 * C::TS + ib * 8 * C::TS + 4 * qg];
          const double c = twp[2 * NBP * 8 * C::TS + ib * 8 * C::TS + 4 * qg];
          A[ib] = a * s0 + b * s1 + c * s2;
          B[ib] = a * r0 + b * r1 + c * r2;
        }
#pragma unroll
        for (int ib = 0; ib < NBP; ++ib)
#pragma unroll
          for (int jb = 0; jb < NBP; ++jb)
            if (!SYM || ib <= jb) dmma884(acc[ib][jb][0], acc[ib][jb][1], A[ib], B[jb]);
      }
#pragma unroll
      for (int ib = 0; ib < NBP; ++ib)
#pragma unroll
        for (int jb = 0; jb < NBP; ++jb) {
          if (SYM && ib > jb) continue;
          const int i = ib * 8 + g, j = jb * 8 + 2 * t4;
          if (SYM && T::PACKED) {
            if (i < NPE && j < NPE && i <= j) Tl[T::pk(i, j)] = acc[ib][jb][0];
            if (i < NPE && j + 1 < NPE && i <= j + 1) Tl[T::pk(i, j + 1)] = acc[ib][jb][1];
          } else if (SYM) {  // full tile: the block and its mirror image
            if (i < NPE && j < NPE) { Tl[i * TLD + j] = acc[ib][jb][0]; Tl[j * TLD + i] = acc[ib][jb][0]; }
            if (i < NPE && j + 1 < NPE) { Tl[i * TLD + j + 1] = acc[ib][jb][1]; Tl[(j + 1) * TLD + i] = acc[ib][jb][1]; }
          } else {
            if (i < NPE && j < NPE) Tl[i * TLD + j] = acc[ib][jb][0];
            if (i < NPE && j + 1 < NPE) Tl[i * TLD + j + 1] = acc[ib][jb][1];
          }
        }
    };
    // Emit the NPE record rows of variable block i in order.  Row (i, a), column block k, lane = b:
    //   k == i:  (l1 + l2) P_ii[a][b] + l2 ((P_00 + P_11) + P_22)[a][b]
    //   k >  i:  l1 P_ik[a][b] + l2 P_ik[b][a]            k < i:  l2 P_ki[a][b] + l1 P_ki[b][a]
    // and Fe_(i,a) = f_i m_a behind the last block (lane NPE of the third store).  tk[k] = the tile that holds P_ik / P_ki.
    auto emit = [&](auto iv, const double* t_a, const double* t_b) {
      constexpr int i = decltype(iv)::value;
      if (lane > NPE) return;
      const int b = lane < NPE ? lane : 0;
      const double fi = i == 0 ? f0 : (i == 1 ? f1 : f2);
#pragma unroll 3
      for (int a = 0; a < NPE; ++a) {
        double* row = out + (size_t)(i * NPE + a) * LDR;
        const int lo = a < b ? a : b, hi = a < b ? b : a;
        const int pidx = T::PACKED ? T::pk(lo, hi) : a * TLD + b;
        const double p0 = diag[pidx], p1 = diag[T::PK + pidx], p2 = diag[2 * T::PK + pidx];
        const double tr = lam2 * ((p0 + p1) + p2);
        const double pi_ = i == 0 ? p0 : (i == 1 ? p1 : p2);
        int which = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          double v;
          if (k == i) {
            v = (lam1 + lam2) * pi_ + tr;
          } else {
            const double* Tl = which++ == 0 ? t_a : t_b;
            const double p = Tl[a * TLD + b], pt = Tl[b * TLD + a];
            v = i < k ? lam1 * p + lam2 * pt : lam2 * p + lam1 * pt;
          }
          if (lane < NPE) row[k * NPE + lane] = v;
          else if (k == 2) row[ND] = fi * mfe[a];
        }
      }
    };
    double* t0 = off;
    double* t1 = off + T::TILE;
    product(0, 1, t0, std::false_type{});
    product(0, 2, t1, std::false_type{});
    product(0, 0, diag, std::true_type{});
    product(1, 1, diag + T::PK, std::true_type{});
    product(2, 2, diag + 2 * T::PK, std::true_type{});
    __syncwarp();
    emit(std::integral_constant<int, 0>{}, t0, t1);   // columns k = 1: P_01, k = 2: P_02
    __syncwarp();
    product(1, 2, t1, std::false_type{});
    __syncwarp();
    emit(std::integral_constant<int, 1>{}, t0, t1);   // k = 0: P_01 (transposed roles), k = 2: P_12
    __syncwarp();
    product(0, 2, t0, std::false_type{});
    __syncwarp();
    emit(std::integral_constant<int, 2>{}, t0, t1);   // k = 0: P_02, k = 1: P_12
    __syncwarp();
  }
}

template <int NPE, int NQ, int WARPS>
int launch_elasticity(feasm_ctx* ctx, const feasm_params& p, int64_t e0, int64_t e1, double* d_kf, int rs) {
  using C = ElemCfg<NPE, NQ>;
  constexpr int NBP = (NPE + 7) / 8;
  const size_t smem = sizeof(double) * (size_t)(C::TAB_SZ + 3 * NBP * 8 * C::TS + WARPS * ElastTiles<NPE>::PER_WARP);
  auto kern = k_elem_elasticity<NPE, NQ, WARPS>;
  FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // systems_of_equations_ex6.C:136-151
  const double E = p.young_modulus != 0. ? p.young_modulus : 1., nu = p.poisson_ratio != 0. ? p.poisson_ratio : 0.3;
  const double lam1 = (E * nu) / ((1. + nu) * (1. - 2. * nu)), lam2 = E / (2. * (1. + nu));
  double f[3] = {p.body_force[0], p.body_force[1], p.body_force[2]};
  if (f[0] == 0. && f[1] == 0. && f[2] == 0.) f[2] = -1.;
  int64_t blocks = (e1 - e0 + WARPS - 1) / WARPS;
  if (blocks > ctx->sm_count) blocks = ctx->sm_count;
  if (blocks < 1) blocks = 1;
  kern<<<(unsigned)blocks, WARPS * 32, smem, ctx->stream>>>(ctx->d_xyz, ctx->d_conn, ctx->d_tab, lam1, lam2, f[0], f[1], f[2], e0, e1,
                                                           d_kf, rs, ctx->d_flags);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

}  // namespace

// One warp per element of the constrained-element list: the element's records are staged through shared memory (whole
// sectors in, whole sectors out — the constraint pass itself walks columns, which would be one sector per entry in
// global memory), constrained there, and written back.
template <bool BATCH>
__global__ void k_constrain_list(const uint32_t* __restrict__ list, int n_list, int64_t e0, int64_t e1, double* __restrict__ kf, int rs,
                                 int nd, ElemConArgs con) {
  extern __shared__ __align__(16) double s_k[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w = blockIdx.x * (blockDim.x >> 5) + warp;
  if (w >= n_list) return;
  const int64_t e = list[w];
  if (e < e0 || e >= e1) return;
  const int ldr = feasm_ldr(nd), h = ldr >> 1;
  double* S = s_k + (size_t)warp * nd * ldr;
  double* K = kf + (uint64_t)(e - e0) * nd * rs;
  // large elements: eight 16 B loads in flight per lane (the rolled loop waits for every load before the next goes out:
  // 0.234 -> 0.199 ms at 96^3 HEX27); small ones (HEX8: 48 loads per element) are faster rolled
  if (!BATCH) {
    for (int t = lane; t < nd * h; t += 32) {
      const int row = t / h, c = t % h;
      *reinterpret_cast<double2*>(S + row * ldr + 2 * c) = *reinterpret_cast<const double2*>(K + (size_t)row * rs + 2 * c);
    }
  } else
  for (int t0 = 0; t0 < nd * h; t0 += 256) {
    double2 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int t = t0 + 32 * j + lane;
      if (t < nd * h) v[j] = *reinterpret_cast<const double2*>(K + (size_t)(t / h) * rs + 2 * (t % h));
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int t = t0 + 32 * j + lane;
      if (t < nd * h) *reinterpret_cast<double2*>(S + (t / h) * ldr + 2 * (t % h)) = v[j];
    }
  }
  constrain_element(S, ldr, nd, con.elem_dofs + (uint64_t)e * nd, con.con_flag, con.con_val, con.mode, lane);
  for (int t = lane; t < nd * h; t += 32) {
    const int row = t / h, c = t % h;
    *reinterpret_cast<double2*>(K + (size_t)row * rs + 2 * c) = *reinterpret_cast<const double2*>(S + row * ldr + 2 * c);
  }
}

static int launch_element_kernel(feasm_ctx* ctx, const feasm_params& p, const double* d_f, int64_t e0, int64_t e1, double* d_kf, int rs,
                                 int cm, bool* fused) {
  *fused = false;
  if (p.problem == FEASM_POISSON && p.affine_tensor) {
    if (ctx->elem_type == FEASM_HEX27 && ctx->tab_nq == 27) return launch_poisson_affine<27, 27, 16, 12, 2>(ctx, p, d_f, e0, e1, d_kf, rs);
    if (ctx->elem_type == FEASM_HEX8 && ctx->tab_nq == 8) return launch_poisson_affine<8, 8, 16, 16>(ctx, p, d_f, e0, e1, d_kf, rs);
    if (ctx->elem_type == FEASM_HEX8 && ctx->tab_nq == 27) return launch_poisson_affine<8, 27, 16, 16>(ctx, p, d_f, e0, e1, d_kf, rs);
    if (ctx->elem_type == FEASM_TET10 && ctx->tab_nq == 14) return launch_poisson_affine<10, 14, 16, 16>(ctx, p, d_f, e0, e1, d_kf, rs);
  }
  if (p.problem == FEASM_POISSON) {
    if (ctx->elem_type == FEASM_HEX27 && ctx->tab_nq == 27) {
      if (rs == 32) return launch_poisson<27, 27, 12, 32>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);  // 256 B records (uint8 positions)
      return launch_poisson<27, 27, 12>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);
    }
    if (ctx->elem_type == FEASM_HEX8 && ctx->tab_nq == 8) {
      static const bool one_per_warp = getenv("FEASM_HEX8_PACK") && atoi(getenv("FEASM_HEX8_PACK")) == 0;  // A/B knob
      if (!one_per_warp) {
        if (rs == 16) return launch_poisson_hex8x4<16, 16>(ctx, p, d_f, e0, e1, d_kf, rs);
        return launch_poisson_hex8x4<16, 0>(ctx, p, d_f, e0, e1, d_kf, rs);
      }
      if (rs == 16) return launch_poisson<8, 8, 16, 16>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);
      return launch_poisson<8, 8, 16>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);
    }
    if (ctx->elem_type == FEASM_HEX8 && ctx->tab_nq == 27) return launch_poisson<8, 27, 16>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);
    // HEX20 (serendipity): 16 rows of Ke through DMMA, the last 4 through DFMA (ElemCfg::REM)
    if (ctx->elem_type == FEASM_HEX20 && ctx->tab_nq == 27) return launch_poisson<20, 27, 12>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);
    // PRISM18 with the 21-point rule: 16 rows of Ke through DMMA, 2 through DFMA
    if (ctx->elem_type == FEASM_PRISM18 && ctx->tab_nq == 21) return launch_poisson<18, 21, 12>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);
    if (ctx->elem_type == FEASM_TET10 && ctx->tab_nq == 14) {
      static const bool one_per_warp = getenv("FEASM_TET10_PACK") && atoi(getenv("FEASM_TET10_PACK")) == 0;  // A/B knob
      if (!one_per_warp) {
        if (rs == 16) return launch_poisson_tet10x2<16, 16>(ctx, p, d_f, e0, e1, d_kf, rs);
        return launch_poisson_tet10x2<16, 0>(ctx, p, d_f, e0, e1, d_kf, rs);
      }
      if (rs == 16) return launch_poisson<10, 14, 16, 16>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);
      return launch_poisson<10, 14, 16>(ctx, p, d_f, e0, e1, d_kf, rs, cm, fused);
    }
    return feasm_fail(ctx, FEASM_ERR_ARG, "no Poisson kernel for this element type / quadrature order");
  }
  if (p.problem == FEASM_ELASTICITY) {
    if (ctx->n_vars != 3) return feasm_fail(ctx, FEASM_ERR_ARG, "elasticity needs 3 variables per node");
    if (ctx->elem_type == FEASM_HEX27 && ctx->tab_nq == 27) return launch_elasticity<27, 27, FEASM_ELAST_WARPS>(ctx, p, e0, e1, d_kf, rs);
    if (ctx->elem_type == FEASM_HEX20 && ctx->tab_nq == 27) return launch_elasticity<20, 27, 12>(ctx, p, e0, e1, d_kf, rs);
    if (ctx->elem_type == FEASM_PRISM18 && ctx->tab_nq == 21) return launch_elasticity<18, 21, 12>(ctx, p, e0, e1, d_kf, rs);
    if (ctx->elem_type == FEASM_HEX8 && ctx->tab_nq == 8) return launch_elasticity<8, 8, 16>(ctx, p, e0, e1, d_kf, rs);
    if (ctx->elem_type == FEASM_TET10 && ctx->tab_nq == 14) return launch_elasticity<10, 14, 16>(ctx, p, e0, e1, d_kf, rs);
    return feasm_fail(ctx, FEASM_ERR_ARG, "no elasticity kernel for this element type / quadrature order");
  }
  return feasm_fail(ctx, FEASM_ERR_ARG, "problem not implemented");
}

int feasm_launch_element_kernel(feasm_ctx* ctx, const feasm_params& p, const double* d_f, int64_t e0, int64_t e1,
                                double* d_kf, int rs, int constrain_mode) {
  bool fused = false;  // the kernel applied the element-level constraint itself
  int rc = launch_element_kernel(ctx, p, d_f, e0, e1, d_kf, rs, constrain_mode, &fused);
  if (rc) return rc;
  FEASM_CUDA(ctx, cudaEventRecord(ctx->ev[3], ctx->stream));  // feasm_last_timing: element kernel | side loads + constraint pass
  // boundary integrals over the loaded sides (systems_of_equations_ex6.C:240-257), before the constraints like there
  rc = feasm_launch_side_loads(ctx, ctx->tab_qorder, e0, e1, d_kf, rs);  // QGauss(dim-1) of the volume rule's order
  if (rc) return rc;
  // Dirichlet constraints on the few elements that touch a constrained dof, as the reference applies them to Ke/Fe
  // before add_matrix (constrain_mode < 0: leave Ke/Fe unconstrained)
  if (!fused && constrain_mode >= 0 && ctx->n_con_elems > 0) {
    const ElemConArgs con{ctx->d_elem_dofs, ctx->d_con_flag, ctx->d_con_val, constrain_mode};
    const size_t per_warp = sizeof(double) * ctx->nd * feasm_ldr(ctx->nd);
    int warps = (int)(32768 / per_warp);
    warps = warps < 1 ? 1 : (warps > 8 ? 8 : warps);
    const size_t smem = warps * per_warp;
    if (smem > ctx->smem_optin) return feasm_fail(ctx, FEASM_ERR_NOMEM, "element too large for the constraint pass");
    auto kern = ctx->nd * (feasm_ldr(ctx->nd) >> 1) >= 256 ? k_constrain_list<true> : k_constrain_list<false>;
    FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)((ctx->n_con_elems + warps - 1) / warps), warps * 32, smem, ctx->stream>>>(
        ctx->d_con_elems, (int)ctx->n_con_elems, e0, e1, d_kf, rs, ctx->nd, con);
    FEASM_CUDA(ctx, cudaGetLastError());
  }
  return FEASM_OK;
}
