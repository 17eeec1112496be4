// feasm_elem_dev.cuh — device code of the element kernels of feasm_element.cu: configuration, Elem::has_affine_map,
// the ex4 forcing, FEMap::compute_map (geom_phase), the Ke/Fe contraction (elem_compute / elem_store) and the
// element-level Dirichlet constraint.  See feasm_element.cu for the design.
#ifndef FEASM_ELEM_DEV_CUH
#define FEASM_ELEM_DEV_CUH
#include "feasm_internal.h"

namespace {
// What the element-level constraint pass needs.
struct ElemConArgs {
  const uint32_t* elem_dofs;  // [n_elem][nd]
  const uint8_t* con_flag;    // [n_rows]
  const double* con_val;      // [n_rows]
  int mode;
};


__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

constexpr int table_stride(int nqp) {
  // lane (g = lane>>2, t = lane&3) reads T[row0+g][q0+t]: conflict-free 64-bit LDS needs stride = 4 or 12 (mod 16)
  int s = nqp;
  while (s % 16 != 4 && s % 16 != 12) ++s;
  return s;
}

template <int NPE, int NQ>
struct ElemCfg {
  static constexpr int NBF = NPE / 8;               // full 8-row blocks of Ke: DMMA
  static constexpr int REM = NPE - 8 * NBF;         // leftover rows: DFMA
  static constexpr int NB = NBF + (REM ? 1 : 0);    // fragment row blocks a lane generates
  static constexpr int NQP = (NQ + 3) & ~3;         // quadrature points padded to the MMA k step
  static constexpr int QG = NQP / 4;
  static constexpr int TS = table_stride(NQP);      // row stride of the scaled-gradient table
  static constexpr int NBLK = NBF * (NBF + 1) / 2;
  static constexpr int LDR = feasm_ldr(NPE);
  static constexpr int TAB = 4 * NPE * NQ + NQ;     // phi | dxi | deta | dzeta | qw  (global layout, [i][q])
  static constexpr int TAB_SZ = (TAB + 1) & ~1;
  static constexpr int TW_SZ = 3 * NB * 8 * TS;     // sqrt(w_q) dref_c[i][q], zero padded
  static constexpr int X_SZ = (NPE * 3 + 1) & ~1;
  static constexpr int WARP_DBL = X_SZ + 32;        // X | fw
  static constexpr int SMEM_DBL(int warps) { return TAB_SZ + TW_SZ + warps * WARP_DBL; }
  static_assert(NQP <= 32, "lane = quadrature point needs nq <= 32");
  static_assert(NPE <= 32, "lane = shape function needs npe <= 32");
  static_assert(REM <= 4, "leftover rows beyond 4 should become a DMMA block");
};

// Elem::has_affine_map as a list of vector comparisons (a1-a0)*sa ~ (b1-b0)*sb with relative_fuzzy_equals at
// affine_tol = TOLERANCE^2 (include/geom/elem.h:2070, include/utils/fuzzy_equals.h:96-112); one lane per check.
struct AffineCheck { signed char a1, a0, b1, b0; float sa, sb; };
__constant__ AffineCheck c_affine_hex8[4] = {{1, 0, 2, 3, 1, 1}, {1, 0, 5, 4, 1, 1}, {1, 0, 6, 7, 1, 1}, {4, 0, 7, 3, 1, 1}};  // cell_hex8.C:121-135
__constant__ AffineCheck c_affine_hex27[23] = {  // cell_hex27.C:125-166
    {1, 0, 2, 3, 1, 1},     {1, 0, 5, 4, 1, 1},     {1, 0, 6, 7, 1, 1},     {1, 0, 8, 0, .5f, 1},   {1, 0, 10, 3, .5f, 1},
    {1, 0, 16, 4, .5f, 1},  {1, 0, 18, 7, .5f, 1},  {1, 0, 20, 11, .5f, 1}, {1, 0, 21, 12, .5f, 1}, {1, 0, 23, 15, .5f, 1},
    {1, 0, 25, 19, .5f, 1}, {1, 0, 26, 24, .5f, 1}, {4, 0, 7, 3, 1, 1},     {4, 0, 12, 0, .5f, 1},  {4, 0, 13, 1, .5f, 1},
    {4, 0, 14, 2, .5f, 1},  {4, 0, 15, 3, .5f, 1},  {4, 0, 22, 9, .5f, 1},  {4, 0, 24, 11, .5f, 1}, {3, 0, 11, 0, .5f, 1},
    {3, 0, 9, 1, .5f, 1},   {3, 0, 17, 5, .5f, 1},  {3, 0, 19, 4, .5f, 1}};
__constant__ AffineCheck c_affine_hex20[16] = {  // cell_hex20.C:119-153
    {1, 0, 2, 3, 1, 1},    {1, 0, 5, 4, 1, 1},    {1, 0, 6, 7, 1, 1},    {1, 0, 8, 0, .5f, 1},  {1, 0, 10, 3, .5f, 1}, {1, 0, 16, 4, .5f, 1},
    {1, 0, 18, 7, .5f, 1}, {4, 0, 7, 3, 1, 1},    {4, 0, 12, 0, .5f, 1}, {4, 0, 13, 1, .5f, 1}, {4, 0, 14, 2, .5f, 1}, {4, 0, 15, 3, .5f, 1},
    {3, 0, 11, 0, .5f, 1}, {3, 0, 9, 1, .5f, 1},  {3, 0, 17, 5, .5f, 1}, {3, 0, 19, 4, .5f, 1}};
__constant__ AffineCheck c_affine_tet10[6] = {{1, 0, 4, 0, 1, 2}, {2, 1, 5, 1, 1, 2}, {2, 0, 6, 0, 1, 2},  // cell_tet10.C:143-170
                                              {3, 0, 7, 0, 1, 2}, {3, 1, 8, 1, 1, 2}, {3, 2, 9, 2, 1, 2}};

__constant__ AffineCheck c_affine_prism18[14] = {  // cell_prism18.C:122-151
    {3, 0, 4, 1, 1, 1},    {3, 0, 5, 2, 1, 1},    {3, 0, 9, 0, .5f, 1},  {3, 0, 10, 1, .5f, 1}, {3, 0, 11, 2, .5f, 1}, {3, 0, 15, 6, .5f, 1}, {3, 0, 16, 7, .5f, 1},
    {3, 0, 17, 8, .5f, 1}, {1, 0, 6, 0, .5f, 1},  {1, 0, 12, 3, .5f, 1}, {2, 0, 8, 0, .5f, 1},  {2, 0, 14, 3, .5f, 1}, {2, 1, 7, 1, .5f, 1},  {2, 1, 13, 4, .5f, 1}};

template <int NPE>
__device__ __forceinline__ bool affine_check_lane(const double* X, int lane) {
  constexpr int N = NPE == 27 ? 23 : (NPE == 20 ? 16 : (NPE == 18 ? 14 : (NPE == 8 ? 4 : 6)));
  if (lane >= N) return true;
  const AffineCheck c = NPE == 27 ? c_affine_hex27[lane]
                                  : (NPE == 20 ? c_affine_hex20[lane] : (NPE == 18 ? c_affine_prism18[lane] : (NPE == 8 ? c_affine_hex8[lane] : c_affine_tet10[lane])));
  double na = 0., nb = 0., nd = 0.;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double a = (X[3 * c.a1 + d] - X[3 * c.a0 + d]) * (double)c.sa;
    const double b = (X[3 * c.b1 + d] - X[3 * c.b0 + d]) * (double)c.sb;
    na += fabs(a); nb += fabs(b); nd += fabs(a - b);
  }
  return nd <= 1e-12 * (na + nb);
}

// sin (add = 0) or cos (add = 1) of a moderate argument, branch free so that the nine evaluations of one forcing
// value interleave in the FP64 pipe: round-to-nearest quadrant via the 1.5*2^52 magic add, three-term Cody-Waite
// reduction with FMA (pi/2 split into correctly rounded hi/mid/lo), then the classic fdlibm kernel polynomials with
// the coefficient set selected per lane.  Max error 1 ulp against glibc over |a| < 1e3 (checked on the host,
// oracle/check_trig.c); callers guard |a| <= 6e4 and use the library functions beyond.
__device__ __forceinline__ double trig_fast(double a, int add) {
  const double MAGIC = 6755399441055744.0;
  const double q = __fma_rn(a, 0x1.45f306dc9c883p-1, MAGIC);
  const int n = __double2loint(q) + add;
  const double j = __dadd_rn(q, -MAGIC);
  double r = __fma_rn(-j, 0x1.921fb54442d18p+0, a);
  r = __fma_rn(-j, 0x1.1a62633145c07p-54, r);
  r = __fma_rn(-j, -0x1.f1976b7ed8fbcp-110, r);
  const double x2 = __dmul_rn(r, r);
  const bool odd = n & 1;
  double z = odd ? -1.13596475577881948265e-11 : 1.58969099521155010221e-10;
  z = __fma_rn(z, x2, odd ? 2.08757232129817482790e-09 : -2.50507602534068634195e-08);
  z = __fma_rn(z, x2, odd ? -2.75573143513906633035e-07 : 2.75573137070700676789e-06);
  z = __fma_rn(z, x2, odd ? 2.48015872894767294178e-05 : -1.98412698298579493134e-04);
  z = __fma_rn(z, x2, odd ? -1.38888888888741095749e-03 : 8.33333333332248946124e-03);
  z = __fma_rn(z, x2, odd ? 4.16666666666666019037e-02 : -1.66666666666666324348e-01);
  const double w = __fma_rn(z, x2, odd ? -0.5 : 0.0);
  const double res = __fma_rn(w, odd ? x2 : r, odd ? 1.0 : r);
  return (n & 2) ? -res : res;
}

// trig_fast in two halves, so that evaluations whose quadrants have the same parity share ONE selection of the polynomial
// coefficients: the reduction (quadrant n, reduced argument r) ...
struct TrigRed { double r; int n; };
__device__ __forceinline__ TrigRed trig_reduce(double a, int add) {
  const double MAGIC = 6755399441055744.0;
  const double q = __fma_rn(a, 0x1.45f306dc9c883p-1, MAGIC);
  TrigRed t;
  t.n = __double2loint(q) + add;
  const double j = __dadd_rn(q, -MAGIC);
  double r = __fma_rn(-j, 0x1.921fb54442d18p+0, a);
  r = __fma_rn(-j, 0x1.1a62633145c07p-54, r);
  t.r = __fma_rn(-j, -0x1.f1976b7ed8fbcp-110, r);
  return t;
}
// ... and the polynomial for three reduced arguments of one parity (odd = cos kernel, even = sin kernel): the same
// operations on the same values as three trig_fast calls — bitwise the same results, a third of the coefficient selects
// (every select is an issue slot beside the FP64 pipe, profiles/r2/issue_probe.md).
__device__ __forceinline__ void trig_poly3(const TrigRed (&t)[3], const bool odd, double (&out)[3]) {
  const double c6 = odd ? -1.13596475577881948265e-11 : 1.58969099521155010221e-10;
  const double c5 = odd ? 2.08757232129817482790e-09 : -2.50507602534068634195e-08;
  const double c4 = odd ? -2.75573143513906633035e-07 : 2.75573137070700676789e-06;
  const double c3 = odd ? 2.48015872894767294178e-05 : -1.98412698298579493134e-04;
  const double c2 = odd ? -1.38888888888741095749e-03 : 8.33333333332248946124e-03;
  const double c1 = odd ? 4.16666666666666019037e-02 : -1.66666666666666324348e-01;
  const double c0 = odd ? -0.5 : 0.0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double r = t[k].r, x2 = __dmul_rn(r, r);
    double z = __fma_rn(c6, x2, c5);
    z = __fma_rn(z, x2, c4);
    z = __fma_rn(z, x2, c3);
    z = __fma_rn(z, x2, c2);
    z = __fma_rn(z, x2, c1);
    const double w = __fma_rn(z, x2, c0);
    const double res = __fma_rn(w, odd ? x2 : r, odd ? 1.0 : r);
    out[k] = (t[k].n & 2) ? -res : res;
  }
}

// introduction_ex4.C:500-537 with exact_solution of introduction_ex3/exact_solution.C:40-47.  The seven
// exact_solution() calls share nine distinct trig values; every product below is evaluated in the example's order
// ((cos * sin) * cos), so the result equals the literal seven-call form bit for bit up to the last bit of sin/cos.
// Explicit _rn intrinsics: no FMA contraction.  x / eps is evaluated as the correctly rounded quotient through
// q0 = x*(1/eps), r = fma(-q0, eps, x), q = fma(r, 1/eps, q0) (Markstein; checked exhaustively-at-random on the host).
// FAST: the three arguments of one axis (a - h eps, a, a + h eps) lie in quadrants of one parity except within 1.6e-3 of a
// multiple of pi/4; when that holds for all three axes on every lane of the warp, each axis takes trig_poly3, else every
// evaluation selects its own coefficients (trig_fast) — the same values either way.
template <bool FAST>
__device__ __forceinline__ double ex4_forcing_t(double x, double y, double z) {
  const double hp = .5 * 3.141592653589793;  // .5 * acos(-1.)
  const double eps = 1.e-3;
  const double ax[3] = {__dmul_rn(hp, x), __dmul_rn(hp, __dadd_rn(x, -eps)), __dmul_rn(hp, __dadd_rn(x, eps))};
  const double ay[3] = {__dmul_rn(hp, y), __dmul_rn(hp, __dadd_rn(y, -eps)), __dmul_rn(hp, __dadd_rn(y, eps))};
  const double az[3] = {__dmul_rn(hp, z), __dmul_rn(hp, __dadd_rn(z, -eps)), __dmul_rn(hp, __dadd_rn(z, eps))};
  double cx[3], sy[3], cz[3];  // [0] at the point, [1] minus eps, [2] plus eps
  if (FAST) {
    TrigRed tx[3], ty[3], tz[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { tx[k] = trig_reduce(ax[k], 1); ty[k] = trig_reduce(ay[k], 0); tz[k] = trig_reduce(az[k], 1); }
    const int mixed = ((tx[1].n ^ tx[0].n) | (tx[2].n ^ tx[0].n) | (ty[1].n ^ ty[0].n) | (ty[2].n ^ ty[0].n) | (tz[1].n ^ tz[0].n) | (tz[2].n ^ tz[0].n)) & 1;
#ifdef FEASM_DEV_NO_TRIG3  // development build: every evaluation selects its own coefficients
    if (false) {
#else
    if (!__any_sync(0xffffffffu, mixed)) {
#endif
      trig_poly3(tx, tx[0].n & 1, cx);
      trig_poly3(ty, ty[0].n & 1, sy);
      trig_poly3(tz, tz[0].n & 1, cz);
    } else {
#pragma unroll
      for (int k = 0; k < 3; ++k) { cx[k] = trig_fast(ax[k], 1); sy[k] = trig_fast(ay[k], 0); cz[k] = trig_fast(az[k], 1); }
    }
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) { cx[k] = cos(ax[k]); sy[k] = sin(ay[k]); cz[k] = cos(az[k]); }
  }
  const double cx0 = cx[0], cxm = cx[1], cxp = cx[2], sy0 = sy[0], sym = sy[1], syp = sy[2], cz0 = cz[0], czm = cz[1], czp = cz[2];
  auto u = [](double a, double b, double c) { return __dmul_rn(__dmul_rn(a, b), c); };
  auto dv = [eps](double a) {
    if (!FAST) return __ddiv_rn(a, eps);
    const double rc = 1. / eps;  // folded at compile time
    const double q0 = __dmul_rn(a, rc);
    return __fma_rn(__fma_rn(-q0, eps, a), rc, q0);
  };
  auto fd = [&](double m, double p, double u0) { return dv(dv(__dadd_rn(__dadd_rn(m, p), __dmul_rn(-2., u0)))); };
  const double u0 = u(cx0, sy0, cz0);
  const double uxx = fd(u(cxm, sy0, cz0), u(cxp, sy0, cz0), u0);
  const double uyy = fd(u(cx0, sym, cz0), u(cx0, syp, cz0), u0);
  const double uzz = fd(u(cx0, sy0, czm), u(cx0, sy0, czp), u0);
  return -__dadd_rn(__dadd_rn(uxx, uyy), uzz);
}
__device__ __noinline__ double ex4_forcing_lib(double x, double y, double z) { return ex4_forcing_t<false>(x, y, z); }

// Phase A of an element, lane = quadrature point: FEMap::compute_map.  Returns has_affine_map; fills
// S = sqrt(det J) * (dxidx dxidy dxidz | detadx .. | dzetadx ..), JxW and (when want_xyz, or always on the general
// branch) the physical point.  X = node coordinates of the element in shared memory.
template <int NPE, int NQ>
__device__ __forceinline__ bool geom_phase(const double* X, const double* s_tab, int lane, bool want_xyz, int* flags, double (&S)[9],
                                           double& JxW, double& x, double& y, double& z) {
  const double* s_phi = s_tab;
  const double* s_dxi = s_tab + NPE * NQ;
  const double* s_qw = s_tab + 4 * NPE * NQ;
  const bool affine = __all_sync(0xffffffffu, affine_check_lane<NPE>(X, lane));
  {
      x = 0; y = 0; z = 0;
      double xa = 0, ya = 0, za = 0, xb = 0, yb = 0, zb = 0, xc = 0, yc = 0, zc = 0;
      const int q = lane < NQ ? lane : 0;
      if (affine) {
        // compute_affine_map (fe_map.C:1167-1244): the Jacobian at qp 0 only — lane c < 9 sums component c —
        // and xyz at every qp (:1190-1197)
        // The nine Jacobian sums run in the reference's order — node after node, product rounded, then added
        // (Point::add_scaled, fe_map.C:994-1008) — with explicit _rn intrinsics, no FMA contraction: the sum cancels from
        // O(|X|) terms down to O(h), so any other order or rounding differs by eps*|X|/h relative (at 96^3 that alone is
        // 1e-14 of the row maximum in Ke).  The products are independent; only the adds chain.
        double d = 0.;
        if (lane < 9) {
          const double* tabd = s_dxi + (lane / 3) * (NPE * NQ);
          const int cd = lane % 3;
#ifdef FEASM_DEV_JAC3
          double d0 = 0., d1 = 0., d2 = 0.;
#pragma unroll
          for (int i = 0; i + 2 < NPE; i += 3) {
            d0 += tabd[i * NQ] * X[3 * i + cd];
            d1 += tabd[(i + 1) * NQ] * X[3 * (i + 1) + cd];
            d2 += tabd[(i + 2) * NQ] * X[3 * (i + 2) + cd];
          }
#pragma unroll
          for (int i = NPE - NPE % 3; i < NPE; ++i) d0 += tabd[i * NQ] * X[3 * i + cd];
          d = (d0 + d1) + d2;
#else
#pragma unroll
          for (int i = 0; i < NPE; ++i) d = __dadd_rn(d, __dmul_rn(tabd[i * NQ], X[3 * i + cd]));
#endif
        }
        if (want_xyz) {
          double x1 = 0., y1 = 0., z1 = 0., x2 = 0., y2 = 0., z2 = 0.;
#pragma unroll
          for (int i = 0; i + 2 < NPE; i += 3) {
            const double p0 = s_phi[i * NQ + q], p1 = s_phi[(i + 1) * NQ + q], p2 = s_phi[(i + 2) * NQ + q];
            x += p0 * X[3 * i]; y += p0 * X[3 * i + 1]; z += p0 * X[3 * i + 2];
            x1 += p1 * X[3 * i + 3]; y1 += p1 * X[3 * i + 4]; z1 += p1 * X[3 * i + 5];
            x2 += p2 * X[3 * i + 6]; y2 += p2 * X[3 * i + 7]; z2 += p2 * X[3 * i + 8];
          }
#pragma unroll
          for (int i = NPE - NPE % 3; i < NPE; ++i) {
            const double p = s_phi[i * NQ + q];
            x += p * X[3 * i]; y += p * X[3 * i + 1]; z += p * X[3 * i + 2];
          }
          x = (x + x1) + x2; y = (y + y1) + y2; z = (z + z1) + z2;
        }
        xa = __shfl_sync(0xffffffffu, d, 0); ya = __shfl_sync(0xffffffffu, d, 1); za = __shfl_sync(0xffffffffu, d, 2);
        xb = __shfl_sync(0xffffffffu, d, 3); yb = __shfl_sync(0xffffffffu, d, 4); zb = __shfl_sync(0xffffffffu, d, 5);
        xc = __shfl_sync(0xffffffffu, d, 6); yc = __shfl_sync(0xffffffffu, d, 7); zc = __shfl_sync(0xffffffffu, d, 8);
      } else {
        // general branch: compute_single_point_map per qp (fe_map.C:994-1026)
        const double* s_deta = s_dxi + NPE * NQ;
        const double* s_dzeta = s_deta + NPE * NQ;
#pragma unroll 3
        for (int i = 0; i < NPE; ++i) {
          const double px = X[3 * i], py = X[3 * i + 1], pz = X[3 * i + 2];
          const double p = s_phi[i * NQ + q], a = s_dxi[i * NQ + q], b = s_deta[i * NQ + q], c = s_dzeta[i * NQ + q];
          x += p * px; y += p * py; z += p * pz;
          // Jacobian sums: the reference's order and roundings (see the affine branch)
          xa = __dadd_rn(xa, __dmul_rn(a, px)); ya = __dadd_rn(ya, __dmul_rn(a, py)); za = __dadd_rn(za, __dmul_rn(a, pz));
          xb = __dadd_rn(xb, __dmul_rn(b, px)); yb = __dadd_rn(yb, __dmul_rn(b, py)); zb = __dadd_rn(zb, __dmul_rn(b, pz));
          xc = __dadd_rn(xc, __dmul_rn(c, px)); yc = __dadd_rn(yc, __dmul_rn(c, py)); zc = __dadd_rn(zc, __dmul_rn(c, pz));
        }
      }
      // det J in the order of fe_map.C:1046-1048, inverse rows :1056-1068
      const double jac = (xa * (yb * zc - zb * yc) + ya * (zb * xc - xb * zc) + za * (xb * yc - yb * xc));
      const bool bad = !(jac > 0.);
      if (bad && lane < NQ) atomicAdd(&flags[0], 1);  // check_for_degenerate_map, fe_map.C:467-506
      // fe_map.C:1242 / :1052.  JxW[0] / qw[0] is the correctly rounded quotient by q0 = x * rc, q = fma(fma(-q0, d, x), rc, q0)
      // with rc = RN(1 / qw[0]) (Markstein; 0 mismatches against the IEEE division in 5e7 random trials per rule)
      if (affine && q != 0) {
        const double w0 = s_qw[0], rc = 1. / w0, j0 = __dmul_rn(jac, w0), q0 = __dmul_rn(j0, rc);
        JxW = __dmul_rn(__fma_rn(__fma_rn(-q0, w0, j0), rc, q0), s_qw[q]);
      } else {
        JxW = jac * s_qw[q];
      }
      const double s = bad ? 0. : rsqrt(jac);  // sqrt(jac) / jac
      S[0] = (yb * zc - zb * yc) * s; S[1] = (zb * xc - xb * zc) * s; S[2] = (xb * yc - yb * xc) * s;
      S[3] = (za * yc - ya * zc) * s; S[4] = (xa * zc - za * xc) * s; S[5] = (ya * xc - xa * yc) * s;
      S[6] = (ya * zb - za * yb) * s; S[7] = (za * xb - xa * zb) * s; S[8] = (xa * yb - ya * xb) * s;
  }
  return affine;
}

// What one warp holds of its element after the contraction: the DMMA accumulators of the upper-triangular 8x8 blocks, the
// leftover rows, and (lane = shape function) Fe.
template <int NPE, int NQ>
struct ElemRegs {
  using C = ElemCfg<NPE, NQ>;
  double acc[C::NBLK > 0 ? C::NBLK : 1][2];
  double pr[C::NB][C::REM > 0 ? C::REM : 1];  // leftover rows: K[ib*8+g][8*NBF+r] on lanes t4 == r after the reduction
  double Fe;
};

// Phase A of one element by one warp: FEMap::compute_map at quadrature point q = lane and the forcing (fw = JxW f).  X holds the element's node coordinates, fw is a 32-double scratch.
// Returns has_affine_map; S = sqrt(det J) * inverse-Jacobian rows (the same on every lane when affine).
template <int NPE, int NQ>
__device__ __forceinline__ bool elem_geom(const double* X, double* fw, const double* s_tab, int lane, int forcing,
                                          const double* __restrict__ f_tab_e, int* flags, double (&S)[9], double& Fe_out) {
  const double* s_phi = s_tab;
  // FEASM_ABLATE (never defined in product builds; scripts/gpu_ablate.py): 2 = no contraction, 3 = no
  // stores, 4 = no geometry / forcing — timing probes of the phases.
  double JxW, x, y, z;
#if defined(FEASM_ABLATE) && FEASM_ABLATE == 4
  const bool affine = true;
  for (int k = 0; k < 9; ++k) S[k] = X[k] + k;
  JxW = 1.; x = y = z = 0.;
  if (false)
#else
  const bool affine = geom_phase<NPE, NQ>(X, s_tab, lane, forcing == FEASM_FORCING_EX4_FD, flags, S, JxW, x, y, z);
#endif
  {
    double f = 0.;
    if (forcing == FEASM_FORCING_EX4_FD) {
      // every lane evaluates (pad lanes repeat qp 0): no divergence; huge coordinates take the library path
      const bool big = !(fabs(x) + fabs(y) + fabs(z) < 2.e4);
      f = __any_sync(0xffffffffu, big) ? ex4_forcing_lib(x, y, z) : ex4_forcing_t<true>(x, y, z);
    } else if (forcing == FEASM_FORCING_TABULATED && lane < NQ) {
      f = f_tab_e[lane];
    }
    fw[lane] = lane < NQ ? JxW * f : 0.;
  }
  Fe_out = 0.;
  return affine;
}

// Fe(a) on lane a from fw (introduction_ex4.C:536-537); three independent chains
template <int NPE, int NQ>
__device__ __forceinline__ double elem_fe(const double* fw, const double* s_tab, int lane, int forcing) {
  const double* s_phi = s_tab;
  double Fe_out = 0.;
  if (lane < NPE && forcing != FEASM_FORCING_NONE) {
    double Fe = 0., f1 = 0., f2 = 0.;
#pragma unroll
    for (int q = 0; q + 2 < NQ; q += 3) {
      Fe += fw[q] * s_phi[lane * NQ + q];
      f1 += fw[q + 1] * s_phi[lane * NQ + q + 1];
      f2 += fw[q + 2] * s_phi[lane * NQ + q + 2];
    }
#pragma unroll
    for (int q = NQ - NQ % 3; q < NQ; ++q) Fe += fw[q] * s_phi[lane * NQ + q];
    Fe_out = (Fe + f1) + f2;
  }
  return Fe_out;
}

// Phases B+C: Ke = G G^T, G[i][(d,q)] = sqrt(JxW_q) dphi_i/dx_d(q), fragments generated in registers from S and the
// scaled reference-gradient table (twp = s_tw + (lane>>2)*TS + (lane&3)).
template <int NPE, int NQ>
__device__ __forceinline__ void elem_contract(const double (&S)[9], const bool affine, const double* twp, int lane, ElemRegs<NPE, NQ>& R) {
  using C = ElemCfg<NPE, NQ>;
  const int t4 = lane & 3;
#pragma unroll
  for (int b = 0; b < C::NBLK; ++b) { R.acc[b][0] = 0.; R.acc[b][1] = 0.; }
#pragma unroll
  for (int ib = 0; ib < C::NB; ++ib)
#pragma unroll
    for (int r = 0; r < (C::REM > 0 ? C::REM : 1); ++r) R.pr[ib][r] = 0.;
#if defined(FEASM_ABLATE) && FEASM_ABLATE == 2
  if (lane == 99)
#endif
#pragma unroll
  for (int qg = 0; qg < C::QG; ++qg) {
    double Sq[9];
    if (affine) {
#pragma unroll
      for (int k = 0; k < 9; ++k) Sq[k] = S[k];
    } else {
#pragma unroll
      for (int k = 0; k < 9; ++k) Sq[k] = __shfl_sync(0xffffffffu, S[k], 4 * qg + t4);
    }
    double G[C::NB][3];
#pragma unroll
    for (int ib = 0; ib < C::NB; ++ib) {
      const double a = twp[ib * 8 * C::TS + 4 * qg];
      const double b = twp[C::NB * 8 * C::TS + ib * 8 * C::TS + 4 * qg];
      const double c = twp[2 * C::NB * 8 * C::TS + ib * 8 * C::TS + 4 * qg];
      // h1_fe_transformation.C:150-187: dphi/dx_d = dphi/dxi dxi/dx_d + dphi/deta deta/dx_d + dphi/dzeta dzeta/dx_d
      G[ib][0] = a * Sq[0] + b * Sq[3] + c * Sq[6];
      G[ib][1] = a * Sq[1] + b * Sq[4] + c * Sq[7];
      G[ib][2] = a * Sq[2] + b * Sq[5] + c * Sq[8];
    }
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      int blk = 0;
#pragma unroll
      for (int jb = 0; jb < C::NBF; ++jb)
#pragma unroll
        for (int ib = 0; ib <= jb; ++ib) { dmma884(R.acc[blk][0], R.acc[blk][1], G[ib][d], G[jb][d]); ++blk; }
      if (C::REM > 0) {
#pragma unroll
        for (int r = 0; r < C::REM; ++r) {
          const double gr = __shfl_sync(0xffffffffu, G[C::NB - 1][d], 4 * r + t4);
#pragma unroll
          for (int ib = 0; ib < C::NB; ++ib) R.pr[ib][r] += G[ib][d] * gr;
        }
      }
    }
  }
  if (C::REM > 0) {
#pragma unroll
    for (int ib = 0; ib < C::NB; ++ib)
#pragma unroll
      for (int r = 0; r < C::REM; ++r) {
        R.pr[ib][r] += __shfl_xor_sync(0xffffffffu, R.pr[ib][r], 1);
        R.pr[ib][r] += __shfl_xor_sync(0xffffffffu, R.pr[ib][r], 2);
      }
  }
}

template <int NPE, int NQ>
__device__ __forceinline__ void elem_compute(const double* X, double* fw, const double* s_tab, const double* twp, int lane,
                                             int forcing, const double* __restrict__ f_tab_e, int* flags, ElemRegs<NPE, NQ>& R) {
  double S[9];
  const bool affine = elem_geom<NPE, NQ>(X, fw, s_tab, lane, forcing, f_tab_e, flags, S, R.Fe);
#ifdef FEASM_DEV_FE_EARLY
  __syncwarp();
  R.Fe = elem_fe<NPE, NQ>(fw, s_tab, lane, forcing);
  elem_contract<NPE, NQ>(S, affine, twp, lane, R);
#else
  elem_contract<NPE, NQ>(S, affine, twp, lane, R);  // its shuffles order the fw stores before the reads below
  __syncwarp();
  R.Fe = elem_fe<NPE, NQ>(fw, s_tab, lane, forcing);
#endif
}

// Phase D: the warp's fragments into the element's block of NPE records, rs doubles apart (row a = Ke(a,:) then Fe(a)
// at column NPE) — in global memory, or in a shared-memory tile when the element-level constraint follows.
template <int NPE, int NQ>
__device__ __forceinline__ void elem_store(const ElemRegs<NPE, NQ>& R, double* __restrict__ out, const int rs, int lane) {
  using C = ElemCfg<NPE, NQ>;
  const int g = lane >> 2, t4 = lane & 3;
#if defined(FEASM_ABLATE) && FEASM_ABLATE == 3
  if (rs != 99) return;
#endif
  if (lane < NPE) out[lane * rs + NPE] = R.Fe;
  {
    int blk = 0;
#pragma unroll
    for (int jb = 0; jb < C::NBF; ++jb)
#pragma unroll
      for (int ib = 0; ib <= jb; ++ib, ++blk) {
        const int i = ib * 8 + g, j = jb * 8 + 2 * t4;
        *reinterpret_cast<double2*>(out + i * rs + j) = make_double2(R.acc[blk][0], R.acc[blk][1]);
        if (ib != jb) {  // mirror into the lower triangle
          out[j * rs + i] = R.acc[blk][0];
          out[(j + 1) * rs + i] = R.acc[blk][1];
        }
      }
  }
  if (C::REM > 0) {
    constexpr int R0 = 8 * C::NBF;
#pragma unroll
    for (int ib = 0; ib < C::NBF; ++ib) {
      const int i = ib * 8 + g;
#pragma unroll
      for (int r = 0; r < C::REM; ++r)
        if (t4 == r) { out[i * rs + R0 + r] = R.pr[ib][r]; out[(R0 + r) * rs + i] = R.pr[ib][r]; }
    }
    if (g < C::REM) {
#pragma unroll
      for (int r = 0; r < C::REM; ++r)
        if (t4 == r) out[(R0 + g) * rs + R0 + r] = R.pr[C::NB - 1][r];
    }
  }
}

template <int NPE, int NQ>
__device__ __forceinline__ void elem_process(const double* X, double* fw, const double* s_tab, const double* twp, int lane,
                                             int forcing, const double* __restrict__ f_tab_e, int* flags, double* __restrict__ out,
                                             const int rs) {
  ElemRegs<NPE, NQ> R;
  elem_compute<NPE, NQ>(X, fw, s_tab, twp, lane, forcing, f_tab_e, flags, R);
  elem_store<NPE, NQ>(R, out, rs, lane);
}

// Reference tables into shared memory: s_tab = phi | dxi | deta | dzeta | qw ([i][q]: lane = q reads are conflict free),
// s_tw[c][row][TS] = sqrt(w_q) * d phi_row / d xi_c (q), zero padded (fragment reads are conflict free).
template <int NPE, int NQ>
__device__ __forceinline__ void elem_load_tables(double* s_tab, double* s_tw, const double* __restrict__ tab, int tid, int nthreads) {
  using C = ElemCfg<NPE, NQ>;
  for (int t = tid; t < C::TAB; t += nthreads) s_tab[t] = tab[t];
  for (int t = tid; t < C::TW_SZ; t += nthreads) {
    const int c = t / (C::NB * 8 * C::TS), row = (t / C::TS) % (C::NB * 8), q = t % C::TS;
    s_tw[t] = (row < NPE && q < NQ) ? sqrt(tab[4 * NPE * NQ + q]) * tab[(1 + c) * NPE * NQ + row * NQ + q] : 0.;
  }
}
// The same constraint for one variable per node (nd = NPE <= 32) on a shared-memory tile T[NPE][LDR] (row a = Ke(a,:), Fe(a) at
// column NPE), with the element's constraint data already in registers: lane a holds "dof a is constrained" (bit a of
// `mask`) and its value g_a (0 when homogeneous).  Same operations in the same order as constrain_element below.
template <int NPE, int LDR>
__device__ __forceinline__ void constrain_tile(double* T, const unsigned mask, const double g_lane, const int lane) {
  constexpr unsigned FULL = 0xffffffffu;
  __syncwarp();  // the tile was written by other lanes
  const int row = lane < NPE ? lane : 0;
  double fe = T[row * LDR + NPE];
  for (unsigned m = mask; m; m &= m - 1) {  // constrained columns: K(i,j) -> 0, F(i) -= K(i,j) g_j
    const int j = __ffs(m) - 1;
    const double g = __shfl_sync(FULL, g_lane, j);
    const double kij = T[row * LDR + j];
    if (lane < NPE) T[row * LDR + j] = 0.;
    fe -= kij * g;
  }
  __syncwarp();
  for (unsigned m = mask; m; m &= m - 1) {  // constrained rows: K(i,:) = e_i
    const int i = __ffs(m) - 1;
    if (lane < NPE) T[i * LDR + lane] = lane == i ? 1. : 0.;
  }
  if (lane < NPE) T[lane * LDR + NPE] = ((mask >> lane) & 1u) ? g_lane : fe;  // F(i) = g_i on constrained rows
  __syncwarp();
}

// Dirichlet constraints at element level, in place on the element's records right after they were written (called by
// the whole warp for the few elements that touch a constrained dof).  This is
// DofMap::heterogeneously_constrain_element_matrix_and_vector (dof_map_constraints.C:2518-2614) /
// constrain_element_matrix_and_vector (:2034-2093) for C = diag(0 on constrained dofs, 1 else), H = g:
//   unconstrained row i:  F(i) -= sum_{j constrained} K(i,j) g_j (heterogeneous only), K(i,j) = 0
//   constrained row i:    K(i,:) = e_i, F(i) = g_i (heterogeneous) or 0
__device__ __noinline__ void constrain_element(double* K, const int rs, const int nd, const uint32_t* __restrict__ dofs,
                                               const uint8_t* __restrict__ con_flag, const double* __restrict__ con_val,
                                               const int mode, const int lane) {
  constexpr unsigned FULL = 0xffffffffu;
  __syncwarp();  // the rows were written by other lanes
  const bool hetero = mode == FEASM_CONSTRAIN_HETEROGENEOUS;
  for (int j0 = 0; j0 < nd; j0 += 32) {  // constrained columns j0..j0+31
    const int jm = j0 + lane;
    uint32_t dj = 0; int cj = 0; double gj = 0.;
    if (jm < nd) { dj = dofs[jm]; cj = con_flag[dj]; if (cj && hetero) gj = con_val[dj]; }
    uint32_t mask = __ballot_sync(FULL, cj);
    while (mask) {
      const int jl = __ffs(mask) - 1, j = j0 + jl;
      mask &= mask - 1;
      const double g = __shfl_sync(FULL, gj, jl);
      for (int i = lane; i < nd; i += 32) {  // column j of every row (row j itself is overwritten below)
        const double kij = K[i * rs + j];
        K[i * rs + j] = 0.;
        if (hetero) K[i * rs + nd] -= kij * g;
      }
    }
  }
  __syncwarp();
  for (int i0 = 0; i0 < nd; i0 += 32) {  // constrained rows
    const int im = i0 + lane;
    uint32_t di = 0; int ci = 0; double gi = 0.;
    if (im < nd) { di = dofs[im]; ci = con_flag[di]; if (ci && hetero) gi = con_val[di]; }
    uint32_t mask = __ballot_sync(FULL, ci);
    while (mask) {
      const int il = __ffs(mask) - 1, i = i0 + il;
      mask &= mask - 1;
      const double g = __shfl_sync(FULL, gi, il);
      for (int t = lane; t <= nd; t += 32) K[i * rs + t] = t == i ? 1. : (t == nd ? g : 0.);
    }
  }
  __syncwarp();
}
}  // namespace

#endif
