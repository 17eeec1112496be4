// orc_fe.cpp — quadrature rules, Lagrange shape tables and the reference->physical map of the CPU oracle.
// TEST INFRASTRUCTURE (see fe_oracle.h).  Restates, in libMesh's own operation order:
//   src/quadrature/quadrature_gauss_1D.C:28-70, src/quadrature/quadrature.C:283-305,
//   src/quadrature/quadrature_gauss_3D.C:233-290, include/fe/fe_lagrange_shape_1D.h:31-167,
//   src/fe/fe_lagrange_shape_3D.C:72-413 (hex), :961-1009 and :2179-2330 (TET10),
//   src/fe/fe_map.C:958-1077, :1167-1244, :1323-1372, src/fe/h1_fe_transformation.C:150-187,
//   src/geom/cell_hex8.C:121-135, cell_hex27.C:125-166, cell_tet10.C:143-170.
#include <cmath>
#include <cstring>

#include "orc_internal.h"

namespace orc {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
const std::string& get_error() { return g_err; }

// ---- 1D Gauss rules (quadrature_gauss_1D.C:28-70) --------------------------------------------
static int gauss_1d(int qorder, double* x, double* w) {
  switch (qorder) {
    case 0:
    case 1:
      x[0] = 0.;
      w[0] = 2.;
      return 1;
    case 2:
    case 3:
      x[0] = -std::sqrt(3.0) / 3;
      x[1] = -x[0];
      w[0] = 1.;
      w[1] = w[0];
      return 2;
    case 4:
    case 5:
      x[0] = -7.7459666924148337703585307995648e-01;
      x[1] = 0.;
      x[2] = -x[0];
      w[0] = 5.0 / 9;
      w[1] = 8.0 / 9;
      w[2] = w[0];
      return 3;
    default:
      return 0;
  }
}

static int tri_rule_fifth(double* pts /*7*2*/, double* w /*7*/);
static void tri6_shape(int i, double xi, double eta, double& v, double& dxi, double& deta);

bool gauss_rule(int elem_type, int qorder, QRule& q) {
  if (elem_type == ORC_HEX8 || elem_type == ORC_HEX27 || elem_type == ORC_HEX20) {
    double x[3], w[3];
    const int np = gauss_1d(qorder, x, w);
    if (!np) { set_error("unsupported hex quadrature order"); return false; }
    q.nq = np * np * np;
    q.pts.resize(3 * q.nq);
    q.w.resize(q.nq);
    int p = 0;  // QBase::tensor_product_hex, quadrature.C:283-305
    for (int k = 0; k < np; k++)
      for (int j = 0; j < np; j++)
        for (int i = 0; i < np; i++) {
          q.pts[3 * p + 0] = x[i];
          q.pts[3 * p + 1] = x[j];
          q.pts[3 * p + 2] = x[k];
          q.w[p] = w[i] * w[j] * w[k];
          p++;
        }
    return true;
  }
  if (elem_type == ORC_TET10) {
    if (qorder != 5) { set_error("tet quadrature: only FIFTH restated"); return false; }
    // Walkington 14-point rule, quadrature_gauss_3D.C:233-290
    q.nq = 14;
    q.pts.assign(42, 0.);
    q.w.assign(14, 0.);
    const double a[3] = {0.31088591926330060980, 0.092735250310891226402, 0.045503704125649649492};
    const double wt[3] = {0.018781320953002641800, 0.012248840519393658257, 0.0070910034628469110730};
    auto set = [&](int p, double x, double y, double z) {
      q.pts[3 * p] = x; q.pts[3 * p + 1] = y; q.pts[3 * p + 2] = z;
    };
    for (int i = 0; i < 2; ++i) {
      const int off = 4 * i;
      const double b = 1. - 3. * a[i];
      set(off + 0, a[i], a[i], a[i]);
      set(off + 1, a[i], b, a[i]);
      set(off + 2, b, a[i], a[i]);
      set(off + 3, a[i], a[i], b);
      for (int j = 0; j < 4; ++j) q.w[off + j] = wt[i];
    }
    {
      const int off = 8;
      const double b = 0.5 * (1. - 2. * a[2]);
      set(off + 0, b, b, a[2]);
      set(off + 1, b, a[2], a[2]);
      set(off + 2, a[2], a[2], b);
      set(off + 3, a[2], b, a[2]);
      set(off + 4, b, a[2], b);
      set(off + 5, a[2], b, b);
      for (int j = 0; j < 6; ++j) q.w[off + j] = wt[2];
    }
    return true;
  }
  if (elem_type == ORC_PRISM18) {
    // quadrature_gauss_3D.C:507-527: QGauss(1, order) x QGauss(2, order) on TRI3 through QBase::tensor_product_prism
    // (quadrature.C:310-332: the 1D rule outermost, point = (tri.x, tri.y, x_1D), weight = w_2D * w_1D)
    if (qorder != 5) { set_error("prism quadrature: only FIFTH restated"); return false; }
    double x[3], w[3], tp[14], tw[7];
    const int n1 = gauss_1d(qorder, x, w);
    const int n2 = tri_rule_fifth(tp, tw);
    q.nq = n1 * n2;
    q.pts.resize(3 * q.nq);
    q.w.resize(q.nq);
    int p = 0;
    for (int j = 0; j < n1; j++)
      for (int i = 0; i < n2; i++) {
        q.pts[3 * p + 0] = tp[2 * i];
        q.pts[3 * p + 1] = tp[2 * i + 1];
        q.pts[3 * p + 2] = x[j];
        q.w[p] = tw[i] * w[j];
        p++;
      }
    return true;
  }
  set_error("unknown element type");
  return false;
}

// ---- 1D Lagrange (fe_lagrange_shape_1D.h) ----------------------------------------------------
static inline double lin(int i, double xi) { return i == 0 ? .5 * (1. - xi) : .5 * (1. + xi); }
static inline double dlin(int i) { return i == 0 ? -.5 : .5; }
static inline double quad(int i, double xi) {
  switch (i) {
    case 0: return .5 * xi * (xi - 1.);
    case 1: return .5 * xi * (xi + 1);
    default: return (1. - xi * xi);
  }
}
static inline double dquad(int i, double xi) {
  switch (i) {
    case 0: return xi - .5;
    case 1: return xi + .5;
    default: return -2. * xi;
  }
}

// fe_lagrange_shape_3D.C:109-111 and :163-165
static const int H8_I0[8] = {0, 1, 1, 0, 0, 1, 1, 0};
static const int H8_I1[8] = {0, 0, 1, 1, 0, 0, 1, 1};
static const int H8_I2[8] = {0, 0, 0, 0, 1, 1, 1, 1};
static const int H27_I0[27] = {0, 1, 1, 0, 0, 1, 1, 0, 2, 1, 2, 0, 0, 1, 1, 0, 2, 1, 2, 0, 2, 2, 1, 2, 0, 2, 2};
static const int H27_I1[27] = {0, 0, 1, 1, 0, 0, 1, 1, 0, 2, 1, 2, 0, 0, 1, 1, 0, 2, 1, 2, 2, 0, 2, 1, 2, 2, 2};
static const int H27_I2[27] = {0, 0, 0, 0, 1, 1, 1, 1, 0, 0, 0, 0, 2, 2, 2, 2, 1, 1, 1, 1, 0, 2, 2, 2, 2, 1, 2};

void shape_table(int elem_type, int np, const double* pts, ShapeTable& t) {
  const int n = elem_type;
  t.n = n;
  t.np = np;
  t.phi.assign((size_t)n * np, 0.);
  t.dxi.assign((size_t)n * np, 0.);
  t.deta.assign((size_t)n * np, 0.);
  t.dzeta.assign((size_t)n * np, 0.);
  for (int p = 0; p < np; ++p) {
    const double xi = pts[3 * p], eta = pts[3 * p + 1], zeta = pts[3 * p + 2];
    if (elem_type == ORC_HEX8) {
      const double s[3][2] = {{lin(0, xi), lin(1, xi)}, {lin(0, eta), lin(1, eta)}, {lin(0, zeta), lin(1, zeta)}};
      const double d[3][2] = {{dlin(0), dlin(1)}, {dlin(0), dlin(1)}, {dlin(0), dlin(1)}};
      for (int i = 0; i < 8; ++i) {
        t.phi[i * np + p] = s[0][H8_I0[i]] * s[1][H8_I1[i]] * s[2][H8_I2[i]];
        t.dxi[i * np + p] = d[0][H8_I0[i]] * s[1][H8_I1[i]] * s[2][H8_I2[i]];
        t.deta[i * np + p] = s[0][H8_I0[i]] * d[1][H8_I1[i]] * s[2][H8_I2[i]];
        t.dzeta[i * np + p] = s[0][H8_I0[i]] * s[1][H8_I1[i]] * d[2][H8_I2[i]];
      }
    } else if (elem_type == ORC_HEX27) {
      const double s[3][3] = {{quad(0, xi), quad(1, xi), quad(2, xi)},
                              {quad(0, eta), quad(1, eta), quad(2, eta)},
                              {quad(0, zeta), quad(1, zeta), quad(2, zeta)}};
      const double d[3][3] = {{dquad(0, xi), dquad(1, xi), dquad(2, xi)},
                              {dquad(0, eta), dquad(1, eta), dquad(2, eta)},
                              {dquad(0, zeta), dquad(1, zeta), dquad(2, zeta)}};
      for (int i = 0; i < 27; ++i) {
        t.phi[i * np + p] = s[0][H27_I0[i]] * s[1][H27_I1[i]] * s[2][H27_I2[i]];
        t.dxi[i * np + p] = d[0][H27_I0[i]] * s[1][H27_I1[i]] * s[2][H27_I2[i]];
        t.deta[i * np + p] = s[0][H27_I0[i]] * d[1][H27_I1[i]] * s[2][H27_I2[i]];
        t.dzeta[i * np + p] = s[0][H27_I0[i]] * s[1][H27_I1[i]] * d[2][H27_I2[i]];
      }
    } else if (elem_type == ORC_HEX20) {
      // Serendipity functions, fe_lagrange_shape_3D.C:848-925 (values) and :1876-2090 (derivatives), written there as 20
      // cases in x = (xi + 1)/2 etc.  Restated by node class, with the reference's order of operations:
      //   corner with end bits (b0,b1,b2):  f0 f1 f2 L,   f_d = b_d ? c_d : 1 - c_d,   L = sum_d (+-2 c_d) + (1 - 2 popcount),
      //     summed starting from the first axis whose bit is set (the constant first for node 0);
      //   mid-edge node with free axis m:   4 x the factors in axis order, the free axis giving c_m (1 - c_m).
      const double c[3] = {.5 * (xi + 1.), .5 * (eta + 1.), .5 * (zeta + 1.)};
      for (int i = 0; i < 20; ++i) {
        const int b[3] = {H27_I0[i], H27_I1[i], H27_I2[i]};  // 0 / 1: the end of that axis, 2: the middle (nodes 8-19)
        double* out[4] = {&t.phi[i * np + p], &t.dxi[i * np + p], &t.deta[i * np + p], &t.dzeta[i * np + p]};
        if (i < 8) {
          double f[3];
          for (int d = 0; d < 3; ++d) f[d] = b[d] ? c[d] : (1. - c[d]);
          const int first = b[0] ? 0 : (b[1] ? 1 : (b[2] ? 2 : -1));
          const int pop = b[0] + b[1] + b[2];
          double L;
          if (first < 0) {
            L = 1. - 2. * c[0] - 2. * c[1] - 2. * c[2];
          } else {
            L = 2. * c[first];
            for (int d = 0; d < 3; ++d)
              if (d != first) L = b[d] ? L + 2. * c[d] : L - 2. * c[d];
            L = L - (2. * pop - 1.);
          }
          *out[0] = f[0] * f[1] * f[2] * L;
          for (int d = 0; d < 3; ++d) {
            const int o1 = d == 0 ? 1 : 0, o2 = d == 2 ? 1 : 2;
            const double s2 = b[d] ? 2. : -2., s1 = b[d] ? 1. : -1.;
            *out[1 + d] = .5 * f[o1] * f[o2] * (f[d] * s2 + s1 * L);
          }
        } else {
          const int m = b[0] == 2 ? 0 : (b[1] == 2 ? 1 : 2);
          double v = 4.;
          for (int d = 0; d < 3; ++d) v = d == m ? v * c[d] * (1. - c[d]) : v * (b[d] ? c[d] : (1. - c[d]));
          *out[0] = v;
          for (int d = 0; d < 3; ++d) {
            double g = 2.;
            for (int a = 0; a < 3; ++a) {
              if (a == d) continue;
              g = a == m ? g * c[a] * (1. - c[a]) : g * (b[a] ? c[a] : (1. - c[a]));
            }
            if (d == m) g = g * (1. - 2. * c[d]);
            else if (!b[d]) g = g * (-1.);
            *out[1 + d] = g;
          }
        }
      }
    } else if (elem_type == ORC_PRISM18) {
      // fe_lagrange_shape_3D.C:1077-1094 (values), :2478-2510 (derivatives): TRI6 SECOND in (xi, eta) times the 1D quadratic
      // Lagrange function in zeta, with the reference's index tables
      static const int i0[18] = {0, 0, 0, 1, 1, 1, 0, 0, 0, 2, 2, 2, 1, 1, 1, 2, 2, 2};
      static const int i1[18] = {0, 1, 2, 0, 1, 2, 3, 4, 5, 0, 1, 2, 3, 4, 5, 3, 4, 5};
      for (int i = 0; i < 18; ++i) {
        double tv, tdx, tde;
        tri6_shape(i1[i], xi, eta, tv, tdx, tde);
        t.phi[i * np + p] = tv * quad(i0[i], zeta);
        t.dxi[i * np + p] = tdx * quad(i0[i], zeta);
        t.deta[i * np + p] = tde * quad(i0[i], zeta);
        t.dzeta[i * np + p] = tv * dquad(i0[i], zeta);
      }
    } else {  // TET10, fe_lagrange_shape_3D.C:961-1009, :2179-2330
      const double z1 = xi, z2 = eta, z3 = zeta;
      const double z0 = 1. - z1 - z2 - z3;
      double* ph = &t.phi[p];
      ph[0 * np] = z0 * (2. * z0 - 1.);
      ph[1 * np] = z1 * (2. * z1 - 1.);
      ph[2 * np] = z2 * (2. * z2 - 1.);
      ph[3 * np] = z3 * (2. * z3 - 1.);
      ph[4 * np] = 4. * z0 * z1;
      ph[5 * np] = 4. * z1 * z2;
      ph[6 * np] = 4. * z2 * z0;
      ph[7 * np] = 4. * z0 * z3;
      ph[8 * np] = 4. * z1 * z3;
      ph[9 * np] = 4. * z2 * z3;
      // d zeta_k / d(xi,eta,zeta)
      const double dz[3][4] = {{-1., 1., 0., 0.}, {-1., 0., 1., 0.}, {-1., 0., 0., 1.}};
      double* out[3] = {&t.dxi[p], &t.deta[p], &t.dzeta[p]};
      const double z[4] = {z0, z1, z2, z3};
      for (int j = 0; j < 3; ++j) {
        const double* g = dz[j];
        double* o = out[j];
        o[0 * np] = (4. * z[0] - 1.) * g[0];
        o[1 * np] = (4. * z[1] - 1.) * g[1];
        o[2 * np] = (4. * z[2] - 1.) * g[2];
        o[3 * np] = (4. * z[3] - 1.) * g[3];
        o[4 * np] = 4. * (z[0] * g[1] + g[0] * z[1]);
        o[5 * np] = 4. * (z[1] * g[2] + g[1] * z[2]);
        o[6 * np] = 4. * (z[0] * g[2] + g[0] * z[2]);
        o[7 * np] = 4. * (z[0] * g[3] + g[0] * z[3]);
        o[8 * np] = 4. * (z[1] * g[3] + g[1] * z[3]);
        o[9 * np] = 4. * (z[2] * g[3] + g[2] * z[3]);
      }
    }
  }
}

void FEData::resize(int n_, int np_) {
  n = n_;
  np = np_;
  xyz.assign((size_t)np * 3, 0.);
  JxW.assign(np, 0.);
  jinv.assign((size_t)np * 9, 0.);
  dphi.assign((size_t)n * np * 3, 0.);
}

// ---- affine tests ------------------------------------------------------------------------------
namespace {
struct P3 { double x, y, z; };
inline P3 pt(const double* X, int i) { return {X[3 * i], X[3 * i + 1], X[3 * i + 2]}; }
inline P3 sub(P3 a, P3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline P3 scale(P3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline P3 divide(P3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
inline double l1(P3 a) { return std::fabs(a.x) + std::fabs(a.y) + std::fabs(a.z); }
// include/utils/fuzzy_equals.h:96-112 with tol = affine_tol = TOLERANCE*TOLERANCE (include/geom/elem.h:2070)
inline bool rfe(P3 a, P3 b) {
  const double tol = 1e-6 * 1e-6;
  return l1(sub(a, b)) <= tol * (l1(a) + l1(b));
}
}  // namespace

bool has_affine_map(int type, const double* X) {
  auto P = [&](int i) { return pt(X, i); };
  if (type == ORC_HEX8) {  // cell_hex8.C:121-135
    P3 v = sub(P(1), P(0));
    if (!rfe(v, sub(P(2), P(3))) || !rfe(v, sub(P(5), P(4))) || !rfe(v, sub(P(6), P(7)))) return false;
    v = sub(P(4), P(0));
    if (!rfe(v, sub(P(7), P(3)))) return false;
    return true;
  }
  if (type == ORC_HEX27) {  // cell_hex27.C:125-166
    P3 v = sub(P(1), P(0));
    if (!rfe(v, sub(P(2), P(3))) || !rfe(v, sub(P(5), P(4))) || !rfe(v, sub(P(6), P(7)))) return false;
    v = divide(v, 2);
    if (!rfe(v, sub(P(8), P(0))) || !rfe(v, sub(P(10), P(3))) || !rfe(v, sub(P(16), P(4))) ||
        !rfe(v, sub(P(18), P(7))) || !rfe(v, sub(P(20), P(11))) || !rfe(v, sub(P(21), P(12))) ||
        !rfe(v, sub(P(23), P(15))) || !rfe(v, sub(P(25), P(19))) || !rfe(v, sub(P(26), P(24))))
      return false;
    v = sub(P(4), P(0));
    if (!rfe(v, sub(P(7), P(3)))) return false;
    v = divide(v, 2);
    if (!rfe(v, sub(P(12), P(0))) || !rfe(v, sub(P(13), P(1))) || !rfe(v, sub(P(14), P(2))) ||
        !rfe(v, sub(P(15), P(3))) || !rfe(v, sub(P(22), P(9))) || !rfe(v, sub(P(24), P(11))))
      return false;
    v = divide(sub(P(3), P(0)), 2);
    if (!rfe(v, sub(P(11), P(0))) || !rfe(v, sub(P(9), P(1))) || !rfe(v, sub(P(17), P(5))) ||
        !rfe(v, sub(P(19), P(4))))
      return false;
    return true;
  }
  if (type == ORC_HEX20) {  // cell_hex20.C:119-153
    P3 v = sub(P(1), P(0));
    if (!rfe(v, sub(P(2), P(3))) || !rfe(v, sub(P(5), P(4))) || !rfe(v, sub(P(6), P(7)))) return false;
    v = divide(v, 2);
    if (!rfe(v, sub(P(8), P(0))) || !rfe(v, sub(P(10), P(3))) || !rfe(v, sub(P(16), P(4))) || !rfe(v, sub(P(18), P(7)))) return false;
    v = sub(P(4), P(0));
    if (!rfe(v, sub(P(7), P(3)))) return false;
    v = divide(v, 2);
    if (!rfe(v, sub(P(12), P(0))) || !rfe(v, sub(P(13), P(1))) || !rfe(v, sub(P(14), P(2))) || !rfe(v, sub(P(15), P(3)))) return false;
    v = divide(sub(P(3), P(0)), 2);
    if (!rfe(v, sub(P(11), P(0))) || !rfe(v, sub(P(9), P(1))) || !rfe(v, sub(P(17), P(5))) || !rfe(v, sub(P(19), P(4)))) return false;
    return true;
  }
  if (type == ORC_PRISM18) {  // cell_prism18.C:122-151
    P3 v = sub(P(3), P(0));
    if (!rfe(v, sub(P(4), P(1))) || !rfe(v, sub(P(5), P(2)))) return false;
    v = divide(v, 2);
    if (!rfe(v, sub(P(9), P(0))) || !rfe(v, sub(P(10), P(1))) || !rfe(v, sub(P(11), P(2))) || !rfe(v, sub(P(15), P(6))) ||
        !rfe(v, sub(P(16), P(7))) || !rfe(v, sub(P(17), P(8))))
      return false;
    v = divide(sub(P(1), P(0)), 2);
    if (!rfe(v, sub(P(6), P(0))) || !rfe(v, sub(P(12), P(3)))) return false;
    v = divide(sub(P(2), P(0)), 2);
    if (!rfe(v, sub(P(8), P(0))) || !rfe(v, sub(P(14), P(3)))) return false;
    v = divide(sub(P(2), P(1)), 2);
    if (!rfe(v, sub(P(7), P(1))) || !rfe(v, sub(P(13), P(4)))) return false;
    return true;
  }
  // TET10, cell_tet10.C:143-170
  static const int e[6][3] = {{1, 0, 4}, {2, 1, 5}, {2, 0, 6}, {3, 0, 7}, {3, 1, 8}, {3, 2, 9}};
  for (auto& t : e) {
    P3 v = sub(P(t[0]), P(t[1]));
    if (!rfe(v, scale(sub(P(t[2]), P(t[1])), 2))) return false;
  }
  return true;
}

// ---- compute_single_point_map, 3D case (fe_map.C:958-1077) --------------------------------------
static bool single_point_map(int n, const double* X, const ShapeTable& t, const double* qw, int p, FEData& fe) {
  const int np = t.np;
  double x[3] = {0, 0, 0}, dxi[3] = {0, 0, 0}, deta[3] = {0, 0, 0}, dzeta[3] = {0, 0, 0};
  for (int i = 0; i < n; ++i) {
    const double* P = &X[3 * i];
    const double ph = t.phi[i * np + p], a = t.dxi[i * np + p], b = t.deta[i * np + p], c = t.dzeta[i * np + p];
    for (int d = 0; d < 3; ++d) {
      x[d] += ph * P[d];
      dxi[d] += a * P[d];
      deta[d] += b * P[d];
      dzeta[d] += c * P[d];
    }
  }
  for (int d = 0; d < 3; ++d) fe.xyz[3 * p + d] = x[d];
  const double dx_dxi = dxi[0], dy_dxi = dxi[1], dz_dxi = dxi[2];
  const double dx_deta = deta[0], dy_deta = deta[1], dz_deta = deta[2];
  const double dx_dzeta = dzeta[0], dy_dzeta = dzeta[1], dz_dzeta = dzeta[2];
  const double jac = (dx_dxi * (dy_deta * dz_dzeta - dz_deta * dy_dzeta) +
                      dy_dxi * (dz_deta * dx_dzeta - dx_deta * dz_dzeta) +
                      dz_dxi * (dx_deta * dy_dzeta - dy_deta * dx_dzeta));
  if (jac <= 0.) {  // jacobian_tolerance default 0 (fe_map.C:70, check :467)
    set_error("Jacobian at or under tolerance (degenerate map)");
    return false;
  }
  fe.JxW[p] = jac * qw[p];
  const double inv_jac = 1. / jac;
  double* J = &fe.jinv[9 * p];
  J[0] = (dy_deta * dz_dzeta - dz_deta * dy_dzeta) * inv_jac;
  J[1] = (dz_deta * dx_dzeta - dx_deta * dz_dzeta) * inv_jac;
  J[2] = (dx_deta * dy_dzeta - dy_deta * dx_dzeta) * inv_jac;
  J[3] = (dz_dxi * dy_dzeta - dy_dxi * dz_dzeta) * inv_jac;
  J[4] = (dx_dxi * dz_dzeta - dz_dxi * dx_dzeta) * inv_jac;
  J[5] = (dy_dxi * dx_dzeta - dx_dxi * dy_dzeta) * inv_jac;
  J[6] = (dy_dxi * dz_deta - dz_dxi * dy_deta) * inv_jac;
  J[7] = (dz_dxi * dx_deta - dx_dxi * dz_deta) * inv_jac;
  J[8] = (dx_dxi * dy_deta - dy_dxi * dx_deta) * inv_jac;
  return true;
}

int fe_reinit(int type, const double* X, const ShapeTable& t, const double* qw, FEData& fe) {
  const int n = t.n, np = t.np;
  if (fe.n != n || fe.np != np) fe.resize(n, np);
  fe.tab = &t;
  int affine = 0;
  if (has_affine_map(type, X)) {  // FEMap::compute_affine_map, fe_map.C:1167-1244
    affine = 1;
    if (!single_point_map(n, X, t, qw, 0, fe)) return -1;
    for (int p = 1; p < np; ++p) {
      double x[3] = {0, 0, 0};
      for (int i = 0; i < n; ++i)
        for (int d = 0; d < 3; ++d) x[d] += t.phi[i * np + p] * X[3 * i + d];
      for (int d = 0; d < 3; ++d) fe.xyz[3 * p + d] = x[d];
      std::memcpy(&fe.jinv[9 * p], &fe.jinv[0], 9 * sizeof(double));
      fe.JxW[p] = fe.JxW[0] / qw[0] * qw[p];
    }
  } else {  // FEMap::compute_map general branch, fe_map.C:1370-1371
    for (int p = 0; p < np; ++p)
      if (!single_point_map(n, X, t, qw, p, fe)) return -1;
  }
  // H1FETransformation::map_dphi, h1_fe_transformation.C:150-187
  for (int i = 0; i < n; ++i)
    for (int p = 0; p < np; ++p) {
      const double a = t.dxi[i * np + p], b = t.deta[i * np + p], c = t.dzeta[i * np + p];
      const double* J = &fe.jinv[9 * p];
      double* g = &fe.dphi[((size_t)i * np + p) * 3];
      g[0] = (a * J[0] + b * J[3] + c * J[6]);
      g[1] = (a * J[1] + b * J[4] + c * J[7]);
      g[2] = (a * J[2] + b * J[5] + c * J[8]);
    }
  return affine;
}

// ---- sides: FE<3,LAGRANGE>::reinit(elem, side) ------------------------------------------------------------------
// src/fe/fe_boundary.C:109-240: the side's own Gauss rule (QBase::tensor_product_quad, quadrature.C:256-277), the
// side's Lagrange map psi (init_face_shape_functions :458-660; QUAD4/QUAD9 tensor shapes fe_lagrange_shape_2D.C:355-356,
// :488-489), FEMap::compute_face_map 3D (:885-1017: JxW = sqrt(g11 g22 - g12 g21) w), FE::side_map (:347-400: parent
// reference points = sum_i psi_i refnode[local_side_node(s, i)]) and the parent's shapes at those points.
static const int H27_SIDE[6][9] = {{0, 3, 2, 1, 11, 10, 9, 8, 20},  {0, 1, 5, 4, 8, 13, 16, 12, 21},   // cell_hex27.C:37-45
                                   {1, 2, 6, 5, 9, 14, 17, 13, 22}, {2, 3, 7, 6, 10, 15, 18, 14, 23},
                                   {3, 0, 4, 7, 11, 12, 19, 15, 24}, {4, 5, 6, 7, 16, 17, 18, 19, 25}};
static const int H8_SIDE[6][4] = {{0, 3, 2, 1}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {3, 0, 4, 7}, {4, 5, 6, 7}};  // cell_hex8.C:42-50
static const int Q4_I0[4] = {0, 1, 1, 0}, Q4_I1[4] = {0, 0, 1, 1};
static const int Q9_I0[9] = {0, 1, 1, 0, 2, 1, 2, 0, 2}, Q9_I1[9] = {0, 0, 1, 1, 0, 2, 1, 2, 2};

// TET10: TRI6 sides (cell_tet10.C:37-43), the TRI6 Lagrange map (fe_lagrange_shape_2D.C:499-530, derivatives :920-975),
// the 7-point degree-5 triangle rule (quadrature_gauss_2D.C:234-257 through QGauss::dunavant_rule2,
// quadrature_gauss.C:195-265: centroid, then (a,a), (a,1-2a), (1-2a,a) per orbit), reference nodes fe_abstract.C:488-501.
static const int T10_SIDE[4][6] = {{0, 2, 1, 6, 5, 4}, {0, 1, 3, 4, 8, 7}, {1, 2, 3, 5, 9, 8}, {2, 0, 3, 6, 7, 9}};
static const double T10_REF[10][3] = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {.5, 0, 0}, {.5, .5, 0}, {0, .5, 0}, {0, 0, .5}, {.5, 0, .5}, {0, .5, .5}};

static int tri_rule_fifth(double* pts /*7*2*/, double* w /*7*/) {
  const double wts[3] = {9. / 80, 31. / 480 + std::sqrt(15.) / 2400, 31. / 480 - std::sqrt(15.) / 2400};
  const double a[3] = {0., 2. / 7 + std::sqrt(15.) / 21, 2. / 7 - std::sqrt(15.) / 21};
  int q = 0;
  pts[0] = 1. / 3; pts[1] = 1. / 3; w[q++] = wts[0];
  for (int p = 1; p < 3; ++p) {
    pts[2 * q] = a[p]; pts[2 * q + 1] = a[p]; w[q++] = wts[p];
    pts[2 * q] = a[p]; pts[2 * q + 1] = 1 - 2 * a[p]; w[q++] = wts[p];
    pts[2 * q] = 1 - 2 * a[p]; pts[2 * q + 1] = a[p]; w[q++] = wts[p];
  }
  return q;
}

static void tri6_shape(int i, double xi, double eta, double& v, double& dxi, double& deta) {
  const double z1 = xi, z2 = eta, z0 = 1. - z1 - z2;
  const double d0x = -1., d1x = 1., d2x = 0., d0e = -1., d1e = 0., d2e = 1.;
  switch (i) {
    case 0: v = 2. * z0 * (z0 - 0.5); dxi = (4. * z0 - 1.) * d0x; deta = (4. * z0 - 1.) * d0e; break;
    case 1: v = 2. * z1 * (z1 - 0.5); dxi = (4. * z1 - 1.) * d1x; deta = (4. * z1 - 1.) * d1e; break;
    case 2: v = 2. * z2 * (z2 - 0.5); dxi = (4. * z2 - 1.) * d2x; deta = (4. * z2 - 1.) * d2e; break;
    case 3: v = 4. * z0 * z1; dxi = 4. * z1 * d0x + 4. * z0 * d1x; deta = 4. * z1 * d0e + 4. * z0 * d1e; break;
    case 4: v = 4. * z1 * z2; dxi = 4. * z2 * d1x + 4. * z1 * d2x; deta = 4. * z2 * d1e + 4. * z1 * d2e; break;
    default: v = 4. * z2 * z0; dxi = 4. * z2 * d0x + 4 * z0 * d2x; deta = 4. * z2 * d0e + 4 * z0 * d2e; break;
  }
}

// QUAD8 (the side of a HEX20), fe_lagrange_shape_2D.C:434-471 (values) and :795-879 (derivatives), expression for expression
static void quad8_shape(int i, double xi, double eta, double& v, double& dxi, double& deta) {
  switch (i) {
    case 0: v = .25 * (1. - xi) * (1. - eta) * (-1. - xi - eta);
      dxi = .25 * (1. - eta) * ((1. - xi) * (-1.) + (-1.) * (-1. - xi - eta));
      deta = .25 * (1. - xi) * ((1. - eta) * (-1.) + (-1.) * (-1. - xi - eta)); break;
    case 1: v = .25 * (1. + xi) * (1. - eta) * (-1. + xi - eta);
      dxi = .25 * (1. - eta) * ((1. + xi) * (1.) + (1.) * (-1. + xi - eta));
      deta = .25 * (1. + xi) * ((1. - eta) * (-1.) + (-1.) * (-1. + xi - eta)); break;
    case 2: v = .25 * (1. + xi) * (1. + eta) * (-1. + xi + eta);
      dxi = .25 * (1. + eta) * ((1. + xi) * (1.) + (1.) * (-1. + xi + eta));
      deta = .25 * (1. + xi) * ((1. + eta) * (1.) + (1.) * (-1. + xi + eta)); break;
    case 3: v = .25 * (1. - xi) * (1. + eta) * (-1. - xi + eta);
      dxi = .25 * (1. + eta) * ((1. - xi) * (-1.) + (-1.) * (-1. - xi + eta));
      deta = .25 * (1. - xi) * ((1. + eta) * (1.) + (1.) * (-1. - xi + eta)); break;
    case 4: v = .5 * (1. - xi * xi) * (1. - eta); dxi = .5 * (-2. * xi) * (1. - eta); deta = .5 * (1. - xi * xi) * (-1.); break;
    case 5: v = .5 * (1. + xi) * (1. - eta * eta); dxi = .5 * (1.) * (1. - eta * eta); deta = .5 * (1. + xi) * (-2. * eta); break;
    case 6: v = .5 * (1. - xi * xi) * (1. + eta); dxi = .5 * (-2. * xi) * (1. + eta); deta = .5 * (1. - xi * xi) * (1.); break;
    default: v = .5 * (1. - xi) * (1. - eta * eta); dxi = .5 * (-1.) * (1. - eta * eta); deta = .5 * (1. - xi) * (-2. * eta); break;
  }
}

int side_nodes(int elem_type, int side, int* nodes) {
  if (side < 0 || side >= 6) return 0;
  if (elem_type == ORC_HEX27) { for (int i = 0; i < 9; ++i) nodes[i] = H27_SIDE[side][i]; return 9; }
  if (elem_type == ORC_HEX20) { for (int i = 0; i < 8; ++i) nodes[i] = H27_SIDE[side][i]; return 8; }  // cell_hex20.C:37-45
  if (elem_type == ORC_HEX8) { for (int i = 0; i < 4; ++i) nodes[i] = H8_SIDE[side][i]; return 4; }
  if (elem_type == ORC_TET10 && side < 4) { for (int i = 0; i < 6; ++i) nodes[i] = T10_SIDE[side][i]; return 6; }
  return 0;
}

int side_reinit(int elem_type, const double* X, int side, int qorder, SideData& sd) {
  int sn[9];
  const int ns = side_nodes(elem_type, side, sn);
  if (!ns) { set_error("side_reinit: HEX8/HEX20/HEX27/TET10 sides only"); return -1; }
  const bool q2 = ns == 9, tri = ns == 6, q8 = ns == 8;
  double gx[3], gw[3];
  const int np = tri ? 0 : gauss_1d(qorder, gx, gw);
  if (tri && qorder != 5 && qorder != 0) { set_error("TRI6 sides: QGauss FIFTH only"); return -1; }
  if (!tri && !np) { set_error("unsupported side quadrature order"); return -1; }
  const int nqf = tri ? 7 : np * np;
  sd.nqf = nqf;
  std::vector<double> pts(2 * nqf), qw(nqf);
  if (tri) {
    tri_rule_fifth(pts.data(), qw.data());
  } else {
    int q = 0;
    for (int j = 0; j < np; j++)
      for (int i = 0; i < np; i++) {
        pts[2 * q] = gx[i];
        pts[2 * q + 1] = gx[j];
        qw[q] = gw[i] * gw[j];
        q++;
      }
  }
  // psi_map, dpsidxi_map, dpsideta_map [i][p]
  std::vector<double> psi(ns * nqf), dxi(ns * nqf), deta(ns * nqf);
  for (int i = 0; i < ns; ++i)
    for (int p = 0; p < nqf; ++p) {
      const double xi = pts[2 * p], eta = pts[2 * p + 1];
      if (tri) {
        tri6_shape(i, xi, eta, psi[i * nqf + p], dxi[i * nqf + p], deta[i * nqf + p]);
      } else if (q8) {
        quad8_shape(i, xi, eta, psi[i * nqf + p], dxi[i * nqf + p], deta[i * nqf + p]);
      } else if (q2) {
        psi[i * nqf + p] = quad(Q9_I0[i], xi) * quad(Q9_I1[i], eta);
        dxi[i * nqf + p] = dquad(Q9_I0[i], xi) * quad(Q9_I1[i], eta);
        deta[i * nqf + p] = quad(Q9_I0[i], xi) * dquad(Q9_I1[i], eta);
      } else {
        psi[i * nqf + p] = lin(Q4_I0[i], xi) * lin(Q4_I1[i], eta);
        dxi[i * nqf + p] = dlin(Q4_I0[i]) * lin(Q4_I1[i], eta);
        deta[i * nqf + p] = lin(Q4_I0[i], xi) * dlin(Q4_I1[i]);
      }
    }
  // compute_face_map, 3D case
  sd.xyz.assign(3 * nqf, 0.);
  sd.JxW.assign(nqf, 0.);
  sd.normal.assign(3 * nqf, 0.);
  std::vector<double> dxyzdxi(3 * nqf, 0.), dxyzdeta(3 * nqf, 0.);
  for (int i = 0; i < ns; ++i) {
    const double* sp = &X[3 * sn[i]];
    for (int p = 0; p < nqf; ++p)
      for (int d = 0; d < 3; ++d) {
        sd.xyz[3 * p + d] += psi[i * nqf + p] * sp[d];
        dxyzdxi[3 * p + d] += dxi[i * nqf + p] * sp[d];
        dxyzdeta[3 * p + d] += deta[i * nqf + p] * sp[d];
      }
  }
  for (int p = 0; p < nqf; ++p) {
    const double* a = &dxyzdxi[3 * p];
    const double* b = &dxyzdeta[3 * p];
    const double n[3] = {a[1] * b[2] - a[2] * b[1], -a[0] * b[2] + a[2] * b[0], a[0] * b[1] - a[1] * b[0]};  // TypeVector::cross
    const double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    for (int d = 0; d < 3; ++d) sd.normal[3 * p + d] = n[d] / len;
    const double g11 = (a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
    const double g12 = (a[0] * b[0] + a[1] * b[1] + a[2] * b[2]);
    const double g21 = g12;
    const double g22 = (b[0] * b[0] + b[1] * b[1] + b[2] * b[2]);
    const double the_jac = std::sqrt(g11 * g22 - g12 * g21);
    if (!(the_jac > 0.)) { set_error("degenerate side map"); return -1; }
    sd.JxW[p] = the_jac * qw[p];
  }
  // side_map: the quadrature points in the parent's reference space (refspace nodes: fe_abstract.C get_refspace_nodes)
  static const double REF[3] = {-1., 1., 0.};
  sd.ref.assign(3 * nqf, 0.);
  for (int i = 0; i < ns; ++i) {
    const int a = sn[i];
    double node[3];
    if (tri) { node[0] = T10_REF[a][0]; node[1] = T10_REF[a][1]; node[2] = T10_REF[a][2]; }
    else { const bool h27 = q2 || q8; node[0] = REF[h27 ? H27_I0[a] : H8_I0[a]]; node[1] = REF[h27 ? H27_I1[a] : H8_I1[a]]; node[2] = REF[h27 ? H27_I2[a] : H8_I2[a]]; }
    for (int p = 0; p < nqf; ++p)
      for (int d = 0; d < 3; ++d) sd.ref[3 * p + d] += psi[i * nqf + p] * node[d];
  }
  // this->reinit(elem, &qp): the parent's shape functions at the mapped points
  ShapeTable t;
  shape_table(elem_type, nqf, sd.ref.data(), t);
  sd.phi = t.phi;
  // ... and its map: reinit(elem, &qp) runs FEMap::compute_map on the PARENT at the mapped points (fe.C:333-338), so what
  // get_xyz() returns after reinit(elem, side) is sum_i phi_i(qp) x_i over the parent's nodes (fe_map.C:994-1000, :1190-1198) —
  // the side map's own points above agree with it to rounding only
  sd.xyz.assign(3 * nqf, 0.);
  for (int i = 0; i < t.n; ++i)
    for (int p = 0; p < nqf; ++p)
      for (int d = 0; d < 3; ++d) sd.xyz[3 * p + d] += t.phi[i * nqf + p] * X[3 * i + d];
  return nqf;
}

// examples/introduction/introduction_ex3/exact_solution.C:40-47
double exact_solution(double x, double y, double z) {
  static const double pi = std::acos(-1.);
  return std::cos(.5 * pi * x) * std::sin(.5 * pi * y) * std::cos(.5 * pi * z);
}

// examples/introduction/introduction_ex4/introduction_ex4.C:500-537 (dim == 3)
double ex4_forcing(double x, double y, double z) {
  const double eps = 1.e-3;
  const double uxx = (exact_solution(x - eps, y, z) + exact_solution(x + eps, y, z) + -2. * exact_solution(x, y, z)) / eps / eps;
  const double uyy = (exact_solution(x, y - eps, z) + exact_solution(x, y + eps, z) + -2. * exact_solution(x, y, z)) / eps / eps;
  const double uzz = (exact_solution(x, y, z - eps) + exact_solution(x, y, z + eps) + -2. * exact_solution(x, y, z)) / eps / eps;
  return -(uxx + uyy + uzz);
}

}  // namespace orc

namespace orc { const std::string& get_error(); }

extern "C" {

const char* orc_last_error(void) { return orc::get_error().c_str(); }
double orc_exact_solution(double x, double y, double z) { return orc::exact_solution(x, y, z); }

int orc_quadrature(int elem_type, int qorder, int cap, double* points, double* weights) {
  orc::QRule q;
  if (!orc::gauss_rule(elem_type, qorder, q)) return -1;
  const int m = q.nq < cap ? q.nq : cap;
  if (points) std::memcpy(points, q.pts.data(), sizeof(double) * 3 * m);
  if (weights) std::memcpy(weights, q.w.data(), sizeof(double) * m);
  return q.nq;
}

int orc_shapes(int elem_type, int np, const double* ref_points, double* phi, double* dxi, double* deta, double* dzeta) {
  orc::ShapeTable t;
  orc::shape_table(elem_type, np, ref_points, t);
  const size_t b = sizeof(double) * t.n * np;
  if (phi) std::memcpy(phi, t.phi.data(), b);
  if (dxi) std::memcpy(dxi, t.dxi.data(), b);
  if (deta) std::memcpy(deta, t.deta.data(), b);
  if (dzeta) std::memcpy(dzeta, t.dzeta.data(), b);
  return t.n;
}

int orc_side_reinit(int elem_type, const double* node_xyz, int side, int qorder, int cap, double* ref_points, double* xyz,
                    double* JxW, double* normals, double* phi) {
  orc::SideData sd;
  const int nqf = orc::side_reinit(elem_type, node_xyz, side, qorder, sd);
  if (nqf < 0) return nqf;
  if (nqf > cap) return -2;
  if (ref_points) std::memcpy(ref_points, sd.ref.data(), sizeof(double) * 3 * nqf);
  if (xyz) std::memcpy(xyz, sd.xyz.data(), sizeof(double) * 3 * nqf);
  if (JxW) std::memcpy(JxW, sd.JxW.data(), sizeof(double) * nqf);
  if (normals) std::memcpy(normals, sd.normal.data(), sizeof(double) * 3 * nqf);
  if (phi) std::memcpy(phi, sd.phi.data(), sizeof(double) * elem_type * nqf);
  return nqf;
}

int orc_has_affine_map(int elem_type, const double* node_xyz) { return orc::has_affine_map(elem_type, node_xyz) ? 1 : 0; }

int orc_fe_reinit(int elem_type, const double* node_xyz, int np, const double* ref_points, const double* weights,
                  double* xyz, double* JxW, double* phi, double* dphi) {
  orc::ShapeTable t;
  orc::shape_table(elem_type, np, ref_points, t);
  orc::FEData fe;
  const int rc = orc::fe_reinit(elem_type, node_xyz, t, weights, fe);
  if (rc < 0) return rc;
  if (xyz) std::memcpy(xyz, fe.xyz.data(), sizeof(double) * 3 * np);
  if (JxW) std::memcpy(JxW, fe.JxW.data(), sizeof(double) * np);
  if (phi) std::memcpy(phi, t.phi.data(), sizeof(double) * t.n * np);
  if (dphi) std::memcpy(dphi, fe.dphi.data(), sizeof(double) * 3 * t.n * np);
  return rc;
}

}  // extern "C"
