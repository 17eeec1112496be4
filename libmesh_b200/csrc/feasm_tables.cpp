// feasm_tables.cpp — reference-element quadrature and Lagrange shape tables (host, built once per element type).
//
// Product code.  Covers what libMesh's QGauss (src/quadrature/quadrature_gauss_1D.C:28-70,
// quadrature.C:283-305, quadrature_gauss_3D.C:233-290) and FE<3,LAGRANGE> (src/fe/fe_lagrange_shape_3D.C:72-413,
// :961-1009, :2179-2330) produce for HEX8/Q1, HEX27/Q2 and TET10/P2.  The tables are generated from the
// reference-node coordinates of each element: a hex basis function is the product of the 1D Lagrange
// polynomials attached to the node's (xi,eta,zeta) in {-1,0,+1}; a TET10 function is built from barycentric
// coordinates and the vertex pair of each mid-edge node.
#include <cmath>

#include "feasm_internal.h"

namespace {

// Reference coordinates of the nodes in libMesh's element-local order (include/geom/cell_hex27.h diagram):
// 8 vertices, 12 mid-edge, 6 mid-face, 1 centre.  Values are -1, 0, +1.
const signed char kHexNode[27][3] = {
    {-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}, {-1, -1, 1}, {1, -1, 1}, {1, 1, 1},  {-1, 1, 1},  {0, -1, -1},
    {1, 0, -1},   {0, 1, -1},  {-1, 0, -1}, {-1, -1, 0}, {1, -1, 0},  {1, 1, 0},  {-1, 1, 0}, {0, -1, 1},  {1, 0, 1},
    {0, 1, 1},    {-1, 0, 1},  {0, 0, -1}, {0, -1, 0},  {1, 0, 0},   {0, 1, 0},  {-1, 0, 0}, {0, 0, 1},   {0, 0, 0}};

// 1D Lagrange polynomial of degree `deg` attached to reference abscissa c in {-1,0,1}; value and derivative.
struct V1 { double v, d; };
V1 lagrange1d(int deg, int c, double x) {
  if (deg == 1) return c < 0 ? V1{.5 * (1. - x), -.5} : V1{.5 * (1. + x), .5};
  if (c < 0) return {.5 * x * (x - 1.), x - .5};
  if (c > 0) return {.5 * x * (x + 1.), x + .5};
  return {1. - x * x, -2. * x};
}

int gauss1d(int qorder, double x[3], double w[3]) {
  if (qorder <= 1) { x[0] = 0.; w[0] = 2.; return 1; }
  if (qorder <= 3) {
    x[0] = -std::sqrt(3.0) / 3; x[1] = -x[0];
    w[0] = w[1] = 1.;
    return 2;
  }
  if (qorder <= 5) {
    x[0] = -7.7459666924148337703585307995648e-01; x[1] = 0.; x[2] = -x[0];
    w[0] = w[2] = 5.0 / 9; w[1] = 8.0 / 9;
    return 3;
  }
  return 0;
}

// Sides of a hex in libMesh's side order (0 = zeta-min, 1 = eta-min, 2 = xi-max, 3 = eta-max, 4 = xi-min, 5 = zeta-max):
// parent-local node of side node i, for i in the QUAD4/QUAD9 node order (4 corners, 4 mid-edges, centre).
// src/geom/cell_hex27.C:37-45; the first four columns are cell_hex8.C:42-50.
const unsigned char kHexSide[6][9] = {{0, 3, 2, 1, 11, 10, 9, 8, 20},  {0, 1, 5, 4, 8, 13, 16, 12, 21}, {1, 2, 6, 5, 9, 14, 17, 13, 22},
                                      {2, 3, 7, 6, 10, 15, 18, 14, 23}, {3, 0, 4, 7, 11, 12, 19, 15, 24}, {4, 5, 6, 7, 16, 17, 18, 19, 25}};
// Reference coordinates of the QUAD9 nodes (QUAD4 = the first four)
const signed char kQuadNode[9][2] = {{-1, -1}, {1, -1}, {1, 1}, {-1, 1}, {0, -1}, {1, 0}, {0, 1}, {-1, 0}, {0, 0}};

// TET10: vertex pair of each mid-edge node 4..9
const int kTetEdge[6][2] = {{0, 1}, {1, 2}, {0, 2}, {0, 3}, {1, 3}, {2, 3}};

}  // namespace

bool feasm_make_tables(int elem_type, int qorder, FeasmTables& t, std::string& err) {
  t = FeasmTables();
  t.npe = elem_type;
  if (elem_type == FEASM_HEX8 || elem_type == FEASM_HEX27 || elem_type == FEASM_HEX20) {
    double x[3], w[3];
    const int n1 = gauss1d(qorder, x, w);
    if (!n1) { err = "unsupported Gauss order for hex"; return false; }
    t.nq = n1 * n1 * n1;
    t.qp.resize(3 * t.nq);
    t.qw.resize(t.nq);
    for (int q = 0; q < t.nq; ++q) {  // q = i + n1*(j + n1*k), i fastest
      const int i = q % n1, j = (q / n1) % n1, k = q / (n1 * n1);
      t.qp[3 * q] = x[i]; t.qp[3 * q + 1] = x[j]; t.qp[3 * q + 2] = x[k];
      t.qw[q] = w[i] * w[j] * w[k];
    }
    const int deg = elem_type == FEASM_HEX8 ? 1 : 2;
    const size_t sz = (size_t)t.npe * t.nq;
    t.phi.resize(sz); t.dxi.resize(sz); t.deta.resize(sz); t.dzeta.resize(sz);
    if (elem_type == FEASM_HEX20) {
      // The 20-node serendipity basis (Lagrange SECOND on HEX20, fe_lagrange_shape_3D.C:848-925 / :1876-2090), in the
      // closed form over the node's reference coordinates (x0, y0, z0) in {-1, 0, 1}:
      //   corner:    (1 + x0 xi)(1 + y0 eta)(1 + z0 zeta)(x0 xi + y0 eta + z0 zeta - 2) / 8
      //   mid-edge:  (1 - c_m^2)(1 + a0 a)(1 + b0 b) / 4      (m the axis with coordinate 0, a and b the other two)
      for (int a = 0; a < 20; ++a)
        for (int q = 0; q < t.nq; ++q) {
          const double c[3] = {t.qp[3 * q], t.qp[3 * q + 1], t.qp[3 * q + 2]};
          const double n0[3] = {(double)kHexNode[a][0], (double)kHexNode[a][1], (double)kHexNode[a][2]};
          const size_t o = (size_t)a * t.nq + q;
          double* g[3] = {&t.dxi[o], &t.deta[o], &t.dzeta[o]};
          if (a < 8) {
            const double f[3] = {1. + n0[0] * c[0], 1. + n0[1] * c[1], 1. + n0[2] * c[2]};
            const double L = n0[0] * c[0] + n0[1] * c[1] + n0[2] * c[2] - 2.;
            t.phi[o] = .125 * f[0] * f[1] * f[2] * L;
            for (int d = 0; d < 3; ++d) {
              const double others = f[(d + 1) % 3] * f[(d + 2) % 3];
              *g[d] = .125 * n0[d] * others * (L + f[d]);
            }
          } else {
            const int m = n0[0] == 0. ? 0 : (n0[1] == 0. ? 1 : 2), u = (m + 1) % 3, v = (m + 2) % 3;
            const double bub = 1. - c[m] * c[m], fu = 1. + n0[u] * c[u], fv = 1. + n0[v] * c[v];
            t.phi[o] = .25 * bub * fu * fv;
            *g[m] = -.5 * c[m] * fu * fv;
            *g[u] = .25 * bub * n0[u] * fv;
            *g[v] = .25 * bub * fu * n0[v];
          }
        }
      return true;
    }
    for (int a = 0; a < t.npe; ++a)
      for (int q = 0; q < t.nq; ++q) {
        const V1 X = lagrange1d(deg, kHexNode[a][0], t.qp[3 * q]);
        const V1 Y = lagrange1d(deg, kHexNode[a][1], t.qp[3 * q + 1]);
        const V1 Z = lagrange1d(deg, kHexNode[a][2], t.qp[3 * q + 2]);
        const size_t o = (size_t)a * t.nq + q;
        t.phi[o] = X.v * Y.v * Z.v;
        t.dxi[o] = X.d * Y.v * Z.v;
        t.deta[o] = X.v * Y.d * Z.v;
        t.dzeta[o] = X.v * Y.v * Z.d;
      }
    return true;
  }
  if (elem_type == FEASM_TET10) {
    if (qorder != 5 && qorder != 4) { err = "TET10: only the FIFTH-order Gauss rule is built"; return false; }
    // 14-point degree-5 rule (Walkington): two vertex-symmetric orbits of 4 points and one edge orbit of 6.
    const double a[3] = {0.31088591926330060980, 0.092735250310891226402, 0.045503704125649649492};
    const double wt[3] = {0.018781320953002641800, 0.012248840519393658257, 0.0070910034628469110730};
    t.nq = 14;
    t.qp.assign(42, 0.);
    t.qw.assign(14, 0.);
    int q = 0;
    for (int o = 0; o < 2; ++o) {
      const double b = 1. - 3. * a[o];
      const double P[4][3] = {{a[o], a[o], a[o]}, {a[o], b, a[o]}, {b, a[o], a[o]}, {a[o], a[o], b}};
      for (auto& p : P) { t.qp[3 * q] = p[0]; t.qp[3 * q + 1] = p[1]; t.qp[3 * q + 2] = p[2]; t.qw[q++] = wt[o]; }
    }
    {
      const double c = a[2], b = 0.5 * (1. - 2. * c);
      const double P[6][3] = {{b, b, c}, {b, c, c}, {c, c, b}, {c, b, c}, {b, c, b}, {c, b, b}};
      for (auto& p : P) { t.qp[3 * q] = p[0]; t.qp[3 * q + 1] = p[1]; t.qp[3 * q + 2] = p[2]; t.qw[q++] = wt[2]; }
    }
    const size_t sz = (size_t)10 * t.nq;
    t.phi.resize(sz); t.dxi.resize(sz); t.deta.resize(sz); t.dzeta.resize(sz);
    // barycentric gradients wrt (xi, eta, zeta): L0 = 1-xi-eta-zeta, L1 = xi, L2 = eta, L3 = zeta
    const double dL[4][3] = {{-1, -1, -1}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int p = 0; p < t.nq; ++p) {
      const double L[4] = {1. - t.qp[3 * p] - t.qp[3 * p + 1] - t.qp[3 * p + 2], t.qp[3 * p], t.qp[3 * p + 1], t.qp[3 * p + 2]};
      double* out[3] = {t.dxi.data(), t.deta.data(), t.dzeta.data()};
      for (int v = 0; v < 4; ++v) {
        t.phi[(size_t)v * t.nq + p] = L[v] * (2. * L[v] - 1.);
        for (int d = 0; d < 3; ++d) out[d][(size_t)v * t.nq + p] = (4. * L[v] - 1.) * dL[v][d];
      }
      for (int e = 0; e < 6; ++e) {
        const int u = kTetEdge[e][0], v = kTetEdge[e][1];
        t.phi[(size_t)(4 + e) * t.nq + p] = 4. * L[u] * L[v];
        for (int d = 0; d < 3; ++d) out[d][(size_t)(4 + e) * t.nq + p] = 4. * (L[u] * dL[v][d] + dL[u][d] * L[v]);
      }
    }
    return true;
  }
  if (elem_type == FEASM_PRISM18) {
    // QGauss(3, FIFTH) on a prism (quadrature_gauss_3D.C:507-527): the 7-point degree-5 triangle rule (centroid and two
    // orbits (a,a), (a,1-2a), (1-2a,a)) times the 3-point Gauss rule, the 1D rule outermost (quadrature.C:310-332).
    if (qorder != 5 && qorder != 4) { err = "PRISM18: only the FIFTH-order Gauss rule is built"; return false; }
    double gx[3], gw[3];
    gauss1d(5, gx, gw);
    const double r15 = std::sqrt(15.);
    const double ta[2] = {2. / 7 + r15 / 21, 2. / 7 - r15 / 21}, tw[2] = {31. / 480 + r15 / 2400, 31. / 480 - r15 / 2400};
    double tp[7][2], tww[7];
    tp[0][0] = tp[0][1] = 1. / 3; tww[0] = 9. / 80;
    for (int o = 0; o < 2; ++o) {
      const double P[3][2] = {{ta[o], ta[o]}, {ta[o], 1 - 2 * ta[o]}, {1 - 2 * ta[o], ta[o]}};
      for (int i = 0; i < 3; ++i) { tp[1 + 3 * o + i][0] = P[i][0]; tp[1 + 3 * o + i][1] = P[i][1]; tww[1 + 3 * o + i] = tw[o]; }
    }
    t.nq = 21;
    t.qp.resize(63); t.qw.resize(21);
    for (int j = 0; j < 3; ++j)
      for (int i = 0; i < 7; ++i) {
        const int q = 7 * j + i;
        t.qp[3 * q] = tp[i][0]; t.qp[3 * q + 1] = tp[i][1]; t.qp[3 * q + 2] = gx[j];
        t.qw[q] = tww[i] * gw[j];
      }
    // PRISM18 = TRI6 x EDGE3 (fe_lagrange_shape_3D.C:1077-1094, :2478-2510), written from where the node sits: in the
    // triangle a vertex v (L_v (2 L_v - 1)) or the middle of the edge (u, v) (4 L_u L_v), along the axis at -1, +1 or 0.
    // Node order of cell_prism18.h: bottom vertices, top vertices, bottom edges, vertical edges, top edges, mid layer.
    static const signed char kTri[18][2] = {{0, -1}, {1, -1}, {2, -1}, {0, -1}, {1, -1}, {2, -1}, {0, 1}, {1, 2}, {0, 2}, {0, -1}, {1, -1}, {2, -1},
                                            {0, 1}, {1, 2}, {0, 2}, {0, 1}, {1, 2}, {0, 2}};
    static const signed char kAx[18] = {-1, -1, -1, 1, 1, 1, -1, -1, -1, 0, 0, 0, 1, 1, 1, 0, 0, 0};
    const double dL[3][2] = {{-1, -1}, {1, 0}, {0, 1}};  // barycentric gradients wrt (xi, eta): L0 = 1 - xi - eta, L1 = xi, L2 = eta
    const size_t sz = (size_t)18 * t.nq;
    t.phi.resize(sz); t.dxi.resize(sz); t.deta.resize(sz); t.dzeta.resize(sz);
    for (int a = 0; a < 18; ++a)
      for (int q = 0; q < t.nq; ++q) {
        const double L[3] = {1. - t.qp[3 * q] - t.qp[3 * q + 1], t.qp[3 * q], t.qp[3 * q + 1]};
        double tv, tdx, tde;
        const int u = kTri[a][0], v = kTri[a][1];
        if (v < 0) { tv = L[u] * (2. * L[u] - 1.); tdx = (4. * L[u] - 1.) * dL[u][0]; tde = (4. * L[u] - 1.) * dL[u][1]; }
        else { tv = 4. * L[u] * L[v]; tdx = 4. * (L[u] * dL[v][0] + dL[u][0] * L[v]); tde = 4. * (L[u] * dL[v][1] + dL[u][1] * L[v]); }
        const V1 Z = lagrange1d(2, kAx[a], t.qp[3 * q + 2]);
        const size_t o = (size_t)a * t.nq + q;
        t.phi[o] = tv * Z.v; t.dxi[o] = tdx * Z.v; t.deta[o] = tde * Z.v; t.dzeta[o] = tv * Z.d;
      }
    return true;
  }
  err = "unknown element type";
  return false;
}

// Side tables of a hex: the side element's Gauss rule (tensor rule, i fastest: quadrature.C:256-277), its Lagrange map
// functions and their derivatives at those points (FEMap::init_face_shape_functions, fe_boundary.C:458-660), and for
// each of the six sides the parent's shape functions at the side points carried into the parent's reference space
// (FE::side_map, fe_boundary.C:347-400: sum over the side nodes of psi_i times the reference node).
// TET10: TRI6 sides (Tet10::side_nodes_map, cell_tet10.C:37-43), the degree-5 triangle rule of QGauss(2, FIFTH)
// (quadrature_gauss_2D.C:234-257 via dunavant_rule2: centroid, then the orbits (a,a), (a,1-2a), (1-2a,a)), the TRI6 map
// functions (fe_lagrange_shape_2D.C:499-530, :920-975) and the parent's shapes at the side points carried into the
// reference tet (reference nodes fe_abstract.C:488-501).  Tables are laid out for six sides (sides 4, 5 stay zero) so
// the kernels address them like a hex's.
// Value of the HEX20 serendipity function of node a at a reference point (the closed form of feasm_make_tables)
static double hex20_value(int a, const double c[3]) {
  const double n0[3] = {(double)kHexNode[a][0], (double)kHexNode[a][1], (double)kHexNode[a][2]};
  if (a < 8) return .125 * (1. + n0[0] * c[0]) * (1. + n0[1] * c[1]) * (1. + n0[2] * c[2]) * (n0[0] * c[0] + n0[1] * c[1] + n0[2] * c[2] - 2.);
  const int m = n0[0] == 0. ? 0 : (n0[1] == 0. ? 1 : 2), u = (m + 1) % 3, v = (m + 2) % 3;
  return .25 * (1. - c[m] * c[m]) * (1. + n0[u] * c[u]) * (1. + n0[v] * c[v]);
}
// QUAD8, the side of a HEX20 (fe_lagrange_shape_2D.C:434-471, :795-879), in the closed form over the node's coordinates:
//   corner (x0, y0):  (1 + x0 xi)(1 + y0 eta)(x0 xi + y0 eta - 1) / 4;   mid-side: (1 - xi^2)(1 + y0 eta) / 2 or (1 + x0 xi)(1 - eta^2) / 2
static void quad8(int a, double xi, double eta, double& v, double& dx, double& de) {
  const double x0 = kQuadNode[a][0], y0 = kQuadNode[a][1];
  if (a < 4) {
    const double fx = 1. + x0 * xi, fy = 1. + y0 * eta, L = x0 * xi + y0 * eta - 1.;
    v = .25 * fx * fy * L;
    dx = .25 * x0 * fy * (L + fx);
    de = .25 * y0 * fx * (L + fy);
  } else if (x0 == 0.) {
    v = .5 * (1. - xi * xi) * (1. + y0 * eta); dx = -xi * (1. + y0 * eta); de = .5 * (1. - xi * xi) * y0;
  } else {
    v = .5 * (1. + x0 * xi) * (1. - eta * eta); dx = .5 * x0 * (1. - eta * eta); de = -(1. + x0 * xi) * eta;
  }
}

static bool make_tet10_side_tables(int qorder, FeasmSideTables& t, std::string& err) {
  if (qorder != 5) { err = "TET10 sides: only the FIFTH-order rule is built"; return false; }
  static const int kSide[4][6] = {{0, 2, 1, 6, 5, 4}, {0, 1, 3, 4, 8, 7}, {1, 2, 3, 5, 9, 8}, {2, 0, 3, 6, 7, 9}};
  static const double kRefNode[10][3] = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {.5, 0, 0}, {.5, .5, 0}, {0, .5, 0}, {0, 0, .5}, {.5, 0, .5}, {0, .5, .5}};
  t.npe = 10; t.ns = 6; t.nqf = 7;
  const double wts[3] = {9. / 80, 31. / 480 + std::sqrt(15.) / 2400, 31. / 480 - std::sqrt(15.) / 2400};
  const double aa[3] = {0., 2. / 7 + std::sqrt(15.) / 21, 2. / 7 - std::sqrt(15.) / 21};
  double pts[7][2];
  t.qw.resize(7);
  int q = 0;
  pts[0][0] = 1. / 3; pts[0][1] = 1. / 3; t.qw[q++] = wts[0];
  for (int o = 1; o < 3; ++o) {
    pts[q][0] = aa[o]; pts[q][1] = aa[o]; t.qw[q++] = wts[o];
    pts[q][0] = aa[o]; pts[q][1] = 1 - 2 * aa[o]; t.qw[q++] = wts[o];
    pts[q][0] = 1 - 2 * aa[o]; pts[q][1] = aa[o]; t.qw[q++] = wts[o];
  }
  t.psi.resize(42); t.dxi.resize(42); t.deta.resize(42);
  for (int p = 0; p < 7; ++p) {
    const double z1 = pts[p][0], z2 = pts[p][1], z0 = 1. - z1 - z2;
    const double z[3] = {z0, z1, z2}, dx[3] = {-1., 1., 0.}, de[3] = {-1., 0., 1.};
    for (int v = 0; v < 3; ++v) {
      t.psi[(size_t)v * 7 + p] = 2. * z[v] * (z[v] - 0.5);
      t.dxi[(size_t)v * 7 + p] = (4. * z[v] - 1.) * dx[v];
      t.deta[(size_t)v * 7 + p] = (4. * z[v] - 1.) * de[v];
    }
    const int eu[3] = {0, 1, 2}, ev[3] = {1, 2, 0};  // mid-side nodes 3, 4, 5: 4 z0 z1, 4 z1 z2, 4 z2 z0
    for (int e = 0; e < 3; ++e) {
      const int u = eu[e], v = ev[e];
      t.psi[(size_t)(3 + e) * 7 + p] = 4. * z[u] * z[v];
      t.dxi[(size_t)(3 + e) * 7 + p] = 4. * z[v] * dx[u] + 4. * z[u] * dx[v];
      t.deta[(size_t)(3 + e) * 7 + p] = 4. * z[v] * de[u] + 4. * z[u] * de[v];
    }
  }
  t.side_nodes.assign(6 * t.ns, 0);
  t.phi.assign((size_t)6 * t.npe * t.nqf, 0.);
  for (int s = 0; s < 4; ++s) {
    for (int a = 0; a < 6; ++a) t.side_nodes[s * t.ns + a] = kSide[s][a];
    for (int p = 0; p < 7; ++p) {
      double r[3] = {0., 0., 0.};
      for (int a = 0; a < 6; ++a)
        for (int d = 0; d < 3; ++d) r[d] += t.psi[(size_t)a * 7 + p] * kRefNode[kSide[s][a]][d];
      const double L[4] = {1. - r[0] - r[1] - r[2], r[0], r[1], r[2]};
      for (int v = 0; v < 4; ++v) t.phi[((size_t)s * 10 + v) * 7 + p] = L[v] * (2. * L[v] - 1.);
      for (int e = 0; e < 6; ++e) t.phi[((size_t)s * 10 + 4 + e) * 7 + p] = 4. * L[kTetEdge[e][0]] * L[kTetEdge[e][1]];
    }
  }
  return true;
}

bool feasm_make_side_tables(int elem_type, int qorder, FeasmSideTables& t, std::string& err) {
  t = FeasmSideTables();
  if (elem_type == FEASM_TET10) return make_tet10_side_tables(qorder, t, err);
  if (elem_type != FEASM_HEX8 && elem_type != FEASM_HEX27 && elem_type != FEASM_HEX20) { err = "side terms: HEX8 / HEX20 / HEX27 / TET10 sides only"; return false; }
  const int deg = elem_type == FEASM_HEX8 ? 1 : 2;
  const bool ser = elem_type == FEASM_HEX20;  // QUAD8 sides: the first eight QUAD9 nodes (cell_hex20.C:37-45 = the first eight columns of cell_hex27.C:37-45)
  t.npe = elem_type;
  t.ns = deg == 1 ? 4 : (ser ? 8 : 9);
  double x[3], w[3];
  const int n1 = gauss1d(qorder, x, w);
  if (!n1) { err = "unsupported Gauss order for a side"; return false; }
  t.nqf = n1 * n1;
  t.qw.resize(t.nqf);
  t.psi.resize((size_t)t.ns * t.nqf); t.dxi.resize(t.psi.size()); t.deta.resize(t.psi.size());
  for (int p = 0; p < t.nqf; ++p) {
    const int i = p % n1, j = p / n1;
    t.qw[p] = w[i] * w[j];
    for (int a = 0; a < t.ns; ++a) {
      if (ser) {
        quad8(a, x[i], x[j], t.psi[(size_t)a * t.nqf + p], t.dxi[(size_t)a * t.nqf + p], t.deta[(size_t)a * t.nqf + p]);
        continue;
      }
      const V1 X = lagrange1d(deg, kQuadNode[a][0], x[i]), Y = lagrange1d(deg, kQuadNode[a][1], x[j]);
      t.psi[(size_t)a * t.nqf + p] = X.v * Y.v;
      t.dxi[(size_t)a * t.nqf + p] = X.d * Y.v;
      t.deta[(size_t)a * t.nqf + p] = X.v * Y.d;
    }
  }
  t.side_nodes.resize(6 * t.ns);
  t.phi.assign((size_t)6 * t.npe * t.nqf, 0.);
  for (int s = 0; s < 6; ++s) {
    for (int a = 0; a < t.ns; ++a) t.side_nodes[s * t.ns + a] = kHexSide[s][a];
    for (int p = 0; p < t.nqf; ++p) {
      double r[3] = {0., 0., 0.};
      for (int a = 0; a < t.ns; ++a)
        for (int d = 0; d < 3; ++d) r[d] += t.psi[(size_t)a * t.nqf + p] * (double)kHexNode[kHexSide[s][a]][d];
      for (int a = 0; a < t.npe; ++a)
        t.phi[((size_t)s * t.npe + a) * t.nqf + p] =
            ser ? hex20_value(a, r)
                : lagrange1d(deg, kHexNode[a][0], r[0]).v * lagrange1d(deg, kHexNode[a][1], r[1]).v * lagrange1d(deg, kHexNode[a][2], r[2]).v;
    }
  }
  return true;
}
