// ref_shim.cpp — C entry points over the REFERENCE's own shape-function (and quadrature) sources, compiled where they lie
// under /root/reference into oracle/_ref/libref_fe.so by oracle/Makefile (target _ref; stub headers: oracle/ref_stubs/).
// Test infrastructure: the tests use it to check the oracle's restatement against the reference code itself; nothing under
// libmesh_b200/ may touch it.  Element types and orders are libMesh's enum values (include/enums/enum_elem_type.h,
// enum_order.h).
#include "libmesh/fe.h"
#include "libmesh/fe_map.h"
#include "libmesh/fe_interface.h"
#include "libmesh/enum_to_string.h"
#include "libmesh/quadrature_gauss.h"

using namespace libMesh;

// QBase::clone() names the rule factory (src/quadrature/quadrature_build.C, which would pull in every rule class of the
// library); nothing here clones a rule.
std::unique_ptr<QBase> QBase::build(const QuadratureType, const unsigned int, const Order) { libmesh_not_implemented(); }
// FEMap's side maps live in src/fe/fe_boundary.C together with explicit instantiations of FE<Dim, T>::reinit(elem, side) for
// every family of the library — not compilable as a whole.  The DEFINITIONS of FEMap::init_face_shape_functions<Dim> and
// FEMap::compute_face_map are cut out of it at build time (oracle/ref_excerpt.py, set "fe"), together with
// FEAbstract::get_refspace_nodes (src/fe/fe_abstract.C) and the side_nodes_map tables of the element classes (src/geom/cell_*.C),
// and compiled here against the stub declarations of ref_stubs/.  Edge maps are not on the path.
namespace libMesh {
#include "ref_fe_excerpts.inc"
template void FEMap::init_face_shape_functions<3>(const std::vector<Point>&, const Elem*);
}  // namespace libMesh
void FEMap::compute_edge_map(int, const std::vector<Real>&, const Elem*) { libmesh_not_implemented(); }

// examples/introduction/introduction_ex3/exact_solution.C (the manufactured solution of ex3 / ex4), compiled where it lies
Real exact_solution(const Real x, const Real y, const Real z);

extern "C" {

double ref_exact_solution(double x, double y, double z) { return exact_solution(x, y, z); }

// phi[i][p], dxi[i][p], deta[i][p], dzeta[i][p] of the LAGRANGE functions i < n_shape of `order` on `elem_type` at np
// reference points (x, y, z): FE<dim, LAGRANGE>::shape / shape_deriv (src/fe/fe_lagrange_shape_{2,3}D.C).  Returns 0, or -1
// with a message in err when the reference code throws.
int ref_lagrange_shapes(int dim, int elem_type, int order, int n_shape, int np, const double* pts, double* phi, double* dxi,
                        double* deta, double* dzeta, char* err, int err_len) {
  try {
    const Elem elem(static_cast<ElemType>(elem_type), 0);
    const Order o = static_cast<Order>(order);
    for (int i = 0; i < n_shape; ++i)
      for (int p = 0; p < np; ++p) {
        const Point pt(pts[3 * p], pts[3 * p + 1], pts[3 * p + 2]);
        const size_t k = (size_t)i * np + p;
        if (dim == 3) {
          if (phi) phi[k] = FE<3, LAGRANGE>::shape(&elem, o, i, pt, false);
          if (dxi) dxi[k] = FE<3, LAGRANGE>::shape_deriv(&elem, o, i, 0, pt, false);
          if (deta) deta[k] = FE<3, LAGRANGE>::shape_deriv(&elem, o, i, 1, pt, false);
          if (dzeta) dzeta[k] = FE<3, LAGRANGE>::shape_deriv(&elem, o, i, 2, pt, false);
        } else {
          if (phi) phi[k] = FE<2, LAGRANGE>::shape(&elem, o, i, pt, false);
          if (dxi) dxi[k] = FE<2, LAGRANGE>::shape_deriv(&elem, o, i, 0, pt, false);
          if (deta) deta[k] = FE<2, LAGRANGE>::shape_deriv(&elem, o, i, 1, pt, false);
        }
      }
    return 0;
  } catch (const std::exception& e) {
    if (err && err_len > 0) {
      int n = 0;
      for (const char* c = e.what(); *c && n < err_len - 1; ++c) err[n++] = *c;
      err[n] = 0;
    }
    return -1;
  }
}

// QGauss(dim, order).init(elem_type): the reference's Gauss rule for that element type (src/quadrature/quadrature_gauss_{1,2,3}D.C,
// quadrature.C tensor products, quadrature_gauss.C Dunavant / Keast tables, quadrature_conical*.C).  Returns the number of
// points (the first `cap` are copied out), or -1 when the reference code throws.
int ref_gauss_rule(int dim, int elem_type, int order, int cap, double* pts, double* w, char* err, int err_len) {
  try {
    QGauss q(dim, static_cast<Order>(order));
    q.init(static_cast<ElemType>(elem_type), 0, /*simple_type_only=*/true);
    const int n = (int)q.n_points();
    for (int p = 0; p < n && p < cap; ++p) {
      if (pts) { pts[3 * p] = q.qp(p)(0); pts[3 * p + 1] = q.qp(p)(1); pts[3 * p + 2] = q.qp(p)(2); }
      if (w) w[p] = q.w(p);
    }
    return n;
  } catch (const std::exception& e) {
    if (err && err_len > 0) {
      int n = 0;
      for (const char* c = e.what(); *c && n < err_len - 1; ++c) err[n++] = *c;
      err[n] = 0;
    }
    return -1;
  }
}

// Elem::has_affine_map() of the element class (Hex8 / Hex20 / Hex27 / Tet10 / Prism18::has_affine_map, src/geom/cell_*.C, the
// reference's own definitions): 1 / 0, or -1 for a type without one here
int ref_has_affine_map(int elem_type, int n_nodes, const double* node_xyz) {
  const ElemType t = static_cast<ElemType>(elem_type);
  if (t != HEX8 && t != HEX20 && t != HEX27 && t != TET10 && t != PRISM18) return -1;
  Elem elem(t, 0);
  elem.set_nodes(node_xyz, (unsigned int)n_nodes);
  return elem.has_affine_map() ? 1 : 0;
}

// FEMap::init_reference_to_physical_map<dim> + FEMap::compute_map (src/fe/fe_map.C:109-444, :1323-1372 with
// compute_single_point_map :448-1077 or compute_affine_map :1167-1244) for one element given by its node coordinates — a 3D
// cell (dim 3) or a 2D element embedded in space (dim 2: the surface-Jacobian branch sqrt(g11 g22 - g12 g21), :810-860) — at
// the nq reference points `pts` with weights `qw`.  `affine` plays Elem::has_affine_map() (the element classes cannot be
// compiled stand-alone; the caller forces either branch).  Out: xyz[nq][3], JxW[nq], jac[nq] and
// inv[nq][9] = dxidx, dxidy, dxidz, detadx, detady, detadz, dzetadx, dzetady, dzetadz.  Returns 0, or -1 when the reference
// code throws (DegenerateMap: the message goes to err).
int ref_fe_map(int dim, int elem_type, int n_nodes, const double* node_xyz, int affine, int nq, const double* pts, const double* qw,
               double* xyz, double* JxW, double* jac, double* inv, char* err, int err_len) {
  try {
    Elem elem(static_cast<ElemType>(elem_type), 0);
    elem.set_nodes(node_xyz, (unsigned int)n_nodes);
    if (affine >= 0) elem.force_affine(affine != 0);  // < 0: the element class's own has_affine_map decides
    std::vector<Point> qp;
    std::vector<Real> w(qw, qw + nq);
    for (int p = 0; p < nq; ++p) qp.emplace_back(pts[3 * p], pts[3 * p + 1], pts[3 * p + 2]);
    FEMap map;
    const std::vector<Point>& m_xyz = map.get_xyz();  // the getters switch the calculations on (fe_map.h:216-380)
    const std::vector<Real>& m_JxW = map.get_JxW();
    const std::vector<Real>& m_jac = map.get_jacobian();
    const std::vector<Real>* m_inv[9] = {&map.get_dxidx(),  &map.get_dxidy(),  &map.get_dxidz(),  &map.get_detadx(),  &map.get_detady(),
                                         &map.get_detadz(), &map.get_dzetadx(), &map.get_dzetady(), &map.get_dzetadz()};
    if (dim == 3) map.init_reference_to_physical_map<3>(qp, &elem);
    else map.init_reference_to_physical_map<2>(qp, &elem);
    map.compute_map((unsigned int)dim, w, &elem, /*calculate_d2phi=*/false);
    for (int p = 0; p < nq; ++p) {
      if (xyz) { xyz[3 * p] = m_xyz[p](0); xyz[3 * p + 1] = m_xyz[p](1); xyz[3 * p + 2] = m_xyz[p](2); }
      if (JxW) JxW[p] = m_JxW[p];
      if (jac) jac[p] = (std::size_t)p < m_jac.size() ? m_jac[p] : 0.;
      if (inv) for (int k = 0; k < 9; ++k) inv[9 * p + k] = (std::size_t)p < m_inv[k]->size() ? (*m_inv[k])[p] : 0.;  // (a 2D map sizes no zeta entries)
    }
    return 0;
  } catch (const std::exception& e) {
    if (err && err_len > 0) {
      int n = 0;
      for (const char* c = e.what(); *c && n < err_len - 1; ++c) err[n++] = *c;
      err[n] = 0;
    }
    return -1;
  }
}

}  // extern "C"
