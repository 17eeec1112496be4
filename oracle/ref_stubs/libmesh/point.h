// ref_stubs/libmesh/point.h — STUB (see libmesh_common.h in this directory): a three-component point with the accessors the
// reference's shape-function and quadrature sources use (include/geom/point.h, include/numerics/type_vector.h).
#ifndef REF_STUB_POINT_H
#define REF_STUB_POINT_H
#include "libmesh/libmesh_common.h"
namespace libMesh {
class Point {
public:
  Point(Real x = 0., Real y = 0., Real z = 0.) { c[0] = x; c[1] = y; c[2] = z; }
  Real operator()(unsigned int i) const { return c[i]; }
  Real& operator()(unsigned int i) { return c[i]; }
  Real& slice(unsigned int i) { return c[i]; }                                                 // TypeVector::slice (type_vector.h)
  Real slice(unsigned int i) const { return c[i]; }
  Point operator-() const { return Point(-c[0], -c[1], -c[2]); }
  Point operator-(const Point& o) const { return Point(c[0] - o.c[0], c[1] - o.c[1], c[2] - o.c[2]); }
  Point operator+(const Point& o) const { return Point(c[0] + o.c[0], c[1] + o.c[1], c[2] + o.c[2]); }
  Point operator*(Real s) const { return Point(c[0] * s, c[1] * s, c[2] * s); }
  Real operator*(const Point& o) const { return c[0] * o.c[0] + c[1] * o.c[1] + c[2] * o.c[2]; }
  void zero() { c[0] = c[1] = c[2] = 0.; }                                                     // TypeVector::zero
  void add(const Point& p) { c[0] += p.c[0]; c[1] += p.c[1]; c[2] += p.c[2]; }                 // TypeVector::add
  void add_scaled(const Point& p, Real f) { c[0] += f * p.c[0]; c[1] += f * p.c[1]; c[2] += f * p.c[2]; }  // TypeVector::add_scaled (type_vector.h:629-644)
  Real norm_sq() const { return c[0] * c[0] + c[1] * c[1] + c[2] * c[2]; }
  bool operator==(const Point& o) const { return c[0] == o.c[0] && c[1] == o.c[1] && c[2] == o.c[2]; }
  Point cross(const Point& o) const { return Point(c[1] * o.c[2] - c[2] * o.c[1], c[2] * o.c[0] - c[0] * o.c[2], c[0] * o.c[1] - c[1] * o.c[0]); }
  // TypeVector::l1_norm (type_vector.h:961-971) and relative_fuzzy_equals (include/utils/fuzzy_equals.h:95-112): |a - b|_1 <= tol (|a|_1 + |b|_1)
  Real l1_norm() const { Real ret{}; for (int i = 0; i < 3; ++i) ret += std::abs(c[i]); return ret; }
  bool absolute_fuzzy_equals(const Point& rhs, Real tol = TOLERANCE) const { return (*this - rhs).l1_norm() <= tol; }
  bool relative_fuzzy_equals(const Point& rhs, Real tol = TOLERANCE) const { return absolute_fuzzy_equals(rhs, tol * (l1_norm() + rhs.l1_norm())); }
  Point unit() const { const Real l = norm(); return Point(c[0] / l, c[1] / l, c[2] / l); }       // TypeVector::unit (type_vector.h)
  Real norm() const { return std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]); }
  Point& operator+=(const Point& o) { c[0] += o.c[0]; c[1] += o.c[1]; c[2] += o.c[2]; return *this; }
  Point& operator*=(Real s) { c[0] *= s; c[1] *= s; c[2] *= s; return *this; }
  Point& operator/=(Real s) { c[0] /= s; c[1] /= s; c[2] /= s; return *this; }
  Point operator/(Real s) const { return Point(c[0] / s, c[1] / s, c[2] / s); }
private:
  Real c[3];
};
inline Point operator*(Real s, const Point& p) { return p * s; }
inline Real triple_product(const Point& a, const Point& b, const Point& c) { return a * b.cross(c); }  // type_vector.h
inline std::ostream& operator<<(std::ostream& os, const Point& p) { return os << "(" << p(0) << ", " << p(1) << ", " << p(2) << ")"; }
// include/numerics/type_tensor.h, as far as the polyhedron case label of the shape sources names it (never executed here)
class RealTensor {
public:
  RealTensor(Real a = 0., Real b = 0., Real c = 0., Real d = 0., Real e = 0., Real f = 0., Real g = 0., Real h = 0., Real i = 0.) : m{a, b, c, d, e, f, g, h, i} {}
  Real& operator()(unsigned int r, unsigned int c) { return m[3 * r + c]; }
  Point operator*(const Point& v) const { return Point(m[0] * v(0) + m[1] * v(1) + m[2] * v(2), m[3] * v(0) + m[4] * v(1) + m[5] * v(2), m[6] * v(0) + m[7] * v(1) + m[8] * v(2)); }
  RealTensor operator-() const { return RealTensor(-m[0], -m[1], -m[2], -m[3], -m[4], -m[5], -m[6], -m[7], -m[8]); }
  void solve(const Point&, Point&) const { libmesh_not_implemented(); }  // only inverse_map (never called here) uses it
  Real operator()(unsigned int r, unsigned int c) const { return m[3 * r + c]; }
  Real det() const { return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) + m[2] * (m[3] * m[7] - m[4] * m[6]); }
private:
  Real m[9];
};
typedef RealTensor RealTensorValue;
class Node : public Point {  // include/geom/node.h: a point with an id
public:
  Node(const Point& p = Point(), dof_id_type id = 0) : Point(p), _id(id) {}
  dof_id_type id() const { return _id; }
private:
  dof_id_type _id;
};
typedef Point RealGradient;
typedef Point RealVectorValue;
}  // namespace libMesh
#endif
