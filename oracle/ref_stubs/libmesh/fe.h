// ref_stubs/libmesh/fe.h — STUB (see libmesh_common.h in this directory): the DECLARATION of FE<Dim, T> that the reference's
// shape-function sources define members of (include/fe/fe.h:129-560: shape, shape_deriv, shape_second_deriv in their three
// overloads, the vectorised all_shapes / shapes / shape_derivs / all_shape_derivs and their default_* loops).  No shape
// arithmetic lives here: every number comes out of the reference's own src/fe/fe_lagrange_shape_{2,3}D.C and
// include/fe/fe_lagrange_shape_1D.h, compiled where they lie.
#ifndef REF_STUB_FE_H
#define REF_STUB_FE_H
#include "libmesh/elem.h"
#include "libmesh/enum_fe_family.h"
#include "libmesh/fe_type.h"
namespace libMesh {
// include/fe/fe_interface.h finite-difference fall-backs: only PYRAMID18 (not on this path) reaches them
template <typename F> inline Real fe_fdm_deriv(const ElemType, const Order, const Elem*, const unsigned int, const unsigned int, const Point&, F) { libmesh_not_implemented(); }
template <typename F> inline Real fe_fdm_second_deriv(const ElemType, const Order, const Elem*, const unsigned int, const unsigned int, const Point&, F) { libmesh_not_implemented(); }

// include/fe/fe_abstract.h: DECLARATION only; the definition (the reference-element node coordinates of every element type) is
// the reference's own, cut out of src/fe/fe_abstract.C at build time
class FEAbstract {
public:
  static void get_refspace_nodes(const ElemType t, std::vector<Point>& nodes);
};

template <unsigned int Dim, FEFamily T>
class FE {
public:
  typedef Real OutputShape;
  static Real shape(const ElemType t, const Order o, const unsigned int i, const Point& p);
  static Real shape(const Elem* elem, const Order o, const unsigned int i, const Point& p, const bool add_p_level = true);
  static Real shape(const FEType fet, const Elem* elem, const unsigned int i, const Point& p, const bool add_p_level = true);
  static Real shape_deriv(const ElemType t, const Order o, const unsigned int i, const unsigned int j, const Point& p);
  static Real shape_deriv(const Elem* elem, const Order o, const unsigned int i, const unsigned int j, const Point& p, const bool add_p_level = true);
  static Real shape_deriv(const FEType fet, const Elem* elem, const unsigned int i, const unsigned int j, const Point& p, const bool add_p_level = true);
  static Real shape_second_deriv(const ElemType t, const Order o, const unsigned int i, const unsigned int j, const Point& p);
  static Real shape_second_deriv(const Elem* elem, const Order o, const unsigned int i, const unsigned int j, const Point& p, const bool add_p_level = true);
  static Real shape_second_deriv(const FEType fet, const Elem* elem, const unsigned int i, const unsigned int j, const Point& p, const bool add_p_level = true);
  static void all_shapes(const Elem* elem, const Order o, const std::vector<Point>& p, std::vector<std::vector<OutputShape>>& v, const bool add_p_level = true);
  static void shapes(const Elem* elem, const Order o, const unsigned int i, const std::vector<Point>& p, std::vector<OutputShape>& v, const bool add_p_level = true);
  static void shape_derivs(const Elem* elem, const Order o, const unsigned int i, const unsigned int j, const std::vector<Point>& p, std::vector<OutputShape>& v, const bool add_p_level = true);
  static void all_shape_derivs(const Elem* elem, const Order o, const std::vector<Point>& p, std::vector<std::vector<OutputShape>>* comps[3], const bool add_p_level = true);
  // the plain loops the reference falls back to (include/fe/fe.h: default_*): one call of shape / shape_deriv per entry
  static void default_all_shapes(const Elem* elem, const Order o, const std::vector<Point>& p, std::vector<std::vector<OutputShape>>& v, const bool add_p_level) {
    for (std::size_t i = 0; i < v.size(); ++i) default_shapes(elem, o, (unsigned int)i, p, v[i], add_p_level);
  }
  static void default_shapes(const Elem* elem, const Order o, const unsigned int i, const std::vector<Point>& p, std::vector<OutputShape>& v, const bool add_p_level) {
    for (std::size_t q = 0; q < p.size(); ++q) v[q] = shape(elem, o, i, p[q], add_p_level);
  }
  static void default_shape_derivs(const Elem* elem, const Order o, const unsigned int i, const unsigned int j, const std::vector<Point>& p, std::vector<OutputShape>& v, const bool add_p_level) {
    for (std::size_t q = 0; q < p.size(); ++q) v[q] = shape_deriv(elem, o, i, j, p[q], add_p_level);
  }
  static void default_all_shape_derivs(const Elem* elem, const Order o, const std::vector<Point>& p, std::vector<std::vector<OutputShape>>* comps[3], const bool add_p_level) {
    for (unsigned int d = 0; d < Dim; ++d)
      if (comps[d])
        for (std::size_t i = 0; i < comps[d]->size(); ++i) default_shape_derivs(elem, o, (unsigned int)i, d, p, (*comps[d])[i], add_p_level);
  }
};
}  // namespace libMesh

// include/fe/fe.h:1738-1788
#define LIBMESH_DEFAULT_VECTORIZED_FE(MyDim, MyType)                                                                                           \
  template <> void FE<MyDim, MyType>::all_shapes(const Elem* elem, const Order o, const std::vector<Point>& p,                                \
                                                  std::vector<std::vector<OutputShape>>& v, const bool add_p_level) {                        \
    FE<MyDim, MyType>::default_all_shapes(elem, o, p, v, add_p_level);                                                                         \
  }                                                                                                                                            \
  template <> void FE<MyDim, MyType>::shapes(const Elem* elem, const Order o, const unsigned int i, const std::vector<Point>& p,              \
                                              std::vector<OutputShape>& v, const bool add_p_level) {                                          \
    FE<MyDim, MyType>::default_shapes(elem, o, i, p, v, add_p_level);                                                                          \
  }                                                                                                                                            \
  template <> void FE<MyDim, MyType>::shape_derivs(const Elem* elem, const Order o, const unsigned int i, const unsigned int j,               \
                                                    const std::vector<Point>& p, std::vector<OutputShape>& v, const bool add_p_level) {       \
    FE<MyDim, MyType>::default_shape_derivs(elem, o, i, j, p, v, add_p_level);                                                                 \
  }                                                                                                                                            \
  template <> void FE<MyDim, MyType>::all_shape_derivs(const Elem* elem, const Order o, const std::vector<Point>& p,                          \
                                                        std::vector<std::vector<OutputShape>>* comps[3], const bool add_p_level) {            \
    FE<MyDim, MyType>::default_all_shape_derivs(elem, o, p, comps, add_p_level);                                                               \
  }
#endif
