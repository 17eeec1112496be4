// ref_stubs/libmesh/fe_interface.h — STUB (see libmesh_common.h in this directory): FEInterface's run-time dispatch
// (src/fe/fe_interface.C) reduced to the LAGRANGE family — every call lands in the reference's own FE<Dim, LAGRANGE>
// members (src/fe/fe_lagrange_shape_{2,3}D.C).
#ifndef REF_STUB_FE_INTERFACE_H
#define REF_STUB_FE_INTERFACE_H
#include "libmesh/fe.h"
namespace libMesh {
class FEInterface {
public:
  typedef Real (*shape_ptr)(const FEType fet, const Elem* elem, const unsigned int i, const Point& p, const bool add_p_level);
  typedef Real (*shape_deriv_ptr)(const FEType fet, const Elem* elem, const unsigned int i, const unsigned int j, const Point& p, const bool add_p_level);
  typedef Real (*shape_second_deriv_ptr)(const FEType fet, const Elem* elem, const unsigned int i, const unsigned int j, const Point& p, const bool add_p_level);
  // Lagrange maps of default order: one mapping function per node (FE<Dim, LAGRANGE>::n_shape_functions, fe_lagrange.C)
  static unsigned int n_shape_functions(const FEType&, const int, const Elem* elem) { return elem->n_nodes(); }
  static shape_ptr shape_function(const FEType&, const Elem* elem) { return elem->dim() == 3 ? (shape_ptr)&FE<3, LAGRANGE>::shape : (shape_ptr)&FE<2, LAGRANGE>::shape; }
  static shape_deriv_ptr shape_deriv_function(const FEType&, const Elem* elem) { return elem->dim() == 3 ? (shape_deriv_ptr)&FE<3, LAGRANGE>::shape_deriv : (shape_deriv_ptr)&FE<2, LAGRANGE>::shape_deriv; }
  static shape_second_deriv_ptr shape_second_deriv_function(const FEType&, const Elem* elem) {
    return elem->dim() == 3 ? (shape_second_deriv_ptr)&FE<3, LAGRANGE>::shape_second_deriv : (shape_second_deriv_ptr)&FE<2, LAGRANGE>::shape_second_deriv;
  }
  static void all_shapes(const unsigned int dim, const FEType& fet, const Elem* elem, const std::vector<Point>& p, std::vector<std::vector<Real>>& v, const bool add_p_level) {
    if (dim == 3) FE<3, LAGRANGE>::all_shapes(elem, fet.order, p, v, add_p_level);
    else FE<2, LAGRANGE>::all_shapes(elem, fet.order, p, v, add_p_level);
  }
  static void all_shape_derivs(const unsigned int dim, const FEType& fet, const Elem* elem, const std::vector<Point>& p, std::vector<std::vector<Real>>* comps[3], const bool add_p_level) {
    if (dim == 3) FE<3, LAGRANGE>::all_shape_derivs(elem, fet.order, p, comps, add_p_level);
    else FE<2, LAGRANGE>::all_shape_derivs(elem, fet.order, p, comps, add_p_level);
  }
  static bool is_InfFE_elem(const ElemType) { return false; }
};
}  // namespace libMesh
#endif
