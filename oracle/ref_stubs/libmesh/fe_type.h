// ref_stubs/libmesh/fe_type.h — STUB (see libmesh_common.h in this directory): the order / family pair of include/fe/fe_type.h.
#ifndef REF_STUB_FE_TYPE_H
#define REF_STUB_FE_TYPE_H
#include "libmesh/enum_fe_family.h"
#include "libmesh/enum_order.h"
namespace libMesh {
struct OrderWrapper {  // include/fe/fe_type.h:47-120
  OrderWrapper(Order o) : _order(static_cast<int>(o)) {}
  OrderWrapper(int o) : _order(o) {}
  operator Order() const { return static_cast<Order>(_order); }
  int get_order() const { return _order; }
  int _order;
};
inline OrderWrapper operator+(const OrderWrapper& o, unsigned int p) { return OrderWrapper(o._order + (int)p); }
inline OrderWrapper operator+(const OrderWrapper& o, int p) { return OrderWrapper(o._order + p); }
class FEType {  // include/fe/fe_type.h:196
public:
  FEType(int o = 1, FEFamily f = LAGRANGE) : order(o), family(f) {}
  OrderWrapper order;
  FEFamily family;
  // include/fe/fe_type.h:414-418
  Order default_quadrature_order() const { return static_cast<Order>(2 * static_cast<unsigned int>(order.get_order()) + 1); }
};
}  // namespace libMesh
#endif
