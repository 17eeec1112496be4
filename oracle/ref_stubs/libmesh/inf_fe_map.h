// ref_stubs/libmesh/inf_fe_map.h — STUB: infinite elements are outside this path (Elem::infinite() is false here).
#ifndef REF_STUB_INF_FE_MAP_H
#define REF_STUB_INF_FE_MAP_H
#include "libmesh/elem.h"
namespace libMesh {
struct InfFEMap {
  static Point inverse_map(const unsigned int, const Elem*, const Point&, const Real = 0., const bool = true) { libmesh_not_implemented(); }
  static void inverse_map(const unsigned int, const Elem*, const std::vector<Point>&, std::vector<Point>&, const Real = 0., const bool = true) { libmesh_not_implemented(); }
  static Point map(const unsigned int, const Elem*, const Point&) { libmesh_not_implemented(); }
};
}  // namespace libMesh
#endif
