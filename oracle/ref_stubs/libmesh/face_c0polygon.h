// ref_stubs/libmesh/face_c0polygon.h — STUB: arbitrary polygons are outside this path (see cell_c0polyhedron.h).
#ifndef REF_STUB_C0POLYGON_H
#define REF_STUB_C0POLYGON_H
#include <array>
#include <tuple>
#include "libmesh/elem.h"
namespace libMesh {
class C0Polygon : public Elem {
public:
  std::tuple<unsigned int, Real, Real> subtriangle_coordinates(const Point&, Real = 0.) const { libmesh_not_implemented(); }
  std::array<int, 3> subtriangle(unsigned int) const { libmesh_not_implemented(); }
  std::array<Point, 3> master_subtriangle(unsigned int) const { libmesh_not_implemented(); }
  unsigned int n_subtriangles() const { return 0; }
  unsigned int n_nodes() const { return 0; }
  unsigned int n_sides() const { return 0; }
  Point master_point(unsigned int) const { return Point(); }
  const Point& point(unsigned int) const { static Point p; return p; }
};
}  // namespace libMesh
#endif
