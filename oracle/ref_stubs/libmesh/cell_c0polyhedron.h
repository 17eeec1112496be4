// ref_stubs/libmesh/cell_c0polyhedron.h — STUB: arbitrary polyhedra are outside this path; the reference's case label
// compiles against this declaration and throws if it were ever reached.
#ifndef REF_STUB_C0POLYHEDRON_H
#define REF_STUB_C0POLYHEDRON_H
#include <array>
#include <tuple>
#include "libmesh/elem.h"
namespace libMesh {
class C0Polyhedron : public Elem {
public:
  std::tuple<unsigned int, Real, Real, Real> subelement_coordinates(const Point&, Real = 0.) const { libmesh_not_implemented(); }
  std::array<int, 4> subelement(unsigned int) const { libmesh_not_implemented(); }
  std::array<Point, 4> master_subelement(unsigned int) const { libmesh_not_implemented(); }
  unsigned int n_subelements() const { return 0; }
  std::vector<unsigned int> nodes_on_side(unsigned int) const { return {}; }
  unsigned int n_nodes() const { return 0; }
  unsigned int n_sides() const { return 0; }
  Point master_point(unsigned int) const { return Point(); }
  const Point& point(unsigned int) const { static Point p; return p; }
};
}  // namespace libMesh
#endif
