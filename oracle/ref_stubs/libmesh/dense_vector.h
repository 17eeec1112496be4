// ref_stubs/libmesh/dense_vector.h — STUB: a plain vector with the two calls the 2D-in-3D second-derivative branch of
// FEMap::compute_single_point_map names (never executed on this path).
#ifndef REF_STUB_DENSE_VECTOR_H
#define REF_STUB_DENSE_VECTOR_H
#include "libmesh/libmesh_common.h"
namespace libMesh {
template <typename T> class DenseVector {
public:
  explicit DenseVector(unsigned int n = 0) : v(n, T(0)) {}
  T& operator()(unsigned int i) { return v[i]; }
  const T& operator()(unsigned int i) const { return v[i]; }
  unsigned int size() const { return (unsigned int)v.size(); }
  void resize(unsigned int n) { v.assign(n, T(0)); }
  void add(T f, const DenseVector<T>& o) { for (std::size_t i = 0; i < v.size(); ++i) v[i] += f * o.v[i]; }
  std::vector<T> v;
};
}  // namespace libMesh
#endif
