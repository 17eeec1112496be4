// ref_stubs/libmesh/dense_matrix.h — STUB: see dense_vector.h in this directory.
#ifndef REF_STUB_DENSE_MATRIX_H
#define REF_STUB_DENSE_MATRIX_H
#include "libmesh/dense_vector.h"
namespace libMesh {
template <typename T> class DenseMatrix {
public:
  DenseMatrix(unsigned int m = 0, unsigned int n = 0) : _m(m), _n(n), a((std::size_t)m * n, T(0)) {}
  T& operator()(unsigned int i, unsigned int j) { return a[(std::size_t)i * _n + j]; }
  const T& operator()(unsigned int i, unsigned int j) const { return a[(std::size_t)i * _n + j]; }
  void vector_mult(DenseVector<T>& dest, const DenseVector<T>& arg) const {
    dest.resize(_m);
    for (unsigned int i = 0; i < _m; ++i) for (unsigned int j = 0; j < _n; ++j) dest(i) += (*this)(i, j) * arg(j);
  }
private:
  unsigned int _m, _n;
  std::vector<T> a;
};
}  // namespace libMesh
#endif
