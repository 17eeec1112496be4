// gpu_numeric_vector.h — libMesh-side adaptor: a NumericVector<Number> whose entries live in B200 memory behind the
// device-vector entry points of the libfeasm C ABI (include/feasm.h, feasm_vec_*).  SURVEY §8b route (2): with
// system.rhs (and the solution / Krylov work vectors) of this type, the batched assemble hook hands the RHS over on the
// device, and NumericVector::add_vector(v, A) — the product SparseMatrix::vector_mult resolves to — runs on the device
// matrix of a GpuSparseMatrix.
//
// Storage is the local part (local_size entries from first_local_index, as DistributedVector,
// include/numerics/distributed_vector.h); on several ranks the reductions finish with the communicator's sum / max / min
// in rank order (DistributedVector::dot, distributed_vector.C), localize gathers.  The Krylov operations (dot, add(a,v),
// scale, norms, pointwise ops) are device kernels with fixed-order reductions; entry access (operator(), set/add of single
// entries) goes through single-entry transfers and is meant for set-up code, not loops.  add_vector (the product with a
// GpuSparseMatrix) works on several ranks as well: libfeasm fetches the entries of v other ranks own (feasm_build_halo, set
// up by the assemble hook) inside feasm_vector_mult.
#ifndef FEASM_GPU_NUMERIC_VECTOR_H
#define FEASM_GPU_NUMERIC_VECTOR_H

#include <cmath>
#include <cstdlib>
#include <memory>

#include "gpu_sparse_matrix.h"

namespace libMesh {

// the context the vectors of this process run on (one per process: device FEASM_DEVICE, else the rank of the first vector
// made — one process per GPU, as the matrix contexts)
inline feasm_ctx* feasm_vector_context(int rank = 0) {
  struct Holder {
    feasm_ctx* ctx = nullptr;
    explicit Holder(int rank) {
      const char* dev = std::getenv("FEASM_DEVICE");
      if (feasm_create(dev ? std::atoi(dev) : rank, &ctx)) libmesh_error_msg("libfeasm: " << feasm_last_error(nullptr));
    }
    ~Holder() { if (ctx) feasm_destroy(ctx); }
  };
  static Holder h(rank);
  return h.ctx;
}

template <typename T>
class GpuNumericVector : public NumericVector<T> {
 public:
  explicit GpuNumericVector(const Parallel::Communicator& comm, const ParallelType ptype = AUTOMATIC) : NumericVector<T>(comm, ptype) {}
  GpuNumericVector(const Parallel::Communicator& comm, const numeric_index_type n, const ParallelType ptype = AUTOMATIC)
      : NumericVector<T>(comm, ptype) { this->init(n, n, false, ptype); }
  GpuNumericVector(const Parallel::Communicator& comm, const numeric_index_type n, const numeric_index_type n_local, const ParallelType ptype = AUTOMATIC)
      : NumericVector<T>(comm, ptype) { this->init(n, n_local, false, ptype); }  // numeric_vector.h:99
  ~GpuNumericVector() override { release(); }

  // ---- lifetime ---------------------------------------------------------------------------------------------------
  void init(const numeric_index_type n, const numeric_index_type n_local, const bool = false, const ParallelType ptype = AUTOMATIC) override {
    release();
    _n = n_local; _n_global = n; _first = 0;
    if (this->n_processors() > 1 && ptype != SERIAL) {  // DistributedVector::init, distributed_vector.h:457-480
      std::vector<numeric_index_type> sizes;
      this->comm().allgather(n_local, sizes);
      numeric_index_type tot = 0;
      for (processor_id_type q = 0; q < this->n_processors(); ++q) { if (q == this->processor_id()) _first = tot; tot += sizes[q]; }
      if (tot != n) libmesh_error_msg("GpuNumericVector::init: local sizes sum to " << tot << ", not " << n);
    } else if (n != n_local) {
      libmesh_error_msg("GpuNumericVector::init: n != n_local on one rank");
    }
    _d = feasm_vec_alloc(ctx(), _n ? _n : 1);  // zero-initialised (fast == true would allow garbage; zero is always valid)
    if (!_d) libmesh_error_msg("GpuNumericVector: device allocation of " << _n << " entries failed");
    this->_type = ptype == AUTOMATIC ? PARALLEL : ptype;
    this->_is_initialized = true;
    this->_is_closed = true;
  }
  void init(const numeric_index_type n, const bool fast = false, const ParallelType ptype = AUTOMATIC) override { this->init(n, n, fast, ptype); }
  void init(const numeric_index_type n, const numeric_index_type n_local, const std::vector<numeric_index_type>&, const bool fast = false,
            const ParallelType ptype = AUTOMATIC) override { this->init(n, n_local, fast, ptype); }
  void init(const NumericVector<T>& other, const bool fast = false) override { this->init(other.size(), other.local_size(), fast, other.type()); }
  void close() override { this->_is_closed = true; }  // every operation below is complete when it returns
  void zero() override { check(feasm_vec_fill(ctx(), _d, _n, 0.)); }
  std::unique_ptr<NumericVector<T>> zero_clone() const override {
    auto v = std::make_unique<GpuNumericVector<T>>(this->comm(), this->type());
    v->init(*this);
    return v;
  }
  std::unique_ptr<NumericVector<T>> clone() const override {
    auto v = std::make_unique<GpuNumericVector<T>>(this->comm(), this->type());
    v->init(*this);
    *v = *this;
    return v;
  }

  // ---- assignment -------------------------------------------------------------------------------------------------
  // `*g = *this` between two GpuNumericVectors must not pick an implicit member-wise copy (it would share _d between
  // the two and free it twice): deep copy, as NumericVector::operator= (numeric_vector.h:279) means it
  GpuNumericVector(const GpuNumericVector&) = delete;
  GpuNumericVector& operator=(const GpuNumericVector& v) {
    this->operator=(static_cast<const NumericVector<T>&>(v));
    return *this;
  }
  NumericVector<T>& operator=(const NumericVector<T>& v) override {
    if (v.size() != _n_global) libmesh_error_msg("GpuNumericVector::operator=: size mismatch");
    if (const auto* g = dynamic_cast<const GpuNumericVector<T>*>(&v)) {
      if (g->_n != _n) libmesh_error_msg("GpuNumericVector::operator=: different local sizes");
      check(feasm_vec_copy(ctx(), FEASM_COPY_D2D, _d, g->_d, _n));
    } else { std::vector<T> h; v.localize(h); *this = h; }  // the whole vector on every rank; the local part is kept
    return *this;
  }
  NumericVector<T>& operator=(const T s) override { check(feasm_vec_fill(ctx(), _d, _n, s)); return *this; }
  NumericVector<T>& operator=(const std::vector<T>& v) override {
    // numeric_vector.h:295: v holds either the local part or the whole vector
    if (v.size() == _n_global) check(feasm_vec_copy(ctx(), FEASM_COPY_H2D, _d, v.data() + _first, _n));
    else if (v.size() == _n) check(feasm_vec_copy(ctx(), FEASM_COPY_H2D, _d, v.data(), _n));
    else libmesh_error_msg("GpuNumericVector::operator=(std::vector): size mismatch");
    return *this;
  }
  // the batched assemble hook: take n entries that are already on the device (feasm_device_rhs of the matrix context)
  void assign_device(const double* d_src, numeric_index_type n) {
    if (n != _n) libmesh_error_msg("GpuNumericVector::assign_device: size mismatch");
    check(feasm_vec_copy(ctx(), FEASM_COPY_D2D, _d, d_src, _n));
  }
  const double* device_ptr() const { return _d; }
  double* device_ptr() { return _d; }

  // ---- reductions ---------------------------------------------------------------------------------------------------
  Real min() const override { double r = _n ? reduce(FEASM_VEC_MIN) : HUGE_VAL; if (distributed()) this->comm().min(r); return r; }
  Real max() const override { double r = _n ? reduce(FEASM_VEC_MAX) : -HUGE_VAL; if (distributed()) this->comm().max(r); return r; }
  T sum() const override { double r = reduce(FEASM_VEC_SUM); if (distributed()) this->comm().sum(r); return r; }
  Real l1_norm() const override { double r = reduce(FEASM_VEC_L1); if (distributed()) this->comm().sum(r); return r; }
  Real l2_norm() const override { return std::sqrt(this->dot(*this)); }
  Real linfty_norm() const override { double r = reduce(FEASM_VEC_LINF); if (distributed()) this->comm().max(r); return r; }
  T dot(const NumericVector<T>& v) const override {
    double r = 0.;
    if (_n) check(feasm_vec_reduce(ctx(), FEASM_VEC_DOT, _d, on_device(v).ptr, _n, &r));
    if (distributed()) this->comm().sum(r);
    return r;
  }

  // ---- sizes --------------------------------------------------------------------------------------------------------
  numeric_index_type size() const override { return _n_global; }
  numeric_index_type local_size() const override { return _n; }
  numeric_index_type first_local_index() const override { return _first; }
  numeric_index_type last_local_index() const override { return _first + _n; }
  std::size_t max_allowed_id() const override { return 0xffffffffu; }

  // ---- entries ------------------------------------------------------------------------------------------------------
  T operator()(const numeric_index_type i) const override {
    double v = 0.;
    check(feasm_vec_copy(ctx(), FEASM_COPY_D2H, &v, _d + local(i), 1));
    return v;
  }
  void set(const numeric_index_type i, const T value) override { check(feasm_vec_set_entry(ctx(), _d, local(i), value, 0)); this->_is_closed = false; }
  void add(const numeric_index_type i, const T value) override { check(feasm_vec_set_entry(ctx(), _d, local(i), value, 1)); this->_is_closed = false; }

  // ---- algebra ------------------------------------------------------------------------------------------------------
  NumericVector<T>& operator+=(const NumericVector<T>& v) override { this->add(T(1), v); return *this; }
  NumericVector<T>& operator-=(const NumericVector<T>& v) override { this->add(T(-1), v); return *this; }
  NumericVector<T>& operator*=(const NumericVector<T>& v) override { check(feasm_vec_pointwise(ctx(), FEASM_VEC_MULT, _d, _d, on_device(v).ptr, 0., _n)); return *this; }
  NumericVector<T>& operator/=(const NumericVector<T>& v) override { check(feasm_vec_pointwise(ctx(), FEASM_VEC_DIV, _d, _d, on_device(v).ptr, 0., _n)); return *this; }
  void reciprocal() override { check(feasm_vec_pointwise(ctx(), FEASM_VEC_RECIPROCAL, _d, _d, nullptr, 0., _n)); }
  void conjugate() override {}  // real numbers
  void add(const T s) override { check(feasm_vec_pointwise(ctx(), FEASM_VEC_ADD_SCALAR, _d, _d, nullptr, s, _n)); }
  void add(const NumericVector<T>& v) override { this->add(T(1), v); }
  void add(const T a, const NumericVector<T>& v) override { check(feasm_vec_axpy(ctx(), _d, a, on_device(v).ptr, _n)); }
  void scale(const T factor) override { check(feasm_vec_pointwise(ctx(), FEASM_VEC_SCALE, _d, _d, nullptr, factor, _n)); }
  void abs() override { check(feasm_vec_pointwise(ctx(), FEASM_VEC_ABS, _d, _d, nullptr, 0., _n)); }
  void pointwise_mult(const NumericVector<T>& a, const NumericVector<T>& b) override {
    const OnDevice da = on_device(a), db = on_device(b);
    check(feasm_vec_pointwise(ctx(), FEASM_VEC_MULT, _d, da.ptr, db.ptr, 0., _n));
  }
  void pointwise_divide(const NumericVector<T>& a, const NumericVector<T>& b) override {
    const OnDevice da = on_device(a), db = on_device(b);
    check(feasm_vec_pointwise(ctx(), FEASM_VEC_DIV, _d, da.ptr, db.ptr, 0., _n));
  }
  // this += A v on the device matrix (numeric_vector.h:561; what SparseMatrix::vector_mult / vector_mult_add call)
  void add_vector(const NumericVector<T>& v, const SparseMatrix<T>& A) override {
    auto* G = dynamic_cast<GpuSparseMatrix<T>*>(const_cast<SparseMatrix<T>*>(&A));
    if (!G) libmesh_error_msg("GpuNumericVector::add_vector: the matrix is not a GpuSparseMatrix");
    // on several ranks v and this are the owned parts; feasm_vector_mult fetches the entries of v the other ranks own
    if (A.n() != v.size() || A.row_stop() - A.row_start() != _n || v.local_size() != _n) libmesh_error_msg("GpuNumericVector::add_vector: size mismatch");
    const OnDevice dv = on_device(v);
    double* tmp = feasm_vec_alloc(ctx(), _n);
    if (!tmp) libmesh_error_msg("GpuNumericVector::add_vector: device allocation failed");
    int rc = feasm_vector_mult(G->ctx(), dv.ptr, tmp);
    if (!rc) rc = feasm_synchronize(G->ctx());  // the product ran on the matrix context's stream
    if (!rc) rc = feasm_vec_axpy(ctx(), _d, 1., tmp, _n);
    feasm_vec_free(ctx(), tmp);
    if (rc) libmesh_error_msg("libfeasm error " << rc << " in add_vector: " << feasm_last_error(G->ctx()));
  }
  // this += A^T v (numeric_vector.h): the transposed product through SparseMatrix::get_transpose on the device matrix
  void add_vector_transpose(const NumericVector<T>& v, const SparseMatrix<T>& A) override {
    auto* G = dynamic_cast<const GpuSparseMatrix<T>*>(&A);
    if (!G) libmesh_error_msg("GpuNumericVector::add_vector_transpose: the matrix is not a GpuSparseMatrix");
    GpuSparseMatrix<T> At(A.comm());
    G->get_transpose(At);
    this->add_vector(v, At);
  }

  // ---- localize: the whole vector on every rank (numeric_vector.h:637); on one rank every flavour is a copy ------------
  void localize(std::vector<T>& v_local) const override {
    v_local.resize(_n);
    if (_n) check(feasm_vec_copy(ctx(), FEASM_COPY_D2H, v_local.data(), _d, _n));
    if (distributed()) this->comm().allgather(v_local);  // local parts in rank order = global order
  }
  void localize(NumericVector<T>& v_local) const override {
    if (auto* g = dynamic_cast<GpuNumericVector<T>*>(&v_local)) { if (g->_n == _n) { *g = *this; return; } }
    std::vector<T> h;
    this->localize(h);
    v_local = h;
  }
  void localize(NumericVector<T>& v_local, const std::vector<numeric_index_type>&) const override { this->localize(v_local); }
  void localize(std::vector<T>& v_local, const std::vector<numeric_index_type>& indices) const override {
    std::vector<T> h;
    this->localize(h);
    v_local.resize(indices.size());
    for (std::size_t k = 0; k < indices.size(); ++k) v_local[k] = h[indices[k]];
  }
  void localize(const numeric_index_type, const numeric_index_type, const std::vector<numeric_index_type>&) override {}
  void localize_to_one(std::vector<T>& v_local, const processor_id_type = 0) const override { this->localize(v_local); }

 private:
  // a device view of any NumericVector: our own buffer, or a temporary upload of a host vector
  struct OnDevice {
    const double* ptr = nullptr;
    double* owned = nullptr;
    feasm_ctx* ctx = nullptr;
    OnDevice() = default;
    OnDevice(const OnDevice&) = delete;
    OnDevice(OnDevice&& o) noexcept : ptr(o.ptr), owned(o.owned), ctx(o.ctx) { o.owned = nullptr; }
    ~OnDevice() { if (owned) feasm_vec_free(ctx, owned); }
  };
  OnDevice on_device(const NumericVector<T>& v) const {
    if (v.size() != _n_global) libmesh_error_msg("GpuNumericVector: size mismatch (" << v.size() << " vs " << _n_global << ")");
    OnDevice d;
    if (const auto* g = dynamic_cast<const GpuNumericVector<T>*>(&v)) {
      if (g->_n != _n) libmesh_error_msg("GpuNumericVector: different local sizes");
      d.ptr = g->_d;
      return d;
    }
    std::vector<T> h;
    v.localize(h);  // the whole vector; the local part is uploaded
    d.ctx = ctx();
    d.owned = feasm_vec_alloc(d.ctx, _n ? _n : 1);
    if (!d.owned) libmesh_error_msg("GpuNumericVector: device allocation failed");
    if (_n) check(feasm_vec_copy(d.ctx, FEASM_COPY_H2D, d.owned, h.data() + (h.size() == _n_global ? _first : 0), _n));
    d.ptr = d.owned;
    return d;
  }
  Real reduce(int op) const {
    double r = 0.;
    if (_n) check(feasm_vec_reduce(ctx(), op, _d, nullptr, _n, &r));
    return r;
  }
  feasm_ctx* ctx() const { return feasm_vector_context((int)this->processor_id()); }
  void check(int rc) const { if (rc) libmesh_error_msg("libfeasm error " << rc << ": " << feasm_last_error(ctx())); }
  bool distributed() const { return this->n_processors() > 1 && this->_type != SERIAL; }
  numeric_index_type local(numeric_index_type i) const {
    if (i < _first || i >= _first + _n) libmesh_error_msg("GpuNumericVector: index " << i << " is not local to this rank");
    return i - _first;
  }
  void release() {
    if (_d) feasm_vec_free(ctx(), _d);
    _d = nullptr; _n = 0; _n_global = 0; _first = 0;
    this->_is_initialized = false;
  }
  double* _d = nullptr;
  numeric_index_type _n = 0, _n_global = 0, _first = 0;  // local size, global size, first local index
};

}  // namespace libMesh
#endif
