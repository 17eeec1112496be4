// gpu_sparse_matrix.h — libMesh-side adaptor: a SparseMatrix<Number> whose storage and assembly live on a B200
// behind the libfeasm C ABI (include/feasm.h).
//
// How a libMesh application uses it (INTEGRATION.md has the full recipe):
//     system.add_matrix<libMesh::GpuSparseMatrix>("System Matrix");     // include/systems/system.h:1904, :2646-2665
//     system.attach_assemble_function(feasm_libmesh::assemble_poisson_ex4);   // system.h:1470
// ImplicitSystem::add_matrices then points ImplicitSystem::matrix at this object (src/systems/implicit_system.C:113-128);
// DofMap::attach_matrix calls attach_dof_map, update_sparsity_pattern (because need_full_sparsity_pattern() is true,
// src/base/dof_map.C:280-287) and init() (dof_map.C:240-254).
//
// Compiled against the real headers in a libMesh build, against libmesh_mock.h (-DFEASM_SHIM_MOCK) in this repo's
// build check.  Errors of the C ABI become libmesh_error_msg exceptions, like every other back-end.
#ifndef FEASM_GPU_SPARSE_MATRIX_H
#define FEASM_GPU_SPARSE_MATRIX_H

#ifdef FEASM_SHIM_MOCK
#include "libmesh_mock.h"
#else
#include "libmesh/dense_matrix.h"
#include "libmesh/dof_map.h"
#include "libmesh/numeric_vector.h"
#include "libmesh/sparse_matrix.h"
#endif

#include <algorithm>
#include <cmath>
#include <mutex>

#include "feasm.h"

namespace libMesh {

template <typename T>
class GpuSparseMatrix : public SparseMatrix<T> {
 public:
  explicit GpuSparseMatrix(const Parallel::Communicator& comm) : SparseMatrix<T>(comm) {}
  ~GpuSparseMatrix() override { this->clear(); }

  SolverPackage solver_package() override { return INVALID_SOLVER_PACKAGE; }  // storage only: a solver copies it out
  bool need_full_sparsity_pattern() const override { return true; }           // as LaspackMatrix, laspack_matrix.h:96

  // SparseMatrix::update_sparsity_pattern (sparse_matrix.h:175): keep the Graph as CSR until the context has the dofs
  void update_sparsity_pattern(const SparsityPattern::Graph& g) override {
    _row_ptr.assign(1, 0);
    _col_idx.clear();
    for (const auto& row : g) {
      _col_idx.insert(_col_idx.end(), row.begin(), row.end());  // rows of the Graph are sorted and unique
      _row_ptr.push_back(_col_idx.size());
    }
    _pattern_sent = false;
    _uploaded = Uploaded();  // the dof set behind the Graph may have changed: the hook flattens again
  }

  void init(const numeric_index_type m, const numeric_index_type n, const numeric_index_type m_l, const numeric_index_type n_l,
            const numeric_index_type = 30, const numeric_index_type = 10, const numeric_index_type = 1) override {
    _m = m; _n = n; _row_start = 0; _row_stop = m_l; (void)n_l;
    this->_is_initialized = true;
  }
  void init(ParallelType = PARALLEL) override {
    if (!this->_dof_map) libmesh_error_msg("GpuSparseMatrix::init: no DofMap attached");
    _m = _n = this->_dof_map->n_dofs();
    _row_start = this->_dof_map->first_dof();
    _row_stop = this->_dof_map->end_dof();
    this->_is_initialized = true;
  }
  void clear() override {
    if (_ctx) feasm_destroy(_ctx);
    _ctx = nullptr;
    _row_ptr.clear(); _col_idx.clear(); _host_values.clear();
    _pattern_sent = false; _closed = false; this->_is_initialized = false;
    _uploaded = Uploaded();  // the context is gone: nothing is uploaded any more
    _host_valid = false;
    _ghosts.clear(); _comm_ready = false;
  }
  void zero() override { if (_ctx && _pattern_sent) check(feasm_zero(_ctx)); _host_valid = false; }
  // same communicator, dof map and sparsity pattern, all values zero (sparse_matrix.h:219); its context is created on
  // first use
  std::unique_ptr<SparseMatrix<T>> zero_clone() const override {
    auto Z = std::make_unique<GpuSparseMatrix<T>>(this->comm());
    Z->_dof_map = this->_dof_map;
    Z->_row_ptr = _row_ptr; Z->_col_idx = _col_idx;
    Z->_m = _m; Z->_n = _n; Z->_row_start = _row_start; Z->_row_stop = _row_stop;
    Z->_is_initialized = this->_is_initialized;
    return Z;
  }
  // sparse_matrix.h:226: zero_clone plus the values, copied on the device
  std::unique_ptr<SparseMatrix<T>> clone() const override {
    auto Z = this->zero_clone();
    auto* G = static_cast<GpuSparseMatrix<T>*>(Z.get());
    if (_ctx && _pattern_sent) {
      G->ensure_matrix();
      check(feasm_synchronize(_ctx));
      check(feasm_vec_copy(G->ctx(), FEASM_COPY_D2D, feasm_device_values(G->ctx()), feasm_device_values(_ctx), (int64_t)_col_idx.size()));
    }
    return Z;
  }
  void close() override { if (_ctx) check(feasm_synchronize(_ctx)); _closed = true; }
  bool closed() const override { return _closed; }

  numeric_index_type m() const override { return _m; }
  numeric_index_type n() const override { return _n; }
  numeric_index_type row_start() const override { return _row_start; }
  numeric_index_type row_stop() const override { return _row_stop; }
  numeric_index_type col_start() const override { return _row_start; }
  numeric_index_type col_stop() const override { return _row_stop; }

  void set(const numeric_index_type i, const numeric_index_type j, const T value) override {  // sparse_matrix.h:295
    std::lock_guard<std::mutex> lock(_mutex);
    ensure_matrix();
    const uint32_t r = local_row(i), c = j;
    check(feasm_set_matrix(ctx(), &value, &r, 1, &c, 1));
    _host_valid = false; _closed = false;
  }
  void add(const numeric_index_type i, const numeric_index_type j, const T value) override {
    std::lock_guard<std::mutex> lock(_mutex);
    ensure_matrix();
    const uint32_t r = local_row(i), c = j;
    check(feasm_add_matrix(ctx(), &value, &r, 1, &c, 1));
    _host_valid = false; _closed = false;
  }
  // sparse_matrix.h:312 — the per-element route stays functional (one small kernel per call); the batched hook
  // (gpu_assemble.C) is the fast path.  Thread safe like PetscMatrix::add_matrix (src/numerics/petsc_matrix.C:747).
  void add_matrix(const DenseMatrix<T>& dm, const std::vector<numeric_index_type>& rows,
                  const std::vector<numeric_index_type>& cols) override {
    std::lock_guard<std::mutex> lock(_mutex);
    if (dm.m() != rows.size() || dm.n() != cols.size()) libmesh_error_msg("add_matrix: size mismatch");
    ensure_matrix();  // SparseMatrix::init + add_matrix works without the batched hook
    std::vector<uint32_t> lr(rows.size());
    for (std::size_t k = 0; k < rows.size(); ++k) lr[k] = local_row(rows[k]);
    check(feasm_add_matrix(ctx(), dm.get_values().data(), lr.data(), (int)lr.size(), cols.data(), (int)cols.size()));
    _host_valid = false; _closed = false;
  }
  void add_matrix(const DenseMatrix<T>& dm, const std::vector<numeric_index_type>& dof_indices) override {
    this->add_matrix(dm, dof_indices, dof_indices);
  }
  // this += a X for a matrix with the same sparsity pattern (sparse_matrix.h:344), on the device
  void add(const T a, const SparseMatrix<T>& X) override {
    std::lock_guard<std::mutex> lock(_mutex);
    auto* G = dynamic_cast<GpuSparseMatrix<T>*>(const_cast<SparseMatrix<T>*>(&X));
    if (!G) libmesh_error_msg("GpuSparseMatrix::add(a, X): X is not a GpuSparseMatrix");
    if (G->_row_ptr != _row_ptr || G->_col_idx != _col_idx) libmesh_error_msg("GpuSparseMatrix::add(a, X): X has a different sparsity pattern");
    ensure_matrix();
    G->ensure_matrix();
    check(feasm_synchronize(G->ctx()));  // X's values are complete before this context's stream reads them
    check(feasm_vec_axpy(ctx(), feasm_device_values(ctx()), a, feasm_device_values(G->ctx()), (int64_t)_col_idx.size()));
    _host_valid = false; _closed = false;
  }

  T operator()(const numeric_index_type i, const numeric_index_type j) const override {
    mirror();
    const uint64_t lo = _row_ptr[i - _row_start], hi = _row_ptr[i - _row_start + 1];
    for (uint64_t k = lo; k < hi; ++k) if (_col_idx[k] == j) return _host_values[k];
    return 0.;
  }
  Real l1_norm() const override {  // max column sum (sparse_matrix.h:379), one rank
    if (this->n_processors() > 1) libmesh_not_implemented();
    mirror();
    std::vector<Real> col(_n, 0.);
    for (std::size_t k = 0; k < _col_idx.size(); ++k) col[_col_idx[k]] += std::abs(_host_values[k]);
    Real mx = 0.;
    for (Real v : col) mx = std::max(mx, v);
    return mx;
  }
  Real linfty_norm() const override {  // max row sum (sparse_matrix.h:391)
    mirror();
    Real mx = 0.;
    for (std::size_t r = 0; r + 1 < _row_ptr.size(); ++r) {
      Real s = 0.;
      for (uint64_t k = _row_ptr[r]; k < _row_ptr[r + 1]; ++k) s += std::abs(_host_values[k]);
      mx = std::max(mx, s);
    }
    if (this->n_processors() > 1) this->comm().max(mx);  // rows are whole on their owner
    return mx;
  }
  void print_personal(std::ostream& os = libMesh::out) const override {
    mirror();
    for (std::size_t r = 0; r + 1 < _row_ptr.size(); ++r)
      for (uint64_t k = _row_ptr[r]; k < _row_ptr[r + 1]; ++k) os << r + _row_start << " " << _col_idx[k] << " " << _host_values[k] << "\n";
  }
  void get_diagonal(NumericVector<T>& dest) const override {
    mirror();
    for (std::size_t r = 0; r + 1 < _row_ptr.size(); ++r) dest.set(r + _row_start, (*this)(r + _row_start, r + _row_start));
    dest.close();
  }
  // sparse_matrix.h:598: dest = this^T, a square matrix on one rank; the transposed CSR is built on the host (a setup-time
  // operation in the reference too) and uploaded as a matrix-only context
  void get_transpose(SparseMatrix<T>& dest) const override {
    auto* D = dynamic_cast<GpuSparseMatrix<T>*>(&dest);
    if (!D) libmesh_error_msg("GpuSparseMatrix::get_transpose: dest is not a GpuSparseMatrix");
    if (_m != _n || _row_start != 0 || _row_stop != _m) libmesh_not_implemented();  // one rank, square
    mirror();
    const std::size_t n = _row_ptr.size() - 1;
    std::vector<uint64_t> rp(n + 1, 0);
    for (uint32_t c : _col_idx) ++rp[c + 1];
    for (std::size_t r = 0; r < n; ++r) rp[r + 1] += rp[r];
    std::vector<uint32_t> ci(_col_idx.size());
    std::vector<T> val(_col_idx.size());
    std::vector<uint64_t> cur(rp.begin(), rp.end() - 1);
    for (std::size_t r = 0; r < n; ++r)  // rows ascending: every transposed row comes out with ascending columns
      for (uint64_t k = _row_ptr[r]; k < _row_ptr[r + 1]; ++k) { const uint64_t q = cur[_col_idx[k]]++; ci[q] = (uint32_t)r; val[q] = _host_values[k]; }
    if (D == this) libmesh_error_msg("GpuSparseMatrix::get_transpose: dest must be another matrix");
    D->clear();
    D->_dof_map = this->_dof_map;
    D->_row_ptr = rp; D->_col_idx = ci;
    D->_m = _m; D->_n = _n; D->_row_start = _row_start; D->_row_stop = _row_stop;
    D->_is_initialized = true;
    D->ensure_matrix();
    D->check(feasm_set_values(D->ctx(), val.data(), nullptr));
    D->_closed = true;
  }
  void get_row(numeric_index_type i, std::vector<numeric_index_type>& indices, std::vector<T>& values) const override {
    mirror();  // ascending columns, as tests/numerics/sparse_matrix_test.h:88-149 expects
    const uint64_t lo = _row_ptr[i - _row_start], hi = _row_ptr[i - _row_start + 1];
    indices.assign(_col_idx.begin() + lo, _col_idx.begin() + hi);
    values.assign(_host_values.begin() + lo, _host_values.begin() + hi);
  }

  // ---- the step after assembly, on the device matrix -----------------------------------------
  // y = A x: what NumericVector::add_vector(v, A) (numeric_vector.h:561) asks of a back-end; host vectors of the
  // local size.  A GpuNumericVector subclass would call feasm_vector_mult with its device pointers instead.
  void multiply(std::vector<T>& y, const std::vector<T>& x) {
    y.resize(x.size());
    check(feasm_vector_mult_host(ctx(), x.data(), y.data()));
  }
  // A x = b with the rhs the batched hook assembled: CG + Jacobi on the device (LinearSolver::solve analogue)
  std::pair<unsigned int, Real> solve_cg(std::vector<T>& x, Real rel_tol, unsigned int max_its) {
    int its = 0;
    double rel = 0.;
    x.resize(_row_stop - _row_start);  // the owned part (the whole vector on one rank); collective on several ranks
    const int rc = feasm_cg_solve(ctx(), rel_tol, (int)max_its, nullptr, x.data(), &its, &rel);
    if (rc < 0) check(rc);  // FEASM_NOT_CONVERGED (> 0) is reported through the returned residual, like LinearSolver::solve
    return {(unsigned int)its, rel};
  }

  // ---- what the batched hook needs ---------------------------------------------------------------------------
  feasm_ctx* ctx() {
    if (!_ctx) {
      const char* dev = std::getenv("FEASM_DEVICE");
      if (feasm_create(dev ? std::atoi(dev) : (int)this->processor_id(), &_ctx)) libmesh_error_msg("libfeasm: " << feasm_last_error(nullptr));
    }
    return _ctx;
  }
  // The per-element virtuals without the batched hook: the stored Graph becomes a matrix-only context
  void ensure_matrix() {
    if (_pattern_sent) return;
    if (_row_ptr.empty()) libmesh_error_msg("GpuSparseMatrix: update_sparsity_pattern was never called");
    const int64_t n_local = (int64_t)_row_ptr.size() - 1;
    check(feasm_init_matrix(ctx(), n_local, n_local, _row_ptr.data(), _col_idx.data()));
    _pattern_sent = true;
  }
  // ---- several ranks (one process per GPU) -----------------------------------------------------------------------
  // Rows of the context: the owned rows [row_start, row_stop) first, then the ghost rows (dofs of local elements that
  // another rank owns) in ascending global id — the numbering feasm_build_exchange expects.
  void set_ghosts(std::vector<dof_id_type> g) { _ghosts = std::move(g); }
  const std::vector<dof_id_type>& ghosts() const { return _ghosts; }
  uint32_t local_row(numeric_index_type i) const {
    if (i >= _row_start && i < _row_stop) return i - _row_start;
    const auto it = std::lower_bound(_ghosts.begin(), _ghosts.end(), i);
    if (it == _ghosts.end() || *it != i) libmesh_error_msg("GpuSparseMatrix: row " << i << " is neither owned by nor a ghost of this rank");
    return (uint32_t)(_row_stop - _row_start + (it - _ghosts.begin()));
  }
  // The NCCL communicator of the context, made once per matrix: rank 0 draws the bootstrap id, the libMesh communicator
  // (MPI) broadcasts its bytes, every rank joins (feasm_comm_init is collective).
  void ensure_communicator() {
    if (_comm_ready || this->n_processors() == 1) return;
    std::vector<char> id(FEASM_COMM_ID_BYTES, 0);
    int rc = 0;
    if (this->processor_id() == 0) rc = feasm_comm_unique_id(id.data());
    this->comm().broadcast(id, 0);
    if (rc) libmesh_error_msg("libfeasm: " << feasm_last_error(nullptr));
    check(feasm_comm_init(ctx(), (int)this->n_processors(), (int)this->processor_id(), id.data()));
    _comm_ready = true;
  }
  // After feasm_build_pattern + feasm_build_exchange: the context's pattern is the local one (owned rows merged with what
  // the peers sent, then the ghost rows).  Its owned rows must be the Graph libMesh built and synchronised itself
  // (SparsityPattern::Build::parallel_sync) — checked here, entry by entry.
  void adopt_device_pattern() {
    const std::size_t n_local = _row_stop - _row_start, n_rows = n_local + _ghosts.size();
    std::vector<uint64_t> rp(n_rows + 1);
    std::vector<uint32_t> ci((std::size_t)feasm_nnz(ctx()));
    check(feasm_get_pattern(ctx(), rp.data(), ci.data()));
    if (_row_ptr.size() != n_local + 1) libmesh_error_msg("GpuSparseMatrix: the Graph has " << _row_ptr.size() - 1 << " rows, the dof map " << n_local);
    for (std::size_t r = 0; r <= n_local; ++r)
      if (rp[r] != _row_ptr[r]) libmesh_error_msg("GpuSparseMatrix: row " << r + _row_start << " of the merged device pattern differs in length from the Graph");
    if (!std::equal(_col_idx.begin(), _col_idx.end(), ci.begin())) libmesh_error_msg("GpuSparseMatrix: merged device pattern differs from the Graph");
    _pattern_sent = true;
  }

  // what the batched hook has uploaded into the context (reset by clear() and update_sparsity_pattern())
  struct Uploaded {
    const void* mesh = nullptr;
    std::size_t n_elem = 0, n_dofs = 0, n_local = 0, first = 0;
    uint64_t xyz_sum = 0;  // checksum of the node coordinates last uploaded
  };
  Uploaded& uploaded() { return _uploaded; }
  void pattern_invalidated() { _pattern_sent = false; }  // feasm_set_dofs dropped the pattern with the old dof set
  // after feasm_set_mesh / feasm_set_dofs: hand the stored Graph to the context (host-pattern route)
  void send_pattern() {
    if (_pattern_sent) return;
    if (_row_ptr.empty()) libmesh_error_msg("GpuSparseMatrix: update_sparsity_pattern was never called");
    check(feasm_set_pattern(ctx(), _row_ptr.data(), _col_idx.data()));
    _pattern_sent = true;
  }
  void check(int rc) const { if (rc) libmesh_error_msg("libfeasm error " << rc << ": " << feasm_last_error(_ctx)); }
  void invalidate_host_copy() { _host_valid = false; _closed = false; }
  std::mutex& mutex() { return _mutex; }
  const std::vector<uint64_t>& row_ptr() const { return _row_ptr; }
  const std::vector<uint32_t>& col_idx() const { return _col_idx; }
  // host copy of the CSR values (what a CPU LinearSolver reads after close())
  const std::vector<T>& host_values() const { mirror(); return _host_values; }

 private:
  void mirror() const {
    if (_host_valid) return;
    if (!_ctx || !_pattern_sent) libmesh_error_msg("GpuSparseMatrix: matrix has no values yet");
    _host_values.resize(std::max<std::size_t>(_col_idx.size(), (std::size_t)feasm_nnz(_ctx)));  // owned rows first; ghost rows follow on several ranks
    check(feasm_get_values(_ctx, _host_values.data(), nullptr));
    _host_valid = true;
  }
  feasm_ctx* _ctx = nullptr;
  numeric_index_type _m = 0, _n = 0, _row_start = 0, _row_stop = 0;
  std::vector<uint64_t> _row_ptr;
  std::vector<uint32_t> _col_idx;
  mutable std::vector<T> _host_values;
  mutable bool _host_valid = false;
  bool _pattern_sent = false, _closed = false;
  Uploaded _uploaded;
  std::vector<dof_id_type> _ghosts;
  bool _comm_ready = false;
  std::mutex _mutex;
};

}  // namespace libMesh
#endif
