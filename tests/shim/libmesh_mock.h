// libmesh_mock.h — MOCK of the libMesh declarations the adaptor classes use, so that gpu_sparse_matrix.h /
// gpu_assemble.C compile — and shim_test.C runs them end to end — in a container where libMesh itself cannot be built
// (its TIMPI and MetaPhysicL submodules are empty).  Every declaration cites the reference header it mirrors; the
// bodies are minimal data holders (a test fills the public underscore members), not libMesh's implementations.
// A real build defines nothing of this: it includes the libmesh/ headers instead (see gpu_sparse_matrix.h).
#ifndef FEASM_LIBMESH_MOCK_H
#define FEASM_LIBMESH_MOCK_H

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <iostream>
#include <map>
#include <set>
#include <tuple>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace libMesh {
typedef double Real;                    // include/base/libmesh_common.h:144
typedef Real Number;                    // libmesh_common.h:217 (real build)
typedef uint32_t numeric_index_type;    // include/base/id_types.h:99
typedef uint32_t dof_id_type;           // id_types.h:66-76
typedef uint32_t processor_id_type;     // id_types.h:104
typedef int8_t boundary_id_type;
static std::ostream& out = std::cout;
struct LogicError : std::logic_error { using std::logic_error::logic_error; };
struct NotImplemented : std::logic_error { using std::logic_error::logic_error; };
#define libmesh_error_msg(msg) do { std::ostringstream s__; s__ << msg; throw libMesh::LogicError(s__.str()); } while (0)
#define libmesh_not_implemented() throw libMesh::NotImplemented(std::string("not implemented: ") + __func__)
#define libmesh_assert(x) ((void)0)

enum ParallelType { AUTOMATIC = 0, SERIAL, PARALLEL, GHOSTED };     // include/enums/enum_parallel_type.h
enum SolverPackage { PETSC_SOLVERS = 0, TRILINOS_SOLVERS, LASPACK_SOLVERS, SLEPC_SOLVERS, EIGEN_SOLVERS, NLOPT_SOLVERS, INVALID_SOLVER_PACKAGE };
enum ElemType { TET10 = 9, HEX8 = 10, HEX20 = 11, HEX27 = 12, PRISM18 = 15, INVALID_ELEM = 99 };  // include/enums/enum_elem_type.h

// include/parallel/parallel.h -> TIMPI's Communicator (timpi/communicator.h): rank / size and the few collectives the
// adaptor uses (broadcast, allgather, sum, max, min).  One rank by default; a multi-process test sets FEASM_MOCK_RANK,
// FEASM_MOCK_SIZE and FEASM_MOCK_DIR, and the collectives then run over files in that directory (every rank writes
// <sequence>.<rank>, renames it into place and reads the others' — enough to stand in for MPI in a two-process test).
namespace Parallel {
struct Communicator {
  Communicator() {
    const char* r = std::getenv("FEASM_MOCK_RANK");
    const char* n = std::getenv("FEASM_MOCK_SIZE");
    const char* d = std::getenv("FEASM_MOCK_DIR");
    if (r && n && d) { _rank = (processor_id_type)std::atoi(r); _size = (processor_id_type)std::atoi(n); _dir = d; }
  }
  processor_id_type rank() const { return _rank; }
  processor_id_type size() const { return _size; }
  // every rank's bytes, in rank order (all the same length)
  std::vector<std::vector<char>> exchange(const std::vector<char>& mine) const {
    std::vector<std::vector<char>> all(_size);
    all[_rank] = mine;
    if (_size == 1) return all;
    const unsigned seq = _seq++;
    const std::string base = _dir + "/" + std::to_string(seq) + ".";
    {
      const std::string tmp = base + std::to_string(_rank) + ".tmp", fin = base + std::to_string(_rank);
      FILE* f = std::fopen(tmp.c_str(), "wb");
      if (!f) throw std::runtime_error("mock communicator: cannot write " + tmp);
      if (!mine.empty()) std::fwrite(mine.data(), 1, mine.size(), f);
      std::fclose(f);
      std::rename(tmp.c_str(), fin.c_str());
    }
    for (processor_id_type q = 0; q < _size; ++q) {
      if (q == _rank) continue;
      const std::string fin = base + std::to_string(q);
      FILE* f = nullptr;
      for (int tries = 0; tries < 60000 && !(f = std::fopen(fin.c_str(), "rb")); ++tries) std::this_thread::sleep_for(std::chrono::milliseconds(1));
      if (!f) throw std::runtime_error("mock communicator: rank " + std::to_string(q) + " never wrote " + fin);
      char buf[4096];
      std::size_t got;
      while ((got = std::fread(buf, 1, sizeof(buf), f)) > 0) all[q].insert(all[q].end(), buf, buf + got);
      std::fclose(f);
    }
    return all;
  }
  template <typename T> void broadcast(std::vector<T>& data, const unsigned int root_id = 0) const {  // timpi: broadcast(T&, root)
    std::vector<char> mine;
    if (_rank == root_id) mine.assign((const char*)data.data(), (const char*)(data.data() + data.size()));
    const auto all = exchange(mine);
    data.resize(all[root_id].size() / sizeof(T));
    std::memcpy(data.data(), all[root_id].data(), all[root_id].size());
  }
  template <typename T> void allgather(const T& send, std::vector<T>& recv) const {                 // timpi: allgather(const T&, vector<T>&)
    const auto all = exchange(std::vector<char>((const char*)&send, (const char*)(&send + 1)));
    recv.resize(_size);
    for (processor_id_type q = 0; q < _size; ++q) std::memcpy(&recv[q], all[q].data(), sizeof(T));
  }
  template <typename T> void allgather(std::vector<T>& r, const bool = false) const {               // timpi: in place, variable lengths
    const auto all = exchange(std::vector<char>((const char*)r.data(), (const char*)(r.data() + r.size())));
    r.clear();
    for (processor_id_type q = 0; q < _size; ++q) r.insert(r.end(), (const T*)all[q].data(), (const T*)(all[q].data() + all[q].size()));
  }
  template <typename T> void sum(T& x) const { std::vector<T> a; allgather(x, a); x = T(0); for (const T& v : a) x += v; }   // rank order: deterministic
  template <typename T> void max(T& x) const { std::vector<T> a; allgather(x, a); for (const T& v : a) x = v > x ? v : x; }
  template <typename T> void min(T& x) const { std::vector<T> a; allgather(x, a); for (const T& v : a) x = v < x ? v : x; }
  void barrier() const { int one = 1; sum(one); }
  processor_id_type _rank = 0, _size = 1;
  std::string _dir;
  mutable unsigned _seq = 0;
};
}  // namespace Parallel
struct ParallelObject {                                              // include/parallel/parallel_object.h
  explicit ParallelObject(const Parallel::Communicator& c) : _communicator(c) {}
  const Parallel::Communicator& comm() const { return _communicator; }
  processor_id_type n_processors() const { return _communicator.size(); }
  processor_id_type processor_id() const { return _communicator.rank(); }
  const Parallel::Communicator& _communicator;
};

struct Point { Real c[3]; Real operator()(unsigned i) const { return c[i]; } };   // include/geom/point.h
struct Node : Point { dof_id_type _id; dof_id_type id() const { return _id; } };  // include/geom/node.h
struct Elem {                                                                      // include/geom/elem.h
  ElemType _type; std::vector<const Node*> _nodes; dof_id_type _id; processor_id_type _pid;
  std::vector<const Elem*> _neighbors;                                             // one per side; nullptr = boundary
  ElemType type() const { return _type; }
  unsigned int n_sides() const { return (unsigned)_neighbors.size(); }
  std::vector<unsigned short> side_index_range() const {                           // elem.h:2710 (an IntRange there)
    std::vector<unsigned short> r(_neighbors.size());
    for (std::size_t i = 0; i < r.size(); ++i) r[i] = (unsigned short)i;
    return r;
  }
  const Elem* neighbor_ptr(unsigned int side) const { return _neighbors[side]; }   // elem.h:2598
  unsigned int n_nodes() const { return (unsigned)_nodes.size(); }
  const Node* node_ptr(unsigned i) const { return _nodes[i]; }
  dof_id_type node_id(unsigned i) const { return _nodes[i]->id(); }
  dof_id_type id() const { return _id; }
  processor_id_type processor_id() const { return _pid; }
};
struct BoundaryInfo {                                                              // include/mesh/boundary_info.h
  std::set<std::tuple<const Elem*, unsigned short, boundary_id_type>> _sides;
  bool has_boundary_id(const Elem* e, unsigned short side, boundary_id_type id) const {  // boundary_info.h:521
    return _sides.count(std::make_tuple(e, side, id)) != 0;
  }
};
struct MeshBase {                                                                  // include/mesh/mesh_base.h
  std::vector<const Elem*> _local;
  BoundaryInfo _bi;
  const BoundaryInfo& get_boundary_info() const { return _bi; }                    // mesh_base.h:165
  const std::vector<const Elem*>& active_local_element_ptr_range() const { return _local; }  // mesh_base.h:1484
  dof_id_type n_nodes() const { return 0; }
  dof_id_type max_node_id() const { return 0; }
};

template <typename T> struct DenseMatrix {                                         // include/numerics/dense_matrix.h
  unsigned _m = 0, _n = 0; std::vector<T> _val;
  unsigned m() const { return _m; } unsigned n() const { return _n; }
  const std::vector<T>& get_values() const { return _val; }                        // dense_matrix.h:382 (row-major, :953-978)
};
namespace SparsityPattern {                                                        // include/base/sparsity_pattern.h:55-57
typedef std::vector<dof_id_type> Row;
typedef std::vector<Row> Graph;
}
typedef std::map<dof_id_type, Number> DofConstraintValueMap;                       // include/base/dof_map.h:123
typedef std::map<dof_id_type, Real> DofConstraintRow;                              // dof_map.h:97-100
typedef std::map<dof_id_type, DofConstraintRow> DofConstraints;                    // dof_map.h:108-112

class DofMap {                                                                     // include/base/dof_map.h
 public:
  dof_id_type n_dofs() const { return _n_dofs; }                                   // dof_map_base.h
  // dof_map_base.h:185-197: first_dof(proc) / end_dof(proc), proc defaulting to this rank.  _first empty = one rank.
  dof_id_type first_dof(processor_id_type p = (processor_id_type)-1) const { return _first.empty() ? 0 : _first[p == (processor_id_type)-1 ? _rank : p]; }
  dof_id_type end_dof(processor_id_type p = (processor_id_type)-1) const { return _first.empty() ? _n_dofs : _first[(p == (processor_id_type)-1 ? _rank : p) + 1]; }
  processor_id_type n_processors() const { return _first.empty() ? 1 : (processor_id_type)(_first.size() - 1); }
  processor_id_type processor_id() const { return _rank; }
  dof_id_type n_local_dofs() const { return end_dof() - first_dof(); }
  void dof_indices(const Elem* e, std::vector<dof_id_type>& di) const {            // dof_map.h:2201 / src/base/dof_map.C:2201
    di.assign(_elem_dofs.begin() + (std::size_t)e->id() * _nd, _elem_dofs.begin() + (std::size_t)(e->id() + 1) * _nd);
  }
  bool is_constrained_dof(dof_id_type d) const { return _vals.count(d) != 0 || _rows.count(d) != 0; }  // dof_map.h:2426
  DofConstraints::const_iterator constraint_rows_begin() const { return _rows.begin(); }  // dof_map.h:1164
  DofConstraints::const_iterator constraint_rows_end() const { return _rows.end(); }      // dof_map.h:1170
  DofConstraintValueMap& get_primal_constraint_values() { return _vals; }          // dof_map.h:1263
  unsigned int n_variables() const { return 1; }
  DofConstraintValueMap _vals;
  DofConstraints _rows;                 // non-empty constraint rows (hanging nodes); Dirichlet dofs live in _vals only
  dof_id_type _n_dofs = 0;
  unsigned _nd = 0;
  std::vector<dof_id_type> _elem_dofs;  // [n_elem][_nd], element id major
  std::vector<dof_id_type> _first;      // [n_processors + 1] dof ranges (empty: one rank owns everything)
  processor_id_type _rank = 0;
};

template <typename T> class SparseMatrix;
// include/numerics/numeric_vector.h — every virtual the reference declares pure (cited by line) is declared here with a
// host std::vector body, so that (a) `override` in GpuNumericVector checks each signature and (b) the mock itself is a
// usable plain vector for the route-(1) test (system.rhs as an ordinary NumericVector).
template <typename T> class NumericVector : public ParallelObject {
 public:
  explicit NumericVector(const Parallel::Communicator& c, const ParallelType ptype = AUTOMATIC) : ParallelObject(c), _type(ptype) {}  // :84
  virtual ~NumericVector() = default;
  virtual bool initialized() const { return _is_initialized; }                                    // :152
  virtual bool closed() const { return _is_closed; }                                              // :183
  ParallelType type() const { return _type; }                                                     // :157
  virtual NumericVector<T>& operator=(const NumericVector<T>& v) { _v = v._v; return *this; }      // :119
  virtual void close() { _is_closed = true; }                                                     // :216
  virtual void zero() { std::fill(_v.begin(), _v.end(), T(0)); }                                   // :227
  virtual std::unique_ptr<NumericVector<T>> zero_clone() const { auto p = std::make_unique<NumericVector<T>>(comm()); p->_v.assign(_v.size(), T(0)); return p; }  // :235
  virtual std::unique_ptr<NumericVector<T>> clone() const { auto p = std::make_unique<NumericVector<T>>(comm()); p->_v = _v; return p; }  // :242
  virtual void init(const numeric_index_type n, const numeric_index_type, const bool = false, const ParallelType = AUTOMATIC) { _v.assign(n, T(0)); _is_initialized = true; }  // :254
  virtual void init(const numeric_index_type n, const bool = false, const ParallelType = AUTOMATIC) { _v.assign(n, T(0)); _is_initialized = true; }  // :262
  virtual void init(const numeric_index_type n, const numeric_index_type, const std::vector<numeric_index_type>&, const bool = false,
                    const ParallelType = AUTOMATIC) { _v.assign(n, T(0)); _is_initialized = true; }  // :270
  virtual void init(const NumericVector<T>& other, const bool = false) { _v.assign(other.size(), T(0)); _is_initialized = true; }  // :280
  virtual NumericVector<T>& operator=(const T s) { std::fill(_v.begin(), _v.end(), s); return *this; }  // :288
  virtual NumericVector<T>& operator=(const std::vector<T>& v) { _v = v; return *this; }          // :295
  virtual Real min() const { Real m = _v.empty() ? 0 : _v[0]; for (T x : _v) m = x < m ? x : m; return m; }  // :301
  virtual Real max() const { Real m = _v.empty() ? 0 : _v[0]; for (T x : _v) m = x > m ? x : m; return m; }  // :307
  virtual T sum() const { T s = 0; for (T x : _v) s += x; return s; }                              // :312
  virtual Real l1_norm() const { Real s = 0; for (T x : _v) s += x < 0 ? -x : x; return s; }       // :318
  virtual Real l2_norm() const { Real s = 0; for (T x : _v) s += x * x; return std::sqrt(s); }     // :324
  virtual Real linfty_norm() const { Real s = 0; for (T x : _v) s = (x < 0 ? -x : x) > s ? (x < 0 ? -x : x) : s; return s; }  // :330
  virtual numeric_index_type size() const { return (numeric_index_type)_v.size(); }               // :372
  virtual numeric_index_type local_size() const { return (numeric_index_type)_v.size(); }         // :377
  virtual numeric_index_type first_local_index() const { return 0; }                              // :385
  virtual numeric_index_type last_local_index() const { return (numeric_index_type)_v.size(); }   // :393
  virtual T operator()(const numeric_index_type i) const { return _v[i]; }                         // :398
  virtual NumericVector<T>& operator+=(const NumericVector<T>& v) { add(T(1), v); return *this; }  // :430
  virtual NumericVector<T>& operator-=(const NumericVector<T>& v) { add(T(-1), v); return *this; } // :439
  virtual NumericVector<T>& operator*=(const NumericVector<T>& v) { for (std::size_t i = 0; i < _v.size(); ++i) _v[i] *= v(i); return *this; }  // :457
  virtual NumericVector<T>& operator/=(const NumericVector<T>& v) { for (std::size_t i = 0; i < _v.size(); ++i) _v[i] /= v(i); return *this; }  // :474
  virtual void reciprocal() { for (T& x : _v) x = T(1) / x; }                                      // :480
  virtual void conjugate() {}                                                                      // :485
  virtual void set(const numeric_index_type i, const T v) { if (i >= _v.size()) _v.resize(i + 1); _v[i] = v; }  // :491
  virtual void add(const numeric_index_type i, const T v) { _v[i] += v; }                          // :497
  virtual void add(const T s) { for (T& x : _v) x += s; }                                          // :503
  virtual void add(const NumericVector<T>& v) { add(T(1), v); }                                    // :510
  virtual void add(const T a, const NumericVector<T>& v) { for (std::size_t i = 0; i < _v.size(); ++i) _v[i] += a * v(i); }  // :517
  virtual void add_vector(const NumericVector<T>&, const SparseMatrix<T>&) { libmesh_not_implemented(); }            // :561
  virtual void add_vector_transpose(const NumericVector<T>&, const SparseMatrix<T>&) { libmesh_not_implemented(); }  // :577
  virtual void scale(const T f) { for (T& x : _v) x *= f; }                                        // :618
  virtual void abs() { for (T& x : _v) x = x < 0 ? -x : x; }                                       // :623
  virtual T dot(const NumericVector<T>& v) const { T s = 0; for (std::size_t i = 0; i < _v.size(); ++i) s += _v[i] * v(i); return s; }  // :631
  virtual void localize(std::vector<T>& v_local) const { v_local = _v; }                           // :637
  virtual void localize(NumericVector<T>& v_local) const { v_local = _v; }                         // :643
  virtual void localize(NumericVector<T>& v_local, const std::vector<numeric_index_type>&) const { v_local = _v; }  // :649
  virtual void localize(std::vector<T>& v_local, const std::vector<numeric_index_type>& indices) const {            // :681
    v_local.resize(indices.size());
    for (std::size_t k = 0; k < indices.size(); ++k) v_local[k] = _v[indices[k]];
  }
  virtual void localize(const numeric_index_type, const numeric_index_type, const std::vector<numeric_index_type>&) {}  // :688
  virtual void localize_to_one(std::vector<T>& v_local, const processor_id_type = 0) const { v_local = _v; }            // :697
  virtual void pointwise_mult(const NumericVector<T>& a, const NumericVector<T>& b) { for (std::size_t i = 0; i < _v.size(); ++i) _v[i] = a(i) * b(i); }    // :729
  virtual void pointwise_divide(const NumericVector<T>& a, const NumericVector<T>& b) { for (std::size_t i = 0; i < _v.size(); ++i) _v[i] = a(i) / b(i); }  // :737
  virtual std::size_t max_allowed_id() const { return 0xffffffffu; }                               // :847
  std::vector<T> _v;
 protected:
  bool _is_closed = false;                                                                         // :854
  bool _is_initialized = false;                                                                    // :860
  ParallelType _type;                                                                              // :865
};

template <typename T> class SparseMatrix : public ParallelObject {                  // include/numerics/sparse_matrix.h
 public:
  explicit SparseMatrix(const Parallel::Communicator& c) : ParallelObject(c) {}     // :78
  virtual ~SparseMatrix() = default;
  virtual SolverPackage solver_package() = 0;                                       // :127
  virtual bool initialized() const { return _is_initialized; }                      // :133
  void attach_dof_map(const DofMap& d) { _dof_map = &d; }                           // :138
  virtual bool need_full_sparsity_pattern() const { return false; }                 // :162
  virtual void update_sparsity_pattern(const SparsityPattern::Graph&) {}            // :175
  virtual void init(const numeric_index_type m, const numeric_index_type n, const numeric_index_type m_l,
                    const numeric_index_type n_l, const numeric_index_type nnz = 30, const numeric_index_type noz = 10,
                    const numeric_index_type blocksize = 1) = 0;                    // :189
  virtual void init(ParallelType type = PARALLEL) = 0;                              // :201
  virtual void clear() = 0;                                                         // :206
  virtual void zero() = 0;                                                          // :211
  virtual std::unique_ptr<SparseMatrix<T>> zero_clone() const = 0;                  // :219
  virtual std::unique_ptr<SparseMatrix<T>> clone() const = 0;                       // :226
  virtual void close() = 0;                                                         // :237
  virtual numeric_index_type m() const = 0;                                         // :249
  virtual numeric_index_type n() const = 0;                                         // :264
  virtual numeric_index_type row_start() const = 0;                                 // :270
  virtual numeric_index_type row_stop() const = 0;                                  // :276
  virtual numeric_index_type col_start() const = 0;                                 // :282
  virtual numeric_index_type col_stop() const = 0;                                  // :288
  virtual void set(const numeric_index_type i, const numeric_index_type j, const T value) = 0;   // :295
  virtual void add(const numeric_index_type i, const numeric_index_type j, const T value) = 0;   // :304
  virtual void add_matrix(const DenseMatrix<T>& dm, const std::vector<numeric_index_type>& rows,
                          const std::vector<numeric_index_type>& cols) = 0;         // :312
  virtual void add_matrix(const DenseMatrix<T>& dm, const std::vector<numeric_index_type>& dof_indices) = 0;  // :320
  virtual void add(const T a, const SparseMatrix<T>& X) = 0;                        // :344
  virtual T operator()(const numeric_index_type i, const numeric_index_type j) const = 0;       // :368
  virtual Real l1_norm() const = 0;                                                 // :379
  virtual Real linfty_norm() const = 0;                                             // :391
  virtual bool closed() const = 0;                                                  // :401
  virtual void print_personal(std::ostream& os = libMesh::out) const = 0;           // :448
  virtual void get_diagonal(NumericVector<T>& dest) const = 0;                      // :592
  virtual void get_transpose(SparseMatrix<T>& dest) const = 0;                      // :598
  virtual void get_row(numeric_index_type i, std::vector<numeric_index_type>& indices, std::vector<T>& values) const = 0;  // :607
 protected:
  const DofMap* _dof_map = nullptr;                                                 // :676
  bool _is_initialized = false;                                                     // :687
};

class System {                                                                      // include/systems/system.h
 public:
  const MeshBase& get_mesh() const { return *_mesh; }                               // :2359
  DofMap& get_dof_map() { return *_dof; }                                           // :2375
  const MeshBase* _mesh = nullptr; DofMap* _dof = nullptr;
};
class ExplicitSystem : public System { public: NumericVector<Number>* rhs = nullptr; };          // include/systems/explicit_system.h:124
class ImplicitSystem : public ExplicitSystem { public: SparseMatrix<Number>* matrix = nullptr; };  // include/systems/implicit_system.h:311
class LinearImplicitSystem : public ImplicitSystem {};
class EquationSystems {                                                             // include/systems/equation_systems.h
 public:
  template <typename S> S& get_system(const std::string&) { return *static_cast<S*>(_sys); }    // :757
  System* _sys = nullptr;
};
}  // namespace libMesh
#endif
