// ref_stubs/libmesh/libmesh_common.h — STUB written for this repository (test infrastructure, not reference code).
// The reference's own libmesh_common.h needs the generated libmesh_config.h and TIMPI, neither of which exists here
// (SURVEY 8c).  These stubs supply only what the reference's shape-function and quadrature sources name — Real, the
// assertion / error macros, a three-component Point — so that those sources can be compiled WHERE THEY LIE under
// /root/reference (oracle/Makefile, target _ref) and called by the tests as the checker of the oracle's restatement.
#ifndef REF_STUB_LIBMESH_COMMON_H
#define REF_STUB_LIBMESH_COMMON_H
#include <cmath>
#include <complex>
#include <iomanip>
#include <type_traits>
#include <cstddef>
#include <iostream>
#include <memory>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#define LIBMESH_DIM 3
#define LIBMESH_ENABLE_SECOND_DERIVATIVES 1
#define LIBMESH_DEFAULT_DOUBLE_PRECISION 1
#define LIBMESH_USE_REAL_NUMBERS 1

namespace libMesh {
typedef double Real;  // include/base/libmesh_common.h:144 (default configuration)
typedef Real Number;
typedef std::complex<Real> Complex;
// libmesh_common.h:189-250, :717-718
template <typename T> inline T libmesh_real(T a) { return a; }
template <typename T> inline T libmesh_conj(T a) { return a; }
template <typename T> inline T libmesh_real(std::complex<T> a) { return std::real(a); }
template <typename T> inline std::complex<T> libmesh_conj(std::complex<T> a) { return std::conj(a); }
template <typename T> inline bool isnan(std::complex<T> a) { return std::isnan(std::real(a)) || std::isnan(std::imag(a)); }
template <typename T> inline bool isinf(std::complex<T> a) { return std::isinf(std::real(a)) || std::isinf(std::imag(a)); }
template <class T> constexpr std::false_type always_false{};
typedef unsigned int dof_id_type;
struct LogicError : std::runtime_error { using std::runtime_error::runtime_error; };
struct NotImplemented : std::runtime_error { using std::runtime_error::runtime_error; };
struct DegenerateMap : std::runtime_error { using std::runtime_error::runtime_error; };        // libmesh_exceptions.h
struct ConvergenceFailure : std::runtime_error { using std::runtime_error::runtime_error; };
namespace Threads {  // include/parallel/threads.h: one thread here
struct spin_mutex { struct scoped_lock { explicit scoped_lock(spin_mutex&) {} }; };
}
template <typename... Args> inline void libmesh_ignore(const Args&...) {}
template <typename Tnew, typename Told> inline Tnew cast_ptr(Told* p) { return static_cast<Tnew>(p); }
template <typename Tnew, typename Told> inline Tnew cast_int(Told v) { return static_cast<Tnew>(v); }
inline constexpr Real operator""_R(long double x) { return static_cast<Real>(x); }  // libmesh_common.h: Real literals
static constexpr Real TOLERANCE = 1.e-6;  // libmesh_common.h
static std::ostream& out = std::cout;  // libMesh::out / libMesh::err (include/base/libmesh_base.h)
static std::ostream& err = std::cerr;
static const unsigned int invalid_uint = static_cast<unsigned int>(-1);
// include/utils/int_range.h: the index loops `for (auto i : index_range(v))` / `make_range(n)` of the reference's sources
template <typename T>
struct IntRange {
  struct iterator {
    T i;
    T operator*() const { return i; }
    iterator& operator++() { ++i; return *this; }
    bool operator!=(const iterator& o) const { return i != o.i; }
  };
  IntRange(T b, T e) : _b(b), _e(e) {}
  iterator begin() const { return {_b}; }
  iterator end() const { return {_e}; }
  T _b, _e;
};
template <typename V> inline IntRange<std::size_t> index_range(const V& v) { return IntRange<std::size_t>(0, v.size()); }
template <typename T> inline IntRange<T> make_range(T b, T e) { return IntRange<T>(b, e); }
template <typename T> inline IntRange<T> make_range(T e) { return IntRange<T>(T(0), e); }
}  // namespace libMesh

#define libmesh_error_msg(msg) do { std::ostringstream ref_os__; ref_os__ << msg; throw libMesh::LogicError(ref_os__.str()); } while (0)
#define libmesh_error_msg_if(cond, msg) do { if (cond) libmesh_error_msg(msg); } while (0)
#define libmesh_error() libmesh_error_msg("libmesh_error")
#define libmesh_not_implemented() throw libMesh::NotImplemented("not implemented")
#define libmesh_assert(x) ((void)0)
#define libmesh_assert_msg(x, m) ((void)0)
#define libmesh_assert_less(a, b) ((void)0)
#define libmesh_assert_less_equal(a, b) ((void)0)
#define libmesh_assert_equal_to(a, b) ((void)0)
#define libmesh_assert_greater(a, b) ((void)0)
#define libmesh_assert_greater_equal(a, b) ((void)0)
#define libmesh_assert_not_equal_to(a, b) ((void)0)
#define libmesh_degenerate_mapping_msg(msg) do { std::stringstream ref_ms__; ref_ms__ << msg << '\n'; throw libMesh::DegenerateMap(ref_ms__.str()); } while (0)
#define libmesh_try try
#define libmesh_catch(e) catch (e)
#define libmesh_here() ((void)0)
#define LIBMESH_EXPORT
#define LOG_SCOPE(a, b) ((void)0)
#define libmesh_dbg_var(var)  /* NDEBUG form, libmesh_common.h:307-312 */
#define libmesh_fallthrough() [[fallthrough]]
#define libmesh_experimental() ((void)0)
#define libmesh_deprecated() ((void)0)
#endif
