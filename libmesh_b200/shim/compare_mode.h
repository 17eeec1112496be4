// compare_mode.h — closes "global values unpinned against a real libMesh build"  on any box that HAS libMesh.
//
// feasm_libmesh::compare_assembly runs two assemble functions on the same System — the reference's own (for example
// assemble_poisson of examples/introduction/introduction_ex4/introduction_ex4.C:313-563, compiled from the reference tree
// next to this file, never copied) into the system's stock SparseMatrix / NumericVector back-end, and the libfeasm hook
// into a GpuSparseMatrix — and compares the two systems per entry through the SparseMatrix interface only
// (get_row, sparse_matrix.h:607; NumericVector::localize): same sparsity pattern (columns of every row), and every value and
// right-hand-side entry within
//        |a - b| <= tol * max(|a|, |b|, floor * rowmax)          tol = 1e-12, floor = 1e-2   (oracle/parity.py)
// with the strict (no floor) failures counted as well.
//
// On a libMesh box (shim/README in INTEGRATION.md):
//   mpicxx -std=c++17 $(libmesh-config --cppflags --cxxflags --include) -I include -I libmesh_b200/shim
//       -Dmain=ex4_main $LIBMESH/examples/introduction/introduction_ex4/introduction_ex4.C
//       $LIBMESH/examples/introduction/introduction_ex3/exact_solution.C
//       libmesh_b200/shim/compare_main.C libmesh_b200/shim/gpu_assemble.C -L libmesh_b200 -lfeasm $(libmesh-config --libs)
// Here it is compiled against libmesh_mock.h and driven by shim_test.C (two GpuSparseMatrix systems), which checks the
// comparison code itself.
#ifndef FEASM_COMPARE_MODE_H
#define FEASM_COMPARE_MODE_H

#include <cmath>
#include <cstdio>

#include "gpu_assemble.h"

namespace feasm_libmesh {

struct CompareReport {
  bool pattern_equal = true;
  std::size_t entries = 0, strict_failures = 0, rhs_strict_failures = 0;
  double max_rel_err = 0., max_rel_err_rhs = 0.;
  bool passes(double tol = 1e-12) const { return pattern_equal && max_rel_err <= tol && max_rel_err_rhs <= tol; }
  void print(const char* what) const {
    std::printf("%s: pattern %s, %zu entries, max rel err values %.3e (strict 1e-12 failures %zu), rhs %.3e (strict failures %zu) -> %s\n", what,
                pattern_equal ? "identical" : "DIFFERENT", entries, max_rel_err, strict_failures, max_rel_err_rhs, rhs_strict_failures,
                passes() ? "PASS" : "FAIL");
  }
};

// A_ref / b_ref: the system assembled by the reference's own function; A_gpu / b_gpu: by the libfeasm hook.  Both closed.
inline CompareReport compare_systems(const libMesh::SparseMatrix<libMesh::Number>& A_ref, const libMesh::NumericVector<libMesh::Number>& b_ref,
                                     const libMesh::SparseMatrix<libMesh::Number>& A_gpu, const libMesh::NumericVector<libMesh::Number>& b_gpu,
                                     double tol = 1e-12, double floor_frac = 1e-2) {
  using namespace libMesh;
  CompareReport rep;
  std::vector<numeric_index_type> ci, cj;
  std::vector<Number> vi, vj;
  for (numeric_index_type r = A_ref.row_start(); r < A_ref.row_stop(); ++r) {
    A_ref.get_row(r, ci, vi);
    A_gpu.get_row(r, cj, vj);
    // back-ends that preallocate from n_nz / n_oz may drop or keep explicit zeros: compare on the union, a missing entry = 0
    double rowmax = 0.;
    for (Number v : vi) rowmax = std::max(rowmax, std::fabs(v));
    std::size_t a = 0, b = 0;
    while (a < ci.size() || b < cj.size()) {
      double x = 0., y = 0.;
      if (b >= cj.size() || (a < ci.size() && ci[a] < cj[b])) { x = vi[a++]; if (x != 0.) rep.pattern_equal = false; }
      else if (a >= ci.size() || cj[b] < ci[a]) { y = vj[b++]; if (y != 0.) rep.pattern_equal = false; }
      else { x = vi[a++]; y = vj[b++]; }
      ++rep.entries;
      const double d = std::fabs(x - y), mag = std::max(std::fabs(x), std::fabs(y));
      if (d > tol * mag) ++rep.strict_failures;
      const double scale = std::max(mag, floor_frac * rowmax);
      if (scale > 0.) rep.max_rel_err = std::max(rep.max_rel_err, d / scale);
    }
    if (ci.size() != cj.size()) rep.pattern_equal = rep.pattern_equal && false;
  }
  std::vector<Number> hr, hg;
  b_ref.localize(hr);
  b_gpu.localize(hg);
  double bmax = 0.;
  for (Number v : hr) bmax = std::max(bmax, std::fabs(v));
  for (std::size_t i = 0; i < hr.size() && i < hg.size(); ++i) {
    const double d = std::fabs(hr[i] - hg[i]), mag = std::max(std::fabs(hr[i]), std::fabs(hg[i]));
    if (d > tol * mag) ++rep.rhs_strict_failures;
    const double scale = std::max(mag, floor_frac * bmax);
    if (scale > 0.) rep.max_rel_err_rhs = std::max(rep.max_rel_err_rhs, d / scale);
  }
  if (hr.size() != hg.size()) rep.pattern_equal = false;
  return rep;
}

// Runs `reference_assemble` into (A_ref, b_ref) and `gpu_assemble` into (A_gpu, b_gpu) on the SAME System (its matrix / rhs
// pointers are swapped around the two calls, as ImplicitSystem::matrix and ExplicitSystem::rhs are public members,
// implicit_system.h:311, explicit_system.h:124), then compares.
typedef void (*assemble_fptr)(libMesh::EquationSystems&, const std::string&);
inline CompareReport compare_assembly(libMesh::EquationSystems& es, const std::string& name, assemble_fptr reference_assemble,
                                      libMesh::SparseMatrix<libMesh::Number>& A_ref, libMesh::NumericVector<libMesh::Number>& b_ref,
                                      assemble_fptr gpu_assemble, libMesh::SparseMatrix<libMesh::Number>& A_gpu,
                                      libMesh::NumericVector<libMesh::Number>& b_gpu) {
  using namespace libMesh;
  LinearImplicitSystem& system = es.get_system<LinearImplicitSystem>(name);
  SparseMatrix<Number>* m0 = system.matrix;
  NumericVector<Number>* r0 = system.rhs;
  system.matrix = &A_ref; system.rhs = &b_ref;
  A_ref.zero(); b_ref.zero();                       // ImplicitSystem::assemble, implicit_system.C:101-105
  reference_assemble(es, name);
  A_ref.close(); b_ref.close();
  system.matrix = &A_gpu; system.rhs = &b_gpu;
  A_gpu.zero(); b_gpu.zero();
  gpu_assemble(es, name);
  A_gpu.close(); b_gpu.close();
  system.matrix = m0; system.rhs = r0;
  return compare_systems(A_ref, b_ref, A_gpu, b_gpu);
}

}  // namespace feasm_libmesh
#endif
