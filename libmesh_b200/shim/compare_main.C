// compare_main.C — the compare-mode program for a box that has a built libMesh (see compare_mode.h for the
// build line).  NOT compiled in this repository's build: libMesh cannot be built here (empty TIMPI / MetaPhysicL
// submodules), and the mock has no Mesh / FE classes.  It sets up introduction_ex4's problem with libMesh's own classes
// (mesh, dof map, Dirichlet boundary), lets the reference's assemble_poisson fill the stock matrix and the libfeasm hook
// fill a GpuSparseMatrix, and prints the per-entry comparison.
//
//   ./compare_ex4 -n 20            # BASELINE config 1: build_cube 20^3 HEX27
#ifndef FEASM_SHIM_MOCK
#include "libmesh/dirichlet_boundaries.h"
#include "libmesh/equation_systems.h"
#include "libmesh/getpot.h"
#include "libmesh/libmesh.h"
#include "libmesh/linear_implicit_system.h"
#include "libmesh/mesh.h"
#include "libmesh/mesh_generation.h"
#include "libmesh/wrapped_function.h"

#include "compare_mode.h"

using namespace libMesh;

// the reference's own functions, from introduction_ex4.C / exact_solution.C compiled next to this file (-Dmain=ex4_main)
void assemble_poisson(EquationSystems& es, const std::string& system_name);
Real exact_solution(const Real x, const Real y, const Real z = 0.);
static void exact_solution_wrapper(DenseVector<Number>& output, const Point& p, const Real) { output(0) = exact_solution(p(0), p(1), p(2)); }

int main(int argc, char** argv) {
  LibMeshInit init(argc, argv);
  GetPot cl(argc, argv);
  const int n = cl.follow(20, "-n");
  Mesh mesh(init.comm());
  MeshTools::Generation::build_cube(mesh, n, n, n, -1., 1., -1., 1., -1., 1., HEX27);  // introduction_ex4.C:180-215
  EquationSystems es(mesh);
  LinearImplicitSystem& system = es.add_system<LinearImplicitSystem>("Poisson");
  const unsigned int u_var = system.add_variable("u", SECOND, LAGRANGE);
  system.add_matrix<GpuSparseMatrix>("feasm");       // shares the DofMap's sparsity Graph (need_full_sparsity_pattern)
  std::set<boundary_id_type> ids = {0, 1, 2, 3, 4, 5};
  std::vector<unsigned int> vars = {u_var};
  WrappedFunction<Number> g(system, exact_solution_wrapper);
  system.get_dof_map().add_dirichlet_boundary(DirichletBoundary(ids, vars, g));  // introduction_ex4.C:242-274
  es.init();
  auto& A_gpu = static_cast<GpuSparseMatrix<Number>&>(system.get_matrix("feasm"));
  GpuNumericVector<Number> b_gpu(init.comm(), system.n_dofs());
  auto b_ref = system.rhs->zero_clone();
  const auto rep = feasm_libmesh::compare_assembly(es, "Poisson", assemble_poisson, *system.matrix, *b_ref,
                                                   feasm_libmesh::assemble_poisson_ex4, A_gpu, b_gpu);
  rep.print("introduction_ex4, reference assemble_poisson vs libfeasm");
  return rep.passes() ? 0 : 1;
}
#endif
