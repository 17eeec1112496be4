// gpu_assemble.h — the batched assemble hooks a libMesh application attaches with
// System::attach_assemble_function (include/systems/system.h:1470; signature void(EquationSystems&, const std::string&),
// src/systems/system.C:1959-1972).  Each replaces the whole per-element loop of the named example with one call into
// libfeasm.  The system's matrix must be a GpuSparseMatrix (see gpu_sparse_matrix.h).
#ifndef FEASM_GPU_ASSEMBLE_H
#define FEASM_GPU_ASSEMBLE_H

#include <string>

#include "gpu_numeric_vector.h"
#include "gpu_sparse_matrix.h"
#ifndef FEASM_SHIM_MOCK
#include "libmesh/equation_systems.h"
#include "libmesh/linear_implicit_system.h"
#include "libmesh/boundary_info.h"
#endif

namespace feasm_libmesh {
// examples/introduction/introduction_ex4/introduction_ex4.C:313-563 (assemble_poisson)
void assemble_poisson_ex4(libMesh::EquationSystems& es, const std::string& system_name);
// the same loop with f = 0 (config 4)
void assemble_laplace(libMesh::EquationSystems& es, const std::string& system_name);
// A boundary integral  Fe_var[i](a) += JxW_face[qp] * (g[i] * phi_face[a][qp])  over the sides that have no neighbour
// and carry boundary id `id` (the traction loop of systems_of_equations_ex6.C:240-257); g has one entry per variable.
struct SideLoad {
  libMesh::boundary_id_type id;
  std::vector<libMesh::Real> g;
};
// Dirichlet data by the penalty method on every side without a neighbour (introduction_ex3.C:386-455):
//   Ke(i,j) += JxW_face*penalty*phi_face[i]*phi_face[j],  Fe(i) += JxW_face*penalty*value*phi_face[i];
// kind = FEASM_PENALTY_CONSTANT (value) or FEASM_PENALTY_EX4_EXACT (the exact solution at the side point).
struct SidePenalty {
  libMesh::Real penalty;
  int kind;
  libMesh::Real value;
};
// examples/systems_of_equations/systems_of_equations_ex6/systems_of_equations_ex6.C:156-264: volume terms and the
// traction (0, 0, -1) on BOUNDARY_ID_MAX_X (= 2, :79)
void assemble_elasticity_ex6(libMesh::EquationSystems& es, const std::string& system_name);
// generic entry: feasm_params as in include/feasm.h; side_load may be null
void assemble(libMesh::EquationSystems& es, const std::string& system_name, const feasm_params& params,
              const SideLoad* side_load = nullptr, const SidePenalty* side_penalty = nullptr);
}  // namespace feasm_libmesh
#endif
