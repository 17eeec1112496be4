// ref_ex6_glue.cpp — the reference's own examples/systems_of_equations/systems_of_equations_ex6/systems_of_equations_ex6.C
// compiled where it lies (this file is one #include of it; oracle/Makefile puts its directory on the include path) over the stub
// world of ref_app/ref_world.h, plus one entry point that runs its LinearElasticity::assemble().  Test infrastructure.
// The class LinearElasticity is local to that file, which is why the entry point lives in the same translation unit.
#define main ref_ex6_main
#include "systems_of_equations_ex6.C"
#undef main

// LinearElasticity::assemble() set up the way the example's main() does (:380-440: "Elasticity" system, assemble object attached,
// variables u, v, w, ZeroFunction DirichletBoundary in LOCAL_VARIABLE_ORDER), on whatever mesh was loaded (main() asks for HEX8 FIRST;
// BASELINE's config 5 is HEX27 SECOND) and with Dirichlet conditions on the boundary ids of `dirichlet_mask` only — the node and edge
// ids main() adds (NODE_BOUNDARY_ID, EDGE_BOUNDARY_ID) are not part of the benchmark configuration.  The traction term is the
// example's own: sides with BOUNDARY_ID_MAX_X.
extern "C" int ref_ex6_assemble(int order, unsigned int dirichlet_mask, char* err, int err_len) {
  try {
    const RefWorldData& w = ref_world();
    const char* none[1] = {"ref_ex6_assemble"};
    LibMeshInit init(1, none);
    Mesh mesh(init.comm(), 3);
    MeshTools::Generation::build_cube(mesh, w.nx, w.ny, w.nz, w.bounds[0], w.bounds[1], w.bounds[2], w.bounds[3], w.bounds[4], w.bounds[5],
                                      static_cast<ElemType>(w.elem_type));
    EquationSystems equation_systems(mesh);
    LinearImplicitSystem& system = equation_systems.add_system<LinearImplicitSystem>("Elasticity");
    LinearElasticity le(equation_systems);
    system.attach_assemble_object(le);
    unsigned int u_var = system.add_variable("u", static_cast<Order>(order), LAGRANGE);
    unsigned int v_var = system.add_variable("v", static_cast<Order>(order), LAGRANGE);
    unsigned int w_var = system.add_variable("w", static_cast<Order>(order), LAGRANGE);
    std::set<boundary_id_type> ids;
    for (int b = 0; b < 8; ++b) if (dirichlet_mask & (1u << b)) ids.insert((boundary_id_type)b);
    ZeroFunction<> zf;
    DirichletBoundary dirichlet_bc(ids, {u_var, v_var, w_var}, zf, LOCAL_VARIABLE_ORDER);
    system.get_dof_map().add_dirichlet_boundary(dirichlet_bc);
    equation_systems.init();
    system.solve();
    return 0;
  } catch (const std::exception& e) {
    if (err && err_len > 0) {
      int n = 0;
      for (const char* c = e.what(); *c && n < err_len - 1; ++c) err[n++] = *c;
      err[n] = 0;
    }
    return -1;
  }
}
