#!/bin/sh
# build_ref.sh <reference root> <suffix> <optimisation flags> — builds oracle/_ref/libref_fe<suffix>.so and libref_app<suffix>.so
# (called by oracle/Makefile; TEST INFRASTRUCTURE).  The reference's sources are compiled WHERE THEY LIE under <reference root>:
#   _ref/inc/libmesh/<name>.h, _ref/inc_app/libmesh/<name>.h   one-line forwarders to the reference's own self-contained headers
#                                                              (or to the stub world); no reference source is copied
#   $GEN/                                                  a few function / table DEFINITIONS cut by ref_excerpt.py out of files
#                                                              that cannot be compiled whole: a build intermediate, removed below
# suffix ""      -O2 -ffp-contract=off   the bitwise checker of the oracle (operation order as written, no FMA contraction)
# suffix "_fast" -O3 -march=x86-64-v3    the same code as a timed data point beside the CPU baseline (bench.py)
set -e
REF="$1"; SFX="$2"; OPT="$3"
CXX="${CXX_REF:-/usr/bin/g++}"
cd "$(dirname "$0")"
OBJ="_ref/obj$SFX"
GEN="_ref/gen$SFX"
mkdir -p _ref/inc/libmesh _ref/inc_app/libmesh "$GEN" "$OBJ"
trap 'rm -rf "$GEN" "$OBJ"' EXIT

REF_HDR="enums/enum_elem_type.h enums/enum_order.h enums/enum_fe_family.h enums/enum_quadrature_type.h fe/fe_lagrange_shape_1D.h
 quadrature/quadrature.h quadrature/quadrature_gauss.h quadrature/quadrature_conical.h quadrature/quadrature_gm.h
 quadrature/quadrature_jacobi.h fe/fe_map.h"
REF_SRC="fe/fe_lagrange_shape_2D.C fe/fe_lagrange_shape_3D.C fe/fe_map.C quadrature/quadrature.C quadrature/quadrature_gauss.C
 quadrature/quadrature_gauss_1D.C quadrature/quadrature_gauss_2D.C quadrature/quadrature_gauss_3D.C quadrature/quadrature_conical.C
 quadrature/quadrature_conical_2D.C quadrature/quadrature_conical_3D.C quadrature/quadrature_gm.C quadrature/quadrature_gm_2D.C
 quadrature/quadrature_gm_3D.C quadrature/quadrature_jacobi.C quadrature/quadrature_jacobi_1D.C"
APP_REF_HDR="numerics/dense_matrix.h numerics/dense_matrix_impl.h numerics/dense_matrix_base.h numerics/dense_matrix_base_impl.h
 numerics/dense_vector.h numerics/dense_vector_base.h numerics/dense_submatrix.h numerics/dense_subvector.h numerics/tensor_tools.h
 utils/compare_types.h enums/enum_solver_package.h"
APP_WORLD_HDR="libmesh.h mesh.h mesh_generation.h exodusII_io.h gnuplot_io.h linear_implicit_system.h equation_systems.h dof_map.h
 sparse_matrix.h numeric_vector.h perf_log.h dirichlet_boundaries.h analytic_function.h string_to_enum.h getpot.h boundary_info.h
 zero_function.h solver_configuration.h petsc_linear_solver.h petsc_macro.h utility.h"
APP_REF_SRC="dense_matrix dense_matrix_base dense_matrix_blas_lapack dense_submatrix dense_subvector"

for h in $REF_HDR; do echo "#include \"$REF/include/$h\"" > _ref/inc/libmesh/$(basename $h); done
for h in $APP_REF_HDR; do echo "#include \"$REF/include/$h\"" > _ref/inc_app/libmesh/$(basename $h); done
for h in $APP_WORLD_HDR; do echo '#include "ref_world.h"' > _ref/inc_app/libmesh/$h; done
echo '#include "../../../ref_app/libmesh_config.h"' > _ref/inc_app/libmesh/libmesh_config.h

# ---- libref_fe: shape functions, quadrature, FEMap (volume and face maps), exact_solution
python3 ref_excerpt.py "$REF" fe $GEN/ref_fe_excerpts.inc
SRCS=""
for s in $REF_SRC; do SRCS="$SRCS $REF/src/$s"; done
$CXX -std=c++17 $OPT -fPIC -shared -I_ref/inc -Iref_stubs -I$GEN -o _ref/libref_fe$SFX.so ref_shim.cpp $SRCS \
    "$REF/examples/introduction/introduction_ex3/exact_solution.C"

# ---- libref_app: the examples (introduction_ex4.C: assemble_poisson and main; systems_of_equations_ex6.C: LinearElasticity::assemble),
# the reference's DenseMatrix sources and its DofMap constraint functions over the stub world of ref_app/ref_world.h
APP="$CXX -std=c++17 $OPT -fPIC -c -I_ref/inc_app -Iref_app -I$GEN -I_ref/inc -Iref_stubs"
python3 ref_excerpt.py "$REF" app $GEN/ref_excerpts.inc
$APP -o $OBJ/app_shim.o ref_app_shim.cpp
$APP -Dmain=ref_ex4_main -o $OBJ/ex4.o "$REF/examples/introduction/introduction_ex4/introduction_ex4.C"
$APP -I"$REF/examples/systems_of_equations/systems_of_equations_ex6" -o $OBJ/ex6.o ref_ex6_glue.cpp
OBJS="$OBJ/app_shim.o $OBJ/ex4.o $OBJ/ex6.o"
for s in $APP_REF_SRC; do $APP -o $OBJ/$s.o "$REF/src/numerics/$s.C"; OBJS="$OBJS $OBJ/$s.o"; done
$CXX -shared -o _ref/libref_app$SFX.so $OBJS -L_ref -lref_fe$SFX -Wl,-rpath,'$ORIGIN'

# ---- libref_mesh (checker build only): the reference's own MeshTools::Generation::build_cube over the recording mesh of
# ref_mesh/ref_mesh_world.h
if [ -z "$SFX" ]; then
  python3 ref_excerpt.py "$REF" mesh_idx $GEN/ref_mesh_idx.inc
  python3 ref_excerpt.py "$REF" mesh_so $GEN/ref_mesh_so_excerpts.inc
  python3 ref_excerpt.py "$REF" mesh $GEN/ref_mesh_excerpts.inc
  $CXX -std=c++17 $OPT -fPIC -shared -I_ref/inc -Iref_stubs -Iref_mesh -I$GEN -o _ref/libref_mesh.so ref_mesh_shim.cpp
fi
