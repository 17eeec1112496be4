// ref_app/ref_world.h — STUB world (test infrastructure, written for this repository; not reference code).
//
// Purpose: compile the reference's own APPLICATION sources for the hot path — examples/introduction/introduction_ex4/
// introduction_ex4.C (assemble_poisson and main) — WHERE THEY LIE under /root/reference and run them, so that the oracle's
// global matrix and right-hand side are checked against numbers produced by the reference's assembly function itself
// (oracle/Makefile target _ref/libref_app.so, tests/test_oracle_ref_app.py).  libMesh as a whole cannot be built here
// (SURVEY 8c); this header supplies, under the reference's header names (one-line forwarders generated in _ref/inc_app/),
// the classes the example names — LibMeshInit, Mesh, EquationSystems, LinearImplicitSystem, DofMap, FEBase, SparseMatrix,
// NumericVector, PerfLog, DirichletBoundary ... — as data holders over arrays the test loads (mesh, dof numbering, sparsity
// pattern: the integer structures, which come from the oracle and are pinned by the reference's published run facts).
//
// What is the reference's own code in that library, and what is glue:
//   reference, compiled unchanged   introduction_ex4.C (the element loop, the finite-difference forcing, main's control flow),
//                                   exact_solution.C, QGauss and its tables, the Lagrange shape functions, FEMap
//                                   (init_reference_to_physical_map, compute_map), DenseMatrix / DenseVector
//                                   (include/numerics/dense_*.h: left_multiply_transpose, right_multiply, vector_mult ...);
//                                   and, excerpted by the Makefile at build time from the files where they lie (function
//                                   bodies, into a build intermediate that is deleted after the compile):
//                                   DofMap::build_constraint_matrix(_and_vector), DofMap::constrain_element_matrix_and_vector,
//                                   DofMap::heterogeneously_constrain_element_matrix_and_vector
//                                   (src/base/dof_map_constraints.C) and H1FETransformation::map_dphi
//                                   (src/fe/h1_fe_transformation.C)
//   glue, restated here             the order of calls inside FE::reinit (src/fe/fe.C:196-368), dof_indices as a table lookup,
//                                   SparseMatrix::add_matrix / NumericVector::add_vector as the (i, j) loop of
//                                   src/numerics/laspack_matrix.C:259-275 / numeric_vector.C:690-697 into a CSR array,
//                                   DirichletBoundary on LAGRANGE variables as nodal evaluation of the boundary function
#ifndef REF_WORLD_H
#define REF_WORLD_H
#include <cstdint>
#include <cstring>
#include <map>
#include <set>

#include "libmesh/libmesh_common.h"
#include "libmesh/elem.h"
#include "libmesh/fe.h"
#include "libmesh/fe_map.h"
#include "libmesh/quadrature.h"
#include "libmesh/dense_matrix.h"  // the reference's own headers (forwarded from _ref/inc_app)
#include "libmesh/dense_vector.h"
#include "libmesh/enum_solver_package.h"

#define LIBMESH_ENABLE_DIRICHLET 1
#define LIBMESH_ENABLE_CONSTRAINTS 1
#define libmesh_example_requires(cond, msg) do { if (!(cond)) { libMesh::out << "requires " << msg << std::endl; return 77; } } while (0)
#define LOG_SCOPE_WITH(a, b, log) ((void)0)
#define LOG_SCOPE_IF(a, b, c) ((void)0)

namespace libMesh {

typedef short boundary_id_type;  // include/base/id_types.h (default LIBMESH_BOUNDARY_ID_BYTES = 2)
static const Real pi = static_cast<Real>(3.1415926535897932384626433832795029L);  // include/base/libmesh.h
static const Number zero = 0.;

class EquationSystems;
class MeshBase;
// include/numerics/vector_value.h, tensor_value.h over this tree's three-component Point / RealTensor
template <typename T> class VectorValue : public Point {
public:
  using Point::Point;
  VectorValue(const Point& p) : Point(p) {}
};
template <typename T> class TensorValue : public RealTensor {
public:
  using RealTensor::RealTensor;
};
namespace Utility {
template <int N, typename T> inline T pow(const T& x) { T r = 1; for (int i = 0; i < N; ++i) r *= x; return r; }  // include/utils/utility.h
}

// ---- what the test loads before the example runs (ref_app_shim.cpp: ref_world_load) -------------------------------------------------
struct RefWorldData {
  // mesh: what MeshTools::Generation::build_cube is asked for, and the result the oracle's generator gave for it
  int elem_type = 0, npe = 0;
  unsigned int nx = 0, ny = 0, nz = 0;
  double bounds[6] = {0, 0, 0, 0, 0, 0};
  std::vector<double> xyz;               // [n_nodes][3]
  std::vector<uint32_t> conn;            // [n_elem][npe]
  std::vector<uint8_t> affine;           // [n_elem]  255: Elem::has_affine_map() is the element class's own test (the default); 0 / 1 force a branch
  std::vector<uint8_t> node_bmask;       // [n_nodes] bit b: the node lies on a side with boundary id b
  int n_sides = 0;
  std::vector<int16_t> side_bid;         // [n_elem][n_sides] BoundaryInfo's id of a boundary side, -1 for an interior side (empty: none loaded)
  // dof numbering: dof of (variable, node), variable-major as DofMap::dof_indices returns them
  int n_vars = 0;
  uint64_t n_dofs = 0;
  std::vector<uint32_t> node_dof;        // [n_vars][n_nodes]
  // sparsity pattern
  std::vector<uint64_t> row_ptr;
  std::vector<uint32_t> col_idx;
  // constraint rows given as such (hanging nodes, and Dirichlet rows as empty rows with a value): u_dof = sum coef u_col + val;
  // what DofMap::create_dof_constraints would have computed from the mesh's refinement tree, which this world does not hold
  std::vector<uint32_t> crow_dof, crow_col;
  std::vector<uint64_t> crow_ptr;
  std::vector<double> crow_coef, crow_val;
  // results
  std::vector<double> values, rhs;
  std::vector<uint8_t> touched;          // [nnz] 1 once SparseMatrix::add_matrix has added into the entry (zeros count: sparse_matrix.h:300-303)
  std::vector<uint32_t> con_dof;         // the constrained dofs and their values, as EquationSystems::init left them
  std::vector<double> con_val;
  std::vector<std::string> args;         // the command line LibMeshInit saw
  double solve_seconds = 0.;
  uint64_t n_nodes() const { return xyz.size() / 3; }
  uint64_t n_elem() const { return npe ? conn.size() / npe : 0; }
};
RefWorldData& ref_world();

// ---- include/base/libmesh.h ----------------------------------------------------------------------------------------------------------
namespace Parallel { class Communicator {}; }
class LibMeshInit {
public:
  LibMeshInit(int argc, const char* const* argv) { ref_world().args.assign(argv, argv + argc); }
  const Parallel::Communicator& comm() const { return _comm; }
private:
  Parallel::Communicator _comm;
};
inline SolverPackage default_solver_package() { return LASPACK_SOLVERS; }
// command_line_next (src/base/libmesh.C): the value after the named option, or the default
template <typename T>
inline T command_line_next(const std::string& name, T dflt) {
  const std::vector<std::string>& a = ref_world().args;
  for (std::size_t i = 0; i + 1 < a.size(); ++i)
    if (a[i] == name) { std::istringstream is(a[i + 1]); T v = dflt; is >> v; return v; }
  return dflt;
}

// ---- include/mesh/*.h ------------------------------------------------------------------------------------------------------------------
class BoundaryInfo {  // include/mesh/boundary_info.h, as far as an assembly loop asks
public:
  bool has_boundary_id(const Elem* const elem, const unsigned short int side, const boundary_id_type id) const { return elem->side_boundary_id(side) == id; }
  // node and edge ids: set by main() of systems_of_equations_ex6.C, which is compiled but not run here
  void add_node(const Node*, const boundary_id_type) { libmesh_not_implemented(); }
  void add_edge(const Elem*, const unsigned short int, const boundary_id_type) { libmesh_not_implemented(); }
  void regenerate_id_sets() {}
};
class MeshBase {
public:
  virtual ~MeshBase() = default;
  unsigned int mesh_dimension() const { return 3; }
  const std::vector<const Elem*>& active_local_element_ptr_range() const { return _ptrs; }
  const std::vector<const Elem*>& element_ptr_range() const { return _ptrs; }
  const BoundaryInfo& get_boundary_info() const { return _boundary_info; }
  BoundaryInfo& get_boundary_info() { return _boundary_info; }
  BoundaryInfo _boundary_info;
  void print_info() const {}
  std::vector<Elem> _elems;
  std::vector<const Elem*> _ptrs;
};
class UnstructuredMesh : public MeshBase {};
class Mesh : public UnstructuredMesh {
public:
  explicit Mesh(const Parallel::Communicator&, unsigned char = 1) {}
};
namespace MeshTools { namespace Generation {
// hands out the loaded mesh after checking that it IS the mesh asked for (include/mesh/mesh_generation.h:70-86)
void build_cube(UnstructuredMesh& mesh, const unsigned int nx, const unsigned int ny = 0, const unsigned int nz = 0, const Real xmin = 0., const Real xmax = 1.,
                const Real ymin = 0., const Real ymax = 1., const Real zmin = 0., const Real zmax = 1., const ElemType type = INVALID_ELEM,
                const bool gauss_lobatto_grid = false);
} }
class GnuPlotIO {
public:
  enum PlottingProperties { GRID_ON = 1 };
  GnuPlotIO(const MeshBase&, const std::string&, int) {}
  void write_equation_systems(const std::string&, const EquationSystems&) {}
};

// ---- include/numerics/sparse_matrix.h, numeric_vector.h: the two calls of the element loop ------------------------------------------
template <typename T>
class SparseMatrix {
public:
  // LaspackMatrix::add_matrix, src/numerics/laspack_matrix.C:259-275: for i, for j: add(rows[i], cols[j], dm(i, j))
  void add_matrix(const DenseMatrix<T>& dm, const std::vector<dof_id_type>& dof_indices) {
    RefWorldData& w = ref_world();
    const unsigned int n = (unsigned int)dof_indices.size();
    for (unsigned int i = 0; i < n; ++i)
      for (unsigned int j = 0; j < n; ++j) {
        const uint64_t r = dof_indices[i], b = w.row_ptr[r], e = w.row_ptr[r + 1];
        const uint32_t* p = std::lower_bound(w.col_idx.data() + b, w.col_idx.data() + e, (uint32_t)dof_indices[j]);
        libmesh_error_msg_if(p == w.col_idx.data() + e || *p != dof_indices[j], "entry (" << r << ", " << dof_indices[j] << ") is not in the sparsity pattern");
        w.values[p - w.col_idx.data()] += dm(i, j);
        w.touched[p - w.col_idx.data()] = 1;
      }
  }
  void zero() { std::fill(ref_world().values.begin(), ref_world().values.end(), 0.); }
};
template <typename T>
class NumericVector {
public:
  // NumericVector::add_vector, src/numerics/numeric_vector.C:690-697: for i: add(dof_indices[i], v(i))
  void add_vector(const DenseVector<T>& v, const std::vector<dof_id_type>& dof_indices) {
    for (std::size_t i = 0; i < dof_indices.size(); ++i) ref_world().rhs[dof_indices[i]] += v((unsigned int)i);
  }
  void zero() { std::fill(ref_world().rhs.begin(), ref_world().rhs.end(), 0.); }
  // named by compute_stresses() of systems_of_equations_ex6.C (post-processing, not on the path): compiled, not run
  dof_id_type first_local_index() const { return 0; }
  dof_id_type last_local_index() const { return 0; }
  void set(const dof_id_type, const T) { libmesh_not_implemented(); }
  void close() {}
};

class PerfLog {  // include/utils/perf_log.h: timing only
public:
  explicit PerfLog(const std::string& = "") {}
  void push(const char*, const char* = "") {}
  void pop(const char*, const char* = "") {}
};

// ---- include/numerics/function_base.h, analytic_function.h, zero_function.h; include/base/dirichlet_boundaries.h -------------------
template <typename Output = Number>
class FunctionBase {
public:
  virtual ~FunctionBase() = default;
  virtual void operator()(const Point& p, const Real time, DenseVector<Output>& output) = 0;
  virtual std::unique_ptr<FunctionBase<Output>> clone() const = 0;
};
template <typename Output = Number>
class AnalyticFunction : public FunctionBase<Output> {
public:
  typedef void (*VectorFunctionPointer)(DenseVector<Output>& output, const Point& p, const Real time);
  AnalyticFunction(VectorFunctionPointer f) : _vector_fptr(f) {}
  void operator()(const Point& p, const Real time, DenseVector<Output>& output) override { _vector_fptr(output, p, time); }  // analytic_function.h
  std::unique_ptr<FunctionBase<Output>> clone() const override { return std::make_unique<AnalyticFunction<Output>>(_vector_fptr); }
private:
  VectorFunctionPointer _vector_fptr;
};
template <typename Output = Number>
class ZeroFunction : public FunctionBase<Output> {
public:
  void operator()(const Point&, const Real, DenseVector<Output>& output) override { output.zero(); }
  std::unique_ptr<FunctionBase<Output>> clone() const override { return std::make_unique<ZeroFunction<Output>>(); }
};
enum VariableIndexing { SYSTEM_VARIABLE_ORDER = 0, LOCAL_VARIABLE_ORDER };
class DirichletBoundary {
public:
  DirichletBoundary(std::set<boundary_id_type> b_in, std::vector<unsigned int> variables_in, const FunctionBase<Number>& f_in,
                    VariableIndexing type = SYSTEM_VARIABLE_ORDER)
      : b(std::move(b_in)), variables(std::move(variables_in)), f(f_in.clone()), f_system_order(type == SYSTEM_VARIABLE_ORDER) {}
  DirichletBoundary(const DirichletBoundary& o) : b(o.b), variables(o.variables), f(o.f->clone()), f_system_order(o.f_system_order) {}
  std::set<boundary_id_type> b;
  std::vector<unsigned int> variables;
  std::unique_ptr<FunctionBase<Number>> f;
  bool f_system_order;
};

// ---- include/base/dof_map.h ----------------------------------------------------------------------------------------------------------
class DofMap {
public:
  typedef std::map<dof_id_type, Real> DofConstraintRow;                 // dof_map.h:97-123 (ordered by dof id)
  typedef std::map<dof_id_type, DofConstraintRow> DofConstraints;
  typedef std::map<dof_id_type, Number> DofConstraintValueMap;
  typedef std::map<unsigned int, DofConstraintValueMap> AdjointDofConstraintValues;

  const FEType& variable_type(const unsigned int v) const { return _var_types.at(v); }
  // a table lookup here: the numbering itself is the oracle's (pinned by the published n_dofs / bandwidth figures)
  void dof_indices(const Elem* const elem, std::vector<dof_id_type>& di) const;
  void dof_indices(const Elem* const elem, std::vector<dof_id_type>& di, const unsigned int vn) const;
  bool is_constrained_dof(const dof_id_type dof) const { return _dof_constraints.count(dof) != 0; }  // dof_map.h (inline)
  void add_dirichlet_boundary(const DirichletBoundary& d) { _dirichlet.emplace_back(d); }
  // DofMap::create_dof_constraints for DirichletBoundary objects on LAGRANGE variables: the boundary function at the nodes
  void create_dof_constraints();

  // DECLARATIONS only (dof_map.h:1351-1396, :2027-2050).  The DEFINITIONS are the reference's own, excerpted from
  // src/base/dof_map_constraints.C at build time (oracle/Makefile) and compiled in ref_app_shim.cpp.
  void build_constraint_matrix(DenseMatrix<Number>& C, std::vector<dof_id_type>& elem_dofs, const bool called_recursively = false) const;
  void build_constraint_matrix_and_vector(DenseMatrix<Number>& C, DenseVector<Number>& H, std::vector<dof_id_type>& elem_dofs, int qoi_index = -1,
                                          const bool called_recursively = false) const;
  void constrain_element_matrix_and_vector(DenseMatrix<Number>& matrix, DenseVector<Number>& rhs, std::vector<dof_id_type>& elem_dofs,
                                           bool asymmetric_constraint_rows = true) const;
  void heterogeneously_constrain_element_matrix_and_vector(DenseMatrix<Number>& matrix, DenseVector<Number>& rhs, std::vector<dof_id_type>& elem_dofs,
                                                           bool asymmetric_constraint_rows = true, int qoi_index = -1) const;
  void heterogenously_constrain_element_matrix_and_vector(DenseMatrix<Number>& matrix, DenseVector<Number>& rhs, std::vector<dof_id_type>& elem_dofs,
                                                          bool asymmetric_constraint_rows = true, int qoi_index = -1) const {
    return this->heterogeneously_constrain_element_matrix_and_vector(matrix, rhs, elem_dofs, asymmetric_constraint_rows, qoi_index);  // dof_map.h:1388-1396
  }

  DofConstraints _dof_constraints;
  DofConstraintValueMap _primal_constraint_values;
  AdjointDofConstraintValues _adjoint_constraint_values;
  std::vector<FEType> _var_types;
  std::vector<DirichletBoundary> _dirichlet;
};

// ---- include/fe/fe_base.h, h1_fe_transformation.h ---------------------------------------------------------------------------------------
template <typename OutputShape> class FEGenericBase;
template <typename OutputShape>
class H1FETransformation {
public:
  // DECLARATION only (h1_fe_transformation.h:88-95); the definition is the reference's, excerpted from
  // src/fe/h1_fe_transformation.C:63-192 at build time
  void map_dphi(const unsigned int dim, const Elem* const elem, const std::vector<Point>& qp, const FEGenericBase<OutputShape>& fe,
                std::vector<std::vector<typename FEGenericBase<OutputShape>::OutputGradient>>& dphi, std::vector<std::vector<OutputShape>>& dphidx,
                std::vector<std::vector<OutputShape>>& dphidy, std::vector<std::vector<OutputShape>>& dphidz) const;
};
template <typename OutputShape>
class FEGenericBase {
public:
  typedef RealGradient OutputGradient;
  static std::unique_ptr<FEGenericBase<OutputShape>> build(const unsigned int dim, const FEType& type);
  FEGenericBase(unsigned int dim, const FEType& t);
  void attach_quadrature_rule(QBase* q) { qrule = q; }
  const std::vector<Real>& get_JxW() const { return _fe_map.get_JxW(); }
  const std::vector<Point>& get_xyz() const { return _fe_map.get_xyz(); }
  const std::vector<std::vector<OutputShape>>& get_phi() const { return phi; }
  const std::vector<std::vector<OutputGradient>>& get_dphi() const { return dphi; }
  const std::vector<std::vector<OutputShape>>& get_dphidxi() const { return dphidxi; }
  const std::vector<std::vector<OutputShape>>& get_dphideta() const { return dphideta; }
  const std::vector<std::vector<OutputShape>>& get_dphidzeta() const { return dphidzeta; }
  const FEMap& get_fe_map() const { return _fe_map; }
  FEType get_fe_type() const { return fe_type; }
  // FE<Dim, T>::reinit(elem), src/fe/fe.C:196-368, in the order of its calls: qrule->init, init_reference_to_physical_map,
  // init_shape_functions (the tables by FE<3, LAGRANGE>::all_shapes / all_shape_derivs), compute_map, map_dphi
  void reinit(const Elem* elem);
  // FE<Dim, T>::reinit(elem, side), src/fe/fe_boundary.C:109-238, in the order of its calls: build_side_ptr, qrule->init(side),
  // init_face_shape_functions, compute_face_map, side_map (:347-400), reinit(elem, &qp), JxW of the side put back
  void reinit(const Elem* elem, const unsigned int side);
  const std::vector<Point>& get_normals() const { return _fe_map.get_normals(); }
  const std::vector<Point>& side_reference_points() const { return _side_ref_points; }  // what side_map gave (for the tests)
private:
  void init_shape_functions(const std::vector<Point>& qp, const Elem* elem);
  unsigned int dim;
  FEType fe_type;
  ElemType _elem_type = INVALID_ELEM, last_side = INVALID_ELEM;
  QBase* qrule = nullptr;
  FEMap _fe_map;
  std::unique_ptr<H1FETransformation<OutputShape>> _fe_trans;  // fe_base.h: std::unique_ptr<FETransformationBase<OutputType>>
  std::vector<std::vector<OutputShape>> phi, dphidxi, dphideta, dphidzeta, dphidx, dphidy, dphidz;
  std::vector<std::vector<OutputGradient>> dphi;
  std::vector<Point> _side_ref_points;
};
typedef FEGenericBase<Real> FEBase;

// ---- include/systems/*.h ---------------------------------------------------------------------------------------------------------------
class System {
public:
  class Assembly {
  public:
    virtual ~Assembly() = default;
    virtual void assemble() = 0;
  };
  System(EquationSystems& es, const std::string& name) : _es(es), _name(name) {}
  virtual ~System() = default;
  unsigned int add_variable(const std::string& var, const Order order = FIRST, const FEFamily family = LAGRANGE) {
    _var_names.push_back(var);
    _dof_map._var_types.emplace_back((int)order, family);
    return (unsigned int)_var_names.size() - 1;
  }
  unsigned short variable_number(const std::string& var) const {
    for (std::size_t i = 0; i < _var_names.size(); ++i) if (_var_names[i] == var) return (unsigned short)i;
    libmesh_error_msg("no variable " << var);
  }
  unsigned int n_vars() const { return (unsigned int)_var_names.size(); }
  const DofMap& get_dof_map() const { return _dof_map; }
  DofMap& get_dof_map() { return _dof_map; }
  const std::string& name() const { return _name; }
  Number current_solution(const dof_id_type) const { libmesh_not_implemented(); }  // post-processing only
  void update() {}
  NumericVector<Number>* solution = &_solution;
  void attach_assemble_function(void fptr(EquationSystems& es, const std::string& name)) { _assemble_fptr = fptr; }
  void attach_assemble_object(Assembly& a) { _assemble_object = &a; }
  // System::user_assembly, src/systems/system.C
  void user_assembly() {
    if (_assemble_fptr) _assemble_fptr(_es, _name);
    else if (_assemble_object) _assemble_object->assemble();
  }
protected:
  EquationSystems& _es;
  std::string _name;
  std::vector<std::string> _var_names;
  DofMap _dof_map;
  NumericVector<Number> _solution;
  void (*_assemble_fptr)(EquationSystems&, const std::string&) = nullptr;
  Assembly* _assemble_object = nullptr;
};
class ExplicitSystem : public System {
public:
  using System::System;
  NumericVector<Number>* rhs = &_rhs;  // include/systems/explicit_system.h:124
private:
  NumericVector<Number> _rhs;
};
class ImplicitSystem : public ExplicitSystem {
public:
  using ExplicitSystem::ExplicitSystem;
  SparseMatrix<Number>& get_system_matrix() { return _matrix; }
  // ImplicitSystem::assemble, src/systems/implicit_system.C:101-105: zero the matrix and the rhs, then the user's function
  void assemble() { _matrix.zero(); rhs->zero(); this->user_assembly(); }
private:
  SparseMatrix<Number> _matrix;
};
class LinearImplicitSystem : public ImplicitSystem {
public:
  using ImplicitSystem::ImplicitSystem;
  void solve() { this->assemble(); }  // the linear solve itself is not on the path
};

class EquationSystems {
public:
  explicit EquationSystems(MeshBase& m) : _mesh(m) {}
  template <typename T> T& add_system(const std::string& name) {
    _systems.emplace_back(std::make_unique<T>(*this, name));
    return *static_cast<T*>(_systems.back().get());
  }
  template <typename T> T& get_system(const std::string& name) {
    for (auto& s : _systems) if (s->name() == name) return *cast_ptr<T*>(s.get());
    libmesh_error_msg("no system " << name);
  }
  const MeshBase& get_mesh() const { return _mesh; }
  MeshBase& get_mesh() { return _mesh; }
  void init();  // distributes dofs (the loaded numbering) and creates the constraints of the Dirichlet boundaries
  void print_info() const {}
private:
  MeshBase& _mesh;
  std::vector<std::unique_ptr<System>> _systems;
};

namespace Utility {
template <typename T> T string_to_enum(const std::string& s);  // include/utils/string_to_enum.h
template <> inline Order string_to_enum<Order>(const std::string& s) {
  if (s == "FIRST") return FIRST;
  if (s == "SECOND") return SECOND;
  libmesh_error_msg("order " << s);
}
template <> inline FEFamily string_to_enum<FEFamily>(const std::string& s) {
  if (s == "LAGRANGE") return LAGRANGE;
  libmesh_error_msg("family " << s);
}
}  // namespace Utility

}  // namespace libMesh
#endif
