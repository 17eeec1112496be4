"""ctypes access to oracle/_ref/libref_fe.so — the REFERENCE's own shape-function and Gauss-quadrature sources
(src/fe/fe_lagrange_shape_{2,3}D.C, include/fe/fe_lagrange_shape_1D.h, src/quadrature/quadrature*.C), compiled where they
lie under /root/reference by oracle/Makefile.  TEST INFRASTRUCTURE: used by tests/test_oracle_ref.py and
tests/golden/make_ref_golden.py to check the oracle's restatement against the reference code itself; the library exists only
where /root/reference does (this container), the golden tables it wrote are committed and travel.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_fe.so")

# libMesh's enum values (include/enums/enum_elem_type.h, enum_order.h)
ELEM = {"TRI6": 4, "QUAD4": 5, "QUAD8": 6, "QUAD9": 7, "TET10": 9, "HEX8": 10, "HEX20": 11, "HEX27": 12, "PRISM18": 15}
ELEM_ALL = dict(ELEM, TET4=8)
N_NODES = {"TRI6": 6, "QUAD4": 4, "QUAD8": 8, "QUAD9": 9, "TET10": 10, "HEX8": 8, "HEX20": 20, "HEX27": 27, "PRISM18": 18}
DIM = {"TRI6": 2, "QUAD4": 2, "QUAD8": 2, "QUAD9": 2, "TET10": 3, "HEX8": 3, "HEX20": 3, "HEX27": 3, "PRISM18": 3}
FE_ORDER = {"TRI6": 2, "QUAD4": 1, "QUAD8": 2, "QUAD9": 2, "TET10": 2, "HEX8": 1, "HEX20": 2, "HEX27": 2, "PRISM18": 2}

_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_lagrange_shapes.restype = C.c_int
        L.ref_lagrange_shapes.argtypes = [C.c_int] * 5 + [C.c_void_p] * 5 + [C.c_char_p, C.c_int]
        L.ref_gauss_rule.restype = C.c_int
        L.ref_gauss_rule.argtypes = [C.c_int] * 4 + [C.c_void_p] * 2 + [C.c_char_p, C.c_int]
        L.ref_fe_map.restype = C.c_int
        L.ref_fe_map.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_char_p, C.c_int]
        L.ref_exact_solution.restype = C.c_double
        L.ref_exact_solution.argtypes = [C.c_double] * 3
        _lib = L
    return _lib


def exact_solution(x: float, y: float, z: float) -> float:
    """examples/introduction/introduction_ex3/exact_solution.C."""
    return lib().ref_exact_solution(x, y, z)


def shapes(elem: str, ref_points):
    """(phi, dxi, deta, dzeta)[i][p] of FE<dim, LAGRANGE>::shape / shape_deriv for the element's own order (dzeta is zero in
    2D)."""
    pts = np.ascontiguousarray(ref_points, dtype=np.float64).reshape(-1, 3)
    n, npnt = N_NODES[elem], pts.shape[0]
    out = [np.zeros((n, npnt)) for _ in range(4)]
    err = C.create_string_buffer(512)
    rc = lib().ref_lagrange_shapes(DIM[elem], ELEM[elem], FE_ORDER[elem], n, npnt, pts.ctypes.data, *[o.ctypes.data for o in out], err, 512)
    if rc:
        raise RuntimeError(err.value.decode())
    return out


def gauss_rule(elem: str, order: int):
    """(points [nq][3], weights [nq]) of QGauss(dim, order).init(elem type)."""
    pts, w = np.zeros((1024, 3)), np.zeros(1024)
    err = C.create_string_buffer(512)
    n = lib().ref_gauss_rule(DIM[elem], ELEM[elem], int(order), 1024, pts.ctypes.data, w.ctypes.data, err, 512)
    if n < 0:
        raise RuntimeError(err.value.decode())
    return pts[:n].copy(), w[:n].copy()


def has_affine_map(elem: str, node_xyz) -> bool:
    """Hex8 / Hex20 / Hex27 / Tet10 / Prism18::has_affine_map of the reference (src/geom/cell_*.C)."""
    L = lib()
    L.ref_has_affine_map.restype = C.c_int
    L.ref_has_affine_map.argtypes = [C.c_int, C.c_int, C.c_void_p]
    X = np.ascontiguousarray(node_xyz, dtype=np.float64).reshape(N_NODES[elem], 3)
    rc = L.ref_has_affine_map(ELEM[elem], N_NODES[elem], X.ctypes.data)
    if rc < 0:
        raise RuntimeError(f"no has_affine_map for {elem}")
    return bool(rc)


def fe_map(elem: str, node_xyz, ref_points, weights, affine: bool):
    """FEMap::init_reference_to_physical_map + FEMap::compute_map (src/fe/fe_map.C) on one element — a 3D cell, or a 2D
    element embedded in space (QUAD4 / QUAD8 / QUAD9 / TRI6: the surface Jacobian) — dict(xyz [nq][3], JxW [nq], jac [nq],
    inv [nq][3][3] with inv[q][a][d] = d(xi_a)/d(x_d)).  `affine` stands in for Elem::has_affine_map()."""
    X = np.ascontiguousarray(node_xyz, dtype=np.float64).reshape(N_NODES[elem], 3)
    pts = np.ascontiguousarray(ref_points, dtype=np.float64).reshape(-1, 3)
    w = np.ascontiguousarray(weights, dtype=np.float64)
    nq = pts.shape[0]
    xyz, JxW, jac, inv = np.zeros((nq, 3)), np.zeros(nq), np.zeros(nq), np.zeros((nq, 9))
    err = C.create_string_buffer(1024)
    rc = lib().ref_fe_map(DIM[elem], ELEM[elem], N_NODES[elem], X.ctypes.data, int(bool(affine)), nq, pts.ctypes.data, w.ctypes.data,
                          xyz.ctypes.data, JxW.ctypes.data, jac.ctypes.data, inv.ctypes.data, err, 1024)
    if rc:
        raise RuntimeError(err.value.decode())
    return dict(xyz=xyz, JxW=JxW, jac=jac, inv=inv.reshape(nq, 3, 3))


# ---- oracle/_ref/libref_app.so: the reference's own APPLICATION source of the path (introduction_ex4.C) over the stub world of
# oracle/ref_app/ref_world.h, plus the constraint functions of src/base/dof_map_constraints.C (see that header for what is the
# reference's code and what is glue) -------------------------------------------------------------------------------------------------
APP_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_app.so")
_ORC_TO_LIBMESH = {8: "HEX8", 20: "HEX20", 27: "HEX27", 10: "TET10", 18: "PRISM18"}
APP_FAST_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_app_fast.so")
_apps = {}
_fast = False


def app_available(fast: bool = False) -> bool:
    return os.path.exists(APP_FAST_LIB_PATH if fast else APP_LIB_PATH)


def use_fast(on: bool) -> None:
    """Route the calls below to libref_app_fast.so — the same sources at -O3 -march=x86-64-v3, for timing (bench.py) — instead of
    the -O2 -ffp-contract=off build that is the bitwise checker."""
    global _fast
    _fast = bool(on)


def app():
    if _fast not in _apps:
        L = C.CDLL(APP_FAST_LIB_PATH if _fast else APP_LIB_PATH)
        L.ref_world_load.restype = C.c_int
        L.ref_world_load.argtypes = [C.c_int, C.c_int, C.c_uint, C.c_uint, C.c_uint, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_ex4_run.restype = C.c_int
        L.ref_ex4_run.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.c_char_p, C.c_int]
        L.ref_ex4_assemble.restype = C.c_int
        L.ref_ex4_assemble.argtypes = [C.c_int, C.c_uint, C.c_char_p, C.c_int]
        L.ref_ex6_assemble.restype = C.c_int
        L.ref_ex6_assemble.argtypes = [C.c_int, C.c_uint, C.c_char_p, C.c_int]
        L.ref_world_set_sides.restype = C.c_int
        L.ref_world_set_sides.argtypes = [C.c_int, C.c_void_p]
        L.ref_world_solve_seconds.restype = C.c_double
        L.ref_world_solve_seconds.argtypes = []
        L.ref_world_untouched_entries.restype = C.c_int64
        L.ref_world_untouched_entries.argtypes = []
        L.ref_world_set_constraint_rows.restype = C.c_int
        L.ref_world_set_constraint_rows.argtypes = [C.c_int64] + [C.c_void_p] * 5
        L.ref_world_results.restype = C.c_int64
        L.ref_world_results.argtypes = [C.c_void_p] * 4
        L.ref_side_reinit.restype = C.c_int
        L.ref_side_reinit.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.c_char_p, C.c_int]
        L.ref_constrain_element.restype = C.c_int
        L.ref_constrain_element.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int]
        _apps[_fast] = L
    return _apps[_fast]


def load_constraint_rows(dofs, row_ptr, cols, coefs, values):
    """Constraint rows for the loaded world, given as such (the argument list of feasm_set_constraint_rows /
    oracle.constraints.ConstraintRows.arrays()): hanging nodes, and Dirichlet rows as empty rows with a value."""
    a = [np.ascontiguousarray(dofs, dtype=np.uint32), np.ascontiguousarray(row_ptr, dtype=np.uint64), np.ascontiguousarray(cols, dtype=np.uint32),
         np.ascontiguousarray(coefs, dtype=np.float64), np.ascontiguousarray(values, dtype=np.float64)]
    pad = [np.zeros(1, dtype=x.dtype) if x.size == 0 else x for x in a]
    app().ref_world_set_constraint_rows(len(a[0]), *[x.ctypes.data for x in pad])


def load_world(mesh, system, n, bounds, affine=None, pattern=None):
    """Hands the stub world the integer structures of an oracle Mesh / System (oracle.py): the mesh that build_cube(n, n, n,
    bounds) is going to be asked for, the dof of every (variable, node), and the sparsity pattern (the system's own, or the
    one given: with constraint rows the Graph holds the couplings DofMap::find_connected_dofs adds)."""
    xyz, conn, mask = mesh.arrays()
    ed, _, _ = system.dofs()
    npe, nv, nn = mesh.npe, system.n_vars, mesh.n_nodes
    node_dof = np.zeros((nv, nn), dtype=np.uint32)
    for v in range(nv):
        node_dof[v, conn.reshape(-1)] = ed[:, v * npe:(v + 1) * npe].reshape(-1)
    if affine is None:  # the reference's own Hex27::has_affine_map ... decides, element by element
        affine = np.full(mesh.n_elem, 255, dtype=np.uint8)
    affine = np.ascontiguousarray(affine, dtype=np.uint8)
    rp, ci = system.sparsity() if pattern is None else pattern
    rp, ci = np.ascontiguousarray(rp, dtype=np.uint64), np.ascontiguousarray(ci, dtype=np.uint32)
    b = np.ascontiguousarray(bounds, dtype=np.float64)
    nx, ny, nz = (n, n, n) if np.isscalar(n) else n
    app().ref_world_load(ELEM[_ORC_TO_LIBMESH[mesh.elem_type]], npe, nx, ny, nz, b.ctypes.data, nn, xyz.ctypes.data, mesh.n_elem, conn.ctypes.data,
                         affine.ctypes.data, mask.ctypes.data, nv, system.n_dofs, node_dof.ctypes.data, rp.ctypes.data, ci.ctypes.data)
    return len(ci), system.n_dofs


def load_world_arrays(elem: str, dims, bounds, xyz, conn, node_bmask, node_dof, row_ptr, col_idx, affine=None):
    """The same from plain arrays — e.g. the mesh of build_cube() and the numbering of distribute_dofs() below, so that nothing the
    world is loaded with comes from the oracle: node_dof [n_vars][n_nodes], node_bmask [n_nodes] (bit b: on boundary id b)."""
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    conn = np.ascontiguousarray(conn, dtype=np.uint32)
    mask = np.ascontiguousarray(node_bmask, dtype=np.uint8)
    node_dof = np.ascontiguousarray(node_dof, dtype=np.uint32).reshape(-1, xyz.shape[0])
    affine = np.full(conn.shape[0], 255, dtype=np.uint8) if affine is None else np.ascontiguousarray(affine, dtype=np.uint8)
    rp, ci = np.ascontiguousarray(row_ptr, dtype=np.uint64), np.ascontiguousarray(col_idx, dtype=np.uint32)
    b = np.ascontiguousarray(bounds, dtype=np.float64)
    n_dofs = len(rp) - 1
    app().ref_world_load(ELEM[elem], conn.shape[1], dims[0], dims[1], dims[2], b.ctypes.data, xyz.shape[0], xyz.ctypes.data, conn.shape[0], conn.ctypes.data,
                         affine.ctypes.data, mask.ctypes.data, node_dof.shape[0], n_dofs, node_dof.ctypes.data, rp.ctypes.data, ci.ctypes.data)
    return len(ci), n_dofs


def _results(nnz, n_dofs):
    values, rhs = np.zeros(nnz), np.zeros(n_dofs)
    nc = app().ref_world_results(None, None, None, None)
    cd, cv = np.zeros(nc, dtype=np.uint32), np.zeros(nc)
    app().ref_world_results(values.ctypes.data, rhs.ctypes.data, cd.ctypes.data, cv.ctypes.data)
    return dict(values=values, rhs=rhs, con_dof=cd, con_val=cv, untouched_entries=int(app().ref_world_untouched_entries()))


def assemble_ex4(order, nnz, n_dofs, boundary_ids=(0, 1, 2, 3, 4, 5)):
    """assemble_poisson of introduction_ex4.C on the loaded mesh, set up as its main() does (a "Poisson" system, one LAGRANGE
    variable of `order`, DirichletBoundary with the exact solution on `boundary_ids`) — for the meshes main() does not build."""
    mask = 0
    for b in boundary_ids:
        mask |= 1 << b
    err = C.create_string_buffer(1024)
    if app().ref_ex4_assemble(int(order), mask, err, 1024) != 0:
        raise RuntimeError(err.value.decode())
    return _results(nnz, n_dofs)


def load_sides(mesh):
    """BoundaryInfo's side ids of a build_cube mesh (oracle Mesh.boundary_sides) for the loaded world: what
    Elem::neighbor_ptr(side) == nullptr and BoundaryInfo::has_boundary_id answer from."""
    ns = {8: 6, 20: 6, 27: 6, 10: 4, 18: 5}[mesh.elem_type]
    bid = np.full((mesh.n_elem, ns), -1, dtype=np.int16)
    for b in range(6):
        e, sd = mesh.boundary_sides(b)
        bid[e, sd] = b
    app().ref_world_set_sides(ns, bid.ctypes.data)


def assemble_ex6(order, nnz, n_dofs, dirichlet_ids=(4,)):
    """LinearElasticity::assemble() of systems_of_equations_ex6.C on the loaded mesh, set up as its main() does (three LAGRANGE
    variables of `order`, ZeroFunction DirichletBoundary on `dirichlet_ids`; traction on the sides with BOUNDARY_ID_MAX_X as the
    example has it)."""
    mask = 0
    for b in dirichlet_ids:
        mask |= 1 << b
    err = C.create_string_buffer(1024)
    if app().ref_ex6_assemble(int(order), mask, err, 1024) != 0:
        raise RuntimeError(err.value.decode())
    return _results(nnz, n_dofs)


def last_solve_seconds() -> float:
    """Wall-clock seconds of the system.solve() (zero + the example's assembly function) of the last assemble_ex4."""
    return float(app().ref_world_solve_seconds())


def run_ex4(args, nnz, n_dofs):
    """main() of examples/introduction/introduction_ex4/introduction_ex4.C with the command line `args` ("-d 3 -n 4 ...") over the
    loaded world: dict(values, rhs, con_dof, con_val) as the run left them (system.solve() assembles; the linear solve is not on
    the path)."""
    argv = [b"introduction_ex4"] + [str(a).encode() for a in args]
    arr = (C.c_char_p * len(argv))(*argv)
    err = C.create_string_buffer(1024)
    rc = app().ref_ex4_run(len(argv), arr, err, 1024)
    if rc != 0:
        raise RuntimeError(f"introduction_ex4 returned {rc}: {err.value.decode()}")
    return _results(nnz, n_dofs)


def constrain_element(K, F, dofs, rows, values=None, heterogeneous=True, asymmetric=True):
    """DofMap::(heterogeneously_)constrain_element_matrix_and_vector of the reference (src/base/dof_map_constraints.C:2439-2614,
    with build_constraint_matrix(_and_vector) :3433-3680) on one element.  rows: {constrained dof: {dof: coefficient}} (an empty
    dict is a Dirichlet row); values: {constrained dof: value}.  Returns (K', F', dofs') — the dof set grows by the constraining
    dofs that are not the element's."""
    K = np.ascontiguousarray(K, dtype=np.float64)
    F = np.ascontiguousarray(F, dtype=np.float64)
    d = np.ascontiguousarray(dofs, dtype=np.uint32)
    n = len(d)
    cd = np.array(sorted(rows), dtype=np.uint32)
    ptr = np.zeros(len(cd) + 1, dtype=np.uint64)
    rd, rc_ = [], []
    for i, c in enumerate(cd):
        for k in sorted(rows[int(c)]):
            rd.append(k)
            rc_.append(rows[int(c)][k])
        ptr[i + 1] = len(rd)
    rd = np.array(rd + [0], dtype=np.uint32)
    rcoef = np.array(rc_ + [0.0], dtype=np.float64)
    cv = np.array([(values or {}).get(int(c), 0.0) for c in cd] + [0.0], dtype=np.float64)
    cap = n + len(rd) + 8
    do, Ko, Fo = np.zeros(cap, dtype=np.uint32), np.zeros(cap * cap), np.zeros(cap)
    err = C.create_string_buffer(1024)
    m = app().ref_constrain_element(n, K.ctypes.data, F.ctypes.data, d.ctypes.data, len(cd), cd.ctypes.data, ptr.ctypes.data, rd.ctypes.data, rcoef.ctypes.data,
                                    cv.ctypes.data, int(heterogeneous), int(asymmetric), cap, do.ctypes.data, Ko.ctypes.data, Fo.ctypes.data, err, 1024)
    if m < 0:
        raise RuntimeError(err.value.decode())
    return Ko[: m * m].reshape(m, m).copy(), Fo[:m].copy(), do[:m].copy()


def side_reinit(elem: str, node_xyz, side: int, qorder: int, affine: bool):
    """FE::reinit(elem, side) with the reference's own pieces — QGauss(2), FEMap::init_face_shape_functions / compute_face_map
    (cut out of src/fe/fe_boundary.C), FEAbstract::get_refspace_nodes, the side_nodes_map tables of src/geom/cell_*.C, the 3D shape
    functions and FEMap at the mapped points — in the call order of fe_boundary.C:109-238: dict(ref, xyz, JxW, normals, phi)."""
    n = N_NODES[elem]
    X = np.ascontiguousarray(node_xyz, dtype=np.float64).reshape(n, 3)
    cap = 64
    ref, xyz, JxW, nrm, phi = np.zeros((cap, 3)), np.zeros((cap, 3)), np.zeros(cap), np.zeros((cap, 3)), np.zeros(n * cap)
    err = C.create_string_buffer(1024)
    nqf = app().ref_side_reinit(ELEM[elem], n, X.ctypes.data, int(bool(affine)), int(side), FE_ORDER[elem], int(qorder), cap, ref.ctypes.data, xyz.ctypes.data,
                                JxW.ctypes.data, nrm.ctypes.data, phi.ctypes.data, err, 1024)
    if nqf < 0:
        raise RuntimeError(err.value.decode())
    return dict(ref=ref[:nqf], xyz=xyz[:nqf], JxW=JxW[:nqf], normals=nrm[:nqf], phi=phi[: n * nqf].reshape(n, nqf))


# ---- oracle/_ref/libref_mesh.so: the reference's own MeshTools::Generation::build_cube over a recording mesh
# (oracle/ref_mesh/ref_mesh_world.h) -------------------------------------------------------------------------------------------------------
MESH_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_mesh.so")
_mesh_lib = None


def mesh_available() -> bool:
    return os.path.exists(MESH_LIB_PATH)


def build_cube(elem: str, nx: int, ny: int, nz: int, bounds):
    """MeshTools::Generation::build_cube(mesh, nx, ny, nz, bounds, type) of the reference (src/mesh/mesh_generation.C:317-1938)
    followed by its ReplicatedMesh::renumber_nodes_and_elements (src/mesh/replicated_mesh.C:700-857; the numbering step of
    prepare_for_use).  dict with
      before = (xyz [lattice nodes][3], conn [n_elem][npe]) as the generator left them when it called prepare_for_use,
      xyz, conn, elem_id                                     after the renumbering (orphaned lattice nodes removed),
      sides = (element position, side, boundary id)          in the order the generator added them,
      second_order_requests, prepared.
    For TET10 `before` holds the 24 TET4 per cell that the reference's own all_second_order then raises."""
    global _mesh_lib
    if _mesh_lib is None:
        L = C.CDLL(MESH_LIB_PATH)
        L.ref_build_cube.restype = C.c_int
        L.ref_build_cube.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_void_p, C.c_int, C.c_char_p, C.c_int]
        L.ref_mesh_sizes.argtypes = [C.c_void_p]
        L.ref_mesh_get.argtypes = [C.c_void_p] * 6
        L.ref_mesh_before_sizes.argtypes = [C.c_void_p]
        L.ref_mesh_before.argtypes = [C.c_void_p] * 2
        _mesh_lib = L
    b = np.ascontiguousarray(bounds, dtype=np.float64)
    err = C.create_string_buffer(1024)
    if _mesh_lib.ref_build_cube(nx, ny, nz, b.ctypes.data, ELEM_ALL[elem], err, 1024) != 0:
        raise RuntimeError(err.value.decode())
    sz = np.zeros(6, dtype=np.int64)
    _mesh_lib.ref_mesh_sizes(sz.ctypes.data)
    nn, ne, npe, ns = (int(x) for x in sz[:4])
    xyz, conn, eid = np.zeros((nn, 3)), np.zeros((ne, npe), dtype=np.uint32), np.zeros(ne, dtype=np.uint32)
    se, sn, sb = np.zeros(ns, dtype=np.uint32), np.zeros(ns, dtype=np.uint8), np.zeros(ns, dtype=np.int16)
    _mesh_lib.ref_mesh_get(xyz.ctypes.data, conn.ctypes.data, eid.ctypes.data, se.ctypes.data, sn.ctypes.data, sb.ctypes.data)
    bs = np.zeros(2, dtype=np.int64)
    _mesh_lib.ref_mesh_before_sizes(bs.ctypes.data)
    xb, cb = np.zeros((int(bs[0]), 3)), np.zeros(int(bs[1]), dtype=np.uint32)
    _mesh_lib.ref_mesh_before(xb.ctypes.data, cb.ctypes.data)
    return dict(xyz=xyz, conn=conn, elem_id=eid, sides=(se, sn, sb), second_order_requests=int(sz[4]), prepared=bool(sz[5]),
                before=(xb, cb.reshape(ne, len(cb) // ne if ne else 0)))


def distribute_dofs(n_vars: int, n_nodes: int):
    """DofMap::distribute_local_dofs_var_major of the reference (src/base/dof_map.C:1417-1535) on the mesh of the last build_cube,
    for one group of n_vars LAGRANGE variables of the element's order: (n_dofs, dof_base [n_nodes]); the dof of variable v at a
    node is dof_base + v."""
    _mesh_lib.ref_distribute_dofs.restype = C.c_int64
    _mesh_lib.ref_distribute_dofs.argtypes = [C.c_uint, C.c_void_p, C.c_char_p, C.c_int]
    base = np.zeros(n_nodes, dtype=np.uint32)
    err = C.create_string_buffer(1024)
    n = _mesh_lib.ref_distribute_dofs(n_vars, base.ctypes.data, err, 1024)
    if n < 0:
        raise RuntimeError(err.value.decode())
    return int(n), base


def partition_linear(n_parts: int, n_elem: int):
    """LinearPartitioner::partition_range of the reference (src/partitioning/linear_partitioner.C:36-117) over all elements of the
    mesh of the last build_cube: elem_pid [n_elem]."""
    _mesh_lib.ref_partition_linear.restype = C.c_int
    _mesh_lib.ref_partition_linear.argtypes = [C.c_uint, C.c_void_p, C.c_char_p, C.c_int]
    pid = np.zeros(n_elem, dtype=np.int32)
    err = C.create_string_buffer(1024)
    if _mesh_lib.ref_partition_linear(n_parts, pid.ctypes.data, err, 1024) != 0:
        raise RuntimeError(err.value.decode())
    return pid
