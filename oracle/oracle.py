"""ctypes binding of the CPU oracle (oracle/fe_oracle.h).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  The product package (libmesh_b200/) must never import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))

HEX8, HEX27, TET10, HEX20, PRISM18 = 8, 27, 10, 20, 18
POISSON_EX4, ELASTICITY_EX6, LAPLACE = 1, 2, 3
FORCING_FD, FORCING_TABULATED = 0, 1

_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def build(force: bool = False) -> None:
    """Compile oracle/_build/liboracle*.so with the committed Makefile (g++ only)."""
    if force:
        subprocess.check_call(["make", "-C", _DIR, "clean"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", _DIR], stdout=subprocess.DEVNULL)


def _load(fast: bool):
    name = "liboracle_fast.so" if fast else "liboracle.so"
    path = os.path.join(_DIR, "_build", name)
    if not os.path.exists(path):
        build()
    lib = C.CDLL(path)
    vp = C.c_void_p
    sigs = {
        "orc_quadrature": (C.c_int, [C.c_int, C.c_int, C.c_int, vp, vp]),
        "orc_shapes": (C.c_int, [C.c_int, C.c_int, _f64p, vp, vp, vp, vp]),
        "orc_fe_reinit": (C.c_int, [C.c_int, _f64p, C.c_int, _f64p, _f64p, vp, vp, vp, vp]),
        "orc_has_affine_map": (C.c_int, [C.c_int, _f64p]),
        "orc_side_reinit": (C.c_int, [C.c_int, _f64p, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp]),
        "orc_mesh_boundary_sides": (C.c_int64, [vp, C.c_int, C.c_int64, vp, vp]),
        "orc_system_set_side_loads": (None, [vp, C.c_int64, vp, vp, vp]),
        "orc_system_set_side_penalty": (None, [vp, C.c_int64, vp, vp, C.c_double, C.c_int, C.c_double]),
        "orc_mesh_build_cube": (vp, [C.c_int] * 4 + [C.c_double] * 6 + [C.c_int]),
        "orc_mesh_from_arrays": (vp, [C.c_int, C.c_int64, C.c_int64, vp, vp, vp]),
        "orc_mesh_free": (None, [vp]),
        "orc_mesh_n_elem": (C.c_int64, [vp]),
        "orc_mesh_n_nodes": (C.c_int64, [vp]),
        "orc_mesh_nodes_per_elem": (C.c_int, [vp]),
        "orc_mesh_get": (None, [vp, vp, vp, vp]),
        "orc_mesh_skew": (None, [vp, C.c_int, C.c_double, C.c_uint64]),
        "orc_mesh_partition_linear": (None, [vp, C.c_int]),
        "orc_mesh_set_partition": (None, [vp, C.c_int, _i32p]),
        "orc_mesh_get_partition": (None, [vp, vp, vp]),
        "orc_system_create": (vp, [vp, C.c_int]),
        "orc_system_free": (None, [vp]),
        "orc_system_n_dofs": (C.c_int64, [vp]),
        "orc_system_dofs_per_elem": (C.c_int, [vp]),
        "orc_system_get_dofs": (None, [vp, vp, vp, vp]),
        "orc_system_add_dirichlet": (C.c_int64, [vp, C.c_uint32, C.c_int]),
        "orc_system_get_constraints": (None, [vp, vp, vp]),
        "orc_system_build_sparsity": (C.c_int64, [vp]),
        "orc_system_get_sparsity": (None, [vp, vp, vp]),
        "orc_element_matrices": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int64, C.c_int64, C.c_int, vp, vp]),
        "orc_forcing_at_qps": (C.c_int, [vp, C.c_int, vp, vp]),
        "orc_assemble": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, vp, vp, vp]),
        "orc_assemble_range": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int64, C.c_int64, vp, vp]),
        "orc_exact_solution": (C.c_double, [C.c_double] * 3),
        "orc_last_error": (C.c_char_p, []),
    }
    for k, (res, args) in sigs.items():
        fn = getattr(lib, k)
        fn.restype = res
        fn.argtypes = args
    return lib


_LIBS: dict = {}


def lib(fast: bool = False):
    if fast not in _LIBS:
        _LIBS[fast] = _load(fast)
    return _LIBS[fast]


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleError(RuntimeError):
    pass


def quadrature(elem_type: int, qorder: int):
    L = lib()
    nq = L.orc_quadrature(elem_type, qorder, 0, None, None)
    if nq < 0:
        raise OracleError(L.orc_last_error().decode())
    pts = np.zeros((nq, 3))
    w = np.zeros(nq)
    L.orc_quadrature(elem_type, qorder, nq, _ptr(pts), _ptr(w))
    return pts, w


def shapes(elem_type: int, ref_points: np.ndarray):
    pts = np.ascontiguousarray(ref_points, dtype=np.float64).reshape(-1, 3)
    n, npnt = elem_type, pts.shape[0]
    out = [np.zeros((n, npnt)) for _ in range(4)]
    lib().orc_shapes(elem_type, npnt, pts, *[_ptr(o) for o in out])
    return out


def fe_reinit(elem_type: int, node_xyz: np.ndarray, ref_points: np.ndarray, weights: np.ndarray):
    X = np.ascontiguousarray(node_xyz, dtype=np.float64).reshape(elem_type, 3)
    pts = np.ascontiguousarray(ref_points, dtype=np.float64).reshape(-1, 3)
    w = np.ascontiguousarray(weights, dtype=np.float64)
    npnt = pts.shape[0]
    xyz, JxW = np.zeros((npnt, 3)), np.zeros(npnt)
    phi, dphi = np.zeros((elem_type, npnt)), np.zeros((elem_type, npnt, 3))
    rc = lib().orc_fe_reinit(elem_type, X, npnt, pts, w, _ptr(xyz), _ptr(JxW), _ptr(phi), _ptr(dphi))
    if rc < 0:
        raise OracleError(lib().orc_last_error().decode())
    return dict(affine=bool(rc), xyz=xyz, JxW=JxW, phi=phi, dphi=dphi)


def side_reinit(elem_type: int, node_xyz: np.ndarray, side: int, qorder: int):
    """FE::reinit(elem, side): dict(ref, xyz, JxW, normals, phi[n][nqf])."""
    X = np.ascontiguousarray(node_xyz, dtype=np.float64).reshape(elem_type, 3)
    cap = 9
    ref, xyz, JxW, nrm, phi = np.zeros((cap, 3)), np.zeros((cap, 3)), np.zeros(cap), np.zeros((cap, 3)), np.zeros(elem_type * cap)
    nqf = lib().orc_side_reinit(elem_type, X, side, qorder, cap, _ptr(ref), _ptr(xyz), _ptr(JxW), _ptr(nrm), _ptr(phi))
    if nqf < 0:
        raise OracleError(lib().orc_last_error().decode())
    return dict(ref=ref[:nqf], xyz=xyz[:nqf], JxW=JxW[:nqf], normals=nrm[:nqf], phi=phi[: elem_type * nqf].reshape(elem_type, nqf))


def has_affine_map(elem_type: int, node_xyz: np.ndarray) -> bool:
    X = np.ascontiguousarray(node_xyz, dtype=np.float64).reshape(elem_type, 3)
    return bool(lib().orc_has_affine_map(elem_type, X))


def exact_solution(x, y, z):
    return lib().orc_exact_solution(x, y, z)


class Mesh:
    """build_cube mesh (src/mesh/mesh_generation.C)."""

    def __init__(self, elem_type, nx, ny, nz, bounds=(0.0, 1.0, 0.0, 1.0, 0.0, 1.0), renumber_nodes=True, fast=False):
        self.L = lib(fast)
        self.fast = fast
        self.h = self.L.orc_mesh_build_cube(elem_type, nx, ny, nz, *[float(b) for b in bounds], int(renumber_nodes))
        if not self.h:
            raise OracleError(self.L.orc_last_error().decode())
        self.elem_type = elem_type
        self.n_parts = 1

    @classmethod
    def from_arrays(cls, elem_type, xyz, conn, mask=None, fast=False):
        """A mesh given by its arrays (locally refined meshes with hanging nodes: what build_cube cannot make)."""
        self = cls.__new__(cls)
        self.L, self.fast = lib(fast), fast
        xyz = np.ascontiguousarray(xyz, dtype=np.float64)
        conn = np.ascontiguousarray(conn, dtype=np.uint32)
        mask = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
        self.h = self.L.orc_mesh_from_arrays(elem_type, conn.shape[0], xyz.shape[0], _ptr(xyz), _ptr(conn), None if mask is None else _ptr(mask))
        if not self.h:
            raise OracleError(self.L.orc_last_error().decode())
        self.elem_type, self.n_parts = elem_type, 1
        return self

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_mesh_free(self.h)
            self.h = None

    @property
    def n_elem(self):
        return self.L.orc_mesh_n_elem(self.h)

    @property
    def n_nodes(self):
        return self.L.orc_mesh_n_nodes(self.h)

    @property
    def npe(self):
        return self.L.orc_mesh_nodes_per_elem(self.h)

    def arrays(self):
        xyz = np.zeros((self.n_nodes, 3))
        conn = np.zeros((self.n_elem, self.npe), dtype=np.uint32)
        mask = np.zeros(self.n_nodes, dtype=np.uint8)
        self.L.orc_mesh_get(self.h, _ptr(xyz), _ptr(conn), _ptr(mask))
        return xyz, conn, mask

    def boundary_sides(self, boundary_id):
        """(elems, sides) carrying the boundary id (BoundaryInfo after build_cube), element-major order."""
        n = self.L.orc_mesh_boundary_sides(self.h, boundary_id, 0, None, None)
        if n < 0:
            raise OracleError("boundary sides: HEX8/HEX27 only")
        e, s = np.zeros(n, dtype=np.uint32), np.zeros(n, dtype=np.uint8)
        self.L.orc_mesh_boundary_sides(self.h, boundary_id, n, _ptr(e), _ptr(s))
        return e, s

    def skew(self, kind=0, amplitude=0.0, seed=12345):
        self.L.orc_mesh_skew(self.h, kind, float(amplitude), seed)

    def partition_linear(self, n_parts):
        self.L.orc_mesh_partition_linear(self.h, n_parts)
        self.n_parts = n_parts

    def set_partition(self, n_parts, elem_pid):
        self.L.orc_mesh_set_partition(self.h, n_parts, np.ascontiguousarray(elem_pid, dtype=np.int32))
        self.n_parts = n_parts

    def partition(self):
        ep = np.zeros(self.n_elem, dtype=np.int32)
        npid = np.zeros(self.n_nodes, dtype=np.int32)
        self.L.orc_mesh_get_partition(self.h, _ptr(ep), _ptr(npid))
        return ep, npid


class System:
    """DOF numbering + constraints + sparsity + assembly on a Mesh."""

    def __init__(self, mesh: Mesh, n_vars=1):
        self.mesh = mesh
        self.L = mesh.L
        self.h = self.L.orc_system_create(mesh.h, n_vars)
        self.n_vars = n_vars
        self.nnz = None

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_system_free(self.h)
            self.h = None

    @property
    def n_dofs(self):
        return self.L.orc_system_n_dofs(self.h)

    @property
    def nd(self):
        return self.L.orc_system_dofs_per_elem(self.h)

    def dofs(self):
        ed = np.zeros((self.mesh.n_elem, self.nd), dtype=np.uint32)
        nb = np.zeros(self.mesh.n_nodes, dtype=np.uint32)
        fd = np.zeros(self.mesh.n_parts + 1, dtype=np.uint32)
        self.L.orc_system_get_dofs(self.h, _ptr(ed), _ptr(nb), _ptr(fd))
        return ed, nb, fd

    def add_dirichlet(self, boundary_ids=(0, 1, 2, 3, 4, 5), kind=0):
        mask = 0
        for b in boundary_ids:
            mask |= 1 << b
        return self.L.orc_system_add_dirichlet(self.h, mask, kind)

    def constraints(self):
        n = self.L.orc_system_add_dirichlet(self.h, 0, 0)
        dofs = np.zeros(n, dtype=np.uint32)
        vals = np.zeros(n)
        self.L.orc_system_get_constraints(self.h, _ptr(dofs), _ptr(vals))
        return dofs, vals

    def set_side_loads(self, elems, sides, g):
        """systems_of_equations_ex6.C:240-257: traction g (one component per variable) on the listed sides."""
        e = np.ascontiguousarray(elems, dtype=np.uint32)
        sd = np.ascontiguousarray(sides, dtype=np.uint8)
        gg = np.zeros(3)
        gg[: self.n_vars] = np.asarray(g, dtype=np.float64).reshape(-1)[: self.n_vars]
        self.L.orc_system_set_side_loads(self.h, len(e), _ptr(e), _ptr(sd), _ptr(gg))

    def set_side_penalty(self, elems, sides, penalty=1e10, kind=1, value=0.0):
        """introduction_ex3.C:386-455: penalty boundary term on the listed sides (kind 1: exact solution, 0: constant)."""
        e = np.ascontiguousarray(elems, dtype=np.uint32)
        sd = np.ascontiguousarray(sides, dtype=np.uint8)
        self.L.orc_system_set_side_penalty(self.h, len(e), _ptr(e), _ptr(sd), float(penalty), int(kind), float(value))

    def build_sparsity(self):
        self.nnz = self.L.orc_system_build_sparsity(self.h)
        return self.nnz

    def sparsity(self):
        if self.nnz is None:
            self.build_sparsity()
        rp = np.zeros(self.n_dofs + 1, dtype=np.uint64)
        ci = np.zeros(self.nnz, dtype=np.uint32)
        self.L.orc_system_get_sparsity(self.h, _ptr(rp), _ptr(ci))
        return rp, ci

    def element_matrices(self, problem, e0=0, e1=None, constrained=True, qorder=0, forcing=FORCING_FD, f_tab=None):
        e1 = self.mesh.n_elem if e1 is None else e1
        nd = self.nd
        Ke = np.zeros((e1 - e0, nd, nd))
        Fe = np.zeros((e1 - e0, nd))
        rc = self.L.orc_element_matrices(self.h, problem, qorder, forcing, _ptr(f_tab), e0, e1, int(constrained), _ptr(Ke), _ptr(Fe))
        if rc < 0:
            raise OracleError(self.L.orc_last_error().decode())
        return Ke, Fe

    def forcing_at_qps(self, qorder=0, nq=None):
        if nq is None:
            nq = len(quadrature(self.mesh.elem_type, 5 if qorder == 0 else qorder)[1])
        f = np.zeros((self.mesh.n_elem, nq))
        xyz = np.zeros((self.mesh.n_elem, nq, 3))
        rc = self.L.orc_forcing_at_qps(self.h, qorder, _ptr(f), _ptr(xyz))
        if rc < 0:
            raise OracleError(self.L.orc_last_error().decode())
        return f, xyz

    def assemble(self, problem, n_threads=1, qorder=0, forcing=FORCING_FD, f_tab=None):
        if self.nnz is None:
            self.build_sparsity()
        values = np.zeros(self.nnz)
        rhs = np.zeros(self.n_dofs)
        phases = np.zeros(4)
        rc = self.L.orc_assemble(self.h, problem, qorder, forcing, _ptr(f_tab), n_threads, _ptr(values), _ptr(rhs), _ptr(phases))
        if rc < 0:
            raise OracleError(self.L.orc_last_error().decode())
        return values, rhs, phases

    def assemble_range(self, problem, e0, e1, values, rhs, n_threads=1, qorder=0, forcing=FORCING_FD, f_tab=None):
        rc = self.L.orc_assemble_range(self.h, problem, qorder, forcing, _ptr(f_tab), n_threads, e0, e1, _ptr(values), _ptr(rhs))
        if rc < 0:
            raise OracleError(self.L.orc_last_error().decode())
