"""ctypes binding of the libfeasm C ABI (include/feasm.h).

This is the only way Python reaches the product: every call goes through the same ``extern "C"`` entry
points a libMesh-side C++ adaptor binds.  There is no fallback: if ``libfeasm.so`` is
missing or no B200 is present, loading / ``Context()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FEASM_LIB") or os.path.join(_HERE, "libfeasm.so")  # FEASM_LIB: development builds

HEX8, HEX27, TET10, HEX20, PRISM18 = 8, 27, 10, 20, 18
POISSON, ELASTICITY = 1, 2
FORCING_NONE, FORCING_EX4_FD, FORCING_TABULATED = 0, 1, 2
CONSTRAIN_HETEROGENEOUS, CONSTRAIN_HOMOGENEOUS = 0, 1
DIRICHLET_ZERO, DIRICHLET_EX4_EXACT = 0, 1
PENALTY_CONSTANT, PENALTY_EX4_EXACT = 0, 1

# every symbol include/feasm.h declares (tests check the library exports each of them)
SYMBOLS = [
    "feasm_build_halo", "feasm_halo_columns", "feasm_set_halo_plan", "feasm_halo_pack", "feasm_halo_recv_buffer",
    "feasm_halo_unpack", "feasm_get_halo_plan",
    "feasm_create", "feasm_destroy", "feasm_last_error", "feasm_version", "feasm_set_mesh",
    "feasm_update_coordinates", "feasm_set_dofs", "feasm_set_constraints", "feasm_set_pattern",
    "feasm_build_pattern", "feasm_nnz", "feasm_get_pattern", "feasm_prepare", "feasm_assemble",
    "feasm_synchronize", "feasm_element_matrices", "feasm_zero", "feasm_add_matrix", "feasm_add_vector",
    "feasm_get_values", "feasm_device_values", "feasm_device_rhs", "feasm_stream", "feasm_set_exchange_plan",
    "feasm_set_side_loads", "feasm_assemble_begin", "feasm_assemble_end", "feasm_send_buffer", "feasm_pack_ghost", "feasm_recv_buffer", "feasm_unpack_ghost", "feasm_cube_sizes", "feasm_build_cube",
    "feasm_distribute_dofs", "feasm_last_timing", "feasm_measure_fp64_peaks", "feasm_vector_mult", "feasm_vector_mult_host", "feasm_cg_solve",
    "feasm_vec_alloc", "feasm_vec_free", "feasm_vec_copy", "feasm_vec_fill", "feasm_vec_axpy", "feasm_vec_pointwise",
    "feasm_vec_reduce", "feasm_vec_set_entry",
    "feasm_generate_cube", "feasm_distribute_dofs_device", "feasm_generate_dirichlet", "feasm_get_mesh", "feasm_get_dofs",
    "feasm_get_constraints", "feasm_set_side_penalty", "feasm_residual", "feasm_add_matrix_noalloc",
    "feasm_set_communicator", "feasm_comm_unique_id", "feasm_comm_init", "feasm_exchange_begin", "feasm_exchange_end",
    "feasm_exchange_ghost_rows", "feasm_comm_allgather_u64", "feasm_set_matrix", "feasm_init_matrix", "feasm_set_values", "feasm_set_interface_prefix",
    "feasm_build_exchange", "feasm_planner_create", "feasm_planner_destroy", "feasm_planner_error", "feasm_planner_message",
    "feasm_planner_receive", "feasm_planner_finish", "feasm_planner_sizes", "feasm_planner_get", "feasm_apply_plan", "feasm_get_exchange_plan",
    "feasm_set_constraint_rows",
]


class FeasmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libfeasm error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [
        ("problem", C.c_int), ("quadrature_order", C.c_int), ("forcing", C.c_int), ("constrain", C.c_int),
        ("f_host", C.c_void_p), ("f_device", C.c_void_p), ("young_modulus", C.c_double),
        ("poisson_ratio", C.c_double), ("body_force", C.c_double * 3), ("affine_tensor", C.c_int),
    ]


_lib = None


def lib():
    """Load libfeasm.so (in-tree).  Raises if it has not been built: there is no other implementation."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FeasmError(-2, f"{LIB_PATH} not built; run `python -c 'import __graft_entry__ as g; g.build()'`")
    L = C.CDLL(LIB_PATH)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    sig = {
        "feasm_create": (i32, [i32, C.POINTER(vp)]),
        "feasm_destroy": (None, [vp]),
        "feasm_last_error": (C.c_char_p, [vp]),
        "feasm_version": (i32, []),
        "feasm_set_mesh": (i32, [vp, i32, i64, i64, vp, vp]),
        "feasm_update_coordinates": (i32, [vp, vp]),
        "feasm_set_dofs": (i32, [vp, i32, i64, i64, vp, vp]),
        "feasm_set_constraints": (i32, [vp, i64, vp, vp]),
        "feasm_set_constraint_rows": (i32, [vp, i64, vp, vp, vp, vp, vp]),
        "feasm_set_pattern": (i32, [vp, vp, vp]),
        "feasm_build_pattern": (i32, [vp]),
        "feasm_nnz": (i64, [vp]),
        "feasm_get_pattern": (i32, [vp, vp, vp]),
        "feasm_prepare": (i32, [vp]),
        "feasm_assemble": (i32, [vp, C.POINTER(Params)]),
        "feasm_synchronize": (i32, [vp]),
        "feasm_element_matrices": (i32, [vp, C.POINTER(Params), i64, i64, i32, vp, vp]),
        "feasm_zero": (i32, [vp]),
        "feasm_add_matrix": (i32, [vp, vp, vp, i32, vp, i32]),
        "feasm_add_vector": (i32, [vp, vp, vp, i32]),
        "feasm_add_matrix_noalloc": (i32, [vp, vp, vp, i32, vp, i32]),
        "feasm_set_interface_prefix": (i32, [vp, i64]),
        "feasm_set_matrix": (i32, [vp, vp, vp, i32, vp, i32]),
        "feasm_init_matrix": (i32, [vp, i64, i64, vp, vp]),
        "feasm_set_values": (i32, [vp, vp, vp]),
        "feasm_set_communicator": (i32, [vp, vp, i32, i32]),
        "feasm_comm_unique_id": (i32, [vp]),
        "feasm_comm_init": (i32, [vp, i32, i32, vp]),
        "feasm_exchange_begin": (i32, [vp]),
        "feasm_exchange_end": (i32, [vp]),
        "feasm_exchange_ghost_rows": (i32, [vp]),
        "feasm_comm_allgather_u64": (i32, [vp, vp, i32, vp]),
        "feasm_build_exchange": (i32, [vp, vp]),
        "feasm_planner_create": (vp, [i32, i32, vp, C.c_uint64, C.c_uint64, vp, vp, vp]),
        "feasm_planner_destroy": (None, [vp]),
        "feasm_planner_error": (C.c_char_p, [vp]),
        "feasm_planner_message": (C.c_uint64, [vp, i32, C.POINTER(vp)]),
        "feasm_planner_receive": (i32, [vp, i32, vp, C.c_uint64]),
        "feasm_planner_finish": (i32, [vp]),
        "feasm_planner_sizes": (i32, [vp, vp]),
        "feasm_planner_get": (i32, [vp, vp, vp, vp, vp, vp, vp, vp]),
        "feasm_apply_plan": (i32, [vp, vp]),
        "feasm_get_exchange_plan": (i32, [vp, vp, vp, vp, vp, vp]),
        "feasm_get_values": (i32, [vp, vp, vp]),
        "feasm_device_values": (vp, [vp]),
        "feasm_device_rhs": (vp, [vp]),
        "feasm_stream": (vp, [vp]),
        "feasm_set_exchange_plan": (i32, [vp, i32, vp, vp, vp, vp, vp]),
        "feasm_set_side_loads": (i32, [vp, i64, vp, vp, vp, i32]),
        "feasm_assemble_begin": (i32, [vp, C.POINTER(Params)]),
        "feasm_assemble_end": (i32, [vp]),
        "feasm_send_buffer": (vp, [vp, C.POINTER(C.c_uint64)]),
        "feasm_pack_ghost": (vp, [vp, C.POINTER(C.c_uint64)]),
        "feasm_recv_buffer": (vp, [vp, C.POINTER(C.c_uint64)]),
        "feasm_unpack_ghost": (i32, [vp]),
        "feasm_cube_sizes": (i32, [i32, i32, i32, i32, C.POINTER(i64), C.POINTER(i64)]),
        "feasm_build_cube": (i32, [i32, i32, i32, i32, vp, vp, vp, vp]),
        "feasm_distribute_dofs": (i64, [i32, i64, i64, vp, i32, i32, vp, vp, vp, vp]),
        "feasm_last_timing": (i32, [vp, vp, C.POINTER(i32)]),
        "feasm_measure_fp64_peaks": (i32, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "feasm_vector_mult": (i32, [vp, vp, vp]),
        "feasm_vector_mult_host": (i32, [vp, vp, vp]),
        "feasm_cg_solve": (i32, [vp, C.c_double, i32, vp, vp, C.POINTER(i32), C.POINTER(C.c_double)]),
        "feasm_vec_alloc": (vp, [vp, i64]),
        "feasm_vec_free": (None, [vp, vp]),
        "feasm_vec_copy": (i32, [vp, i32, vp, vp, i64]),
        "feasm_vec_fill": (i32, [vp, vp, i64, C.c_double]),
        "feasm_vec_axpy": (i32, [vp, vp, C.c_double, vp, i64]),
        "feasm_vec_pointwise": (i32, [vp, i32, vp, vp, vp, C.c_double, i64]),
        "feasm_vec_reduce": (i32, [vp, i32, vp, vp, i64, C.POINTER(C.c_double)]),
        "feasm_vec_set_entry": (i32, [vp, vp, i64, C.c_double, i32]),
        "feasm_generate_cube": (i32, [vp, i32, i32, i32, i32, vp]),
        "feasm_distribute_dofs_device": (i64, [vp, i32]),
        "feasm_generate_dirichlet": (i64, [vp, C.c_uint, i32, C.c_uint]),
        "feasm_get_mesh": (i32, [vp, vp, vp, vp]),
        "feasm_get_dofs": (i32, [vp, vp, vp]),
        "feasm_get_constraints": (i32, [vp, vp, vp]),
        "feasm_set_side_penalty": (i32, [vp, i64, vp, vp, C.c_double, i32, C.c_double]),
        "feasm_residual": (i32, [vp, C.POINTER(Params), vp, vp]),
        "feasm_build_halo": (i32, [vp, vp]),
        "feasm_halo_columns": (i64, [vp, C.c_uint64, C.c_uint64, vp]),
        "feasm_set_halo_plan": (i32, [vp, C.c_uint64, C.c_uint64, i32, vp, vp, vp, vp, vp]),
        "feasm_halo_pack": (vp, [vp, vp, C.POINTER(C.c_uint64)]),
        "feasm_halo_recv_buffer": (vp, [vp, C.POINTER(C.c_uint64)]),
        "feasm_halo_unpack": (i32, [vp]),
        "feasm_get_halo_plan": (i32, [vp, vp, vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a, dtype):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


def make_params(problem=POISSON, quadrature_order=0, forcing=FORCING_EX4_FD, constrain=CONSTRAIN_HETEROGENEOUS,
                f_host=None, f_device=0, young_modulus=0.0, poisson_ratio=0.0, body_force=(0.0, 0.0, 0.0), affine_tensor=0):
    p = Params()
    p.problem, p.quadrature_order, p.forcing, p.constrain = problem, quadrature_order, forcing, constrain
    p.f_host = None if f_host is None else f_host.ctypes.data
    p.f_device = f_device or None
    p.young_modulus, p.poisson_ratio = young_modulus, poisson_ratio
    for i in range(3):
        p.body_force[i] = body_force[i]
    p.affine_tensor = int(affine_tensor)
    p._keep = f_host
    return p


# ---- host-side generators (no GPU needed) ---------------------------------------------------------
def cube_sizes(elem_type, nx, ny, nz):
    """(n_elem, n_nodes) of build_cube."""
    ne, nn = C.c_int64(), C.c_int64()
    if lib().feasm_cube_sizes(elem_type, nx, ny, nz, C.byref(ne), C.byref(nn)):
        raise FeasmError(-1, "bad build_cube arguments")
    return int(ne.value), int(nn.value)


def build_cube(elem_type, nx, ny, nz, bounds=(0.0, 1.0, 0.0, 1.0, 0.0, 1.0)):
    L = lib()
    ne, nn = C.c_int64(), C.c_int64()
    if L.feasm_cube_sizes(elem_type, nx, ny, nz, C.byref(ne), C.byref(nn)):
        raise FeasmError(-1, "bad build_cube arguments")
    xyz = np.empty((nn.value, 3))
    conn = np.empty((ne.value, elem_type), dtype=np.uint32)
    mask = np.empty(nn.value, dtype=np.uint8)
    b = np.asarray(bounds, dtype=np.float64)
    rc = L.feasm_build_cube(elem_type, nx, ny, nz, _p(b), _p(xyz), _p(conn), _p(mask))
    if rc:
        raise FeasmError(rc, "feasm_build_cube failed")
    return xyz, conn, mask


def distribute_dofs(conn, n_nodes, n_vars=1, n_parts=1, elem_pid=None):
    L = lib()
    conn = _c(conn, np.uint32)
    ne, npe = conn.shape
    base = np.empty(n_nodes, dtype=np.uint32)
    first = np.empty(n_parts + 1, dtype=np.uint32)
    npid = np.empty(n_nodes, dtype=np.int32)
    ep = _c(elem_pid, np.int32)
    n = L.feasm_distribute_dofs(npe, ne, n_nodes, _p(conn), n_vars, n_parts, _p(ep), _p(base), _p(first), _p(npid))
    if n < 0:
        raise FeasmError(int(n), "feasm_distribute_dofs failed")
    return int(n), base, first, npid


def elem_dofs_from_nodes(conn, node_base, n_vars=1):
    """DofMap::dof_indices for every element: variable-major, element-local node order."""
    nb = node_base[conn.astype(np.int64)]
    return np.ascontiguousarray(np.concatenate([nb + v for v in range(n_vars)], axis=1), dtype=np.uint32)


class Context:
    """One libfeasm context = one GPU.  Thin, 1:1 over the C ABI."""

    def __init__(self, device=0):
        self.L = lib()
        h = C.c_void_p()
        rc = self.L.feasm_create(device, C.byref(h))
        if rc:
            raise FeasmError(rc, self.L.feasm_last_error(None).decode())
        self.h = h
        self.n_rows = 0
        self.n_owned = 0
        self.nd = 0
        self.n_elem = 0
        self.has_comm = False
        self.comm_size = 1

    def close(self):
        if getattr(self, "h", None):
            self.L.feasm_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def _ck(self, rc):
        if rc:
            raise FeasmError(rc, self.L.feasm_last_error(self.h).decode())

    def set_mesh(self, elem_type, xyz, conn):
        xyz, conn = _c(xyz, np.float64), _c(conn, np.uint32)
        self.elem_type, self.n_elem, self.n_nodes = elem_type, conn.shape[0], xyz.shape[0]
        self._ck(self.L.feasm_set_mesh(self.h, elem_type, conn.shape[0], xyz.shape[0], _p(xyz), _p(conn)))
        self.n_nodes = xyz.shape[0]

    def update_coordinates(self, xyz):
        self._ck(self.L.feasm_update_coordinates(self.h, _p(xyz)))

    def set_dofs(self, elem_dofs, n_rows, n_vars=1, n_owned=None, global_ids=None):
        ed = _c(elem_dofs, np.uint32)
        gid = _c(global_ids, np.uint32)
        self.n_rows, self.nd = int(n_rows), ed.shape[1]
        self.n_owned = int(n_rows if n_owned is None else n_owned)
        self._ck(self.L.feasm_set_dofs(self.h, n_vars, n_rows, n_rows if n_owned is None else n_owned, _p(ed), _p(gid)))

    # ---- device-side ingest (no host arrays): build_cube, distribute_dofs + dof_indices, DirichletBoundary ----
    def generate_cube(self, elem_type, nx, ny, nz, bounds=(0.0, 1.0, 0.0, 1.0, 0.0, 1.0)):
        b = np.asarray(bounds, dtype=np.float64)
        self._ck(self.L.feasm_generate_cube(self.h, elem_type, nx, ny, nz, _p(b)))
        ne, nn = C.c_int64(), C.c_int64()
        self.L.feasm_cube_sizes(elem_type, nx, ny, nz, C.byref(ne), C.byref(nn))
        self.elem_type, self.n_elem, self.n_nodes = elem_type, ne.value, nn.value

    def distribute_dofs(self, n_vars=1):
        n = int(self.L.feasm_distribute_dofs_device(self.h, n_vars))
        if n < 0:
            self._ck(n)
        self.n_rows, self.nd, self.n_vars = n, self.elem_type * n_vars, n_vars
        self.n_owned = n
        return n

    def generate_dirichlet(self, id_mask=0x3f, kind=1, var_mask=0xff):
        n = int(self.L.feasm_generate_dirichlet(self.h, id_mask, kind, var_mask))
        if n < 0:
            self._ck(n)
        return n

    def get_mesh(self, mask=True):
        xyz = np.empty((self.n_nodes, 3))
        conn = np.empty((self.n_elem, self.elem_type), dtype=np.uint32)
        m = np.empty(self.n_nodes, dtype=np.uint8) if mask else None
        self._ck(self.L.feasm_get_mesh(self.h, _p(xyz), _p(conn), _p(m)))
        return xyz, conn, m

    def get_dofs(self, node_base=True):
        ed = np.empty((self.n_elem, self.nd), dtype=np.uint32)
        nb = np.empty(self.n_nodes, dtype=np.uint32) if node_base else None
        self._ck(self.L.feasm_get_dofs(self.h, _p(ed), _p(nb)))
        return ed, nb

    def get_constraints(self):
        f = np.empty(self.n_rows, dtype=np.uint8)
        v = np.empty(self.n_rows)
        self._ck(self.L.feasm_get_constraints(self.h, _p(f), _p(v)))
        return f, v

    def set_constraints(self, rows, values):
        r, v = _c(rows, np.uint32), _c(values, np.float64)
        self._ck(self.L.feasm_set_constraints(self.h, len(r), _p(r), _p(v)))

    def set_constraint_rows(self, dofs, row_ptr, cols, coefs, values):
        """General constraint rows (hanging nodes): u_dofs[i] = sum_k coefs[k] u_cols[k] + values[i]; replaces set_constraints."""
        d, rp = _c(dofs, np.uint32), _c(row_ptr, np.uint64)
        c, w, v = _c(cols, np.uint32), _c(coefs, np.float64), _c(values, np.float64)
        if len(rp) != len(d) + 1 or len(v) != len(d) or len(c) != len(w):
            raise FeasmError(-1, "inconsistent constraint row arrays")
        self._ck(self.L.feasm_set_constraint_rows(self.h, len(d), _p(d), _p(rp), _p(c), _p(w), _p(v)))

    def set_pattern(self, row_ptr, col_idx):
        rp, ci = _c(row_ptr, np.uint64), _c(col_idx, np.uint32)
        self._ck(self.L.feasm_set_pattern(self.h, _p(rp), _p(ci)))

    def build_pattern(self):
        self._ck(self.L.feasm_build_pattern(self.h))
        return self.nnz

    @property
    def nnz(self):
        return int(self.L.feasm_nnz(self.h))

    def get_pattern(self):
        rp = np.empty(self.n_rows + 1, dtype=np.uint64)
        ci = np.empty(self.nnz, dtype=np.uint32)
        self._ck(self.L.feasm_get_pattern(self.h, _p(rp), _p(ci)))
        return rp, ci

    def prepare(self):
        self._ck(self.L.feasm_prepare(self.h))

    def assemble(self, params, sync=True):
        self._ck(self.L.feasm_assemble(self.h, C.byref(params)))
        if sync:
            self._ck(self.L.feasm_synchronize(self.h))

    def set_side_loads(self, elems, sides, g):
        """Boundary integral Fe_var[i](a) += JxW_face (g[i] phi_face[a]) on the listed (element, side) pairs
        (systems_of_equations_ex6.C:240-257); empty lists clear."""
        e, sd, gg = _c(elems, np.uint32), _c(sides, np.uint8), _c(np.atleast_1d(g), np.float64)
        if len(e) != len(sd):
            raise ValueError("elems and sides differ in length")
        self._ck(self.L.feasm_set_side_loads(self.h, len(e), _p(e), _p(sd), _p(gg), len(gg)))

    def set_side_penalty(self, elems, sides, penalty=1e10, kind=1, value=0.0):
        """introduction_ex3.C:386-455 penalty boundary term on the listed (element, side) pairs; empty lists clear."""
        e, sd = _c(elems, np.uint32), _c(sides, np.uint8)
        self._ck(self.L.feasm_set_side_penalty(self.h, len(e), _p(e), _p(sd), float(penalty), int(kind), float(value)))

    def set_interface_prefix(self, n_elems):
        self._ck(self.L.feasm_set_interface_prefix(self.h, int(n_elems)))

    def assemble_begin(self, params):
        self._ck(self.L.feasm_assemble_begin(self.h, C.byref(params)))

    def assemble_end(self):
        self._ck(self.L.feasm_assemble_end(self.h))

    def send_buffer(self):
        n = C.c_uint64()
        ptr = self.L.feasm_send_buffer(self.h, C.byref(n))
        return ptr, n.value

    def synchronize(self):
        self._ck(self.L.feasm_synchronize(self.h))

    def element_matrices(self, params, e0=0, e1=None, constrained=True):
        e1 = self.n_elem if e1 is None else e1
        Ke = np.empty((e1 - e0, self.nd, self.nd))
        Fe = np.empty((e1 - e0, self.nd))
        self._ck(self.L.feasm_element_matrices(self.h, C.byref(params), e0, e1, int(constrained), _p(Ke), _p(Fe)))
        return Ke, Fe

    def zero(self):
        self._ck(self.L.feasm_zero(self.h))

    def add_matrix(self, dm, rows, cols=None):
        dm = _c(dm, np.float64)
        r = _c(rows, np.uint32)
        c = r if cols is None else _c(cols, np.uint32)
        self._ck(self.L.feasm_add_matrix(self.h, _p(dm), _p(r), len(r), _p(c), len(c)))

    def set_matrix(self, dm, rows, cols=None):
        dm = _c(dm, np.float64)
        r = _c(rows, np.uint32)
        c = r if cols is None else _c(cols, np.uint32)
        self._ck(self.L.feasm_set_matrix(self.h, _p(dm), _p(r), len(r), _p(c), len(c)))

    def init_matrix(self, row_ptr, col_idx, n_owned=None):
        rp, ci = _c(row_ptr, np.uint64), _c(col_idx, np.uint32)
        n = len(rp) - 1
        self._ck(self.L.feasm_init_matrix(self.h, n, n if n_owned is None else n_owned, _p(rp), _p(ci)))
        self.n_rows = n
        self.n_owned = int(n if n_owned is None else n_owned)

    def set_values(self, values=None, rhs=None):
        v, b = _c(values, np.float64), _c(rhs, np.float64)
        self._ck(self.L.feasm_set_values(self.h, _p(v), _p(b)))

    def add_matrix_noalloc(self, dm, rows, cols=None):
        """feasm_add_matrix_noalloc: the same add through the library's fixed device scratch (no device allocation)."""
        dm = _c(dm, np.float64)
        r = _c(rows, np.uint32)
        c = r if cols is None else _c(cols, np.uint32)
        self._ck(self.L.feasm_add_matrix_noalloc(self.h, _p(dm), _p(r), len(r), _p(c), len(c)))

    def add_vector(self, v, rows):
        v, r = _c(v, np.float64), _c(rows, np.uint32)
        self._ck(self.L.feasm_add_vector(self.h, _p(v), _p(r), len(r)))

    def get_values(self, values=True, rhs=True, out_values=None, out_rhs=None):
        v = (np.empty(self.nnz) if out_values is None else out_values) if values else None
        b = (np.empty(self.n_rows) if out_rhs is None else out_rhs) if rhs else None
        self._ck(self.L.feasm_get_values(self.h, _p(v), _p(b)))
        return v, b

    def device_values(self):
        return self.L.feasm_device_values(self.h)

    def device_rhs(self):
        return self.L.feasm_device_rhs(self.h)

    def stream(self):
        return self.L.feasm_stream(self.h)

    def last_timing(self):
        ms = (C.c_float * 4)()
        n = C.c_int()
        self._ck(self.L.feasm_last_timing(self.h, ms, C.byref(n)))
        return [ms[0], ms[1], ms[2], ms[3]], n.value

    def vector_mult(self, d_x, d_y):
        """y = A x with device pointers (ints) of n_owned doubles (= n_rows on one rank)."""
        self._ck(self.L.feasm_vector_mult(self.h, C.c_void_p(d_x), C.c_void_p(d_y)))

    def residual_host(self, params, u):
        """R = K(u) u - F from the element records (feasm_residual), host arrays in and out (several ranks: owned parts)."""
        L, n = self.L, self.n_owned
        uc = _c(u, np.float64)
        out = np.empty(n, np.float64)
        d_u, d_r = L.feasm_vec_alloc(self.h, n), L.feasm_vec_alloc(self.h, n)
        try:
            self._ck(L.feasm_vec_copy(self.h, 0, C.c_void_p(d_u), _p(uc), n))
            self._ck(L.feasm_residual(self.h, C.byref(params), C.c_void_p(d_u), C.c_void_p(d_r)))
            self._ck(L.feasm_vec_copy(self.h, 1, _p(out), C.c_void_p(d_r), n))
        finally:
            L.feasm_vec_free(self.h, C.c_void_p(d_u))
            L.feasm_vec_free(self.h, C.c_void_p(d_r))
        return out

    def vector_mult_host(self, x):
        xc = _c(x, np.float64)
        y = np.empty(self.n_owned, np.float64)
        self._ck(self.L.feasm_vector_mult_host(self.h, _p(xc), _p(y)))
        return y

    def cg_solve(self, rel_tol=1e-12, max_its=10000, x0=None):
        x = np.empty(self.n_owned, np.float64)
        x0c = _c(x0, np.float64)
        its, rel = C.c_int(), C.c_double()
        rc = self.L.feasm_cg_solve(self.h, rel_tol, max_its, _p(x0c), _p(x), C.byref(its), C.byref(rel))
        if rc < 0:  # FEASM_NOT_CONVERGED (+1) is a status: x, its, rel are valid and rel > rel_tol says so
            self._ck(rc)
        return x, its.value, rel.value

    def measure_fp64_peaks(self):
        a, b = C.c_double(), C.c_double()
        self._ck(self.L.feasm_measure_fp64_peaks(self.h, C.byref(a), C.byref(b)))
        return {"dfma_tflops": a.value, "dmma_tflops": b.value}

    def set_exchange_plan(self, peer_rank, send_ptr, send_slots, recv_ptr, recv_slots):
        pr = _c(peer_rank, np.int32)
        sp, ss = _c(send_ptr, np.uint64), _c(send_slots, np.uint64)
        rp, rs = _c(recv_ptr, np.uint64), _c(recv_slots, np.uint64)
        self._ck(self.L.feasm_set_exchange_plan(self.h, len(pr), _p(pr), _p(sp), _p(ss), _p(rp), _p(rs)))

    # -- NCCL below the C ABI ------------------------------------------------------------------------------------------
    @staticmethod
    def comm_unique_id():
        """The bootstrap id rank 0 makes (ncclGetUniqueId); broadcast its bytes to every rank by any means."""
        buf = C.create_string_buffer(128)
        rc = lib().feasm_comm_unique_id(buf)
        if rc:
            raise FeasmError(rc, lib().feasm_last_error(None).decode())
        return bytes(buf.raw)

    def comm_init(self, n_ranks, rank, id_bytes):
        buf = C.create_string_buffer(bytes(id_bytes), 128)
        self._ck(self.L.feasm_comm_init(self.h, n_ranks, rank, buf))
        self.has_comm, self.comm_size = True, n_ranks

    def exchange_begin(self):
        self._ck(self.L.feasm_exchange_begin(self.h))

    def exchange_end(self):
        self._ck(self.L.feasm_exchange_end(self.h))

    def exchange_ghost_rows(self):
        self._ck(self.L.feasm_exchange_ghost_rows(self.h))

    def comm_allgather(self, values):
        v = _c(values, np.uint64)
        out = np.empty((self.comm_size, len(v)), dtype=np.uint64)
        self._ck(self.L.feasm_comm_allgather_u64(self.h, _p(v), len(v), _p(out)))
        return out

    def build_exchange(self, first_dof):
        """Collective setup over the context's communicator (feasm_build_exchange): merged pattern, slot plan, scatter maps."""
        fd = _c(first_dof, np.uint64)
        if len(fd) != self.comm_size + 1:
            raise FeasmError(-1, "first_dof must have n_ranks + 1 entries")
        self._ck(self.L.feasm_build_exchange(self.h, _p(fd)))

    # -- the halo of a distributed vector (feasm_halo.cu) ------------------------------------------------------------------
    def build_halo(self, first_dof):
        """Collective (feasm_build_halo): after it vector_mult / cg_solve / residual work on the distributed system."""
        fd = _c(first_dof, np.uint64)
        if len(fd) != self.comm_size + 1:
            raise FeasmError(-1, "first_dof must have n_ranks + 1 entries")
        self._ck(self.L.feasm_build_halo(self.h, _p(fd)))

    def halo_columns(self, n_global, first):
        n = int(self.L.feasm_halo_columns(self.h, int(n_global), int(first), None))
        if n < 0:
            self._ck(n)
        out = np.zeros(n, dtype=np.uint32)
        if n:
            self.L.feasm_halo_columns(self.h, int(n_global), int(first), _p(out))
        return out

    def set_halo_plan(self, n_global, first, peers, send_ptr, send_rows, recv_ptr, recv_gids):
        pr = _c(peers, np.int32)
        sp, sr = _c(send_ptr, np.uint64), _c(send_rows, np.uint32)
        rp, rg = _c(recv_ptr, np.uint64), _c(recv_gids, np.uint32)
        self._ck(self.L.feasm_set_halo_plan(self.h, int(n_global), int(first), len(pr), _p(pr), _p(sp), _p(sr), _p(rp), _p(rg)))

    def halo_pack(self, d_x):
        n = C.c_uint64()
        ptr = self.L.feasm_halo_pack(self.h, C.c_void_p(d_x), C.byref(n))
        if not ptr:
            raise FeasmError(-3, "no halo plan")
        return ptr, n.value

    def halo_recv_buffer(self):
        n = C.c_uint64()
        ptr = self.L.feasm_halo_recv_buffer(self.h, C.byref(n))
        if not ptr:
            raise FeasmError(-3, "no halo plan")
        return ptr, n.value

    def halo_unpack(self):
        self._ck(self.L.feasm_halo_unpack(self.h))

    def halo_plan(self):
        n = int(self.L.feasm_get_halo_plan(self.h, None, None, None, None, None))
        if n < 0:
            self._ck(n)
        peers, sp, rp = np.zeros(n, dtype=np.int32), np.zeros(n + 1, dtype=np.uint64), np.zeros(n + 1, dtype=np.uint64)
        self.L.feasm_get_halo_plan(self.h, _p(peers), _p(sp), None, _p(rp), None)
        sr, rg = np.zeros(int(sp[-1]), dtype=np.uint32), np.zeros(int(rp[-1]), dtype=np.uint32)
        self.L.feasm_get_halo_plan(self.h, None, None, _p(sr), None, _p(rg))
        return peers, sp, sr, rp, rg

    def apply_plan(self, planner):
        self._ck(self.L.feasm_apply_plan(self.h, planner.h))

    def exchange_plan(self):
        """(peers, send_ptr, send_slots, recv_ptr, recv_slots) of the plan the context holds."""
        n = int(self.L.feasm_get_exchange_plan(self.h, None, None, None, None, None))
        if n < 0:
            self._ck(n)
        peers, sp, rp = np.zeros(n, dtype=np.int32), np.zeros(n + 1, dtype=np.uint64), np.zeros(n + 1, dtype=np.uint64)
        if n:
            self.L.feasm_get_exchange_plan(self.h, _p(peers), _p(sp), None, _p(rp), None)
        ss, rs = np.zeros(int(sp[-1]), dtype=np.uint64), np.zeros(int(rp[-1]), dtype=np.uint64)
        if n and int(self.L.feasm_get_exchange_plan(self.h, None, None, _p(ss), None, _p(rs))) < 0:
            self._ck(-2)
        return peers, sp, ss, rp, rs

    def pack_ghost(self):
        n = C.c_uint64()
        ptr = self.L.feasm_pack_ghost(self.h, C.byref(n))
        return ptr, n.value

    def recv_buffer(self):
        n = C.c_uint64()
        ptr = self.L.feasm_recv_buffer(self.h, C.byref(n))
        return ptr, n.value

    def unpack_ghost(self):
        self._ck(self.L.feasm_unpack_ghost(self.h))


class Planner:
    """feasm_planner_*: the host-only half of the multi-rank setup (no device, no communicator).  Messages are uint32 arrays."""

    def __init__(self, rank, n_ranks, first_dof, n_owned, n_rows, global_ids, row_ptr, col_idx):
        self.L = lib()
        fd, gid = _c(first_dof, np.uint64), _c(global_ids, np.uint32)
        rp, ci = _c(row_ptr, np.uint64), _c(col_idx, np.uint32)
        self.n_rows, self.n_ranks = int(n_rows), int(n_ranks)
        self.h = self.L.feasm_planner_create(rank, n_ranks, _p(fd), int(n_owned), int(n_rows), _p(gid), _p(rp), _p(ci))
        if not self.h:
            raise FeasmError(-1, self.L.feasm_planner_error(None).decode())

    def _ck(self, rc):
        if rc:
            raise FeasmError(rc, self.L.feasm_planner_error(self.h).decode())

    def messages(self):
        """{peer: uint32 words} for every peer this rank has ghost rows of."""
        out = {}
        for q in range(self.n_ranks):
            w = C.c_void_p()
            n = int(self.L.feasm_planner_message(self.h, q, C.byref(w)))
            if n:
                out[q] = np.ctypeslib.as_array(C.cast(w, C.POINTER(C.c_uint32)), shape=(n,)).copy()
        return out

    def receive(self, peer, words):
        w = _c(words, np.uint32)
        self._ck(self.L.feasm_planner_receive(self.h, int(peer), _p(w), len(w)))

    def finish(self):
        """-> dict(row_ptr, col_idx, peers, send_ptr, send_slots, recv_ptr, recv_slots, merged)"""
        self._ck(self.L.feasm_planner_finish(self.h))
        sz = np.zeros(5, dtype=np.uint64)
        self._ck(self.L.feasm_planner_sizes(self.h, _p(sz)))
        nnz, npeer, ns, nr, merged = (int(x) for x in sz)
        o = dict(row_ptr=np.zeros(self.n_rows + 1, dtype=np.uint64), col_idx=np.zeros(nnz, dtype=np.uint32), peers=np.zeros(npeer, dtype=np.int32),
                 send_ptr=np.zeros(npeer + 1, dtype=np.uint64), send_slots=np.zeros(ns, dtype=np.uint64),
                 recv_ptr=np.zeros(npeer + 1, dtype=np.uint64), recv_slots=np.zeros(nr, dtype=np.uint64))
        self._ck(self.L.feasm_planner_get(self.h, _p(o["row_ptr"]), _p(o["col_idx"]), _p(o["peers"]), _p(o["send_ptr"]), _p(o["send_slots"]),
                                          _p(o["recv_ptr"]), _p(o["recv_slots"])))
        o["merged"] = bool(merged)
        return o

    def close(self):
        if self.h:
            self.L.feasm_planner_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
