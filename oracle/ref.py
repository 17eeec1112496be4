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
