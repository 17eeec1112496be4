"""The PRODUCT's host-side table generator against the reference's own tables, without a GPU.

libmesh_b200/csrc/feasm_tables.cpp builds what every element kernel reads from shared memory — the Gauss rule and phi,
dphi/dxi, /deta, /dzeta at its points — and the side tables of feasm_side.cu.  It was written independently of the oracle (node
coordinates and 1D Lagrange polynomials instead of libMesh's index maps), so this is a direct check of product code: compiled for
the host alone (tests/probe/tables_probe.cpp), its tables are compared with tests/golden/reference_fe_tables.npz, which the
REFERENCE's own quadrature and shape-function sources wrote (tests/golden/make_ref_golden.py), and its side tables with the
oracle's side_reinit on the reference element, which is the reference's face machinery bit for bit (tests/test_oracle_ref.py).
Bar: the same points in the same order; every table entry within a few ulp of unity (the two generators multiply in different
orders)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "reference_fe_tables.npz"))
NAME = {8: "HEX8", 20: "HEX20", 27: "HEX27", 10: "TET10", 18: "PRISM18"}
TOL = 8 * np.finfo(float).eps


@pytest.fixture(scope="module")
def probe():
    so = os.path.join(ROOT, "build", "libtables_probe.so")
    src = [os.path.join(ROOT, "tests", "probe", "tables_probe.cpp"), os.path.join(ROOT, "libmesh_b200", "csrc", "feasm_tables.cpp")]
    cuda_inc = next((p for p in ("/usr/local/cuda/include", "/usr/local/cuda/targets/x86_64-linux/include") if os.path.exists(os.path.join(p, "cuda_runtime.h"))), None)
    if cuda_inc is None:
        pytest.skip("cuda_runtime.h (named by feasm_internal.h) not found")
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        os.makedirs(os.path.dirname(so), exist_ok=True)
        subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-I" + os.path.join(ROOT, "include"),
                               "-I" + os.path.join(ROOT, "libmesh_b200", "csrc"), "-I" + cuda_inc, "-o", so] + src)
    L = C.CDLL(so)
    L.probe_table_sizes.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_char_p, C.c_int]
    L.probe_tables.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 6
    L.probe_side_sizes.argtypes = [C.c_int, C.c_int] + [C.POINTER(C.c_int)] * 4
    L.probe_side_tables.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 6
    return L


def _tables(L, etype, qorder):
    npe, nq = C.c_int(), C.c_int()
    err = C.create_string_buffer(256)
    assert L.probe_table_sizes(etype, qorder, npe, nq, err, 256) == 0, err.value.decode()
    qp, qw = np.zeros((nq.value, 3)), np.zeros(nq.value)
    tabs = [np.zeros((npe.value, nq.value)) for _ in range(4)]
    assert L.probe_tables(etype, qorder, qp.ctypes.data, qw.ctypes.data, *[t.ctypes.data for t in tabs]) == 0
    return qp, qw, tabs


@pytest.mark.parametrize("etype,qorder", [(8, 3), (8, 5), (27, 3), (27, 5), (20, 5), (10, 5), (18, 5)])
def test_volume_tables_are_the_references(probe, etype, qorder):
    qp, qw, tabs = _tables(probe, etype, qorder)
    k = f"{NAME[etype]}_q{qorder}"
    assert qp.shape == GOLD[k + "_pts"].shape, "another number of quadrature points"
    assert np.abs(qp - GOLD[k + "_pts"]).max() <= TOL and np.abs(qw - GOLD[k + "_w"]).max() <= TOL, "points / weights, or their order, differ"
    for name, t in zip(("phi", "dxi", "deta", "dzeta"), tabs):
        assert np.abs(t - GOLD[f"{k}_{name}"]).max() <= TOL * max(1.0, np.abs(GOLD[f"{k}_{name}"]).max()), name
    assert abs(tabs[0].sum(axis=0) - 1).max() <= TOL  # partition of unity at every point


HEX_REF = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1], [0, -1, -1],
                    [1, 0, -1], [0, 1, -1], [-1, 0, -1], [-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0], [0, -1, 1], [1, 0, 1],
                    [0, 1, 1], [-1, 0, 1], [0, 0, -1], [0, -1, 0], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0, 0, 1], [0, 0, 0]], dtype=float)
TET_REF = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [.5, 0, 0], [.5, .5, 0], [0, .5, 0], [0, 0, .5], [.5, 0, .5], [0, .5, .5]], dtype=float)


@pytest.mark.parametrize("etype,qorder", [(8, 3), (8, 5), (27, 5), (20, 5), (10, 5)])
def test_side_tables_are_the_references(probe, etype, qorder):
    """Parent shape functions at the side points and the side weights, side by side with FE::reinit(elem, side) on the reference
    element (where JxW_face = w for the hex sides, whose map is the identity up to orientation)."""
    L = probe
    npe, ns, nqf, n_sides = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    assert L.probe_side_sizes(etype, qorder, npe, ns, nqf, n_sides) == 0
    sn = np.zeros((n_sides.value, ns.value), dtype=np.int32)
    qw = np.zeros(nqf.value)
    psi, dxi, deta = (np.zeros((ns.value, nqf.value)) for _ in range(3))
    phi = np.zeros((n_sides.value, npe.value, nqf.value))
    assert L.probe_side_tables(etype, qorder, sn.ctypes.data, qw.ctypes.data, psi.ctypes.data, dxi.ctypes.data, deta.ctypes.data, phi.ctypes.data) == 0
    X = TET_REF if etype == 10 else HEX_REF[:etype]
    for side in range(4 if etype == 10 else 6):  # (the table is laid out for six sides; a tet uses four)
        d = O.side_reinit(etype, X, side, qorder)
        assert d["phi"].shape == phi[side].shape
        assert np.abs(d["phi"] - phi[side]).max() <= TOL, side
        if etype != 10:
            assert np.abs(d["JxW"] - qw).max() <= TOL
    assert abs(psi.sum(axis=0) - 1).max() <= TOL
