"""pytest configuration: the `gpu` marker and shared helpers.

`-m "not gpu"` : oracle vs the reference's known answers, host logic, C-ABI symbol export (no GPU calls).
`-m gpu`       : parity of the CUDA path (through the C ABI) against the oracle on a real B200.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Everything native is built once per session (in-tree; see __graft_entry__.build)."""
    import __graft_entry__ as g

    lib = os.path.join(ROOT, "libmesh_b200", "libfeasm.so")
    orc = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
    if not (os.path.exists(lib) and os.path.exists(orc)):
        g.build()


# ---- the parity rule (stated in DESIGN.md §parity) -------------------------------------------------------
# north_star: "matrix and RHS values must agree within a relative 1e-12 per entry in FP64".  Entries that are
# sums with heavy cancellation (analytically ~0 couplings) have no meaningful relative error, so the scale of an
# entry is floored at 1% of the largest magnitude of its row:
#     |a - b| <= 1e-12 * max(|a|, |b|, 0.01 * rowmax)
PARITY_TOL = 1e-12
FLOOR = 1e-2


def csr_parity_error(v, vo, row_ptr):
    rp = row_ptr.astype(np.int64)
    lens = np.diff(rp)
    nz = lens > 0
    rowmax = np.zeros(len(lens))
    rowmax[nz] = np.maximum.reduceat(np.abs(vo), rp[:-1][nz])
    scale = np.maximum(np.maximum(np.abs(v), np.abs(vo)), FLOOR * np.repeat(rowmax, lens))
    scale[scale == 0] = 1.0
    return float((np.abs(v - vo) / scale).max())


def vec_parity_error(r, ro):
    scale = np.maximum(np.maximum(np.abs(r), np.abs(ro)), FLOOR * np.abs(ro).max())
    scale[scale == 0] = 1.0
    return float((np.abs(r - ro) / scale).max())


def dense_parity_error(K, Ko):
    """Element matrices [ne][nd][nd]: scale floored per row like the CSR rule."""
    rowmax = np.abs(Ko).max(axis=-1, keepdims=True)
    scale = np.maximum(np.maximum(np.abs(K), np.abs(Ko)), FLOOR * rowmax)
    scale[scale == 0] = 1.0
    return float((np.abs(K - Ko) / scale).max())
