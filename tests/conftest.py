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


# ---- the parity rule: ONE definition, oracle/parity.py (stated in BASELINE.md §3 and DESIGN.md §4) --------------
#     |a - b| <= 1e-12 * max(|a|, |b|, 1e-2 * rowmax)
from oracle.parity import FLOOR, TOL as PARITY_TOL, csr_parity_error, dense_parity_error, vec_parity_error  # noqa: E402,F401
