"""C-ABI library checks that need no GPU: every symbol include/feasm.h declares is exported, the host-side
generators agree with the oracle, and compute entry points fail loudly (no CPU fallback) without a device."""
import os
import re

import numpy as np
import pytest

from libmesh_b200 import capi as F
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "feasm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(feasm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_header_symbol():
    L = F.lib()
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(L, s), f"libfeasm.so does not export {s}"
    assert sorted(F.SYMBOLS) == syms, "capi.SYMBOLS is out of date with include/feasm.h"
    assert L.feasm_version() == 1


def test_product_never_references_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "libmesh_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp", "Makefile")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "fe_oracle.h" not in txt and "liboracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


@pytest.mark.parametrize("etype", [F.HEX8, F.HEX27])
def test_build_cube_matches_oracle(etype):
    b = (-1, 1, 0, 0.3, 0, 0.1)
    xyz, conn, mask = F.build_cube(etype, 5, 3, 4, b)
    m = O.Mesh(etype, 5, 3, 4, bounds=b, renumber_nodes=False)
    xo, co, mo = m.arrays()
    assert np.array_equal(xyz, xo) and np.array_equal(conn, co) and np.array_equal(mask, mo)


def test_tet10_cube_matches_oracle():
    b = (-1, 1, -1, 1, -1, 1)
    xyz, conn, mask = F.build_cube(F.TET10, 3, 2, 2, b)
    m = O.Mesh(O.TET10, 3, 2, 2, bounds=b, renumber_nodes=False)  # ids compacted in old-id order, like the product
    xo, co, mo = m.arrays()
    assert conn.shape == co.shape == (24 * 12, 10)
    assert np.array_equal(xyz, xo) and np.array_equal(conn, co) and np.array_equal(mask, mo)


@pytest.mark.parametrize("nv", [1, 3])
def test_distribute_dofs_matches_oracle(nv):
    m = O.Mesh(O.HEX27, 4, 3, 3)
    m.partition_linear(2)
    s = O.System(m, nv)
    _, conn, _ = m.arrays()
    ep, npid = m.partition()
    ed, nb, fd = s.dofs()
    n, base, first, npid2 = F.distribute_dofs(conn, m.n_nodes, nv, 2, ep)
    assert n == s.n_dofs and np.array_equal(base, nb) and np.array_equal(first, fd) and np.array_equal(npid2, npid)
    assert np.array_equal(F.elem_dofs_from_nodes(conn, base, nv), ed)


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the loud-failure path is for GPU-less hosts")
    with pytest.raises(F.FeasmError) as e:
        F.Context(0)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)
