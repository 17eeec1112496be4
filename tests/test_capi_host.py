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


@pytest.mark.parametrize("etype", [F.HEX8, F.HEX27, F.HEX20, F.PRISM18])
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


def test_planner_two_ranks_by_hand():
    """feasm_planner_* on a hand-made 1-D chain of 2-dof elements split between two ranks: 5 dofs, rank 0 owns 0-2, rank 1 owns
    3-4; element (2,3) belongs to rank 1, so rank 1 has ghost row 2.  Columns are global ids."""
    fd = [0, 3, 5]
    # rank 0: elements (0,1), (1,2): rows 0..2, no ghosts
    p0 = F.Planner(0, 2, fd, 3, 3, [0, 1, 2], [0, 2, 5, 7], [0, 1, 0, 1, 2, 1, 2])
    # rank 1: elements (2,3), (3,4): owned rows 3, 4 (local 0, 1), ghost row 2 (local 2)
    p1 = F.Planner(1, 2, fd, 2, 3, [3, 4, 2], [0, 3, 5, 7], [2, 3, 4, 3, 4, 2, 3])
    assert p0.messages() == {}
    m = p1.messages()
    assert list(m) == [0] and m[0].tolist() == [1, 2, 2, 2, 3]  # one row, global id 2, two columns: 2 3
    p0.receive(1, m[0])
    o0, o1 = p0.finish(), p1.finish()
    assert o0["merged"] and o0["row_ptr"].tolist() == [0, 2, 5, 8] and o0["col_idx"].tolist() == [0, 1, 0, 1, 2, 1, 2, 3]
    assert not o1["merged"]
    assert o0["peers"].tolist() == [1] and o1["peers"].tolist() == [0]
    assert o1["send_ptr"].tolist() == [0, 3] and o1["send_slots"].tolist() == [5, 6, 7 + 2]  # two values of local row 2, then its rhs
    assert o0["recv_ptr"].tolist() == [0, 3] and o0["recv_slots"].tolist() == [6, 7, 8 + 2]  # columns 2, 3 of merged row 2, rhs of row 2
    assert o0["send_slots"].size == 0 and o1["recv_slots"].size == 0


def test_planner_rejects_bad_input():
    fd = [0, 3, 5]
    with pytest.raises(F.FeasmError):  # ghost ids must ascend and must not be owned
        F.Planner(1, 2, fd, 2, 3, [3, 4, 3], [0, 1, 2, 3], [3, 4, 3])
    with pytest.raises(F.FeasmError):  # n_owned does not match the dof range
        F.Planner(0, 2, fd, 2, 2, [0, 1], [0, 1, 2], [0, 1])
    p0 = F.Planner(0, 2, fd, 3, 3, [0, 1, 2], [0, 1, 2, 3], [0, 1, 2])
    with pytest.raises(F.FeasmError, match="does not own"):
        p0.receive(1, [1, 4, 1, 4])
    with pytest.raises(F.FeasmError, match="length"):
        p0.receive(1, [1, 2, 2, 3])
    with pytest.raises(F.FeasmError, match="bad peer"):
        p0.receive(0, [1, 2, 1, 3])
