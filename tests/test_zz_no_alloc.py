"""feasm_add_matrix_noalloc (include/feasm.h) against feasm_add_matrix, through the C ABI.

The entry point has no counterpart in the reference: it was added for callers that
cannot call cudaMalloc.  Its staging buffers are a fixed file-scope device array (FEASM_NOALLOC_SCRATCH_BYTES) that keeps
the previous call's bytes, so the test makes calls IN A ROW on one device — a large block first, smaller ones after it,
on two contexts — and compares the matrix after every call with the one the allocating entry point builds from the same
sequence.  Collected last (file name) so that a failure here hides no other test under `pytest -x`; it builds its own
inputs from the product-side generators and passes alone in a fresh process as well as after the other tests.
"""
import numpy as np
import pytest

from conftest import PARITY_TOL, csr_parity_error

pytestmark = pytest.mark.gpu


def _ctx(F, etype, n, n_vars=1):
    xyz, conn, _ = F.build_cube(etype, n, n, n, (-1.0, 1.0, -1.0, 1.0, -1.0, 1.0))
    n_dofs, base, _, _ = F.distribute_dofs(conn, xyz.shape[0], n_vars)
    ed = F.elem_dofs_from_nodes(conn, base, n_vars)
    ctx = F.Context(0)
    ctx.set_mesh(etype, xyz, conn)
    ctx.set_dofs(ed, n_dofs, n_vars=n_vars)
    ctx.build_pattern()
    ctx.zero()
    return ctx, ed


def test_noalloc_add_matrix_matches_add_matrix_on_calls_in_a_row():
    from libmesh_b200 import capi as F

    rng = np.random.default_rng(11)
    a, ed = _ctx(F, 27, 3, n_vars=3)  # 81 dofs per element: the largest block the library's elements produce
    b, _ = _ctx(F, 27, 3, n_vars=3)
    rp, _ = a.get_pattern()
    # a full 81 x 81 block, then smaller and rectangular ones (the scratch still holds the larger call's bytes), then a
    # repeat of an element: the adds accumulate
    calls = [(ed[0], ed[0]), (ed[5][:8], ed[5][:8]), (ed[13][:27], ed[13][40:81]), (ed[0][:3], ed[0][:1]), (ed[26], ed[26]), (ed[0], ed[0])]
    for k, (rows, cols) in enumerate(calls):
        dm = rng.standard_normal((len(rows), len(cols)))
        a.add_matrix(dm, rows, cols)
        b.add_matrix_noalloc(dm, rows, cols)
        va, _ = a.get_values()
        vb, _ = b.get_values()
        assert csr_parity_error(vb, va, rp) < PARITY_TOL, f"call {k}"
        assert np.abs(va).max() > 0
    # sparse_matrix.h:300-303 semantics carried over: zero may be added outside the pattern, a nonzero may not
    far = np.array([ed[0, 0], ed[26, 26]], dtype=np.uint32)
    b.add_matrix_noalloc(np.zeros((2, 2)), far)
    with pytest.raises(F.FeasmError) as e:
        b.add_matrix_noalloc(np.ones((2, 2)), far)
    assert e.value.code == -5
    with pytest.raises(F.FeasmError):  # the entries of the block that ARE in the pattern were added: same on the other route
        a.add_matrix(np.ones((2, 2)), far)
    # ... and the matrix is still usable after the error, again equal to the allocating route
    dm = rng.standard_normal((81, 81))
    a.add_matrix(dm, ed[7])
    b.add_matrix_noalloc(dm, ed[7])
    assert csr_parity_error(b.get_values()[0], a.get_values()[0], rp) < PARITY_TOL
    a.close()
    b.close()


def test_noalloc_rejects_blocks_beyond_the_fixed_scratch():
    from libmesh_b200 import capi as F

    ctx, ed = _ctx(F, 8, 6)
    rows = np.unique(ed)[:129].astype(np.uint32)  # 129 > FEASM_NOALLOC_MAX_DIM
    with pytest.raises(F.FeasmError) as e:
        ctx.add_matrix_noalloc(np.zeros((129, 2)), rows, rows[:2])
    assert e.value.code == -1 and "FEASM_NOALLOC_MAX_DIM" in str(e.value)
    with pytest.raises(F.FeasmError):
        ctx.add_matrix_noalloc(np.zeros((2, 129)), rows[:2], rows)
    ctx.add_matrix_noalloc(np.zeros((128, 128)), rows[:128], rows[:128])  # the maximum itself is accepted (zeros: any slot)
    ctx.close()
