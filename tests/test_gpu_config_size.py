"""GPU parity at BASELINE config sizes (VERDICT r1 "What's weak" #2, #3): the paths that only occur at size — several
waves of the persistent element kernel, blocks > sm_count in the row reduction, uint8 position rows on full-length rows —
and the in-kernel FD forcing measured per entry, not asserted at a hand-picked tolerance.

Oracle passes run on all host threads here (OpenMP atomics: the order of the adds into one slot differs from the serial
loop by rounding only, far inside the 1e-12 bar).
"""
import os

import numpy as np
import pytest

from conftest import PARITY_TOL
from oracle.parity import csr_compare, vec_compare
from test_gpu_parity import _mods, make_case, make_ctx

pytestmark = pytest.mark.gpu
THREADS = os.cpu_count() or 1

# The FD forcing of introduction_ex4.C:500-537 divides a difference of O(eps^2) = 1e-6 by eps^2: a one-ulp difference
# between two libm sin/cos (2.2e-16) is 2.2e-10 in f.  That is the conditioning of the reference's own expression, so
# the RHS of this mode is DEFINED only to ~1e-10 relative; the test measures the per-entry deviation, prints it, and
# holds it to that conditioning limit.  The 1e-12 bar for the RHS is carried by FORCING_TABULATED on the same mesh.
FD_RHS_TOL = 2.2e-16 / (1e-3 * 1e-3)


def _report(tag, cmp_v, cmp_r):
    print(f"\n[{tag}] values: max rel {cmp_v['max_rel_err']:.3e}, strict-1e-12 failures {cmp_v['strict_fail_count']} of {cmp_v['entries']}"
          f" (worst |d|/rowmax {cmp_v['strict_fail_max_abs_over_rowmax']:.2e});"
          f" rhs: max rel {cmp_r['max_rel_err']:.3e}, strict failures {cmp_r['strict_fail_count']} of {cmp_r['entries']}")


SIZE_CASES = [
    ("config1_hex27_20", 27, (20, 20, 20), 0),   # BASELINE config 1: introduction_ex4 -d 3 -n 20, the designated parity config
    ("hex27_32", 27, (32, 32, 32), 0),
    ("hex8_48_q3", 8, (48, 48, 48), 3),
    ("hex8_24_q5", 8, (24, 24, 24), 5),
    ("tet10_12", 10, (12, 12, 12), 5),           # 41 472 tets, irregular rows
    ("hex20_24", 20, (24, 24, 24), 0),           # serendipity hex at size (several waves, uint8 position rows)
    ("prism18_20", 18, (20, 20, 20), 0),         # 16 000 prisms
]


@pytest.mark.parametrize("name,etype,dims,qorder", SIZE_CASES, ids=[c[0] for c in SIZE_CASES])
def test_config_size_parity(name, etype, dims, qorder):
    F, O = _mods()
    m, s = make_case(etype, dims)
    ctx = make_ctx(m, s)
    rp, ci = s.sparsity()
    rp2, ci2 = ctx.get_pattern()
    assert ctx.nnz == s.nnz and np.array_equal(rp, rp2) and np.array_equal(ci, ci2)
    # tabulated forcing: the 1e-12 per-entry bar on matrix and RHS
    f_tab, _ = s.forcing_at_qps(qorder)
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab, n_threads=THREADS)
    ctx.assemble(F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab))
    v, r = ctx.get_values()
    cv, cr = csr_compare(v, vo, rp), vec_compare(r, ro)
    _report(name + " tabulated", cv, cr)
    assert cv["max_rel_err"] < PARITY_TOL and cr["max_rel_err"] < PARITY_TOL
    # the mode bench.py times: FD forcing evaluated in the kernel
    vo2, ro2, _ = s.assemble(O.POISSON_EX4, qorder=qorder, n_threads=THREADS)
    ctx.assemble(F.make_params(F.POISSON, qorder, F.FORCING_EX4_FD))
    v2, r2 = ctx.get_values()
    cv2, cr2 = csr_compare(v2, vo2, rp), vec_compare(r2, ro2)
    _report(name + " FD forcing", cv2, cr2)
    assert np.array_equal(v2, v), "the matrix does not depend on the forcing mode"
    assert cv2["max_rel_err"] < PARITY_TOL
    assert cr2["max_rel_err"] < FD_RHS_TOL
    ctx.close()


def test_config5_elasticity_size():
    """systems_of_equations_ex6 geometry at 16x8x8 HEX27 (1024 elements, 81 dofs per element), traction on max-x."""
    F, O = _mods()
    m, s = make_case(27, (16, 8, 8), bounds=(0, 1, 0, 0.3, 0, 0.1), n_vars=3, dirichlet=(4,), kind=1)
    el, sd = m.boundary_sides(2)
    s.set_side_loads(el, sd, (0., 0., -1.))
    ctx = make_ctx(m, s)
    ctx.set_side_loads(el, sd, (0., 0., -1.))
    rp, ci = s.sparsity()
    rp2, ci2 = ctx.get_pattern()
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2)
    vo, ro, _ = s.assemble(O.ELASTICITY_EX6, n_threads=THREADS)
    ctx.assemble(F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS))
    v, r = ctx.get_values()
    cv, cr = csr_compare(v, vo, rp), vec_compare(r, ro)
    _report("config5 elasticity 16x8x8", cv, cr)
    assert cv["max_rel_err"] < PARITY_TOL and cr["max_rel_err"] < PARITY_TOL
    ctx.close()


def test_real_nccl_two_ranks():
    """The real per-step ncclSend/ncclRecv group on two GPUs: two processes, NCCL, the ordering between
    the exchange stream, feasm_assemble_end, the next step's pack and feasm_unpack_ghost — checked against the oracle's
    global system after several back-to-back asynchronous steps.  Skips on a one-GPU box."""
    import subprocess
    import sys

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "tests", "nccl_worker.py")]
    out = subprocess.run(cmd, cwd=root, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "NCCL_PARITY_OK" in out.stdout
