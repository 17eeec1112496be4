"""GPU: the libMesh adaptor (libmesh_b200/shim: GpuSparseMatrix + the batched assemble hook) run end to end against the
mock libMesh classes, compared with a direct drive of the C ABI (tests/shim/shim_test.C, built by __graft_entry__.build())."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_adaptor_end_to_end():
    exe = os.path.join(ROOT, "build", "shim_test")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "libmesh_b200", "shim")])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "SHIM TEST OK" in r.stdout, r.stdout + r.stderr


def _ranks(world, tmp_path):
    exe = os.path.join(ROOT, "build", "shim_test_ranks")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "libmesh_b200", "shim")])
    procs = []
    for r in range(world):
        env = dict(os.environ, FEASM_MOCK_RANK=str(r), FEASM_MOCK_SIZE=str(world), FEASM_MOCK_DIR=str(tmp_path))
        procs.append(subprocess.Popen([exe], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=300)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    report = "\n".join(f"--- rank {r} (exit {p.returncode}) ---\n{out[-3000:]}" for r, (p, out) in enumerate(zip(procs, outs)))
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"rank {r}: SHIM RANKS OK" in out, report
    return outs


def test_adaptor_ranks_program_on_one_rank(tmp_path):
    """The several-ranks test program with one rank: the same checks against the one-context assembly, no exchange."""
    _ranks(1, tmp_path)


def test_adaptor_two_ranks(tmp_path):
    """Two processes, two GPUs: the C++ hook's multi-rank route (localisation, NCCL bootstrap over the host communicator,
    feasm_build_exchange, begin / exchange / end) against a one-context assembly of the whole mesh."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (one process per GPU)")
    outs = _ranks(2, tmp_path)
    assert "ghost rows" in outs[1]
