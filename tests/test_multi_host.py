"""N > 1 host path on the CPU: world_size 2 and 3 gloo jobs running tests/multi_worker.py (see its docstring)."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_world(world, args, timeout=240):
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "multi_worker.py")] + [str(a) for a in args],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r} failed:\n{out[-3000:]}"
        assert f"rank {r} ok" in out


@pytest.mark.parametrize("world,etype,dims,n_vars,scramble", [
    (2, 27, (4, 3, 4), 1, 0),   # HEX27 slab partition, the config-3 layout in small
    (3, 8, (5, 4, 3), 1, 1),    # HEX8, interleaved partition: several neighbours per rank, remainder elements
    (2, 10, (2, 2, 2), 1, 0),   # TET10
    (2, 27, (3, 2, 2), 3, 1),   # 3-variable elasticity numbering (ex6), interleaved partition
])
def test_multi_rank_host_logic(world, etype, dims, n_vars, scramble):
    run_world(world, [etype, *dims, n_vars, scramble])
