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
