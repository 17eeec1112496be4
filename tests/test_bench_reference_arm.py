"""bench.py --impl reference (the arm the driver times beside ours) runs without a GPU: one JSON line with the contract's keys,
the CPU implementations it ran (the oracle port; the reference's own assemble_poisson from oracle/_ref as one process per core,
where that library exists) and the faster of them as its value.  Under torchrun every rank but 0 exits without work."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra, env=None):
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1", "--n", "8"] + extra,
                       capture_output=True, text=True, timeout=600, env=e, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    return [l for l in p.stdout.splitlines() if l.strip()]


@pytest.mark.parametrize("elem", ["hex27", "hex8"])
def test_reference_arm_prints_one_contract_line(elem):
    lines = _run(["--elem", elem])
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "elements/s" and d["higher_is_better"] is True and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["e2e"] == {"value": d["value"], "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["n_elem"] == 512 and "build_cube 8^3" in d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["value"] == d["value"] and cb["kind"] in ("port", "reference") and cb["cores"] >= 1
    assert cb["port"]["value"] > 0
    from oracle import ref as R

    if R.app_available(fast=True):  # the reference's own code ran too, over all 512 elements, and the arm took the faster one
        rc = cb["reference_code"]
        assert rc is not None and "unavailable" not in rc, rc
        assert rc["kind"] == "reference" and rc["value"] > 0 and "512 elements" in rc["sample"]
        assert d["value"] == max(cb["port"]["value"], rc["value"])


def test_other_ranks_print_nothing():
    assert _run([], env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []


def test_elasticity_reference_arm_is_the_port():
    d = json.loads(_run(["--problem", "elasticity"])[0])
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["reference_code"] is None
