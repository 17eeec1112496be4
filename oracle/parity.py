"""The parity rule, in one place (TEST INFRASTRUCTURE: tests/, __graft_entry__.smoke() and bench.py's checker import this).

north_star: "matrix and RHS values must agree within a relative 1e-12 per entry in FP64".  An entry that is a sum with
heavy cancellation (a coupling that is analytically ~0) has no meaningful relative error of its own, so the scale of an
entry is floored at FLOOR times the largest magnitude of its row:

    |a - b| <= TOL * max(|a|, |b|, FLOOR * rowmax)              TOL = 1e-12, FLOOR = 1e-2

(for a vector: FLOOR * max|reference vector|).  Every comparison also reports the STRICT rule — no floor — as a count of
failing entries and their worst deviation, so the effect of the floor is visible, never assumed.
"""
from __future__ import annotations

import numpy as np

TOL = 1e-12
FLOOR = 1e-2


def csr_compare(v, vo, row_ptr, chunk_rows: int = 1 << 20) -> dict:
    """Per-entry comparison of CSR value arrays (v: under test, vo: oracle) in row chunks (full-size arrays are GBs)."""
    rp = np.asarray(row_ptr).astype(np.int64)
    n_rows = len(rp) - 1
    worst = 0.0          # max |a-b| / max(|a|, |b|, FLOOR * rowmax)
    worst_strict = 0.0   # max |a-b| / max(|a|, |b|) over entries where that is defined
    n_strict_fail = 0
    strict_fail_over_rowmax = 0.0  # the largest |a-b| / rowmax among the entries failing the strict rule
    max_abs = 0.0
    n_bitwise = 0
    for r0 in range(0, n_rows, chunk_rows):
        r1 = min(n_rows, r0 + chunk_rows)
        lo, hi = int(rp[r0]), int(rp[r1])
        if hi == lo:
            continue
        a, b = np.asarray(v[lo:hi]), np.asarray(vo[lo:hi])
        lens = np.diff(rp[r0:r1 + 1])
        nz = lens > 0
        rowmax = np.zeros(r1 - r0)
        rowmax[nz] = np.maximum.reduceat(np.abs(b), (rp[r0:r1] - lo)[nz])
        rm = np.repeat(rowmax, lens)
        d = np.abs(a - b)
        n_bitwise += int((a == b).sum())
        mag = np.maximum(np.abs(a), np.abs(b))
        scale = np.maximum(mag, FLOOR * rm)
        scale[scale == 0] = 1.0
        worst = max(worst, float((d / scale).max()))
        max_abs = max(max_abs, float(d.max()))
        fail = d > TOL * mag
        nf = int(fail.sum())
        if nf:
            n_strict_fail += nf
            with np.errstate(divide="ignore", invalid="ignore"):
                rel = np.where(mag[fail] > 0, d[fail] / mag[fail], np.inf)
            finite = rel[np.isfinite(rel)]
            if finite.size:
                worst_strict = max(worst_strict, float(finite.max()))
            rmf = rm[fail]
            rmf[rmf == 0] = 1.0
            strict_fail_over_rowmax = max(strict_fail_over_rowmax, float((d[fail] / rmf).max()))
    n = int(rp[-1])
    return {"entries": n, "max_rel_err": worst, "passes_1e-12": bool(worst <= TOL), "bitwise_equal_entries": n_bitwise,
            "strict_fail_count": n_strict_fail, "strict_fail_max_rel": worst_strict,
            "strict_fail_max_abs_over_rowmax": strict_fail_over_rowmax, "max_abs_diff": max_abs}


def vec_compare(r, ro) -> dict:
    a, b = np.asarray(r), np.asarray(ro)
    d = np.abs(a - b)
    mag = np.maximum(np.abs(a), np.abs(b))
    vmax = float(np.abs(b).max()) if b.size else 0.0
    scale = np.maximum(mag, FLOOR * vmax)
    scale[scale == 0] = 1.0
    fail = d > TOL * mag
    with np.errstate(divide="ignore", invalid="ignore"):
        rel = np.where(mag > 0, d / mag, 0.0)
    return {"entries": int(a.size), "max_rel_err": float((d / scale).max()) if a.size else 0.0,
            "passes_1e-12": bool(a.size == 0 or (d / scale).max() <= TOL), "bitwise_equal_entries": int((a == b).sum()),
            "strict_fail_count": int(fail.sum()), "strict_fail_max_rel": float(rel[fail].max()) if fail.any() else 0.0,
            "strict_fail_max_abs_over_vecmax": float(d[fail].max() / vmax) if fail.any() and vmax > 0 else 0.0,
            "max_abs_diff": float(d.max()) if a.size else 0.0}


def csr_parity_error(v, vo, row_ptr) -> float:
    return csr_compare(v, vo, row_ptr)["max_rel_err"]


def vec_parity_error(r, ro) -> float:
    return vec_compare(r, ro)["max_rel_err"]


def dense_parity_error(K, Ko) -> float:
    """Element matrices [ne][nd][nd]: scale floored per row like the CSR rule."""
    rowmax = np.abs(Ko).max(axis=-1, keepdims=True)
    scale = np.maximum(np.maximum(np.abs(K), np.abs(Ko)), FLOOR * rowmax)
    scale[scale == 0] = 1.0
    return float((np.abs(K - Ko) / scale).max())
