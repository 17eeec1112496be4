"""Writes tests/golden/reference_fe_tables.npz from the REFERENCE's own code (oracle/_ref/libref_fe.so = the reference's
shape-function and quadrature sources compiled where they lie; oracle/Makefile): for every element type of the path and its
side types, the Gauss rule and what FE::reinit tabulates on it — phi and its reference derivatives at the rule's points — plus
the same functions at 32 seeded random points.  Run in the build container (where /root/reference exists):
    python tests/golden/make_ref_golden.py
and tests/golden/reference_fe_map.npz: FEMap::init_reference_to_physical_map + FEMap::compute_map (src/fe/fe_map.C, compiled
as it is) on elements of plain, sheared and perturbed build_cube meshes — xyz, JxW, the Jacobian and the inverse-map entries
at the rule's points.  (The element coordinates are inputs and come from the oracle's mesh generator; which of the two
branches of compute_map ran — Elem::has_affine_map(), part of the element classes that cannot be compiled stand-alone — is
recorded with them.)
The files are committed; tests/test_oracle_ref.py compares the oracle with them bit for bit wherever the tests run."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402  (inputs only: meshes)
from oracle import ref as R  # noqa: E402

CASES = [("HEX8", 3), ("HEX8", 5), ("HEX27", 3), ("HEX27", 5), ("HEX20", 5), ("TET10", 5), ("PRISM18", 5),
         ("QUAD4", 3), ("QUAD4", 5), ("QUAD9", 5), ("QUAD8", 5), ("TRI6", 5)]


def random_points(elem, rng):
    if elem in ("TET10", "TRI6"):
        p = rng.uniform(0.02, 0.3, size=(32, 3))
    elif elem == "PRISM18":
        p = np.c_[rng.uniform(0.02, 0.45, size=(32, 2)), rng.uniform(-1, 1, size=32)]
    else:
        p = rng.uniform(-1, 1, size=(32, 3))
    if R.DIM[elem] == 2:
        p[:, 2] = 0.0
    return p


def main():
    rng = np.random.default_rng(20261105)
    out = {}
    for elem, order in CASES:
        pts, w = R.gauss_rule(elem, order)
        phi, dxi, deta, dzeta = R.shapes(elem, pts)
        k = f"{elem}_q{order}"
        out[k + "_pts"], out[k + "_w"] = pts, w
        out[k + "_phi"], out[k + "_dxi"], out[k + "_deta"], out[k + "_dzeta"] = phi, dxi, deta, dzeta
        if elem + "_rand_pts" not in out:
            rp = random_points(elem, rng)
            sp = R.shapes(elem, rp)
            out[elem + "_rand_pts"] = rp
            for name, a in zip(("phi", "dxi", "deta", "dzeta"), sp):
                out[f"{elem}_rand_{name}"] = a
    path = os.path.join(ROOT, "tests", "golden", "reference_fe_tables.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path)} bytes")

    # ---- FEMap on whole elements -------------------------------------------------------------------------------
    MAP_CASES = [("HEX8", O.HEX8, 3, 0.05), ("HEX27", O.HEX27, 5, 0.05), ("HEX20", O.HEX20, 5, 0.05), ("TET10", O.TET10, 5, 0.01),
                 ("PRISM18", O.PRISM18, 5, 0.01)]
    mp = {}
    for elem, orc, order, amp in MAP_CASES:
        pts, w = R.gauss_rule(elem, order)
        Xs, aff, xyz, jxw, jac, inv = [], [], [], [], [], []
        for skew in (None, (0, 0.0, 0), (1, amp, 7)):  # plain, SkewFunc (affine), seeded perturbation (general branch)
            m = O.Mesh(orc, 2, 2, 2, bounds=(-1, 1, 0, 0.7, 0, 1.3))
            if skew:
                m.skew(*skew)
            nodes, conn, _ = m.arrays()
            for e in range(0, m.n_elem, max(1, m.n_elem // 4)):
                X = nodes[conn[e]]
                a = bool(O.has_affine_map(orc, X))
                r = R.fe_map(elem, X, pts, w, a)
                Xs.append(X); aff.append(a); xyz.append(r["xyz"]); jxw.append(r["JxW"]); jac.append(r["jac"]); inv.append(r["inv"])
        mp[elem + "_order"] = np.array(order)
        mp[elem + "_X"], mp[elem + "_affine"] = np.array(Xs), np.array(aff)
        mp[elem + "_xyz"], mp[elem + "_JxW"], mp[elem + "_jac"], mp[elem + "_inv"] = np.array(xyz), np.array(jxw), np.array(jac), np.array(inv)
    path = os.path.join(ROOT, "tests", "golden", "reference_fe_map.npz")
    np.savez_compressed(path, **mp)
    print(f"wrote {path}: {len(mp)} arrays, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
