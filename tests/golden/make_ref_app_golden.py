"""Writes tests/golden/reference_ex4_runs.npz, reference_ex6_runs.npz and reference_hanging_runs.npz: the global matrix, right-hand side and Dirichlet constraints that the REFERENCE's
own examples/introduction/introduction_ex4/introduction_ex4.C produces — its main() ("-d 3 -n N [-o FIRST]") or, for the meshes
main() does not ask build_cube for, its assemble_poisson set up as main() does — compiled where it lies under /root/reference
into oracle/_ref/libref_app.so (oracle/Makefile; oracle/ref_app/ref_world.h says what in that library is the reference's code
and what is glue); for reference_ex6_runs.npz the same of LinearElasticity::assemble() of
examples/systems_of_equations/systems_of_equations_ex6/systems_of_equations_ex6.C (three variables, traction on the max-x sides,
homogeneous Dirichlet on the min-x side: BASELINE's config 5 in small); for reference_hanging_runs.npz assemble_poisson on
locally refined HEX8 / HEX27 meshes whose hanging-node rows go through the reference's own
DofMap::heterogeneously_constrain_element_matrix_and_vector / build_constraint_matrix_and_vector.  Run here, where /root/reference exists; the .npz travels to the GPU box, where tests/test_golden_ref_app.py
and tests/test_gpu_golden_ref.py compare the oracle and the CUDA path with it.

    python tests/golden/make_ref_app_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from oracle import ref as R  # noqa: E402

BOUNDS = (-1.0, 1.0, -1.0, 1.0, -1.0, 1.0)
# name, oracle element type, n (build_cube n^3), skew (kind, amplitude, seed) or None, FE order, through main()?
CASES = [
    ("hex27_main_n3", O.HEX27, 3, None, 2, True),
    ("hex8_main_n4", O.HEX8, 4, None, 1, True),
    ("hex27_skew_n2", O.HEX27, 2, (1, 0.03, 12345), 2, False),
    ("hex8_skew_n3", O.HEX8, 3, (1, 0.05, 777), 1, False),
    ("tet10_n2", O.TET10, 2, None, 2, False),
    ("tet10_skew_n2", O.TET10, 2, (1, 0.02, 99), 2, False),
    ("hex20_n2", O.HEX20, 2, None, 2, False),
    ("hex20_skew_n2", O.HEX20, 2, (1, 0.03, 4321), 2, False),
    ("prism18_n2", O.PRISM18, 2, None, 2, False),
    ("prism18_skew_n2", O.PRISM18, 2, (1, 0.03, 2468), 2, False),
]


def oracle_case(etype, n, skew):
    m = O.Mesh(etype, n, n, n, BOUNDS)
    if skew is not None:
        m.skew(*skew)
    s = O.System(m, 1)
    s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    s.build_sparsity()
    return m, s


def reference_run(m, s, n, order, through_main):
    nnz, nd = R.load_world(m, s, n, BOUNDS)
    if through_main:
        return R.run_ex4(["-d", 3, "-n", n] + (["-o", "FIRST"] if order == 1 else []), nnz, nd)
    return R.assemble_ex4(order, nnz, nd)


EX6_BOUNDS = (0.0, 1.0, 0.0, 0.3, 0.0, 0.1)
# name, oracle element type, (nx, ny, nz), skew or None, FE order
CASES6 = [
    ("hex27_ex6", O.HEX27, (2, 2, 2), None, 2),
    ("hex27_ex6_perturbed", O.HEX27, (2, 2, 2), (1, 0.004, 7), 2),
    ("hex8_ex6", O.HEX8, (5, 3, 2), None, 1),
    ("hex8_ex6_perturbed", O.HEX8, (4, 3, 3), (1, 0.004, 11), 1),
]


def oracle_case6(etype, dims, skew):
    m = O.Mesh(etype, *dims, EX6_BOUNDS)
    if skew is not None:
        m.skew(*skew)
    s = O.System(m, 3)
    s.add_dirichlet((4,), 1)                      # BOUNDARY_ID_MIN_X, ZeroFunction
    e, sd = m.boundary_sides(2)                   # BOUNDARY_ID_MAX_X
    s.set_side_loads(e, sd, (0.0, 0.0, -1.0))     # g_vec of systems_of_equations_ex6.C:240
    s.build_sparsity()
    return m, s


def reference_run6(m, s, dims, order):
    nnz, nd = R.load_world(m, s, dims, EX6_BOUNDS)
    R.load_sides(m)
    return R.assemble_ex6(order, nnz, nd, (4,))


def main6():
    out = {"names": np.array([c[0] for c in CASES6])}
    for name, etype, dims, skew, order in CASES6:
        m, s = oracle_case6(etype, dims, skew)
        r = reference_run6(m, s, dims, order)
        out[name + "/spec"] = np.array([etype, *dims, order] + list(skew or (-1, 0.0, 0)), dtype=np.float64)
        for k in ("values", "rhs", "con_dof", "con_val"):
            out[name + "/" + k] = r[k]
        print(f"{name}: {len(r['values'])} entries, {len(r['rhs'])} dofs, {len(r['con_dof'])} constrained")
    path = os.path.join(ROOT, "tests", "golden", "reference_ex6_runs.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


# name, element type, n (coarse cells per side), refined block [lo, hi)^3
CASES_H = [("hex8_hanging", O.HEX8, 4, 1, 3), ("hex27_hanging", O.HEX27, 3, 1, 2)]


def exact(p):
    return np.cos(.5 * np.pi * p[:, 0]) * np.sin(.5 * np.pi * p[:, 1]) * np.cos(.5 * np.pi * p[:, 2])


def hanging_case(etype, n, lo, hi):
    """(xyz, conn, mesh, system, elem dofs, ConstraintRows, pattern): the case of tests/test_gpu_constraints.py — hanging rows of
    the refined block plus Dirichlet rows (the exact solution of ex4) on the boundary."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hanging import refined_cube, refined_cube27
    from oracle import constraints as OC

    make = refined_cube27 if etype == O.HEX27 else refined_cube
    xyz, conn, mask, hang = make(n, lambda i, j, k: lo <= i < hi and lo <= j < hi and lo <= k < hi)
    m = O.Mesh.from_arrays(etype, xyz, conn, mask)
    s = O.System(m, 1)
    ed, nb, _ = s.dofs()
    rows = {int(nb[h]): [(int(nb[v]), c) for v, c in lst] for h, lst in hang.items()}
    vals = {}
    bn = np.nonzero(mask)[0]
    for node, v in zip(bn, exact(xyz[bn])):
        rows[int(nb[node])] = []
        vals[int(nb[node])] = float(v)
    con = OC.ConstraintRows(rows, vals)
    return xyz, conn, m, s, ed, con, OC.sparsity(con, ed, s.n_dofs)


def reference_run_h(m, s, con, pattern, etype):
    nnz, nd = R.load_world(m, s, (0, 0, 0), (0.0,) * 6, pattern=pattern)
    R.load_constraint_rows(*con.arrays())
    return R.assemble_ex4(1 if etype == O.HEX8 else 2, nnz, nd, boundary_ids=())


def main_h():
    out = {"names": np.array([c[0] for c in CASES_H])}
    for name, etype, n, lo, hi in CASES_H:
        xyz, conn, m, s, ed, con, pattern = hanging_case(etype, n, lo, hi)
        r = reference_run_h(m, s, con, pattern, etype)
        assert r["untouched_entries"] == 0
        out[name + "/spec"] = np.array([etype, n, lo, hi], dtype=np.int64)
        for k in ("values", "rhs"):
            out[name + "/" + k] = r[k]
        print(f"{name}: {len(r['values'])} entries, {len(r['rhs'])} dofs, {len(con.rows)} constraint rows")
    path = os.path.join(ROOT, "tests", "golden", "reference_hanging_runs.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


def main():
    out = {"names": np.array([c[0] for c in CASES])}
    for name, etype, n, skew, order, through_main in CASES:
        m, s = oracle_case(etype, n, skew)
        r = reference_run(m, s, n, order, through_main)
        out[name + "/spec"] = np.array([etype, n, order, int(through_main)] + list(skew or (-1, 0.0, 0)), dtype=np.float64)
        for k in ("values", "rhs", "con_dof", "con_val"):
            out[name + "/" + k] = r[k]
        print(f"{name}: {len(r['values'])} entries, {len(r['rhs'])} dofs, {len(r['con_dof'])} constrained")
    path = os.path.join(ROOT, "tests", "golden", "reference_ex4_runs.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
    main6()
    main_h()
