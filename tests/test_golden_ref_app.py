"""The oracle against outputs of the reference itself: tests/golden/reference_ex4_runs.npz holds the global matrix, right-hand
side and Dirichlet constraints that the reference's own introduction_ex4.C (main() or assemble_poisson, compiled where it lies —
tests/golden/make_ref_app_golden.py, oracle/ref_app/ref_world.h) produced on ten small meshes: every element family of the
product, affine and skewed; tests/golden/reference_ex6_runs.npz the same of LinearElasticity::assemble() of its
systems_of_equations_ex6.C (three variables, traction on the max-x sides) on four.  The oracle reproduces them bit for bit.
Needs neither /root/reference nor a GPU."""
import os

import numpy as np
import pytest

from oracle import oracle as O

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_ex4_runs.npz"))
BOUNDS = (-1.0, 1.0, -1.0, 1.0, -1.0, 1.0)


def golden_case(name):
    etype, n, order, through_main, kind, amp, seed = G[name + "/spec"]
    m = O.Mesh(int(etype), int(n), int(n), int(n), BOUNDS)
    if kind >= 0:
        m.skew(int(kind), float(amp), int(seed))
    s = O.System(m, 1)
    s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    s.build_sparsity()
    return m, s


@pytest.mark.parametrize("name", [str(n) for n in G["names"]])
def test_oracle_reproduces_the_references_own_output(name):
    m, s = golden_case(name)
    v, r, _ = s.assemble(O.POISSON_EX4)
    cd, cv = s.constraints()
    assert np.array_equal(cd, G[name + "/con_dof"]) and np.array_equal(cv, G[name + "/con_val"])
    assert np.array_equal(v, G[name + "/values"])
    assert np.array_equal(r, G[name + "/rhs"])


G6 = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_ex6_runs.npz"))
EX6_BOUNDS = (0.0, 1.0, 0.0, 0.3, 0.0, 0.1)


def golden_case6(name):
    etype, nx, ny, nz, order, kind, amp, seed = G6[name + "/spec"]
    m = O.Mesh(int(etype), int(nx), int(ny), int(nz), EX6_BOUNDS)
    if kind >= 0:
        m.skew(int(kind), float(amp), int(seed))
    s = O.System(m, 3)
    s.add_dirichlet((4,), 1)
    e, sd = m.boundary_sides(2)
    s.set_side_loads(e, sd, (0.0, 0.0, -1.0))
    s.build_sparsity()
    return m, s, e, sd


@pytest.mark.parametrize("name", [str(n) for n in G6["names"]])
def test_oracle_reproduces_the_references_own_elasticity_output(name):
    m, s, _, _ = golden_case6(name)
    v, r, _ = s.assemble(O.ELASTICITY_EX6)
    cd, cv = s.constraints()
    assert np.array_equal(cd, G6[name + "/con_dof"]) and np.array_equal(cv, G6[name + "/con_val"])
    assert np.array_equal(v, G6[name + "/values"])
    assert np.array_equal(r, G6[name + "/rhs"])


GH = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_hanging_runs.npz"))


def golden_case_h(name):
    """The locally refined mesh of a hanging-node golden case: (xyz, conn, mesh, system, elem dofs, ConstraintRows, (rp, ci))."""
    import sys

    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from hanging import refined_cube, refined_cube27
    from oracle import constraints as OC

    etype, n, lo, hi = (int(x) for x in GH[name + "/spec"])
    make = refined_cube27 if etype == O.HEX27 else refined_cube
    xyz, conn, mask, hang = make(n, lambda i, j, k: lo <= i < hi and lo <= j < hi and lo <= k < hi)
    m = O.Mesh.from_arrays(etype, xyz, conn, mask)
    s = O.System(m, 1)
    ed, nb, _ = s.dofs()
    rows = {int(nb[h]): [(int(nb[v]), c) for v, c in lst] for h, lst in hang.items()}
    vals = {}
    bn = np.nonzero(mask)[0]
    g = np.cos(.5 * np.pi * xyz[bn, 0]) * np.sin(.5 * np.pi * xyz[bn, 1]) * np.cos(.5 * np.pi * xyz[bn, 2])
    for node, v in zip(bn, g):
        rows[int(nb[node])] = []
        vals[int(nb[node])] = float(v)
    con = OC.ConstraintRows(rows, vals)
    return xyz, conn, m, s, ed, con, OC.sparsity(con, ed, s.n_dofs)


@pytest.mark.parametrize("name", [str(n) for n in GH["names"]])
def test_constraint_restatement_reproduces_the_references_hanging_node_system(name):
    """oracle/constraints.py (what the product's general-constraint kernels are tested against) against the global system the
    reference's own element loop and constraint functions produced on a locally refined mesh."""
    from oracle import constraints as OC

    xyz, conn, m, s, ed, con, (rp, ci) = golden_case_h(name)
    Ku, Fu = s.element_matrices(O.POISSON_EX4, constrained=False)
    vo, ro = OC.assemble(con, ed, Ku, Fu, rp, ci)
    assert len(vo) == len(GH[name + "/values"])
    assert np.abs(vo - GH[name + "/values"]).max() <= 4e-16 * np.abs(vo).max()
    assert np.abs(ro - GH[name + "/rhs"]).max() <= 4e-16 * max(np.abs(ro).max(), 1.0)
