"""The general-constraint restatement (oracle/constraints.py) pinned by construction: with Dirichlet rows only it is the C
oracle's element-level constraint entry for entry; on a hanging-node mesh the constrained system reproduces a linear
field exactly (the patch test a conforming constraint must pass), and its matrix rows have the libMesh form."""
import numpy as np
import pytest

from hanging import refined_cube, refined_cube27
from oracle import constraints as OC
from oracle import oracle as O


def linear(p):
    return 0.3 + p[:, 0] - 0.5 * p[:, 1] + 0.25 * p[:, 2]


def hanging_case(n=4):
    xyz, conn, mask, hang = refined_cube(n, lambda i, j, k: 1 <= i < n - 1 and 1 <= j < n - 1 and 1 <= k < n - 1)
    m = O.Mesh.from_arrays(O.HEX8, xyz, conn, mask)
    s = O.System(m, 1)
    ed, nb, _ = s.dofs()
    rows = {int(nb[h]): [(int(nb[v]), c) for v, c in lst] for h, lst in hang.items()}
    vals = {}
    for node in np.nonzero(mask)[0]:
        assert node not in hang, "refine interior elements only: no hanging node on the boundary"
        rows[int(nb[node])] = []
        vals[int(nb[node])] = float(linear(xyz[node:node + 1])[0])
    return xyz, conn, mask, hang, m, s, ed, nb, OC.ConstraintRows(rows, vals)


def test_dirichlet_rows_only_equal_the_c_oracle():
    m = O.Mesh(O.HEX8, 3, 3, 2, bounds=(-1, 1, -1, 1, -1, 1))
    s = O.System(m, 1)
    s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    s.build_sparsity()
    ed, _, _ = s.dofs()
    cd, cv = s.constraints()
    con = OC.ConstraintRows({int(d): [] for d in cd}, {int(d): float(v) for d, v in zip(cd, cv)})
    Ku, Fu = s.element_matrices(O.POISSON_EX4, constrained=False, qorder=3)
    Kc, Fc = s.element_matrices(O.POISSON_EX4, constrained=True, qorder=3)
    for e in range(m.n_elem):
        K, F, dofs = OC.constrain(con, Ku[e], Fu[e], ed[e])
        assert dofs == [int(d) for d in ed[e]]
        assert np.array_equal(K, Kc[e]) or np.abs(K - Kc[e]).max() <= 1e-15 * np.abs(Kc[e]).max()
        assert np.abs(F - Fc[e]).max() <= 1e-14 * max(np.abs(Fc[e]).max(), 1e-300)
    rp, ci = OC.sparsity(con, ed, s.n_dofs)
    rp0, ci0 = s.sparsity()
    assert np.array_equal(rp, rp0) and np.array_equal(ci, ci0)


def test_hanging_node_mesh_passes_the_patch_test():
    xyz, conn, mask, hang, m, s, ed, nb, con = hanging_case(4)
    assert len(hang) > 0 and m.n_elem == 4 ** 3 - 8 + 64
    Ku, Fu = s.element_matrices(O.LAPLACE, constrained=False, qorder=3)
    rp, ci = OC.sparsity(con, ed, s.n_dofs)
    vals, rhs = OC.assemble(con, ed, Ku, Fu, rp, ci)
    A = np.zeros((s.n_dofs, s.n_dofs))
    for r in range(s.n_dofs):
        A[r, ci[int(rp[r]):int(rp[r + 1])]] = vals[int(rp[r]):int(rp[r + 1])]
    u = np.linalg.solve(A, rhs)
    exact = np.zeros(s.n_dofs)
    exact[nb] = linear(xyz)
    assert np.abs(u - exact).max() < 1e-12   # linear fields are in the constrained space: reproduced to rounding
    # a hanging row has the libMesh form: t on the diagonal, -coef t at its masters (t = elements touching it), rhs 0
    h = next(iter(hang))
    r = int(nb[h])
    t = A[r, r]
    assert t == float((conn == h).sum())
    for v, c in hang[h]:
        assert A[r, int(nb[v])] == -c * t
    assert rhs[r] == 0.0
    # an element that touches a hanging node is widened by the masters it does not hold itself
    e = int(np.nonzero((conn == h).any(axis=1))[0][0])
    K, F, dofs = OC.constrain(con, Ku[e], Fu[e], ed[e])
    assert len(dofs) > 8 and K.shape == (len(dofs), len(dofs))


def test_hex27_hanging_nodes_hold_quadratics():
    """HEX27 with one refined cell: 72 hanging nodes with the quadratic interpolation weights (3/8, 3/4, -1/8 on an edge, their
    products on a face); the constrained Laplace system reproduces a harmonic quadratic at every node (Q2 patch test)."""
    quad = lambda p: 0.3 + p[:, 0] - 0.5 * p[:, 1] + 0.25 * p[:, 2] + p[:, 0] ** 2 - p[:, 1] ** 2 + 0.5 * p[:, 0] * p[:, 2]
    xyz, conn, mask, hang = refined_cube27(3, lambda i, j, k: i == 1 and j == 1 and k == 1)
    assert len(hang) == 72 and all(abs(sum(c for _, c in lst) - 1.0) < 1e-15 for lst in hang.values())  # weights sum to one
    m = O.Mesh.from_arrays(O.HEX27, xyz, conn, mask)
    s = O.System(m, 1)
    ed, nb, _ = s.dofs()
    rows = {int(nb[h]): [(int(nb[v]), c) for v, c in lst] for h, lst in hang.items()}
    vals = {}
    for node in np.nonzero(mask)[0]:
        rows[int(nb[node])] = []
        vals[int(nb[node])] = float(quad(xyz[node:node + 1])[0])
    con = OC.ConstraintRows(rows, vals)
    Ku, Fu = s.element_matrices(O.LAPLACE, constrained=False, qorder=5)
    rp, ci = OC.sparsity(con, ed, s.n_dofs)
    vals_, rhs = OC.assemble(con, ed, Ku, Fu, rp, ci)
    A = np.zeros((s.n_dofs, s.n_dofs))
    for r in range(s.n_dofs):
        A[r, ci[int(rp[r]):int(rp[r + 1])]] = vals_[int(rp[r]):int(rp[r + 1])]
    u = np.linalg.solve(A, rhs)
    exact = np.zeros(s.n_dofs)
    exact[nb] = quad(xyz)
    assert np.abs(u - exact).max() < 1e-12


def test_recursion_follows_new_dofs_only():
    """build_constraint_matrix_and_vector recurses only while NEW dofs appear (dof_map_constraints.C:3595-3600): a master that
    is itself constrained in terms of further dofs is substituted (x = C y + h, y = D z + g), a master with an empty
    (Dirichlet) row stays a column of C — its value reaches the system through its own constrained row."""
    con = OC.ConstraintRows({1: [(5, 0.5), (6, 0.5)], 6: []}, {6: 2.0})
    dofs = [0, 1]
    Cm, H = OC.build_constraint_matrix_and_vector(con, dofs)
    assert dofs == [0, 1, 5, 6] and Cm.shape == (2, 4)
    assert Cm[0, 0] == 1.0 and Cm[1, 1] == 0.0 and Cm[1, 2] == 0.5 and Cm[1, 3] == 0.5 and H[1] == 0.0
    # two levels: 1 = (5 + 6) / 2 with 6 = (7 + 8) / 2 + 0.25: the second level is substituted
    con = OC.ConstraintRows({1: [(5, 0.5), (6, 0.5)], 6: [(7, 0.5), (8, 0.5)]}, {6: 0.25})
    dofs = [0, 1]
    Cm, H = OC.build_constraint_matrix_and_vector(con, dofs)
    assert dofs == [0, 1, 5, 6, 7, 8]
    assert Cm[1].tolist() == [0.0, 0.0, 0.5, 0.0, 0.25, 0.25] and H[1] == 0.125
