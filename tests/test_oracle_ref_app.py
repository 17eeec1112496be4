"""The oracle against the REFERENCE's own application sources of the path: examples/introduction/introduction_ex4/
introduction_ex4.C (assemble_poisson and main), examples/systems_of_equations/systems_of_equations_ex6/systems_of_equations_ex6.C
(LinearElasticity::assemble, traction side term included) and the constraint functions of src/base/dof_map_constraints.C, compiled where
they lie under /root/reference into oracle/_ref/libref_app.so (oracle/Makefile) and run over the stub world of
oracle/ref_app/ref_world.h — which says what in that library is the reference's code and what is glue.  The integer structures
(mesh, dof numbering, pattern) are the oracle's; every floating-point number of the global matrix and right-hand side comes out
of the reference's own loop, quadrature, shape functions, FEMap, DenseMatrix and constraint code.  Runs where /root/reference
exists; tests/test_golden_ref_app.py holds the same comparison against committed outputs for machines without it."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from hanging import refined_cube, refined_cube27  # noqa: E402
from oracle import constraints as OC  # noqa: E402
from oracle import oracle as O  # noqa: E402
from oracle import ref as R  # noqa: E402

pytestmark = pytest.mark.skipif(not R.app_available(), reason="oracle/_ref/libref_app.so exists only where /root/reference does")
BOUNDS = (-1.0, 1.0, -1.0, 1.0, -1.0, 1.0)


def _case(etype, n, skew=None):
    m = O.Mesh(etype, n, n, n, BOUNDS)
    if skew is not None:
        m.skew(*skew)
    s = O.System(m, 1)
    s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    s.build_sparsity()
    return m, s


def _same(s, out):
    v, r, _ = s.assemble(O.POISSON_EX4)
    cd, cv = s.constraints()
    assert np.array_equal(cd, out["con_dof"]) and np.array_equal(cv, out["con_val"]), "Dirichlet constraints differ"
    assert np.array_equal(v, out["values"]), f"matrix values differ by {np.abs(v - out['values']).max()}"
    assert np.array_equal(r, out["rhs"]), f"right-hand side differs by {np.abs(r - out['rhs']).max()}"
    # the sparsity pattern is EXACTLY what the example's element loop adds into: an entry outside it throws in the stub matrix
    # ("not in the sparsity pattern"), and no entry of it stays untouched — what SparsityPattern::Build computes for this system
    assert out["untouched_entries"] == 0


@pytest.mark.parametrize("etype,n,extra", [(O.HEX27, 3, []), (O.HEX27, 6, []), (O.HEX8, 5, ["-o", "FIRST"])], ids=["hex27_n3", "hex27_n6_published_shape", "hex8_n5"])
def test_main_of_introduction_ex4_gives_the_oracles_bits(etype, n, extra, capfd):
    """main(): command line, build_cube request, add_variable, DirichletBoundary with exact_solution_wrapper, init, solve."""
    m, s = _case(etype, n)
    nnz, nd = R.load_world(m, s, n, BOUNDS)
    out = R.run_ex4(["-d", 3, "-n", n] + extra, nnz, nd)
    assert "Running introduction_ex4 -d 3 -n %d" % n in capfd.readouterr().out  # introduction_ex4.C:141-146 ran
    _same(s, out)
    if etype == O.HEX27 and n == 6:
        assert nd == 2197 and len(out["con_dof"]) == 866  # the run facts the example's page publishes (tests/golden/reference_runs.json)


def test_main_refuses_a_mesh_it_did_not_ask_for():
    m, s = _case(O.HEX27, 2)
    nnz, nd = R.load_world(m, s, 2, BOUNDS)
    with pytest.raises(RuntimeError, match="not the mesh that was loaded"):
        R.run_ex4(["-d", 3, "-n", 3], nnz, nd)


ASSEMBLE = [("hex27_skew", O.HEX27, 3, (1, 0.03, 12345), 2), ("hex27_skewfunc", O.HEX27, 3, (0, 0.0, 0), 2), ("hex8_skew", O.HEX8, 4, (1, 0.05, 777), 1),
            ("tet10", O.TET10, 2, None, 2), ("tet10_skew", O.TET10, 2, (1, 0.02, 99), 2), ("hex20", O.HEX20, 3, None, 2),
            ("hex20_skew", O.HEX20, 3, (1, 0.03, 4321), 2), ("prism18", O.PRISM18, 3, None, 2), ("prism18_skew", O.PRISM18, 2, (1, 0.03, 2468), 2)]


@pytest.mark.parametrize("name,etype,n,skew,order", ASSEMBLE, ids=[c[0] for c in ASSEMBLE])
def test_assemble_poisson_on_every_family_and_both_map_branches(name, etype, n, skew, order):
    """assemble_poisson on the meshes main() does not build: each element family of the product, affine (compute_affine_map) and
    skewed (compute_single_point_map per point) — bitwise."""
    m, s = _case(etype, n, skew)
    nnz, nd = R.load_world(m, s, n, BOUNDS)
    _same(s, R.assemble_ex4(order, nnz, nd))


@pytest.mark.parametrize("n", [20, 32])
def test_config_1_at_full_size(n):
    """BASELINE.json configs[0]: 20^3 HEX27, 8 000 elements, 68 921 dofs — every one of the 4.1 M entries bitwise; and 32^3
    (32 768 elements, 17 M entries), the size of the GPU suite's largest HEX27 parity case."""
    m, s = _case(O.HEX27, n)
    nnz, nd = R.load_world(m, s, n, BOUNDS)
    assert nd == (2 * n + 1) ** 3
    _same(s, R.run_ex4(["-d", 3, "-n", n], nnz, nd))


# ---- systems_of_equations_ex6.C: LinearElasticity::assemble() -----------------------------------------------------------------------
EX6_BOUNDS = (0.0, 1.0, 0.0, 0.3, 0.0, 0.1)
EX6 = [("hex8_as_the_example_runs_it", O.HEX8, (8, 4, 2), None, 1), ("hex8_perturbed", O.HEX8, (4, 3, 3), (1, 0.004, 11), 1),
       ("hex27_config5_small", O.HEX27, (4, 3, 2), None, 2), ("hex27_perturbed", O.HEX27, (3, 2, 2), (1, 0.004, 7), 2),
       ("hex27_config5_cube_8", O.HEX27, (8, 8, 8), None, 2)]  # BASELINE config 5's mesh shape (n^3 HEX27 on the ex6 bar) at n = 8: 14 739 dofs


@pytest.mark.parametrize("name,etype,dims,skew,order", EX6, ids=[c[0] for c in EX6])
def test_linear_elasticity_assemble_of_ex6_gives_the_oracles_bits(name, etype, dims, skew, order):
    """The 6-deep tensor loop over DenseSubMatrix blocks (:220-237), the body force, the traction on the max-x sides through
    fe_face->reinit(elem, side) (:240-257) and constrain_element_matrix_and_vector (:259) — matrix and right-hand side bitwise."""
    m = O.Mesh(etype, *dims, EX6_BOUNDS)
    if skew is not None:
        m.skew(*skew)
    s = O.System(m, 3)
    s.add_dirichlet((4,), 1)
    e, sd = m.boundary_sides(2)
    s.set_side_loads(e, sd, (0.0, 0.0, -1.0))
    s.build_sparsity()
    nnz, nd = R.load_world(m, s, dims, EX6_BOUNDS)
    R.load_sides(m)
    out = R.assemble_ex6(order, nnz, nd, (4,))
    v, r, _ = s.assemble(O.ELASTICITY_EX6)
    cd, cv = s.constraints()
    assert np.array_equal(cd, out["con_dof"]) and np.array_equal(cv, out["con_val"])
    assert np.array_equal(v, out["values"]) and np.array_equal(r, out["rhs"]) and out["untouched_entries"] == 0
    # the traction really is in there: without the side list the oracle's right-hand side differs
    s.set_side_loads([], [], (0.0, 0.0, -1.0))
    assert np.abs(s.assemble(O.ELASTICITY_EX6)[1] - out["rhs"]).max() > 1e-4


# ---- DofMap::constrain_element_matrix_and_vector / heterogeneously_... / build_constraint_matrix(_and_vector) ----------------------
def _both(rows, vals, K, F, dofs, hetero=True, asym=True):
    K1, F1, d1 = OC.constrain(OC.ConstraintRows(rows, vals), K.copy(), F.copy(), [int(d) for d in dofs], hetero, asym)
    K2, F2, d2 = R.constrain_element(K, F, dofs, {k: dict(v) for k, v in rows.items()}, vals, hetero, asym)
    assert list(d1) == [int(x) for x in d2], "the widened dof sets differ"
    scale = max(np.abs(K).max(), np.abs(F).max(), 1.0)
    assert np.abs(K1 - K2).max() <= 4e-16 * scale * len(d1) and np.abs(F1 - F2).max() <= 4e-16 * scale * len(d1)
    return K2, F2, d2


@pytest.mark.parametrize("hetero", [True, False])
@pytest.mark.parametrize("asym", [True, False])
def test_constraint_functions_of_the_reference_on_small_cases(hetero, asym):
    rng = np.random.default_rng(11)
    K = rng.standard_normal((4, 4))
    K = K + K.T
    F = rng.standard_normal(4)
    _both({1: [(5, 0.5), (6, 0.5)], 6: []}, {6: 2.0}, K, F, [0, 1, 2, 6], hetero, asym)                 # a master with a Dirichlet row
    K2, F2, d2 = _both({1: [(5, 0.5), (6, 0.5)], 6: [(7, 0.5), (8, 0.5)]}, {6: 0.25}, K, F, [0, 1, 2, 3], hetero, asym)  # two levels
    assert [int(x) for x in d2] == [0, 1, 2, 3, 5, 6, 7, 8]
    _both({1: [(0, 0.5), (2, 0.5)]}, {}, K, F, [0, 1, 2, 3], hetero, asym)                               # constrained by the element's own dofs
    _both({9: [(0, 1.0)]}, {9: 1.0}, K, F, [0, 1, 2, 3], hetero, asym)                                   # nothing of this element is constrained
    _both({0: [], 1: [], 2: [], 3: []}, {0: 1.0, 2: -2.0}, K, F, [0, 1, 2, 3], hetero, asym)             # every dof Dirichlet


def test_dirichlet_rows_the_c_oracle_against_the_reference():
    """The C oracle's element-level constraint (the one inside every assembled value above) against the reference's function, per
    element, bitwise."""
    m, s = _case(O.HEX27, 3)
    ed, _, _ = s.dofs()
    cd, cv = s.constraints()
    rows = {int(d): {} for d in cd}
    vals = {int(d): float(v) for d, v in zip(cd, cv)}
    Ku, Fu = s.element_matrices(O.POISSON_EX4, constrained=False)
    Kc, Fc = s.element_matrices(O.POISSON_EX4, constrained=True)
    for e in range(m.n_elem):
        K, F, d = R.constrain_element(Ku[e], Fu[e], ed[e], rows, vals, True, True)
        assert np.array_equal(d, ed[e]) and np.array_equal(K, Kc[e]) and np.array_equal(F, Fc[e])


@pytest.mark.parametrize("which", ["hex8", "hex27"])
def test_hanging_node_rows_the_restatement_against_the_reference(which):
    """Locally refined meshes: every element through the reference's general path (C with off-diagonal weights, widened dof sets,
    C^T K C, C^T (F - K H)) against oracle/constraints.py, which the product's general-constraint kernels are tested against."""
    if which == "hex8":
        xyz, conn, mask, hang = refined_cube(4, lambda i, j, k: 1 <= i < 3 and 1 <= j < 3 and 1 <= k < 3)
        m = O.Mesh.from_arrays(O.HEX8, xyz, conn, mask)
    else:
        xyz, conn, mask, hang = refined_cube27(3, lambda i, j, k: i == 1 and j == 1 and k == 1)
        m = O.Mesh.from_arrays(O.HEX27, xyz, conn, mask)
    s = O.System(m, 1)
    ed, nb, _ = s.dofs()
    rows = {int(nb[h]): [(int(nb[v]), c) for v, c in lst] for h, lst in hang.items()}
    vals = {}
    for node in np.nonzero(mask)[0]:
        rows[int(nb[node])] = []
        vals[int(nb[node])] = float(0.3 + xyz[node, 0] - 0.5 * xyz[node, 1])
    Ku, Fu = s.element_matrices(O.POISSON_EX4, constrained=False)
    widened = 0
    for e in range(m.n_elem):
        _, _, d = _both(rows, vals, Ku[e], Fu[e], ed[e])
        widened += len(d) > len(ed[e])
    assert widened > 0


@pytest.mark.parametrize("which", ["hex8", "hex27"])
def test_hanging_node_mesh_assembled_by_the_references_own_loop_and_constraint_code(which):
    """assemble_poisson of introduction_ex4.C on a locally refined mesh, with the hanging-node and Dirichlet rows loaded into the
    stub DofMap: every element goes through the reference's own heterogeneously_constrain_element_matrix_and_vector (widened dof
    sets, C^T K C, C^T (F - K H)) and add_matrix.  The global system is oracle/constraints.py's (bitwise for HEX8's weights 1/2,
    1/4; to rounding for HEX27, where numpy's matrix products sum in another order), and the pattern with the couplings
    DofMap::find_connected_dofs adds is exactly the set of entries the loop adds into."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_ref_app_golden as G

    name, etype, n, lo, hi = G.CASES_H[0 if which == "hex8" else 1]
    xyz, conn, m, s, ed, con, (rp, ci) = G.hanging_case(etype, n, lo, hi)
    out = G.reference_run_h(m, s, con, (rp, ci), etype)
    assert out["untouched_entries"] == 0 and len(out["con_dof"]) == len(con.rows)
    Ku, Fu = s.element_matrices(O.POISSON_EX4, constrained=False)
    vo, ro = OC.assemble(con, ed, Ku, Fu, rp, ci)
    assert np.abs(vo - out["values"]).max() <= 4e-16 * np.abs(vo).max()
    assert np.abs(ro - out["rhs"]).max() <= 4e-16 * max(np.abs(ro).max(), 1.0)
    if which == "hex8":
        assert np.array_equal(vo, out["values"])
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_hanging_runs.npz"))
    assert np.array_equal(out["values"], g[name + "/values"]) and np.array_equal(out["rhs"], g[name + "/rhs"])


def test_committed_golden_outputs_are_what_the_reference_gives_today():
    """tests/golden/reference_ex4_runs.npz (written by make_ref_app_golden.py) regenerated and compared."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_ref_app_golden as G

    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_ex4_runs.npz"))
    assert list(g["names"]) == [c[0] for c in G.CASES]
    for name, etype, n, skew, order, through_main in G.CASES:
        m, s = G.oracle_case(etype, n, skew)
        r = G.reference_run(m, s, n, order, through_main)
        for k in ("values", "rhs", "con_dof", "con_val"):
            assert np.array_equal(r[k], g[name + "/" + k]), (name, k)
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_ex6_runs.npz"))
    assert list(g["names"]) == [c[0] for c in G.CASES6]
    for name, etype, dims, skew, order in G.CASES6:
        m, s = G.oracle_case6(etype, dims, skew)
        r = G.reference_run6(m, s, dims, order)
        for k in ("values", "rhs", "con_dof", "con_val"):
            assert np.array_equal(r[k], g[name + "/" + k]), (name, k)
