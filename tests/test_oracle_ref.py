"""The oracle against the REFERENCE'S OWN CODE for the tables of FE::reinit: the Lagrange shape functions and their
reference derivatives (src/fe/fe_lagrange_shape_{2,3}D.C, include/fe/fe_lagrange_shape_1D.h) and the Gauss rules
(src/quadrature/quadrature_gauss_{1,2,3}D.C, quadrature.C, quadrature_gauss.C).  Those reference sources are compiled where
they lie into oracle/_ref/libref_fe.so (oracle/Makefile; stub declarations in oracle/ref_stubs/), and their outputs are
committed as tests/golden/reference_fe_tables.npz (tests/golden/make_ref_golden.py).  The bar is bit-for-bit equality: the
restatement follows the reference's operation order, so nothing may differ."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from oracle import ref as R

GOLD_MAP = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_fe_map.npz"))
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_fe_tables.npz"))
ORC = {"HEX8": O.HEX8, "HEX27": O.HEX27, "HEX20": O.HEX20, "TET10": O.TET10, "PRISM18": O.PRISM18}
VOLUME_CASES = [("HEX8", 3), ("HEX8", 5), ("HEX27", 3), ("HEX27", 5), ("HEX20", 5), ("TET10", 5), ("PRISM18", 5)]
needs_ref = pytest.mark.skipif(not R.available(), reason="oracle/_ref/libref_fe.so is built only where /root/reference exists")


@pytest.mark.parametrize("elem,order", VOLUME_CASES)
def test_quadrature_rules_are_the_reference_bits(elem, order):
    pts, w = O.quadrature(ORC[elem], order)
    k = f"{elem}_q{order}"
    assert pts.shape == GOLD[k + "_pts"].shape
    assert np.array_equal(pts, GOLD[k + "_pts"]) and np.array_equal(w, GOLD[k + "_w"])


@pytest.mark.parametrize("elem,order", VOLUME_CASES)
def test_shape_tables_at_the_rule_points_are_the_reference_bits(elem, order):
    """What FE::reinit tabulates once per element type: phi_i(q), dphi_i/dxi(q), /deta, /dzeta."""
    k = f"{elem}_q{order}"
    got = O.shapes(ORC[elem], GOLD[k + "_pts"])
    for name, a in zip(("phi", "dxi", "deta", "dzeta"), got):
        assert np.array_equal(a, GOLD[f"{k}_{name}"]), name


@pytest.mark.parametrize("elem", sorted(ORC))
def test_shape_functions_at_random_points_are_the_reference_bits(elem):
    got = O.shapes(ORC[elem], GOLD[elem + "_rand_pts"])
    for name, a in zip(("phi", "dxi", "deta", "dzeta"), got):
        assert np.array_equal(a, GOLD[f"{elem}_rand_{name}"]), name


HEX_REF = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1], [0, -1, -1],
                    [1, 0, -1], [0, 1, -1], [-1, 0, -1], [-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0], [0, -1, 1], [1, 0, 1],
                    [0, 1, 1], [-1, 0, 1], [0, 0, -1], [0, -1, 0], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0, 0, 1], [0, 0, 0]], dtype=float)
TET_REF = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [.5, 0, 0], [.5, .5, 0], [0, .5, 0], [0, 0, .5], [.5, 0, .5], [0, .5, .5]], dtype=float)


@pytest.mark.parametrize("elem,side_elem,order", [("HEX8", "QUAD4", 3), ("HEX8", "QUAD4", 5), ("HEX27", "QUAD9", 5), ("HEX20", "QUAD8", 5)])
def test_side_rule_on_the_reference_hex_is_the_reference_quad_rule(elem, side_elem, order):
    """Side 5 (zeta = +1) of the reference hex has its side nodes at (xi, eta) = the QUAD's own reference nodes, so the side
    points are (x_q, y_q, 1) of QGauss(2, order) on that QUAD and JxW_face = w_q (the side map is the identity)."""
    X = HEX_REF[: ORC[elem]]
    d = O.side_reinit(ORC[elem], X, 5, order)
    k = f"{side_elem}_q{order}"
    assert np.abs(d["ref"][:, :2] - GOLD[k + "_pts"][:, :2]).max() < 1e-15 and np.abs(d["ref"][:, 2] - 1).max() < 1e-15
    assert np.abs(d["JxW"] - GOLD[k + "_w"]).max() < 1e-15
    assert np.abs(d["xyz"] - d["ref"]).max() < 1e-15


def test_side_rule_on_the_reference_tet_is_the_reference_triangle_rule():
    """Side 0 of the reference tet (nodes 0, 2, 1: the face zeta = 0 walked clockwise): side point (s, t) sits at
    (xi, eta) = (t, s); JxW_face = w_q (unit right triangle)."""
    d = O.side_reinit(O.TET10, TET_REF, 0, 5)
    q, w = GOLD["TRI6_q5_pts"], GOLD["TRI6_q5_w"]
    assert np.abs(d["ref"][:, 0] - q[:, 1]).max() < 1e-15 and np.abs(d["ref"][:, 1] - q[:, 0]).max() < 1e-15
    assert np.abs(d["ref"][:, 2]).max() < 1e-15 and np.abs(d["JxW"] - w).max() < 1e-15


@needs_ref
def test_golden_tables_are_what_the_compiled_reference_gives_now():
    from golden.make_ref_golden import CASES

    for elem, order in CASES:
        pts, w = R.gauss_rule(elem, order)
        k = f"{elem}_q{order}"
        assert np.array_equal(pts, GOLD[k + "_pts"]) and np.array_equal(w, GOLD[k + "_w"])
        for name, a in zip(("phi", "dxi", "deta", "dzeta"), R.shapes(elem, pts)):
            assert np.array_equal(a, GOLD[f"{k}_{name}"])


@needs_ref
@pytest.mark.parametrize("elem", sorted(ORC))
def test_oracle_equals_the_compiled_reference_on_fresh_points(elem):
    rng = np.random.default_rng(sum(map(ord, elem)))
    if elem == "TET10":
        pts = rng.uniform(0.0, 0.33, size=(500, 3))
    elif elem == "PRISM18":
        pts = np.c_[rng.uniform(0.0, 0.5, size=(500, 2)), rng.uniform(-1, 1, size=500)]
    else:
        pts = rng.uniform(-1, 1, size=(500, 3))
    for a, b in zip(O.shapes(ORC[elem], pts), R.shapes(elem, pts)):
        assert np.array_equal(a, b)


@needs_ref
@pytest.mark.parametrize("elem,side", [("HEX27", 2), ("HEX20", 0), ("HEX8", 4), ("TET10", 3)])
def test_parent_shapes_at_the_side_points_are_the_reference_bits(elem, side):
    """FE::reinit(elem, side): the parent's shape functions evaluated at the side points carried into the parent's reference
    space — through the reference's own shape code at the oracle's mapped points."""
    X = TET_REF if elem == "TET10" else HEX_REF[: ORC[elem]]
    d = O.side_reinit(ORC[elem], X, side, 5)
    assert np.array_equal(d["phi"], R.shapes(elem, d["ref"])[0])


@needs_ref
def test_exact_solution_is_the_example_file():
    """examples/introduction/introduction_ex3/exact_solution.C compiled as it is: the Dirichlet data and the FD forcing of
    ex4 are built from this function."""
    rng = np.random.default_rng(4)
    for x, y, z in rng.uniform(-1.001, 1.001, size=(2000, 3)):
        assert O.exact_solution(x, y, z) == R.exact_solution(x, y, z)


def _map_dphi(sh, inv):
    """H1FETransformation::map_dphi (src/fe/h1_fe_transformation.C:150-187), 3D: dphi[i][q][d] =
    dphidxi[i][q] * dxidx_d[q] + dphideta[i][q] * detadx_d[q] + dphidzeta[i][q] * dzetadx_d[q], in that order."""
    return sh[1][:, :, None] * inv[None, :, 0, :] + sh[2][:, :, None] * inv[None, :, 1, :] + sh[3][:, :, None] * inv[None, :, 2, :]


@pytest.mark.parametrize("elem", sorted(ORC))
def test_fe_reinit_is_the_reference_femap_bit_for_bit(elem):
    """FE::reinit = FEMap::compute_map (affine and general branch) + map_dphi: the oracle's xyz, JxW and physical gradients
    against what the reference's own fe_map.C gave for the same elements (plain, sheared, perturbed)."""
    order = int(GOLD_MAP[elem + "_order"])
    k = f"{elem}_q{order}"
    pts, w = GOLD[k + "_pts"], GOLD[k + "_w"]
    sh = [GOLD[f"{k}_{n}"] for n in ("phi", "dxi", "deta", "dzeta")]
    n_general = 0
    for X, aff, xyz, jxw, inv in zip(GOLD_MAP[elem + "_X"], GOLD_MAP[elem + "_affine"], GOLD_MAP[elem + "_xyz"], GOLD_MAP[elem + "_JxW"], GOLD_MAP[elem + "_inv"]):
        r = O.fe_reinit(ORC[elem], X, pts, w)
        assert r["affine"] == bool(aff)
        n_general += not aff
        assert np.array_equal(r["xyz"], xyz) and np.array_equal(r["JxW"], jxw)
        assert np.array_equal(r["phi"], sh[0]) and np.array_equal(r["dphi"], _map_dphi(sh, inv))
    assert 0 < n_general < len(GOLD_MAP[elem + "_affine"])  # both branches of compute_map are in the fixture


@needs_ref
@pytest.mark.parametrize("elem,amp", [("HEX27", 0.04), ("HEX8", 0.05), ("HEX20", 0.04), ("TET10", 0.008), ("PRISM18", 0.01)])
def test_fe_reinit_equals_the_compiled_femap_on_fresh_meshes(elem, amp):
    orc = ORC[elem]
    pts, w = O.quadrature(orc, 3 if elem == "HEX8" else 5)
    sh = R.shapes(elem, pts)
    for skew in (None, (0, 0.0, 0), (1, amp, 31)):
        m = O.Mesh(orc, 3, 2, 2, bounds=(-1, 1, -1, 1, 0, 1.5))
        if skew:
            m.skew(*skew)
        xyz, conn, _ = m.arrays()
        for e in range(m.n_elem):
            X = xyz[conn[e]]
            a = O.fe_reinit(orc, X, pts, w)
            b = R.fe_map(elem, X, pts, w, a["affine"])
            assert np.array_equal(a["xyz"], b["xyz"]) and np.array_equal(a["JxW"], b["JxW"])
            assert np.array_equal(a["dphi"], _map_dphi(sh, b["inv"]))


@needs_ref
def test_the_reference_throws_on_the_degenerate_map_the_oracle_rejects():
    X = HEX_REF[:8].copy()
    X[:, 2] = 0.0  # flat element: det J = 0 (fe_map.C:467-506)
    pts, w = O.quadrature(O.HEX8, 3)
    with pytest.raises(O.OracleError):
        O.fe_reinit(O.HEX8, X, pts, w)
    with pytest.raises(RuntimeError, match="Jacobian"):
        R.fe_map("HEX8", X, pts, w, False)


def _ex4_element_from_reference_tables(elem, X, pts, w, affine):
    """introduction_ex4.C:464-537 written out in numpy over tables that come from the REFERENCE'S compiled code only
    (FEMap for xyz / JxW / the inverse map, the shape files for phi / dphi, exact_solution.C for the forcing): every
    elementwise numpy operation is one IEEE operation, applied in the example's order —
        Ke(i,j) += JxW[qp] * (dphi[i][qp] * dphi[j][qp])      with TypeVector's a0*b0 + a1*b1 + a2*b2
        Fe(i)   += JxW[qp] * fxy * phi[i][qp]                 fxy = -(uxx + uyy + uzz), the FD Laplacian of exact_solution."""
    sh = R.shapes(elem, pts)
    mp = R.fe_map(elem, X, pts, w, affine)
    dphi = _map_dphi(sh, mp["inv"])  # [i][q][d]
    n, nq = sh[0].shape
    Ke, Fe = np.zeros((n, n)), np.zeros(n)
    for q in range(nq):
        g = dphi[:, q, :]
        dot = (g[:, None, 0] * g[None, :, 0] + g[:, None, 1] * g[None, :, 1]) + g[:, None, 2] * g[None, :, 2]
        Ke += mp["JxW"][q] * dot
    eps = 1.e-3
    ex = R.exact_solution
    for q in range(nq):
        x, y, z = (float(v) for v in mp["xyz"][q])
        uxx = (ex(x - eps, y, z) + ex(x + eps, y, z) + -2. * ex(x, y, z)) / eps / eps
        uyy = (ex(x, y - eps, z) + ex(x, y + eps, z) + -2. * ex(x, y, z)) / eps / eps
        uzz = (ex(x, y, z - eps) + ex(x, y, z + eps) + -2. * ex(x, y, z)) / eps / eps
        fxy = -(uxx + uyy + uzz)
        Fe += mp["JxW"][q] * fxy * sh[0][:, q]
    return Ke, Fe


@needs_ref
@pytest.mark.parametrize("elem,dims,skew", [("HEX27", (2, 2, 2), None), ("HEX27", (2, 2, 1), (1, 0.04, 11)), ("HEX8", (3, 2, 2), (1, 0.05, 5)),
                                            ("HEX20", (2, 2, 1), (1, 0.04, 3)), ("TET10", (1, 1, 1), (1, 0.008, 2)), ("PRISM18", (2, 1, 1), (1, 0.01, 9))])
def test_ex4_element_matrices_from_reference_tables_only(elem, dims, skew):
    """The oracle's unconstrained Ke / Fe of introduction_ex4 against the example's loop evaluated over the reference's own
    tables: bit for bit.  What stays between this and the global system is dof numbering, the element constraint and the
    ordered scatter."""
    orc = ORC[elem]
    m = O.Mesh(orc, *dims, bounds=(-1, 1, -1, 1, -1, 1))
    if skew:
        m.skew(*skew)
    s = O.System(m, 1)
    qorder = 3 if elem == "HEX8" else 5
    Ko, Fo = s.element_matrices(O.POISSON_EX4, constrained=False, qorder=qorder)
    xyz, conn, _ = m.arrays()
    pts, w = O.quadrature(orc, qorder)
    for e in range(0, m.n_elem, max(1, m.n_elem // 6)):
        X = xyz[conn[e]]
        Ke, Fe = _ex4_element_from_reference_tables(elem, X, pts, w, O.has_affine_map(orc, X))
        assert np.array_equal(Ke, Ko[e]), np.abs(Ke - Ko[e]).max()
        assert np.array_equal(Fe, Fo[e]), np.abs(Fe - Fo[e]).max()


@needs_ref
@pytest.mark.parametrize("elem,skew", [("HEX27", (1, 0.004, 7)), ("HEX8", None), ("TET10", None)])
def test_ex6_element_matrices_from_reference_tables_only(elem, skew):
    """systems_of_equations_ex6.C:136-151 (elasticity_tensor) and :220-237 (the six-deep Ke loop, the body force) in numpy
    over the reference's compiled tables, in the example's order ((JxW * C_ijkl) * dphi_j(l)) * dphi_i(j), j then l innermost:
    the oracle's unconstrained 3-variable element matrix and vector, bit for bit."""
    orc = ORC[elem]
    m = O.Mesh(orc, 2, 1, 1, bounds=(0, 1, 0, 0.3, 0, 0.1))
    if skew:
        m.skew(*skew)
    s = O.System(m, 3)
    Ko, Fo = s.element_matrices(O.ELASTICITY_EX6, constrained=False)
    xyz, conn, _ = m.arrays()
    qorder = 3 if elem == "HEX8" else 5
    pts, w = O.quadrature(orc, qorder)
    sh = R.shapes(elem, pts)
    n = sh[0].shape[0]
    nu, E = 0.3, 1.
    lam1 = (E * nu) / ((1. + nu) * (1. - 2. * nu))
    lam2 = E / (2. * (1. + nu))
    kd = lambda a, b: 1. if a == b else 0.
    C = lambda i, j, k, l: lam1 * kd(i, j) * kd(k, l) + lam2 * (kd(i, k) * kd(j, l) + kd(i, l) * kd(j, k))
    f = (0., 0., -1.)
    for e in (0, m.n_elem - 1):
        X = xyz[conn[e]]
        mp = R.fe_map(elem, X, pts, w, O.has_affine_map(orc, X))
        dphi = _map_dphi(sh, mp["inv"])
        Ke, Fe = np.zeros((3 * n, 3 * n)), np.zeros(3 * n)
        for q in range(len(w)):
            g = dphi[:, q, :]
            for i in range(3):
                for j in range(3):
                    for k in range(3):
                        for l in range(3):
                            Ke[i * n:(i + 1) * n, k * n:(k + 1) * n] += ((mp["JxW"][q] * C(i, j, k, l)) * g[None, :, l]) * g[:, None, j]
            for i in range(3):
                Fe[i * n:(i + 1) * n] += mp["JxW"][q] * (f[i] * sh[0][:, q])
        assert np.array_equal(Ke, Ko[e]), np.abs(Ke - Ko[e]).max()
        assert np.array_equal(Fe, Fo[e]), np.abs(Fe - Fo[e]).max()


H27_SIDE = [[0, 3, 2, 1, 11, 10, 9, 8, 20], [0, 1, 5, 4, 8, 13, 16, 12, 21], [1, 2, 6, 5, 9, 14, 17, 13, 22],  # cell_hex27.C:37-45
            [2, 3, 7, 6, 10, 15, 18, 14, 23], [3, 0, 4, 7, 11, 12, 19, 15, 24], [4, 5, 6, 7, 16, 17, 18, 19, 25]]
T10_SIDE = [[0, 2, 1, 6, 5, 4], [0, 1, 3, 4, 8, 7], [1, 2, 3, 5, 9, 8], [2, 0, 3, 6, 7, 9]]                      # cell_tet10.C:37-43


@needs_ref
@pytest.mark.parametrize("elem,side_elem,ns,order,amp", [("HEX27", "QUAD9", 9, 5, 0.03), ("HEX20", "QUAD8", 8, 5, 0.03), ("HEX8", "QUAD4", 4, 3, 0.04),
                                                         ("HEX8", "QUAD4", 4, 5, 0.04), ("TET10", "TRI6", 6, 5, 0.008)])
def test_side_reinit_is_the_references_own_face_machinery_bit_for_bit(elem, side_elem, ns, order, amp):
    """FE::reinit(elem, side).  FEMap::init_face_shape_functions and FEMap::compute_face_map (cut out of src/fe/fe_boundary.C),
    FEAbstract::get_refspace_nodes (fe_abstract.C), the side_nodes_map tables of src/geom/cell_*.C, the reference's 2D rule, its
    QUAD4 / QUAD8 / QUAD9 / TRI6 and 3D shape functions and FEMap::compute_map at the mapped points, called in the order of
    fe_boundary.C:109-238: the oracle's side points in the parent's reference space, physical points, JxW_face, normals and
    phi_face are those bits — on straight and on perturbed (curved-side) elements.  (The surface Jacobian of the
    2D-element-in-space branch of compute_single_point_map, fe_map.C:730-860, is the same number as well.)"""
    if not R.app_available():
        pytest.skip("oracle/_ref/libref_app.so not built")
    orc = ORC[elem]
    pts, w = R.gauss_rule(side_elem, order)
    for amplitude in (0.0, amp):
        m = O.Mesh(orc, 2, 2, 1, bounds=(-1, 1, -1, 1, 0, 1))
        if amplitude:
            m.skew(1, amplitude, 5)
        xyz, conn, _ = m.arrays()
        for e in range(0, m.n_elem, max(1, m.n_elem // 5)):
            X = xyz[conn[e]]
            for side in range(4 if elem == "TET10" else 6):
                d = O.side_reinit(orc, X, side, order)
                r = R.side_reinit(elem, X, side, order, O.has_affine_map(orc, X))
                for k in ("ref", "xyz", "JxW", "normals", "phi"):
                    assert np.array_equal(d[k], r[k]), (k, e, side, np.abs(d[k] - r[k]).max())
                nodes = (T10_SIDE if elem == "TET10" else H27_SIDE)[side][:ns]
                assert np.array_equal(d["JxW"], R.fe_map(side_elem, X[nodes], pts, w, False)["JxW"])


@needs_ref
def test_hanging_node_coefficients_of_the_test_meshes_are_the_reference_side_shapes():
    """lagrange_compute_constraints (src/fe/fe_lagrange.C) sets the coefficient of a hanging node on a parent's side to the
    parent side's shape function at the node's position there.  The refined test meshes of tests/hanging.py use 1/2 (HEX8
    edges), 1/4 (HEX8 faces) and the products of 3/8, 3/4, -1/8 (HEX27): exactly what the reference's compiled QUAD4 / QUAD9
    functions give at the half and quarter points."""
    from hanging import _quad_weights

    # QUAD9 nodes on the bottom edge eta = -1: corner 0 (xi = -1), corner 1 (xi = +1), mid-edge 4 (xi = 0)
    for rel in range(5):
        p = (rel - 2) / 2.0
        phi = R.shapes("QUAD9", [[p, -1.0, 0.0]])[0][:, 0]
        want = _quad_weights(rel)
        assert phi[0] == want.get(0, 0.0) and phi[4] == want.get(2, 0.0) and phi[1] == want.get(4, 0.0)
    # a face-hanging HEX27 node at the quarter point (-1/2, +1/2) of a coarse face: tensor products of the 1D weights
    phi = R.shapes("QUAD9", [[-0.5, 0.5, 0.0]])[0][:, 0]
    wx, wy = _quad_weights(1), _quad_weights(3)
    xi_of = {0: 0, 1: 4, 2: 4, 3: 0, 4: 2, 5: 4, 6: 2, 7: 0, 8: 2}   # QUAD9 node -> coarse lattice coordinate along xi (0, 2, 4)
    eta_of = {0: 0, 1: 0, 2: 4, 3: 4, 4: 0, 5: 2, 6: 4, 7: 2, 8: 2}
    for a in range(9):
        assert phi[a] == wx[xi_of[a]] * wy[eta_of[a]]
    # HEX8: the middle of an edge / of a face of the bilinear QUAD4
    assert np.array_equal(R.shapes("QUAD4", [[0.0, -1.0, 0.0]])[0][:, 0], [0.5, 0.5, 0.0, 0.0])
    assert np.array_equal(R.shapes("QUAD4", [[0.0, 0.0, 0.0]])[0][:, 0], [0.25, 0.25, 0.25, 0.25])


@needs_ref
@pytest.mark.parametrize("elem", sorted(ORC))
def test_has_affine_map_is_the_element_classes_own_test(elem):
    """Hex8 / Hex20 / Hex27 / Tet10 / Prism18::has_affine_map (src/geom/cell_*.C; the reference's own definitions, cut out of the
    element classes and compiled over relative_fuzzy_equals as include/utils/fuzzy_equals.h states it) decide which branch of
    FEMap::compute_map an element takes.  Plain, sheared (SkewFunc: still affine) and perturbed elements, with perturbations on
    either side of the tolerance affine_tol = TOLERANCE^2 = 1e-12."""
    n_affine = n_general = 0
    for skew in (None, (0, 0.0, 0), (1, 0.03, 3), (1, 1e-9, 3), (1, 1e-11, 4), (1, 1e-12, 5), (1, 1e-13, 3), (1, 1e-15, 6)):
        m = O.Mesh(ORC[elem], 2, 2, 2, (-1, 1, -1, 1, -1, 1))
        if skew:
            m.skew(*skew)
        xyz, conn, _ = m.arrays()
        for e in range(m.n_elem):
            a, b = O.has_affine_map(ORC[elem], xyz[conn[e]]), R.has_affine_map(elem, xyz[conn[e]])
            assert a == b, (skew, e)
            n_affine += b
            n_general += not b
    assert n_affine > 0 and n_general > 0
