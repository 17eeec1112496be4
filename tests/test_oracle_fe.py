"""Pins the oracle's element-level arithmetic to the known answers of the reference's own unit tests.

Restated from (paths relative to /root/reference):
  tests/quadrature/quadrature_test.C:591-630 (test3DWeights: exact monomial integrals on hex and tet),
  tests/fe/fe_test.h:67-110, :671-690, :812-813 (partition of unity, value / gradient reproduction of the linear
  and quadratic test functions on build_cube meshes, incl. the SkewFunc-skewed mesh :40-64),
  src/quadrature/quadrature_gauss_1D.C:28-70 and quadrature_gauss_3D.C:257-283 (the literal rule constants).
"""
import itertools
import math

import numpy as np
import pytest

from oracle import oracle as O

TOL = 1e-12  # tests/quadrature/quadrature_test.C quadrature_tolerance = TOLERANCE*sqrt(TOLERANCE) ~ 1e-9; we are stricter


def hex_integral(a, b, c):  # int_{[-1,1]^3} x^a y^b z^c
    f = lambda p: 0.0 if p % 2 else 2.0 / (p + 1)
    return f(a) * f(b) * f(c)


def tet_integral(a, b, c):  # int over the reference tet = a! b! c! / (a+b+c+3)!
    return math.factorial(a) * math.factorial(b) * math.factorial(c) / math.factorial(a + b + c + 3)


@pytest.mark.parametrize("etype,qorder,nq,exact", [(O.HEX8, 3, 8, 3), (O.HEX8, 5, 27, 5), (O.HEX27, 5, 27, 5), (O.TET10, 5, 14, 5)])
def test_quadrature_integrates_monomials_exactly(etype, qorder, nq, exact):
    pts, w = O.quadrature(etype, qorder)
    assert len(w) == nq
    for a, b, c in itertools.product(range(exact + 1), repeat=3):
        if etype == O.TET10 and a + b + c > exact:
            continue
        if etype != O.TET10 and max(a, b, c) > exact:
            continue
        got = float((w * pts[:, 0] ** a * pts[:, 1] ** b * pts[:, 2] ** c).sum())
        want = tet_integral(a, b, c) if etype == O.TET10 else hex_integral(a, b, c)
        assert abs(got - want) < TOL, (a, b, c, got, want)


def test_gauss_constants_and_tensor_ordering():
    # quadrature_gauss_1D.C: 2-pt -+sqrt(3)/3 w=1; 3-pt -+0.7745966692414834, 0, w=5/9, 8/9, 5/9
    pts, w = O.quadrature(O.HEX8, 3)
    g = math.sqrt(3.0) / 3.0
    assert np.allclose(sorted(set(np.round(pts[:, 0], 15))), [-g, g], atol=1e-15)
    assert np.allclose(w, 1.0)
    pts, w = O.quadrature(O.HEX27, 5)
    x1 = np.array([-7.7459666924148337703585307995648e-01, 0.0, 7.7459666924148337703585307995648e-01])
    w1 = np.array([5 / 9, 8 / 9, 5 / 9])
    # QBase::tensor_product_hex (quadrature.C:283-305): q = i + np*(j + np*k), x from i, y from j, z from k
    for k in range(3):
        for j in range(3):
            for i in range(3):
                q = i + 3 * (j + 3 * k)
                assert np.allclose(pts[q], [x1[i], x1[j], x1[k]], atol=1e-16)
                assert abs(w[q] - w1[i] * w1[j] * w1[k]) < 1e-16
    # Walkington 14-point tet rule (quadrature_gauss_3D.C:257-283)
    pts, w = O.quadrature(O.TET10, 5)
    a = [0.31088591926330061, 0.092735250310891226, 0.045503704125649649]
    wt = [0.018781320953002642, 0.012248840519393658, 0.0070910034628469111]
    assert abs(w[:4] - wt[0]).max() < 1e-18 and abs(w[4:8] - wt[1]).max() < 1e-18 and abs(w[8:] - wt[2]).max() < 1e-18
    assert abs(pts[0] - a[0]).max() < 1e-18 and abs(pts[4] - a[1]).max() < 1e-18
    assert abs(w.sum() - 1.0 / 6.0) < 1e-15


REF_NODES = {
    O.HEX8: None,
    O.HEX27: None,
    O.TET10: np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [.5, 0, 0], [.5, .5, 0], [0, .5, 0], [0, 0, .5], [.5, 0, .5], [0, .5, .5]], dtype=float),
    # Prism18: triangle (0,0) (1,0) (0,1) x [-1, 1] (cell_prism18.h node order: bottom, top, bottom edges, verticals, top edges, mid layer)
    O.PRISM18: np.array([[0, 0, -1], [1, 0, -1], [0, 1, -1], [0, 0, 1], [1, 0, 1], [0, 1, 1], [.5, 0, -1], [.5, .5, -1], [0, .5, -1],
                         [0, 0, 0], [1, 0, 0], [0, 1, 0], [.5, 0, 1], [.5, .5, 1], [0, .5, 1], [.5, 0, 0], [.5, .5, 0], [0, .5, 0]], dtype=float),
}


def _hex_ref_nodes(n):
    ref = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1], [0, -1, -1],
                    [1, 0, -1], [0, 1, -1], [-1, 0, -1], [-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0], [0, -1, 1], [1, 0, 1],
                    [0, 1, 1], [-1, 0, 1], [0, 0, -1], [0, -1, 0], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0, 0, 1], [0, 0, 0]], dtype=float)
    return ref[:n]


@pytest.mark.parametrize("etype", [O.HEX8, O.HEX27, O.TET10, O.HEX20, O.PRISM18])
def test_shapes_are_nodal_and_partition_unity(etype):
    nodes = REF_NODES[etype] if etype in (O.TET10, O.PRISM18) else _hex_ref_nodes(etype)
    phi, dxi, deta, dzeta = O.shapes(etype, nodes)
    assert np.abs(phi - np.eye(etype)).max() < 1e-14  # Lagrange: phi_i(node_j) = delta_ij
    rng = np.random.default_rng(7)
    pts = rng.uniform(0.05, 0.3, size=(20, 3)) if etype in (O.TET10, O.PRISM18) else rng.uniform(-1, 1, size=(20, 3))
    phi, dxi, deta, dzeta = O.shapes(etype, pts)
    assert np.abs(phi.sum(axis=0) - 1).max() < 1e-14  # fe_test.h:671-690
    for d in (dxi, deta, dzeta):
        assert np.abs(d.sum(axis=0)).max() < 1e-13
    # derivatives agree with central differences of phi
    h = 1e-6
    for d, tab in enumerate((dxi, deta, dzeta)):
        pp, pm = pts.copy(), pts.copy()
        pp[:, d] += h
        pm[:, d] -= h
        fd = (O.shapes(etype, pp)[0] - O.shapes(etype, pm)[0]) / (2 * h)
        assert np.abs(fd - tab).max() < 1e-8


def linear_test(p):
    return p[..., 0] + 0.25 * p[..., 1] + 0.0625 * p[..., 2]


def quadratic_test(p):
    x, y, z = p[..., 0], p[..., 1], p[..., 2]
    return x * x + 0.5 * y * y + 0.25 * z * z + 0.125 * x * y + 0.0625 * x * z + 0.03125 * y * z


def quadratic_grad(p):
    x, y, z = p[..., 0], p[..., 1], p[..., 2]
    return np.stack([2 * x + 0.125 * y + 0.0625 * z, y + 0.125 * x + 0.03125 * z, 0.5 * z + 0.0625 * x + 0.03125 * y], axis=-1)


def test_hex20_shapes_equal_the_textbook_serendipity_form():
    """The oracle follows the 20-case form of fe_lagrange_shape_3D.C:848-925 in x = (xi + 1) / 2; the closed form over the
    node's reference coordinates is the independent statement of the same basis (Zienkiewicz: corner nodes
    (1 + xi xi_i)(1 + eta eta_i)(1 + zeta zeta_i)(xi xi_i + eta eta_i + zeta zeta_i - 2) / 8, mid-edge nodes
    (1 - xi^2)(1 + eta eta_i)(1 + zeta zeta_i) / 4 and permutations)."""
    nodes = _hex_ref_nodes(20)
    pts = np.random.default_rng(3).uniform(-1, 1, size=(64, 3))
    phi = O.shapes(O.HEX20, pts)[0]
    xi, eta, zeta = pts.T
    for i, (x0, y0, z0) in enumerate(nodes):
        if i < 8:
            want = 0.125 * (1 + xi * x0) * (1 + eta * y0) * (1 + zeta * z0) * (xi * x0 + eta * y0 + zeta * z0 - 2)
        elif x0 == 0:
            want = 0.25 * (1 - xi * xi) * (1 + eta * y0) * (1 + zeta * z0)
        elif y0 == 0:
            want = 0.25 * (1 - eta * eta) * (1 + xi * x0) * (1 + zeta * z0)
        else:
            want = 0.25 * (1 - zeta * zeta) * (1 + xi * x0) * (1 + eta * y0)
        assert np.abs(phi[i] - want).max() < 5e-16


def test_prism18_shapes_equal_the_tensor_form_and_the_rule_is_exact():
    """PRISM18 = TRI6 x EDGE3 (fe_lagrange_shape_3D.C:1077-1094): against the barycentric closed form
    L_a (2 L_a - 1) / 4 L_a L_b in the triangle times xi (xi -+ 1) / 2, 1 - xi^2 along the axis, written from the node's
    position; and the 21-point rule (7-point degree-5 triangle rule x 3-point Gauss, quadrature_gauss_3D.C:507-527,
    quadrature.C:310-332) integrates the monomials x^a y^b z^c, a + b <= 5, c <= 5, exactly over the reference prism."""
    from math import factorial as f

    nodes = REF_NODES[O.PRISM18]
    pts = np.random.default_rng(5).uniform(0.05, 0.45, size=(50, 3)) * [1, 1, 4] - [0, 0, 1]
    phi = O.shapes(O.PRISM18, pts)[0]
    L = np.stack([1 - pts[:, 0] - pts[:, 1], pts[:, 0], pts[:, 1]])
    z = pts[:, 2]
    for i, (x0, y0, z0) in enumerate(nodes):
        lz = {-1.0: 0.5 * z * (z - 1), 1.0: 0.5 * z * (z + 1), 0.0: 1 - z * z}[z0]
        bary = np.array([1 - x0 - y0, x0, y0])
        if np.isclose(bary.max(), 1.0):
            a = int(bary.argmax())
            tri = L[a] * (2 * L[a] - 1)
        else:
            a, b = np.nonzero(bary > 0.25)[0]
            tri = 4 * L[a] * L[b]
        assert np.abs(phi[i] - tri * lz).max() < 5e-16
    q, w = O.quadrature(O.PRISM18, 5)
    assert q.shape == (21, 3) and abs(w.sum() - 1.0) < 1e-15
    assert np.array_equal(q[:7, :2], q[7:14, :2]) and np.all(q[:7, 2] == q[0, 2]) and q[0, 2] < q[7, 2] < q[14, 2]  # 1D rule outermost
    for a in range(6):
        for b in range(6 - a):
            for c in range(6):
                exact = f(a) * f(b) / f(a + b + 2) * (2.0 / (c + 1) if c % 2 == 0 else 0.0)
                assert abs((w * q[:, 0] ** a * q[:, 1] ** b * q[:, 2] ** c).sum() - exact) < 1e-15


def test_prism18_build_cube():
    """mesh_generation.C:1634-1756: two prisms per lattice cell on the HEX27 node lattice, all nodes in use; counts as
    tests/mesh/mesh_generation_test.C checks them (n_elem = 2 nx ny nz, n_nodes = (2nx+1)(2ny+1)(2nz+1))."""
    m = O.Mesh(O.PRISM18, 3, 2, 2, bounds=(-1, 1, 0, 2, 0, 3))
    xyz, conn, _ = m.arrays()
    assert m.n_elem == 2 * 3 * 2 * 2 and xyz.shape[0] == 7 * 5 * 5 and len(np.unique(conn)) == xyz.shape[0]
    # the two prisms of cell 0 share the diagonal face: nodes (i+2,j) (i,j+2) and the mid-diagonal column
    assert set(conn[0]) & set(conn[1]) == {conn[0][1], conn[0][2], conn[0][4], conn[0][5], conn[0][7], conn[0][10], conn[0][11], conn[0][13], conn[0][16]}
    # mid-side nodes are the averages of their vertices (element-local edge table of cell_prism18.C)
    edges = {6: (0, 1), 7: (1, 2), 8: (0, 2), 9: (0, 3), 10: (1, 4), 11: (2, 5), 12: (3, 4), 13: (4, 5), 14: (3, 5)}
    for e in range(m.n_elem):
        X = xyz[conn[e]]
        for mid, (a, b) in edges.items():
            assert np.abs(X[mid] - 0.5 * (X[a] + X[b])).max() < 1e-15
        assert np.abs(X[15] - 0.25 * (X[0] + X[1] + X[3] + X[4])).max() < 1e-15


@pytest.mark.parametrize("etype,order", [(O.HEX8, 1), (O.HEX27, 2), (O.TET10, 2), (O.HEX20, 2), (O.PRISM18, 2)])
@pytest.mark.parametrize("skew", [False, True])
def test_fe_reinit_reproduces_polynomials(etype, order, skew):
    """fe_test.h testU/testGradU on the build_cube unit-cube mesh (fe_lagrange_test.C:23-30), plain and SkewFunc-skewed."""
    m = O.Mesh(etype, 2, 2, 2)
    if skew:
        m.skew(0, 0.0)  # kind 0 = SkewFunc: x += 0.1 y, y += 0.2 z
    xyz, conn, _ = m.arrays()
    pts, w = O.quadrature(etype, 5)
    vol = 0.0
    for e in range(m.n_elem):
        X = xyz[conn[e]]
        r = O.fe_reinit(etype, X, pts, w)
        vol += r["JxW"].sum()
        f = quadratic_test if order == 2 else linear_test
        u_nodes = f(X)
        u_q = r["phi"].T @ u_nodes
        assert np.abs(u_q - f(r["xyz"])).max() < 1e-13
        g_q = np.einsum("iqd,i->qd", r["dphi"], u_nodes)
        want = quadratic_grad(r["xyz"]) if order == 2 else np.broadcast_to([1, 0.25, 0.0625], g_q.shape)
        assert np.abs(g_q - want).max() < 1e-12
        # affine test: every build_cube hex/tet (also after the linear SkewFunc) has an affine map
        assert r["affine"] == O.has_affine_map(etype, X) == True  # noqa: E712
    assert abs(vol - 1.0) < 1e-13  # sum_q JxW = volume; SkewFunc is volume preserving


def test_general_branch_matches_affine_branch_and_detects_nonaffine():
    """compute_affine_map (fe_map.C:1167-1244) vs per-qp compute_single_point_map (:958-1077)."""
    m = O.Mesh(O.HEX27, 2, 2, 2)
    m.skew(1, 0.05, 12345)  # seeded interior perturbation: non-affine elements
    xyz, conn, _ = m.arrays()
    pts, w = O.quadrature(O.HEX27, 5)
    n_general = 0
    vol = 0.0
    for e in range(m.n_elem):
        X = xyz[conn[e]]
        r = O.fe_reinit(O.HEX27, X, pts, w)
        n_general += not r["affine"]
        vol += r["JxW"].sum()
        # an isoparametric map reproduces (only) linears exactly in physical space
        u = linear_test(X)
        assert np.abs(r["phi"].T @ u - linear_test(r["xyz"])).max() < 1e-13
        g = np.einsum("iqd,i->qd", r["dphi"], u)
        assert np.abs(g - [1, 0.25, 0.0625]).max() < 1e-12
    assert n_general == m.n_elem
    assert abs(vol - 1.0) < 1e-12  # interior perturbation keeps the cube's volume


def test_degenerate_map_is_an_error():
    X = _hex_ref_nodes(8).copy()
    X[:, 2] = 0.0  # flat element: det J = 0 -> libMesh throws DegenerateMap (fe_map.C:467-506)
    pts, w = O.quadrature(O.HEX8, 3)
    with pytest.raises(O.OracleError):
        O.fe_reinit(O.HEX8, X, pts, w)
