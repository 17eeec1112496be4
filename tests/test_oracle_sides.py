"""Pins the oracle's side (face) code — FE::reinit(elem, side), src/fe/fe_boundary.C:109-240 — to known answers.

Restated from the reference's own side test (tests/fe/fe_side_test.C:100-190: on every side of every element the
shape functions at the side quadrature points reproduce a linear / quadratic function of the physical point) plus
closed-surface identities that hold for any element, affine or not: sum over the six sides of int n dS = 0 and
int x.n dS = 3 |element| (divergence theorem), which tie JxW_face and the normals to the volume map already pinned in
test_oracle_fe.py; and the published geometry of systems_of_equations_ex6 (the loaded face x = 1 of the box
[0,1] x [0,0.3] x [0,0.1] has area 0.03, systems_of_equations_ex6.C:437-444, :240-257).
"""
import numpy as np
import pytest

from oracle import oracle as O


def _mesh(etype, dims, skew):
    m = O.Mesh(etype, *dims, bounds=(-1, 1, -1, 1, -1, 1))
    if skew == "skewfunc":
        m.skew(0)
    elif skew == "perturbed":
        m.skew(1, 0.07, 12345)
    return m


@pytest.mark.parametrize("skew", [None, "skewfunc", "perturbed"])
@pytest.mark.parametrize("etype,qorder", [(O.HEX8, 3), (O.HEX8, 5), (O.HEX27, 5), (O.HEX20, 5), (O.HEX20, 3)])
def test_side_shapes_reproduce_functions_of_the_side_points(etype, qorder, skew):
    m = _mesh(etype, (2, 2, 2), skew)
    xyz, conn, _ = m.arrays()
    lin = lambda p: p[..., 0] + 0.25 * p[..., 1] + 0.0625 * p[..., 2]  # fe_test.h linear_test
    quad = lambda p: p[..., 0] ** 2 + 0.5 * p[..., 1] ** 2 + 0.25 * p[..., 2] ** 2 + 0.125 * p[..., 0] * p[..., 1]
    for e in range(m.n_elem):
        X = xyz[conn[e]]
        for s in range(6):
            d = O.side_reinit(etype, X, s, qorder)
            nqf = {3: 4, 5: 9}[qorder]
            assert d["phi"].shape == (etype, nqf)
            assert np.abs(d["phi"].sum(0) - 1).max() < 1e-14                 # partition of unity on the side
            assert np.abs(d["phi"].T @ X - d["xyz"]).max() < 1e-14           # the parent map hits the side points
            assert np.abs(d["phi"].T @ lin(X) - lin(d["xyz"])).max() < 1e-13
            if etype in (O.HEX27, O.HEX20) and skew != "perturbed":         # quadratic reproduction needs an affine map
                assert np.abs(d["phi"].T @ quad(X) - quad(d["xyz"])).max() < 1e-13
            # the mapped points lie on the reference face of that side
            axis, val = [(2, -1), (1, -1), (0, 1), (1, 1), (0, -1), (2, 1)][s]
            assert np.abs(d["ref"][:, axis] - val).max() < 1e-15


@pytest.mark.parametrize("skew", [None, "skewfunc", "perturbed"])
@pytest.mark.parametrize("etype,qorder", [(O.HEX8, 3), (O.HEX27, 5), (O.HEX20, 5)])
def test_closed_surface_identities(etype, qorder, skew):
    m = _mesh(etype, (2, 2, 2), skew)
    xyz, conn, _ = m.arrays()
    pts, w = O.quadrature(etype, 5)
    for e in range(m.n_elem):
        X = xyz[conn[e]]
        vol = O.fe_reinit(etype, X, pts, w)["JxW"].sum()
        n_int, flux = np.zeros(3), 0.0
        for s in range(6):
            d = O.side_reinit(etype, X, s, 5)
            n_int += (d["JxW"][:, None] * d["normals"]).sum(0)
            flux += (d["JxW"] * (d["xyz"] * d["normals"]).sum(1)).sum()
        assert np.abs(n_int).max() < 1e-13
        assert abs(flux - 3 * vol) < 1e-12 * max(1.0, abs(vol))


def test_side_rule_is_the_tensor_rule_i_fastest():
    # QBase::tensor_product_quad (quadrature.C:256-277): q = i + np*j, point (x_i, x_j); side 0 of the reference hex is
    # zeta = -1 with side nodes (0, 3, 2, 1): side xi runs along parent eta, side eta along parent xi
    X = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], dtype=float)
    d = O.side_reinit(O.HEX8, X, 0, 3)
    g = np.sqrt(3.0) / 3.0
    want = np.array([[-g, -g, -1], [-g, g, -1], [g, -g, -1], [g, g, -1]])  # q = 1 is side (+g, -g) = parent (xi -g, eta +g)
    assert np.abs(d["ref"] - want).max() < 1e-15
    assert np.abs(d["JxW"] - 1.0).max() < 1e-15 and np.abs(d["normals"] - [0, 0, -1]).max() < 1e-15


@pytest.mark.parametrize("etype,dims", [(O.HEX27, (4, 3, 2)), (O.HEX8, (5, 3, 2)), (O.HEX20, (3, 3, 2))])
def test_ex6_traction_face_and_total_load(etype, dims):
    m = O.Mesh(etype, *dims, bounds=(0, 1, 0, 0.3, 0, 0.1))
    e, sd = m.boundary_sides(2)                       # BOUNDARY_ID_MAX_X, systems_of_equations_ex6.C:79
    assert len(e) == dims[1] * dims[2] and (sd == 2).all()
    xyz, conn, _ = m.arrays()
    area = sum(O.side_reinit(etype, xyz[conn[ee]], 2, 5)["JxW"].sum() for ee in e)
    assert abs(area - 0.03) < 1e-15
    for b, cnt in [(0, dims[0] * dims[1]), (1, dims[0] * dims[2]), (4, dims[1] * dims[2]), (5, dims[0] * dims[1])]:
        assert len(m.boundary_sides(b)[0]) == cnt     # mesh_generation.C:1416-1432: one id per box face
    s0, s1 = O.System(m, 3), O.System(m, 3)
    s1.set_side_loads(e, sd, [0.0, 0.0, -1.0])
    (_, F0), (_, F1) = s0.element_matrices(O.ELASTICITY_EX6, constrained=False), s1.element_matrices(O.ELASTICITY_EX6, constrained=False)
    dF = (F1 - F0).reshape(m.n_elem, 3, etype)
    assert np.abs(dF[:, :2]).max() == 0.0                                 # g = (0, 0, -1)
    assert abs(dF[:, 2].sum() + 0.03) < 1e-15                             # int_Gamma g_z ds
    untouched = np.setdiff1d(np.arange(m.n_elem), e)
    assert np.abs(dF[untouched]).max() == 0.0


def _exact(x, y, z):  # introduction_ex3/exact_solution.C:40-47, independent of the oracle
    return np.cos(.5 * np.pi * x) * np.sin(.5 * np.pi * y) * np.cos(.5 * np.pi * z)


def _all_boundary_sides(m):
    es, ss = [], []
    for b in range(6):
        e, s = m.boundary_sides(b)
        es.append(e)
        ss.append(s)
    return np.concatenate(es), np.concatenate(ss)


@pytest.mark.parametrize("etype,qorder,nv", [(O.HEX27, 5, 1), (O.HEX8, 3, 1), (O.HEX8, 3, 3), (O.HEX20, 5, 1)])
def test_penalty_term_is_the_ex3_loop(etype, qorder, nv):
    """introduction_ex3.C:437-453 on the sides of a 3D element, against an independent numpy evaluation from side_reinit."""
    m = _mesh(etype, (3, 2, 2), "perturbed")
    e, sd = _all_boundary_sides(m)
    xyz, conn, _ = m.arrays()
    prob = O.POISSON_EX4 if nv == 1 else O.ELASTICITY_EX6
    s0, s1 = O.System(m, nv), O.System(m, nv)
    s1.set_side_penalty(e, sd, penalty=1e10, kind=1)
    (K0, F0), (K1, F1) = (s.element_matrices(prob, constrained=False, qorder=qorder) for s in (s0, s1))
    n = etype
    for el in range(m.n_elem):
        dK, dF = np.zeros((n, n)), np.zeros(n)
        for k in np.nonzero(e == el)[0]:
            d = O.side_reinit(etype, xyz[conn[el]], int(sd[k]), qorder)
            val = _exact(d["xyz"][:, 0], d["xyz"][:, 1], d["xyz"][:, 2])
            dK += 1e10 * np.einsum("q,iq,jq->ij", d["JxW"], d["phi"], d["phi"])
            dF += 1e10 * (d["phi"] * (d["JxW"] * val)).sum(1)
        for v in range(nv):
            blk = (K1[el] - K0[el]).reshape(nv * n, nv * n)[v * n:(v + 1) * n, v * n:(v + 1) * n]
            assert np.abs(blk - dK).max() <= 1e-12 * max(1.0, np.abs(dK).max())
            assert np.abs((F1[el] - F0[el]).reshape(nv, n)[v] - dF).max() <= 1e-12 * max(1.0, np.abs(dF).max())
        if nv > 1:  # off-diagonal variable blocks are untouched
            full = (K1[el] - K0[el]).reshape(nv * n, nv * n)
            assert np.abs(full[:n, n:2 * n]).max() <= 1e-6 * max(1.0, np.abs(dK).max()) * 1e-9


def test_penalty_enforces_the_boundary_values_like_ex3():
    """The point of the method (introduction_ex3.C:386-400): with penalty 1e10 the solution takes the exact solution's
    values on the boundary; on the ex4 problem the whole nodal error is then the discretisation error."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla

    errs = []
    for nn in (2, 4):
        m = O.Mesh(O.HEX27, nn, nn, nn, bounds=(-1, 1, -1, 1, -1, 1), renumber_nodes=False)
        s = O.System(m, 1)
        s.set_side_penalty(*_all_boundary_sides(m), penalty=1e10, kind=1)
        s.build_sparsity()
        v, r, _ = s.assemble(O.POISSON_EX4)
        rp, ci = s.sparsity()
        A = sp.csr_matrix((v, ci.astype(np.int64), rp.astype(np.int64)), shape=(s.n_dofs, s.n_dofs))
        u = spla.spsolve(A.tocsc(), r)
        xyz, _, mask = m.arrays()
        _, nb, _ = s.dofs()
        exact = _exact(xyz[:, 0], xyz[:, 1], xyz[:, 2])
        err = np.abs(u[nb] - exact)
        errs.append((err[mask != 0].max(), err.max()))
    # boundary nodes carry the L2 projection of the exact solution onto the Q2 faces: O(h^3); 0.044 at h = 1
    assert errs[0][0] < 0.06 and errs[1][0] < errs[0][0] / 5
    assert errs[1][1] < errs[0][1] / 4


# ---- TET10 sides (TRI6, the 7-point degree-5 rule) --------------------------------------------------------------------
def test_tri_side_rule_is_the_degree_5_dunavant_rule():
    """quadrature_gauss_2D.C:234-257: 7 points, weights sum to the reference triangle's area, exact for x^a y^b, a+b <= 5
    (tests/quadrature/quadrature_test.C checks monomials the same way).  Seen through side 0 of the reference tet, whose
    side map is affine: side (xi, eta) -> parent (eta, xi, 0)."""
    from math import factorial
    X = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [.5, 0, 0], [.5, .5, 0], [0, .5, 0], [0, 0, .5], [.5, 0, .5], [0, .5, .5]], dtype=float)
    d = O.side_reinit(O.TET10, X, 0, 5)
    assert d["JxW"].shape == (7,) and abs(d["JxW"].sum() - 0.5) < 1e-15
    assert np.abs(d["xyz"][:, 2]).max() == 0.0 and np.abs(d["normals"] - [0, 0, -1]).max() < 1e-15   # outward: -z
    assert np.abs(d["xyz"][0] - [1 / 3, 1 / 3, 0]).max() < 1e-15                                        # centroid first
    for a in range(6):
        for b in range(6 - a):
            exact = factorial(a) * factorial(b) / factorial(a + b + 2)
            assert abs((d["JxW"] * d["xyz"][:, 0] ** a * d["xyz"][:, 1] ** b).sum() - exact) < 1e-15


@pytest.mark.parametrize("skew", [None, "perturbed"])
def test_tet_side_shapes_and_closed_surface(skew):
    """fe_side_test.C on TET10: side shapes reproduce linear and quadratic functions of the side points; and the
    divergence theorem over the four sides ties JxW_face and the normals to the volume map."""
    m = _mesh(O.TET10, (2, 1, 1), skew)
    xyz, conn, _ = m.arrays()
    lin = lambda p: p[..., 0] + 0.25 * p[..., 1] + 0.0625 * p[..., 2]
    quad = lambda p: p[..., 0] ** 2 + 0.5 * p[..., 1] ** 2 + 0.25 * p[..., 2] ** 2 + 0.125 * p[..., 0] * p[..., 1]
    pts, w = O.quadrature(O.TET10, 5)
    for e in range(0, m.n_elem, 3):
        X = xyz[conn[e]]
        vol = O.fe_reinit(O.TET10, X, pts, w)["JxW"].sum()
        n_int, flux = np.zeros(3), 0.0
        for s in range(4):
            d = O.side_reinit(O.TET10, X, s, 5)
            assert d["phi"].shape == (10, 7)
            assert np.abs(d["phi"].sum(0) - 1).max() < 1e-14
            assert np.abs(d["phi"].T @ X - d["xyz"]).max() < 1e-14
            assert np.abs(d["phi"].T @ lin(X) - lin(d["xyz"])).max() < 1e-13
            if skew is None:
                assert np.abs(d["phi"].T @ quad(X) - quad(d["xyz"])).max() < 1e-13
            n_int += (d["JxW"][:, None] * d["normals"]).sum(0)
            flux += (d["JxW"] * (d["xyz"] * d["normals"]).sum(1)).sum()
        assert np.abs(n_int).max() < 1e-13
        assert abs(flux - 3 * vol) < 1e-12 * max(1.0, abs(vol))


def tet_boundary_sides(m, bid):
    """(element, side) pairs of a TET10 cube whose six side nodes all lie on box face `bid` (BoundaryInfo of build_cube)."""
    sides = [[0, 2, 1, 6, 5, 4], [0, 1, 3, 4, 8, 7], [1, 2, 3, 5, 9, 8], [2, 0, 3, 6, 7, 9]]  # cell_tet10.C:37-43
    _, conn, mask = m.arrays()
    on = (mask >> bid) & 1
    es, ss = [], []
    for s, nodes in enumerate(sides):
        hit = on[conn[:, nodes]].all(axis=1)
        es.append(np.nonzero(hit)[0])
        ss.append(np.full(hit.sum(), s))
    e, sd = np.concatenate(es), np.concatenate(ss)
    o = np.argsort(e, kind="stable")
    return e[o].astype(np.uint32), sd[o].astype(np.uint8)


def test_tet_traction_face_area():
    m = O.Mesh(O.TET10, 3, 2, 2, bounds=(0, 1, 0, 0.3, 0, 0.1))
    e, sd = tet_boundary_sides(m, 2)
    assert len(e) == 4 * 2 * 2                       # four tets of each boundary cell face
    xyz, conn, _ = m.arrays()
    area = sum(O.side_reinit(O.TET10, xyz[conn[ee]], int(s), 5)["JxW"].sum() for ee, s in zip(e, sd))
    assert abs(area - 0.03) < 1e-15
    s0, s1 = O.System(m, 1), O.System(m, 1)
    s1.set_side_loads(e, sd, [1.7])
    (_, F0), (_, F1) = (s.element_matrices(O.POISSON_EX4, constrained=False) for s in (s0, s1))
    assert abs((F1 - F0).sum() - 1.7 * 0.03) < 1e-15
