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
@pytest.mark.parametrize("etype,qorder", [(O.HEX8, 3), (O.HEX8, 5), (O.HEX27, 5)])
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
            if etype == O.HEX27 and skew != "perturbed":                    # quadratic reproduction needs an affine map
                assert np.abs(d["phi"].T @ quad(X) - quad(d["xyz"])).max() < 1e-13
            # the mapped points lie on the reference face of that side
            axis, val = [(2, -1), (1, -1), (0, 1), (1, 1), (0, -1), (2, 1)][s]
            assert np.abs(d["ref"][:, axis] - val).max() < 1e-15


@pytest.mark.parametrize("skew", [None, "skewfunc", "perturbed"])
@pytest.mark.parametrize("etype,qorder", [(O.HEX8, 3), (O.HEX27, 5)])
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


@pytest.mark.parametrize("etype,dims", [(O.HEX27, (4, 3, 2)), (O.HEX8, (5, 3, 2))])
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
