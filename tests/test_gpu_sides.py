"""GPU parity of the side (face) terms — feasm_set_side_loads — against the CPU oracle: the traction of
systems_of_equations_ex6.C:240-257 on the max-x face, on affine and perturbed meshes, and a Neumann flux on a Poisson
system where corner elements carry two loaded sides."""
import numpy as np
import pytest

from conftest import PARITY_TOL, csr_parity_error, dense_parity_error, vec_parity_error
from test_gpu_parity import _mods, make_case, make_ctx

pytestmark = pytest.mark.gpu

EX6 = [
    ("hex27_ex6", 27, (4, 3, 2), None, (0.0, 0.0, -1.0)),
    ("hex27_ex6_perturbed", 27, (3, 2, 2), (1, 0.004, 7), (0.3, -0.2, -1.0)),
    ("hex8_ex6", 8, (5, 3, 2), None, (0.0, 0.0, -1.0)),
    ("hex8_ex6_perturbed", 8, (4, 3, 3), (1, 0.004, 11), (0.3, -0.2, -1.0)),
    ("hex20_ex6", 20, (4, 3, 2), None, (0.0, 0.0, -1.0)),  # QUAD8 sides of the serendipity hex
    ("hex20_ex6_perturbed", 20, (3, 2, 2), (1, 0.004, 5), (0.3, -0.2, -1.0)),
]


@pytest.mark.parametrize("name,etype,dims,skew,g", EX6, ids=[c[0] for c in EX6])
def test_ex6_with_the_traction_term(name, etype, dims, skew, g):
    F, O = _mods()
    m, s = make_case(etype, dims, bounds=(0, 1, 0, 0.3, 0, 0.1), n_vars=3, skew=skew, dirichlet=(4,), kind=1)
    e, sd = m.boundary_sides(2)  # BOUNDARY_ID_MAX_X
    s.set_side_loads(e, sd, g)
    ctx = make_ctx(m, s)
    ctx.set_side_loads(e, sd, g)
    p = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS)
    vo, ro, _ = s.assemble(O.ELASTICITY_EX6)
    ctx.assemble(p)
    v, r = ctx.get_values()
    rp, _ = s.sparsity()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    Ko, Fo = s.element_matrices(O.ELASTICITY_EX6, constrained=False)
    K, Fv = ctx.element_matrices(p, constrained=False)
    assert dense_parity_error(K, Ko) < PARITY_TOL
    assert vec_parity_error(Fv.reshape(-1), Fo.reshape(-1)) < PARITY_TOL
    # a sub-range that starts inside the loaded elements (records are offset by e0)
    e0 = int(e[len(e) // 2])
    _, F2 = ctx.element_matrices(p, e0=e0, e1=m.n_elem, constrained=True)
    _, Fo2 = s.element_matrices(O.ELASTICITY_EX6, e0=e0, e1=m.n_elem, constrained=True)
    assert vec_parity_error(F2.reshape(-1), Fo2.reshape(-1)) < PARITY_TOL
    # the load really is there, and clearing the list takes it away again
    s.set_side_loads([], [], g)
    _, r0, _ = s.assemble(O.ELASTICITY_EX6)
    assert np.abs(ro - r0).max() > 1e-4
    ctx.set_side_loads([], [], g)
    ctx.assemble(p)
    _, r = ctx.get_values()
    assert vec_parity_error(r, r0) < PARITY_TOL
    ctx.close()


@pytest.mark.parametrize("etype,dims,skew", [(27, (3, 3, 2), None), (27, (3, 2, 2), (1, 0.03, 5)), (8, (4, 4, 3), (1, 0.05, 9)), (20, (3, 3, 2), (1, 0.03, 2))])
def test_neumann_flux_with_two_loaded_sides_per_corner_element(etype, dims, skew):
    F, O = _mods()
    m, s = make_case(etype, dims, skew=skew, dirichlet=(4,))
    e2, s2 = m.boundary_sides(2)
    e3, s3 = m.boundary_sides(3)
    e, sd = np.concatenate([e3, e2]), np.concatenate([s3, s2])  # unsorted on purpose
    assert len(np.intersect1d(e2, e3)) > 0
    s.set_side_loads(e, sd, [1.7])
    ctx = make_ctx(m, s)
    ctx.set_side_loads(e, sd, [1.7])
    qorder = 3 if etype == 8 else 5
    f_tab, _ = s.forcing_at_qps(qorder)
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    ctx.assemble(F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab))
    v, r = ctx.get_values()
    rp, _ = s.sparsity()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    ctx.close()


def test_side_loads_reject_bad_input():
    F, O = _mods()
    m, s = make_case(10, (2, 2, 1))
    ctx = make_ctx(m, s)
    with pytest.raises(F.FeasmError):
        ctx.set_side_loads([0], [4], [1.0])  # a tet has sides 0..3
    ctx.close()
    m, s = make_case(8, (2, 2, 2))
    ctx = make_ctx(m, s)
    with pytest.raises(F.FeasmError):
        ctx.set_side_loads([0], [6], [1.0])
    with pytest.raises(F.FeasmError):
        ctx.set_side_loads([m.n_elem], [0], [1.0])
    ctx.close()


def _all_boundary_sides(m):
    es, ss = zip(*(m.boundary_sides(b) for b in range(6)))
    return np.concatenate(es), np.concatenate(ss)


@pytest.mark.parametrize("etype,dims,skew,qorder", [(27, (4, 3, 2), None, 0), (27, (3, 3, 3), (1, 0.04, 7), 0), (8, (5, 4, 3), (1, 0.05, 3), 3), (20, (3, 3, 2), (1, 0.04, 4), 0)])
def test_penalty_sides_match_the_ex3_loop(etype, dims, skew, qorder):
    """feasm_set_side_penalty (introduction_ex3.C:386-455 on HEX sides) against the oracle: element matrices and the
    assembled system of the penalty form of ex4 (no constraint rows; penalty 1e10 on every boundary side)."""
    from libmesh_b200 import capi as F
    from oracle import oracle as O

    m = O.Mesh(etype, *dims, bounds=(-1, 1, -1, 1, -1, 1))
    if skew:
        m.skew(*skew)
    s = O.System(m, 1)
    e, sd = _all_boundary_sides(m)
    s.set_side_penalty(e, sd, penalty=1e10, kind=1)
    s.build_sparsity()
    xyz, conn, _ = m.arrays()
    ctx = F.Context(0)
    ctx.set_mesh(etype, xyz, conn)
    ctx.set_dofs(s.dofs()[0], s.n_dofs)
    ctx.build_pattern()
    ctx.set_side_penalty(e, sd, penalty=1e10, kind=F.PENALTY_EX4_EXACT)
    f_tab, _ = s.forcing_at_qps(qorder)
    p = F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab)
    Ko, Fo = s.element_matrices(O.POISSON_EX4, constrained=False, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    K, Fv = ctx.element_matrices(p, constrained=False)
    assert dense_parity_error(K, Ko) < PARITY_TOL
    assert vec_parity_error(Fv.ravel(), Fo.ravel()) < PARITY_TOL
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    ctx.assemble(p)
    v, r = ctx.get_values()
    rp, _ = s.sparsity()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    # clearing the list gives the plain volume system back
    ctx.set_side_penalty([], [])
    K2, _ = ctx.element_matrices(p, constrained=False)
    K0, _ = O.System(m, 1).element_matrices(O.POISSON_EX4, constrained=False, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    assert dense_parity_error(K2, K0) < PARITY_TOL


def test_penalty_constant_value_on_every_variable():
    from libmesh_b200 import capi as F
    from oracle import oracle as O

    m = O.Mesh(27, 3, 2, 2, bounds=(0, 1, 0, 0.3, 0, 0.1))
    s = O.System(m, 3)
    e, sd = m.boundary_sides(4)  # clamp x-min by penalty instead of the constraint rows
    s.set_side_penalty(e, sd, penalty=1e8, kind=0, value=0.25)
    xyz, conn, _ = m.arrays()
    ctx = F.Context(0)
    ctx.set_mesh(27, xyz, conn)
    ctx.set_dofs(s.dofs()[0], s.n_dofs, n_vars=3)
    ctx.set_side_penalty(e, sd, penalty=1e8, kind=F.PENALTY_CONSTANT, value=0.25)
    p = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS)
    Ko, Fo = s.element_matrices(O.ELASTICITY_EX6, constrained=False)
    K, Fv = ctx.element_matrices(p, constrained=False)
    assert dense_parity_error(K, Ko) < PARITY_TOL
    assert vec_parity_error(Fv.ravel(), Fo.ravel()) < PARITY_TOL


@pytest.mark.parametrize("skew", [None, (1, 0.02, 5)])
def test_tet10_side_load_and_penalty(skew):
    """TRI6 sides of TET10 (QGauss(2, FIFTH), 7 points): a Neumann flux on box face 2 and the ex3 penalty term on the other
    five faces, element vectors / matrices and the assembled system against the oracle."""
    from libmesh_b200 import capi as F
    from oracle import oracle as O
    from test_oracle_sides import tet_boundary_sides

    m = O.Mesh(O.TET10, 3, 2, 2, bounds=(-1, 1, -1, 1, -1, 1))
    if skew:
        m.skew(*skew)
    s = O.System(m, 1)
    le, ls = tet_boundary_sides(m, 2)
    pe, ps = zip(*(tet_boundary_sides(m, b) for b in (0, 1, 3, 4, 5)))
    pe, ps = np.concatenate(pe), np.concatenate(ps)
    s.set_side_loads(le, ls, [1.7])
    s.set_side_penalty(pe, ps, penalty=1e10, kind=1)
    s.build_sparsity()
    xyz, conn, _ = m.arrays()
    ctx = F.Context(0)
    ctx.set_mesh(F.TET10, xyz, conn)
    ctx.set_dofs(s.dofs()[0], s.n_dofs)
    ctx.build_pattern()
    ctx.set_side_loads(le, ls, [1.7])
    ctx.set_side_penalty(pe, ps, penalty=1e10, kind=F.PENALTY_EX4_EXACT)
    f_tab, _ = s.forcing_at_qps(5)
    p = F.make_params(F.POISSON, 5, F.FORCING_TABULATED, f_host=f_tab)
    Ko, Fo = s.element_matrices(O.POISSON_EX4, constrained=False, qorder=5, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    K, Fv = ctx.element_matrices(p, constrained=False)
    assert dense_parity_error(K, Ko) < PARITY_TOL
    assert vec_parity_error(Fv.ravel(), Fo.ravel()) < PARITY_TOL
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=5, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    ctx.assemble(p)
    v, r = ctx.get_values()
    assert csr_parity_error(v, vo, s.sparsity()[0]) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    with pytest.raises(F.FeasmError):
        ctx.set_side_loads([0], [4], [1.0])  # a tet has four sides
