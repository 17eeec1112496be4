"""GPU: device-side mesh ingest (feasm_generate_cube / feasm_distribute_dofs_device / feasm_generate_dirichlet, SURVEY §8f
rank 2) against the oracle's build_cube, DofMap numbering and Dirichlet set — integer structures and coordinates bit for
bit — and the assembled system of the all-device route against the host-array route and the oracle."""
import numpy as np
import pytest

from conftest import PARITY_TOL, csr_parity_error, vec_parity_error

pytestmark = pytest.mark.gpu


def _mods():
    from libmesh_b200 import capi as F
    from oracle import oracle as O

    return F, O


@pytest.mark.parametrize("etype,dims,bounds", [(8, (5, 3, 4), (-1, 1, 0, 0.3, 0, 0.1)), (27, (5, 3, 4), (-1, 1, 0, 0.3, 0, 0.1)),
                                               (27, (1, 1, 1), (0, 1, 0, 1, 0, 1)), (8, (1, 7, 2), (-3, -1, 2, 5, 0, 1e-3)),
                                               (20, (4, 3, 2), (-1, 1, 0, 0.3, 0, 0.1)), (18, (4, 3, 2), (-1, 1, 0, 0.3, 0, 0.1))])
def test_generated_cube_is_build_cube_bit_for_bit(etype, dims, bounds):
    F, O = _mods()
    ctx = F.Context(0)
    ctx.generate_cube(etype, *dims, bounds=bounds)
    xyz, conn, mask = ctx.get_mesh()
    m = O.Mesh(etype, *dims, bounds=bounds, renumber_nodes=False)
    xo, co, mo = m.arrays()
    assert np.array_equal(xyz, xo), "coordinates differ from the two-step build_cube arithmetic"
    assert np.array_equal(conn, co) and np.array_equal(mask, mo)
    xh, ch, mh = F.build_cube(etype, *dims, bounds)
    assert np.array_equal(xyz, xh) and np.array_equal(conn, ch) and np.array_equal(mask, mh)


@pytest.mark.parametrize("dims,bounds", [((3, 2, 2), (-1, 1, -1, 1, -1, 1)), ((1, 1, 1), (0, 1, 0, 1, 0, 1)), ((4, 5, 3), (-1, 1, 0, 0.3, 0, 0.1))])
def test_generated_tet10_cube_is_the_24_tet_split_bit_for_bit(dims, bounds):
    """build_cube(TET10): 24 tets per cell + all_second_order's first-sight edge nodes, by sorts on the device."""
    F, O = _mods()
    ctx = F.Context(0)
    ctx.generate_cube(F.TET10, *dims, bounds=bounds)
    xyz, conn, mask = ctx.get_mesh()
    xh, ch, mh = F.build_cube(F.TET10, *dims, bounds)
    assert conn.shape == ch.shape == (24 * dims[0] * dims[1] * dims[2], 10)
    assert np.array_equal(conn, ch), "tet connectivity / edge-node numbering differs from the host split"
    assert np.array_equal(xyz, xh) and np.array_equal(mask, mh)
    m = O.Mesh(O.TET10, *dims, bounds=bounds, renumber_nodes=False)
    xo, co, mo = m.arrays()
    assert np.array_equal(xyz, xo) and np.array_equal(conn, co) and np.array_equal(mask, mo)
    # and the dof numbering + Dirichlet set on top of it
    n = ctx.distribute_dofs(1)
    s = O.System(m, 1)
    assert n == s.n_dofs and np.array_equal(ctx.get_dofs()[0], s.dofs()[0])
    ctx.generate_dirichlet(0x3f, F.DIRICHLET_EX4_EXACT)
    s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    rows, vals = s.constraints()
    flag, val = ctx.get_constraints()
    assert np.array_equal(np.nonzero(flag)[0], np.sort(rows)) and np.max(np.abs(val[rows] - vals)) <= 4e-16


@pytest.mark.parametrize("etype,dims,nv", [(27, (4, 3, 3), 1), (27, (4, 3, 3), 3), (8, (6, 2, 5), 1), (8, (3, 3, 3), 3)])
def test_device_dof_numbering_is_first_touch(etype, dims, nv):
    F, O = _mods()
    ctx = F.Context(0)
    ctx.generate_cube(etype, *dims)
    n = ctx.distribute_dofs(nv)
    m = O.Mesh(etype, *dims, renumber_nodes=False)
    s = O.System(m, nv)
    ed_o, nb_o, _ = s.dofs()
    ed, nb = ctx.get_dofs()
    assert n == s.n_dofs
    assert np.array_equal(ed, ed_o), "dof_indices differ from DofMap::distribute_dofs first-touch numbering"
    assert np.array_equal(nb, nb_o)


@pytest.mark.parametrize("skew", [None, (1, 0.02, 99)])
def test_device_dof_numbering_on_an_uploaded_tet_mesh(skew):
    """The numbering kernel takes whatever mesh the context holds: TET10 (irregular connectivity) through feasm_set_mesh."""
    F, O = _mods()
    m = O.Mesh(O.TET10, 3, 2, 2, bounds=(-1, 1, -1, 1, -1, 1))
    if skew:
        m.skew(*skew)
    s = O.System(m, 1)
    xyz, conn, _ = m.arrays()
    ctx = F.Context(0)
    ctx.set_mesh(F.TET10, xyz, conn)
    assert ctx.distribute_dofs(1) == s.n_dofs
    ed, _ = ctx.get_dofs()
    assert np.array_equal(ed, s.dofs()[0])


@pytest.mark.parametrize("etype,nv,ids,kind", [(27, 1, (0, 1, 2, 3, 4, 5), 0), (8, 1, (0, 5), 0), (27, 3, (4,), 1), (8, 3, (1, 2), 1)])
def test_device_dirichlet_set_matches_oracle(etype, nv, ids, kind):
    F, O = _mods()
    dims, b = (4, 3, 2), (-1, 1, -1, 1, -1, 1)
    ctx = F.Context(0)
    ctx.generate_cube(etype, *dims, bounds=b)
    ctx.distribute_dofs(nv)
    mask = sum(1 << i for i in ids)
    n = ctx.generate_dirichlet(mask, F.DIRICHLET_EX4_EXACT if kind == 0 else F.DIRICHLET_ZERO)
    s = O.System(O.Mesh(etype, *dims, bounds=b), nv)
    s.add_dirichlet(ids, kind)
    rows, vals = s.constraints()
    flag, val = ctx.get_constraints()
    assert n == len(rows)
    assert np.array_equal(np.nonzero(flag)[0], np.sort(rows)), "constrained dof set differs"
    # CUDA sin/cos against glibc: last-bit differences only
    assert np.max(np.abs(val[rows] - vals)) <= 4e-16


@pytest.mark.parametrize("etype,dims,qorder", [(27, (5, 4, 3), 0), (8, (6, 5, 4), 3), (20, (4, 4, 3), 0), (18, (4, 3, 3), 0)])
def test_all_device_route_assembles_the_same_system(etype, dims, qorder):
    F, O = _mods()
    b = (-1, 1, -1, 1, -1, 1)
    ctx = F.Context(0)
    ctx.generate_cube(etype, *dims, bounds=b)
    ctx.distribute_dofs(1)
    ctx.generate_dirichlet(0x3f, F.DIRICHLET_EX4_EXACT)
    ctx.build_pattern()
    m = O.Mesh(etype, *dims, bounds=b)
    s = O.System(m, 1)
    s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    s.build_sparsity()
    rp, ci = s.sparsity()
    rp2, ci2 = ctx.get_pattern()
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2)
    f_tab, _ = s.forcing_at_qps(qorder)
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    ctx.assemble(F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab))
    v, r = ctx.get_values()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    # and the host-array route with the same Dirichlet data gives the same bits
    xyz, conn, _ = ctx.get_mesh()
    ed, _ = ctx.get_dofs()
    flag, val = ctx.get_constraints()
    rows = np.nonzero(flag)[0].astype(np.uint32)
    c2 = F.Context(0)
    c2.set_mesh(etype, xyz, conn)
    c2.set_dofs(ed, ctx.n_rows)
    c2.set_constraints(rows, val[rows])
    c2.build_pattern()
    c2.assemble(F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab))
    v2, r2 = c2.get_values()
    assert np.array_equal(v, v2) and np.array_equal(r, r2)


def test_ingest_errors_are_loud():
    F, O = _mods()
    ctx = F.Context(0)
    with pytest.raises(F.FeasmError):
        ctx.generate_cube(14, 2, 2, 2)              # PYRAMID14: not an element type of this path
    with pytest.raises(F.FeasmError):
        ctx.distribute_dofs(1)                      # no mesh yet
    xyz, conn, _ = F.build_cube(F.HEX8, 2, 2, 2)
    ctx.set_mesh(F.HEX8, xyz, conn)
    ctx.distribute_dofs(1)
    with pytest.raises(F.FeasmError):
        ctx.generate_dirichlet(0x3f, F.DIRICHLET_ZERO)  # boundary ids exist for generated meshes only
