"""GPU: general constraint rows (hanging nodes) — feasm_set_constraint_rows against the Python restatement of
DofMap::heterogeneously_constrain_element_matrix_and_vector (oracle/constraints.py) on locally refined HEX8 meshes (weights
1/2, 1/4) and HEX27 meshes (the quadratic interpolation weights 3/8, 3/4, -1/8 and their products)."""
import numpy as np
import pytest

from conftest import PARITY_TOL
from hanging import refined_cube, refined_cube27
from oracle import constraints as OC
from oracle.parity import csr_parity_error, vec_parity_error

pytestmark = pytest.mark.gpu


def _mods():
    from libmesh_b200 import capi as F
    from oracle import oracle as O

    return F, O


def exact(p):
    return np.cos(.5 * np.pi * p[:, 0]) * np.sin(.5 * np.pi * p[:, 1]) * np.cos(.5 * np.pi * p[:, 2])


def linear(p):
    return 0.3 + p[:, 0] - 0.5 * p[:, 1] + 0.25 * p[:, 2]


def case(n, lo, hi, bc, etype=8):
    F, O = _mods()
    make = refined_cube27 if etype == 27 else refined_cube
    xyz, conn, mask, hang = make(n, lambda i, j, k: lo <= i < hi and lo <= j < hi and lo <= k < hi)
    m = O.Mesh.from_arrays(etype, xyz, conn, mask)
    s = O.System(m, 1)
    ed, nb, _ = s.dofs()
    rows = {int(nb[h]): [(int(nb[v]), c) for v, c in lst] for h, lst in hang.items()}
    vals = {}
    bn = np.nonzero(mask)[0]
    g = bc(xyz[bn])
    for node, v in zip(bn, g):
        rows[int(nb[node])] = []
        vals[int(nb[node])] = float(v)
    return xyz, conn, hang, m, s, ed, nb, OC.ConstraintRows(rows, vals)


@pytest.mark.parametrize("n,lo,hi,qorder,etype", [(4, 1, 3, 3, 8), (6, 2, 5, 3, 8), (5, 1, 3, 5, 8), (3, 1, 2, 5, 27), (4, 1, 3, 5, 27)])
def test_hanging_node_assembly_matches_the_reference_algorithm(n, lo, hi, qorder, etype):
    F, O = _mods()
    xyz, conn, hang, m, s, ed, nb, con = case(n, lo, hi, exact, etype)
    assert len(hang) > 0
    f_tab, _ = s.forcing_at_qps(qorder)
    Ku, Fu = s.element_matrices(O.POISSON_EX4, constrained=False, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    rp, ci = OC.sparsity(con, ed, s.n_dofs)
    vo, ro = OC.assemble(con, ed, Ku, Fu, rp, ci)

    ctx = F.Context(0)
    ctx.set_mesh(etype, xyz, conn)
    ctx.set_dofs(ed, s.n_dofs)
    ctx.set_constraint_rows(*con.arrays())
    ctx.build_pattern()   # the device-built Graph plus the couplings find_connected_dofs adds
    rp2, ci2 = ctx.get_pattern()
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2), "pattern differs from SparsityPattern::Build with constraints"
    p = F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab)
    ctx.assemble(p)
    v, r = ctx.get_values()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    ctx.assemble(p)       # fixed-order adds: bitwise reproducible
    v2, r2 = ctx.get_values()
    assert np.array_equal(v, v2) and np.array_equal(r, r2)
    # the host-pattern route (libMesh's own Graph) gives the same system
    ctx2 = F.Context(0)
    ctx2.set_mesh(etype, xyz, conn)
    ctx2.set_dofs(ed, s.n_dofs)
    ctx2.set_constraint_rows(*con.arrays())
    ctx2.set_pattern(rp, ci)
    ctx2.assemble(p)
    v3, r3 = ctx2.get_values()
    assert np.array_equal(v, v3) and np.array_equal(r, r3)
    ctx.close(); ctx2.close()


def test_hanging_node_patch_test_on_the_device_matrix():
    """Laplace with linear Dirichlet data: the constrained system assembled on the GPU reproduces the linear field at every
    node, hanging ones included."""
    F, O = _mods()
    xyz, conn, hang, m, s, ed, nb, con = case(4, 1, 3, linear)
    ctx = F.Context(0)
    ctx.set_mesh(8, xyz, conn)
    ctx.set_dofs(ed, s.n_dofs)
    ctx.set_constraint_rows(*con.arrays())
    ctx.build_pattern()
    ctx.assemble(F.make_params(F.POISSON, 3, F.FORCING_NONE))
    v, r = ctx.get_values()
    rp, ci = ctx.get_pattern()
    A = np.zeros((s.n_dofs, s.n_dofs))
    for row in range(s.n_dofs):
        A[row, ci[int(rp[row]):int(rp[row + 1])]] = v[int(rp[row]):int(rp[row + 1])]
    u = np.linalg.solve(A, r)
    want = np.zeros(s.n_dofs)
    want[nb] = linear(xyz)
    assert np.abs(u - want).max() < 1e-12
    ctx.close()


def harmonic_quadratic(p):
    return 0.3 + p[:, 0] - 0.5 * p[:, 1] + 0.25 * p[:, 2] + p[:, 0] ** 2 - p[:, 1] ** 2 + 0.5 * p[:, 0] * p[:, 2]


def test_hex27_hanging_node_patch_test_with_a_quadratic_field():
    """Q2 elements with the quadratic hanging-node weights hold every quadratic: Laplace with Dirichlet data from a harmonic
    quadratic is solved exactly at every node by the system assembled on the GPU."""
    F, O = _mods()
    xyz, conn, hang, m, s, ed, nb, con = case(3, 1, 2, harmonic_quadratic, 27)
    assert len(hang) == 72 and {round(c, 6) for lst in hang.values() for _, c in lst} >= {0.375, 0.75, -0.125, 0.140625, -0.046875}
    ctx = F.Context(0)
    ctx.set_mesh(27, xyz, conn)
    ctx.set_dofs(ed, s.n_dofs)
    ctx.set_constraint_rows(*con.arrays())
    ctx.build_pattern()
    ctx.assemble(F.make_params(F.POISSON, 5, F.FORCING_NONE))
    v, r = ctx.get_values()
    rp, ci = ctx.get_pattern()
    A = np.zeros((s.n_dofs, s.n_dofs))
    for row in range(s.n_dofs):
        A[row, ci[int(rp[row]):int(rp[row + 1])]] = v[int(rp[row]):int(rp[row + 1])]
    u = np.linalg.solve(A, r)
    want = np.zeros(s.n_dofs)
    want[nb] = harmonic_quadratic(xyz)
    assert np.abs(u - want).max() < 1e-12
    ctx.close()


def test_dirichlet_rows_through_the_general_entry_are_the_fast_path():
    """Empty rows only: feasm_set_constraint_rows is feasm_set_constraints (nothing is widened), bit for bit."""
    F, O = _mods()
    m = O.Mesh(O.HEX8, 5, 4, 3, bounds=(-1, 1, -1, 1, -1, 1))
    s = O.System(m, 1)
    s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    s.build_sparsity()
    xyz, conn, _ = m.arrays()
    ed, _, _ = s.dofs()
    cd, cv = s.constraints()
    out = []
    for general in (False, True):
        ctx = F.Context(0)
        ctx.set_mesh(8, xyz, conn)
        ctx.set_dofs(ed, s.n_dofs)
        if general:
            ctx.set_constraint_rows(cd, np.zeros(len(cd) + 1, dtype=np.uint64), [], [], cv)
        else:
            ctx.set_constraints(cd, cv)
        ctx.build_pattern()
        ctx.assemble(F.make_params(F.POISSON, 3, F.FORCING_EX4_FD))
        out.append(ctx.get_values())
        ctx.close()
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1])


def test_general_rows_errors_are_loud():
    F, O = _mods()
    xyz, conn, hang, m, s, ed, nb, con = case(4, 1, 3, exact)
    ctx = F.Context(0)
    ctx.set_mesh(8, xyz, conn)
    ctx.set_dofs(ed, s.n_dofs)
    d, rp, c, w, v = con.arrays()
    bad = c.copy()
    bad[0] = s.n_dofs
    with pytest.raises(F.FeasmError):
        ctx.set_constraint_rows(d, rp, bad, w, v)          # constraining dof out of range
    ctx.set_constraint_rows(d, rp, c, w, v)
    ctx.build_pattern()
    with pytest.raises(F.FeasmError):
        ctx.assemble_begin(F.make_params(F.POISSON, 3, F.FORCING_NONE))   # one rank, feasm_assemble only
    # a pattern without the widened couplings is refused, not silently missed
    ctx3 = F.Context(0)
    ctx3.set_mesh(8, xyz, conn)
    ctx3.set_dofs(ed, s.n_dofs)
    ctx3.build_pattern()
    rp0, ci0 = ctx3.get_pattern()
    ctx3.set_constraint_rows(d, rp, c, w, v)
    ctx3.set_pattern(rp0, ci0)
    with pytest.raises(F.FeasmError, match="coupling"):
        ctx3.assemble(F.make_params(F.POISSON, 3, F.FORCING_NONE))
    ctx.close(); ctx3.close()
