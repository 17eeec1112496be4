"""Oracle mesh / dof numbering / sparsity / assembly against independent restatements and the reference's goldens.

Golden facts come from tests/golden/reference_runs.json (stdout of real libMesh runs, extracted by
tests/golden/make_golden.py).  Structural checks restate, in plain numpy and independently of oracle/*.cpp:
  src/mesh/mesh_generation.C:117-150, :1285-1432, :1575-1626 (build_cube node ids, element order, boundary ids),
  src/base/dof_map.C:1417-1535 (first-touch dof numbering), src/base/sparsity_pattern.C:96-363 (row unions),
  src/base/dof_map_constraints.C:2518-2614 (Dirichlet element constraint), laspack_matrix.C:259-275 (scatter-add).
Also the reference tests' known answers: tests/base/dof_map_test.C:84-119 (contiguous ownership ranges),
tests/systems/systems_test.C:128-205 (identity-Ke scatter), tests/numerics/sparse_matrix_test.h:88-149.
"""
import json
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_runs.json")))


def brute_sparsity(elem_dofs, n_dofs):
    rows = [set() for _ in range(n_dofs)]
    for ed in elem_dofs:
        s = set(int(d) for d in ed)
        for d in s:
            rows[d] |= s
    rp = np.zeros(n_dofs + 1, dtype=np.uint64)
    ci = []
    for r, s in enumerate(rows):
        c = sorted(s)
        ci.extend(c)
        rp[r + 1] = rp[r] + len(c)
    return rp, np.array(ci, dtype=np.uint32)


def first_touch(conn, n_nodes, n_vars=1):
    base = np.full(n_nodes, -1, dtype=np.int64)
    nxt = 0
    for e in conn:
        for n in e:
            if base[n] < 0:
                base[n] = nxt
                nxt += n_vars
    return base, nxt


def test_golden_introduction_ex4_d3_n6():
    g = GOLD["introduction_ex4_d3_n6"]
    m = O.Mesh(O.HEX27, 6, 6, 6, bounds=(-1, 1, -1, 1, -1, 1))
    s = O.System(m)
    assert (m.n_nodes, m.n_elem, s.n_dofs) == (g["n_nodes"], g["n_elem"], g["n_dofs"])
    assert s.add_dirichlet() == g["n_constrained_dofs"]
    _, cv = s.constraints()
    # libMesh counts a constraint as heterogeneous when its rhs value is nonzero (dof_map_constraints.C)
    assert int(np.count_nonzero(cv)) == g["n_heterogeneous_constraints"]
    nnz = s.build_sparsity()
    assert round(nnz / s.n_dofs, 4) == g["avg_on_proc_bandwidth"]
    rp, _ = s.sparsity()
    assert int(np.diff(rp.astype(np.int64)).max()) == g["max_on_proc_bandwidth"]


def test_golden_systems_of_equations_ex6_40x10x5():
    g = GOLD["systems_of_equations_ex6_40x10x5"]
    m = O.Mesh(O.HEX8, 40, 10, 5, bounds=(0, 1, 0, 0.3, 0, 0.1))
    s = O.System(m, 3)
    assert (m.n_nodes, m.n_elem, s.n_dofs) == (g["n_nodes"], g["n_elem"], g["n_dofs"])
    nnz = s.build_sparsity()
    assert round(nnz / s.n_dofs, 4) == g["avg_on_proc_bandwidth"]
    rp, _ = s.sparsity()
    assert int(np.diff(rp.astype(np.int64)).max()) == g["max_on_proc_bandwidth"]
    # ex6 constrains min-x (id 4) + the max-x/max-y/max-z corner node + the max-x/min-y edge, all three variables
    _, _, mask = m.arrays()
    bit = lambda b: (mask >> b) & 1
    con = bit(4) | (bit(2) & bit(3) & bit(5)) | (bit(2) & bit(1))
    assert 3 * int(con.sum()) == g["n_constrained_dofs"]
    assert s.add_dirichlet((4,), 1) == 3 * int(bit(4).sum())


@pytest.mark.parametrize("etype,n", [(O.HEX8, 3), (O.HEX27, 3), (O.TET10, 2)])
def test_mesh_counts_coordinates_and_boundary_ids(etype, n):
    m = O.Mesh(etype, n, n, n, bounds=(-1, 1, 0, 2, 0, 3), renumber_nodes=False)
    xyz, conn, mask = m.arrays()
    if etype == O.HEX8:
        assert (m.n_elem, m.n_nodes) == (n ** 3, (n + 1) ** 3)
    elif etype == O.HEX27:
        assert (m.n_elem, m.n_nodes) == (n ** 3, (2 * n + 1) ** 3)
    else:
        assert m.n_elem == 24 * n ** 3  # mesh_generation_test.C: 24 tets per hex cell
    # every node of every element is referenced, ids are dense
    assert np.array_equal(np.unique(conn), np.arange(m.n_nodes))
    # boundary bits: 0 z-min, 1 y-min, 2 x-max, 3 y-max, 4 x-min, 5 z-max (mesh_generation.C:1416-1432)
    lo, hi = np.array([-1, 0, 0.0]), np.array([1, 2, 3.0])
    want = ((xyz[:, 2] == lo[2]) * 1 | (xyz[:, 1] == lo[1]) * 2 | (xyz[:, 0] == hi[0]) * 4 | (xyz[:, 1] == hi[1]) * 8
            | (xyz[:, 0] == lo[0]) * 16 | (xyz[:, 2] == hi[2]) * 32)
    assert np.array_equal(mask, want.astype(np.uint8))
    if etype != O.TET10:
        # lattice ids, i fastest (mesh_generation.C:117-150), two-step coordinates (:1289-1296, :1779-1785)
        s = 1 if etype == O.HEX8 else 2
        g = s * n
        ids = np.arange(m.n_nodes)
        i, j, k = ids % (g + 1), (ids // (g + 1)) % (g + 1), ids // (g + 1) ** 2
        assert np.array_equal(xyz[:, 0], i / g * 2.0 + -1.0)
        assert np.array_equal(xyz[:, 1], j / g * 2.0 + 0.0)
        assert np.array_equal(xyz[:, 2], k / g * 3.0 + 0.0)
        # element 0 = corner cell; HEX local nodes 0..7 are the corners in libMesh order
        c0 = xyz[conn[0][:8]]
        h = np.array([2.0, 2.0, 3.0]) / n
        ref = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]])
        assert np.allclose(c0, lo + ref * h, atol=1e-15)
        if etype == O.HEX27:
            assert np.allclose(xyz[conn[0][26]], lo + 0.5 * h, atol=1e-15)  # centre node
            assert np.allclose(xyz[conn[0][20]], lo + np.array([.5, .5, 0]) * h, atol=1e-15)  # z-min face node
            assert np.allclose(xyz[conn[0][8]], lo + np.array([.5, 0, 0]) * h, atol=1e-15)  # edge 0-1
    # total volume via the FE map
    pts, w = O.quadrature(etype, 5)
    vol = sum(O.fe_reinit(etype, xyz[conn[e]], pts, w)["JxW"].sum() for e in range(m.n_elem))
    assert abs(vol - 12.0) < 1e-11


@pytest.mark.parametrize("etype,n,nv", [(O.HEX8, 3, 1), (O.HEX27, 2, 1), (O.HEX27, 2, 3), (O.TET10, 2, 1)])
def test_dof_numbering_is_first_touch_and_variable_major(etype, n, nv):
    m = O.Mesh(etype, n, n, n)
    s = O.System(m, nv)
    _, conn, _ = m.arrays()
    ed, nb, fd = s.dofs()
    base, nd = first_touch(conn, m.n_nodes, nv)
    assert s.n_dofs == nd and np.array_equal(nb, base)
    assert list(fd) == [0, nd]
    # dof_indices: variable-major, element-local node order (dof_map.C:2280-2308); vars interleaved per node
    want = np.concatenate([base[conn] + v for v in range(nv)], axis=1)
    assert np.array_equal(ed, want)


def test_multirank_dof_ranges_are_contiguous_and_owner_is_min_rank():
    """tests/base/dof_map_test.C:84-119 on the 3^3 HEX27 mesh it uses, 3 ranks (the reference CI runs mpirun -np 3)."""
    m = O.Mesh(O.HEX27, 3, 3, 3)
    m.partition_linear(3)
    s = O.System(m)
    _, conn, _ = m.arrays()
    ep, npid = m.partition()
    assert list(np.bincount(ep)) == [9, 9, 9]  # LinearPartitioner: contiguous blocks in element order
    ed, nb, fd = s.dofs()
    own = np.full(m.n_nodes, 99)
    for e, c in enumerate(conn):
        own[c] = np.minimum(own[c], ep[e])  # node.C:99-102
    assert np.array_equal(own, npid)
    assert fd[0] == 0 and fd[-1] == s.n_dofs == m.n_nodes
    for p in range(3):
        mine = nb[npid == p]
        assert mine.min() == fd[p] and mine.max() == fd[p + 1] - 1 and len(mine) == fd[p + 1] - fd[p]
    # product-side generator agrees (host logic of the C ABI, no GPU needed)
    from libmesh_b200 import capi as F

    n2, base2, first2, npid2 = F.distribute_dofs(conn, m.n_nodes, 1, 3, ep)
    assert n2 == s.n_dofs and np.array_equal(base2, nb) and np.array_equal(first2, fd) and np.array_equal(npid2, npid)


@pytest.mark.parametrize("etype,n,nv", [(O.HEX8, 4, 1), (O.HEX27, 3, 1), (O.HEX27, 2, 3), (O.TET10, 2, 1)])
def test_sparsity_is_sorted_unique_row_union(etype, n, nv):
    m = O.Mesh(etype, n, n, n)
    s = O.System(m, nv)
    ed, _, _ = s.dofs()
    nnz = s.build_sparsity()
    rp, ci = s.sparsity()
    rp2, ci2 = brute_sparsity(ed, s.n_dofs)
    assert nnz == len(ci2) and np.array_equal(rp, rp2) and np.array_equal(ci, ci2)
    if etype == O.HEX8:
        assert nnz == nv * nv * (3 * n + 1) ** 3
    if etype == O.HEX27:
        assert nnz == nv * nv * (8 * n + 1) ** 3


def test_poisson_element_matrix_invariants():
    m = O.Mesh(O.HEX27, 2, 2, 2, bounds=(-1, 1, -1, 1, -1, 1))
    m.skew(1, 0.05, 12345)
    s = O.System(m)
    Ke, Fe = s.element_matrices(O.POISSON_EX4, constrained=False)
    assert np.abs(Ke - Ke.transpose(0, 2, 1)).max() < 1e-13
    assert np.abs(Ke.sum(axis=2)).max() < 1e-12  # constants are in the kernel of the Laplacian
    assert np.linalg.eigvalsh(Ke[0]).min() > -1e-12
    # independent restatement of introduction_ex4.C:464-467 / :536-537 from FE::reinit outputs
    xyz, conn, _ = m.arrays()
    pts, w = O.quadrature(O.HEX27, 5)
    f_tab, xq = s.forcing_at_qps()
    for e in (0, 5):
        r = O.fe_reinit(O.HEX27, xyz[conn[e]], pts, w)
        K = np.einsum("q,iqd,jqd->ij", r["JxW"], r["dphi"], r["dphi"])
        assert np.abs(K - Ke[e]).max() < 1e-13
        assert np.abs(xq[e] - r["xyz"]).max() < 1e-15
        F = r["phi"] @ (r["JxW"] * f_tab[e])
        assert np.abs(F - Fe[e]).max() < 1e-13
    # the forcing is -Laplace(u) of u = cos(pi x/2) sin(pi y/2) cos(pi z/2) up to the FD truncation (eps = 1e-3)
    u = np.cos(.5 * np.pi * xq[..., 0]) * np.sin(.5 * np.pi * xq[..., 1]) * np.cos(.5 * np.pi * xq[..., 2])
    assert np.abs(f_tab - 0.75 * np.pi ** 2 * u).max() < 1e-5


def test_elasticity_element_matrix_invariants():
    m = O.Mesh(O.HEX27, 2, 1, 1, bounds=(0, 1, 0, 0.3, 0, 0.1))
    s = O.System(m, 3)
    Ke, Fe = s.element_matrices(O.ELASTICITY_EX6, constrained=False)
    assert Ke.shape == (2, 81, 81)
    assert np.abs(Ke - Ke.transpose(0, 2, 1)).max() < 1e-12
    xyz, conn, _ = m.arrays()
    X = xyz[conn[0]]
    # six rigid-body modes in the null space; layout variable-major [u@nodes | v@nodes | w@nodes]
    modes = []
    for d in range(3):
        t = np.zeros((3, 27))
        t[d] = 1
        modes.append(t.reshape(-1))
    for a, b in ((0, 1), (1, 2), (0, 2)):
        r = np.zeros((3, 27))
        r[a], r[b] = -X[:, b], X[:, a]
        modes.append(r.reshape(-1))
    scale = np.abs(Ke[0]).max()
    for v in modes:
        assert np.abs(Ke[0] @ v).max() < 1e-12 * scale * max(1, np.abs(v).max())
    # body force (0,0,-1): Fe_w sums to -volume, Fe_u = Fe_v = 0 (systems_of_equations_ex6.C:233-236)
    vol = 0.5 * 0.3 * 0.1
    assert abs(Fe[0][54:].sum() + vol) < 1e-14 and np.abs(Fe[0][:54]).max() == 0
    # one entry restated literally from :223-230 with C_ijkl of :136-151
    pts, w = O.quadrature(O.HEX27, 5)
    r = O.fe_reinit(O.HEX27, X, pts, w)
    nu, E = 0.3, 1.0
    l1, l2 = E * nu / ((1 + nu) * (1 - 2 * nu)), E / (2 * (1 + nu))
    dl = np.eye(3)
    C = l1 * np.einsum("ij,kl->ijkl", dl, dl) + l2 * (np.einsum("ik,jl->ijkl", dl, dl) + np.einsum("il,jk->ijkl", dl, dl))
    K = np.einsum("q,ijkl,bql,aqj->iakb", r["JxW"], C, r["dphi"], r["dphi"]).reshape(81, 81)
    assert np.abs(K - Ke[0]).max() < 1e-12 * scale


@pytest.mark.parametrize("problem,nv", [(O.POISSON_EX4, 1), (O.ELASTICITY_EX6, 3)])
def test_assembly_is_scatter_add_of_constrained_element_matrices(problem, nv):
    """Global CSR = sum_e P_e^T Ke P_e in element order (laspack_matrix.C:259-275); Dirichlet per
    dof_map_constraints.C:2518-2614 (ex4, heterogeneous) / :2439-2514 (ex6, homogeneous)."""
    if nv == 1:
        m = O.Mesh(O.HEX27, 3, 3, 3, bounds=(-1, 1, -1, 1, -1, 1))
        s = O.System(m)
        s.add_dirichlet()
    else:
        m = O.Mesh(O.HEX27, 3, 2, 2, bounds=(0, 1, 0, 0.3, 0, 0.1))
        s = O.System(m, 3)
        s.add_dirichlet((4,), 1)
    ed, _, _ = s.dofs()
    cd, cv = s.constraints()
    n = s.n_dofs
    s.build_sparsity()
    rp, ci = s.sparsity()
    v, b, _ = s.assemble(problem)
    A = np.zeros((n, n))
    for r in range(n):
        A[r, ci[int(rp[r]):int(rp[r + 1])]] = v[int(rp[r]):int(rp[r + 1])]
    # independent: unconstrained element matrices + a literal restatement of the constraint algebra
    Ke, Fe = s.element_matrices(problem, constrained=False)
    flag = np.zeros(n, dtype=bool)
    flag[cd] = True
    g = np.zeros(n)
    g[cd] = cv
    A2, b2 = np.zeros((n, n)), np.zeros(n)
    for e in range(m.n_elem):
        d = ed[e].astype(np.int64)
        c = flag[d]
        K, F = Ke[e].copy(), Fe[e].copy()
        H = np.where(c, g[d], 0.0)
        Cm = np.diag((~c).astype(float))
        if problem == O.POISSON_EX4:
            F = Cm.T @ (F - K @ H)
        else:
            F = Cm.T @ F
        K = Cm.T @ K @ Cm
        for i in np.nonzero(c)[0]:
            K[i, i] = 1.0
            F[i] = H[i] if problem == O.POISSON_EX4 else 0.0
        A2[np.ix_(d, d)] += K
        b2[d] += F
    assert np.abs(A - A2).max() < 1e-12 * np.abs(A2).max()
    assert np.abs(b - b2).max() < 1e-12 * max(np.abs(b2).max(), 1)
    # constrained diagonal = number of touching elements, rhs = count * g
    cnt = np.bincount(ed.reshape(-1), minlength=n)
    assert np.array_equal(np.diag(A)[cd], cnt[cd].astype(float))
    if problem == O.POISSON_EX4:
        assert np.abs(b[cd] - cnt[cd] * cv).max() < 1e-13
    # constrained rows/cols are otherwise zero but stay in the pattern
    off = A.copy()
    np.fill_diagonal(off, 0)
    assert np.abs(off[cd]).max() == 0 and np.abs(off[:, cd]).max() == 0
    # threaded baseline variant sums the same contributions
    v8, b8, _ = s.assemble(problem, n_threads=4)
    assert np.abs(v8 - v).max() < 1e-13 * np.abs(v).max() and np.abs(b8 - b).max() < 1e-13 * max(np.abs(b).max(), 1)


def test_identity_element_matrix_scatter_counts_touching_elements():
    """tests/systems/systems_test.C:128-205 idea: scatter Ke = I through dof_indices => diagonal = touch count."""
    m = O.Mesh(O.HEX27, 2, 2, 2)
    s = O.System(m)
    ed, _, _ = s.dofs()
    s.build_sparsity()
    rp, ci = s.sparsity()
    cnt = np.bincount(ed.reshape(-1), minlength=s.n_dofs)
    assert cnt.max() == 8 and cnt.min() == 1
    # a vertex node shared by 8 elements has 125 row entries, a cell-centre node 27
    lens = np.diff(rp.astype(np.int64))
    assert lens.max() == 125 and lens.min() == 27
    for r in range(s.n_dofs):  # ascending columns (sparse_matrix_test.h:88-149 get_row contract)
        c = ci[int(rp[r]):int(rp[r + 1])]
        assert np.all(np.diff(c.astype(np.int64)) > 0)
