"""GPU parity: the CUDA path, driven through the C ABI (include/feasm.h), against the CPU oracle.

Bar (north_star): sparsity pattern and dof maps bit-exact; matrix / RHS entries within 1e-12 relative per entry
(scale floored at 1% of the row maximum for cancellation entries, see conftest.py and DESIGN.md).
"""
import numpy as np
import pytest

from conftest import PARITY_TOL, csr_parity_error, dense_parity_error, vec_parity_error

pytestmark = pytest.mark.gpu


def _mods():
    from libmesh_b200 import capi as F
    from oracle import oracle as O

    return F, O


def make_case(etype, dims, bounds=(-1, 1, -1, 1, -1, 1), n_vars=1, skew=None, dirichlet=(0, 1, 2, 3, 4, 5), kind=0):
    F, O = _mods()
    m = O.Mesh(etype, *dims, bounds=bounds)
    if skew is not None:
        m.skew(*skew)
    s = O.System(m, n_vars)
    if dirichlet:
        s.add_dirichlet(dirichlet, kind)
    s.build_sparsity()
    return m, s


def make_ctx(m, s, host_pattern=False):
    F, O = _mods()
    xyz, conn, _ = m.arrays()
    ed, _, _ = s.dofs()
    cd, cv = s.constraints()
    ctx = F.Context(0)
    ctx.set_mesh(m.elem_type, xyz, conn)
    ctx.set_dofs(ed, s.n_dofs, n_vars=s.n_vars)
    ctx.set_constraints(cd, cv)
    if host_pattern:
        ctx.set_pattern(*s.sparsity())
    else:
        ctx.build_pattern()
    return ctx


CASES = [
    ("hex27_ex4_6", 27, (6, 6, 6), None, 0),       # the published introduction_ex4 -d 3 -n 6 shape
    ("hex27_ragged", 27, (5, 3, 2), None, 0),
    ("hex27_skew", 27, (4, 4, 4), (1, 0.03, 12345), 0),  # non-affine: general Jacobian branch
    ("hex27_skewfunc", 27, (3, 3, 3), (0, 0.0, 0), 0),   # tests/fe/fe_test.h SkewFunc (affine, sheared)
    ("hex8_q3", 8, (7, 5, 6), None, 3),            # QGauss THIRD = default_quadrature_order for FIRST
    ("hex8_q5", 8, (6, 6, 6), None, 5),            # ex4-faithful FIFTH on HEX8 (introduction_ex4.C:352)
    ("hex8_skew", 8, (5, 5, 5), (1, 0.05, 777), 3),
    ("tet10", 10, (3, 2, 2), None, 5),
    ("tet10_skew", 10, (2, 2, 2), (1, 0.02, 99), 5),
    ("hex27_single", 27, (1, 1, 1), None, 0),      # one element: every dof constrained
    ("hex20", 20, (4, 3, 3), None, 0),             # serendipity hex: 16 rows by DMMA, 4 by DFMA
    ("hex20_skew", 20, (3, 3, 3), (1, 0.03, 4321), 0),
    ("hex20_single", 20, (1, 1, 1), None, 0),
    ("prism18", 18, (4, 3, 3), None, 0),           # TRI6 x EDGE3 prisms, two per lattice cell, 21-point rule
    ("prism18_skew", 18, (3, 3, 2), (1, 0.03, 2468), 0),
    ("prism18_skewfunc", 18, (3, 2, 2), (0, 0.0, 0), 0),
    ("prism18_single_cell", 18, (1, 1, 1), None, 0),
]


@pytest.mark.parametrize("name,etype,dims,skew,qorder", CASES, ids=[c[0] for c in CASES])
def test_pattern_bit_exact_and_values_within_1e12(name, etype, dims, skew, qorder):
    F, O = _mods()
    m, s = make_case(etype, dims, skew=skew)
    ctx = make_ctx(m, s)
    rp, ci = s.sparsity()
    rp2, ci2 = ctx.get_pattern()
    assert ctx.nnz == s.nnz
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2), "device-built sparsity pattern differs from SparsityPattern::Build"
    f_tab, _ = s.forcing_at_qps(qorder)
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    ctx.assemble(F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab))
    v, r = ctx.get_values()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    # bitwise reproducible: the scatter is a fixed-order reduction, not atomics
    ctx.assemble(F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab))
    v2, r2 = ctx.get_values()
    assert np.array_equal(v, v2) and np.array_equal(r, r2)
    ctx.close()


@pytest.mark.parametrize("etype,dims,qorder", [(27, (6, 6, 6), 0), (8, (8, 8, 8), 3), (10, (2, 2, 2), 5), (20, (5, 4, 3), 0), (18, (4, 4, 3), 0)])
def test_device_fd_forcing(etype, dims, qorder):
    """introduction_ex4.C:500-537 evaluated on the device.  The FD Laplacian divides an O(1e-6) difference by 1e-6,
    so one ulp of cos/sin becomes ~1e-10 in f: parity of this mode is stated against max|rhs| at 1e-11;
    the 1e-12 per-entry bar is carried by the tabulated-forcing mode above.  The matrix does not depend on f."""
    F, O = _mods()
    m, s = make_case(etype, dims)
    ctx = make_ctx(m, s)
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=qorder)
    ctx.assemble(F.make_params(F.POISSON, qorder, F.FORCING_EX4_FD))
    v, r = ctx.get_values()
    assert csr_parity_error(v, vo, s.sparsity()[0]) < PARITY_TOL
    assert np.abs(r - ro).max() < 1e-11 * np.abs(ro).max()
    ctx.close()


def test_laplace_no_forcing_tet10():
    F, O = _mods()
    m, s = make_case(10, (2, 2, 3))
    ctx = make_ctx(m, s)
    vo, ro, _ = s.assemble(O.LAPLACE, qorder=5)
    ctx.assemble(F.make_params(F.POISSON, 5, F.FORCING_NONE))
    v, r = ctx.get_values()
    assert csr_parity_error(v, vo, s.sparsity()[0]) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    ctx.close()


@pytest.mark.parametrize("etype,dims,skew,qorder", [(27, (3, 3, 3), (1, 0.03, 5), 0), (8, (4, 4, 4), None, 3), (10, (2, 2, 2), None, 5),
                                                    (20, (3, 2, 2), (1, 0.03, 6), 0), (18, (3, 2, 2), (1, 0.03, 8), 0)])
@pytest.mark.parametrize("constrained", [False, True])
def test_element_matrices(etype, dims, skew, qorder, constrained):
    F, O = _mods()
    m, s = make_case(etype, dims, skew=skew)
    ctx = make_ctx(m, s)
    f_tab, _ = s.forcing_at_qps(qorder)
    Ko, Fo = s.element_matrices(O.POISSON_EX4, constrained=constrained, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    K, Fv = ctx.element_matrices(F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab), constrained=constrained)
    assert dense_parity_error(K, Ko) < PARITY_TOL
    assert vec_parity_error(Fv.reshape(-1), Fo.reshape(-1)) < PARITY_TOL
    if not constrained:
        assert np.abs(K - K.transpose(0, 2, 1)).max() < 1e-14 * np.abs(K).max()
    ctx.close()


def test_host_pattern_route_and_no_constraints():
    """SparseMatrix::update_sparsity_pattern route (sparse_matrix.h:175) and a system without Dirichlet dofs."""
    F, O = _mods()
    m, s = make_case(27, (3, 2, 2), dirichlet=None)
    ctx = make_ctx(m, s, host_pattern=True)
    f_tab, _ = s.forcing_at_qps()
    vo, ro, _ = s.assemble(O.POISSON_EX4, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    ctx.assemble(F.make_params(F.POISSON, 0, F.FORCING_TABULATED, f_host=f_tab))
    v, r = ctx.get_values()
    rp, _ = s.sparsity()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL and vec_parity_error(r, ro) < PARITY_TOL
    # singular Neumann operator: every row sums to ~0
    rows = np.add.reduceat(v, rp[:-1].astype(np.int64))
    assert np.abs(rows).max() < 1e-12 * np.abs(v).max()
    ctx.close()


def test_degenerate_element_is_reported():
    """det J <= 0 must fail like libMesh's DegenerateMap (fe_map.C:467-506), not produce numbers."""
    F, O = _mods()
    m, s = make_case(8, (2, 2, 2))
    xyz, conn, _ = m.arrays()
    ed, _, _ = s.dofs()
    conn = conn.copy()
    conn[3, [0, 1]] = conn[3, [1, 0]]  # swap two nodes: inverted element
    ctx = F.Context(0)
    ctx.set_mesh(8, xyz, conn)
    ctx.set_dofs(ed, s.n_dofs)
    ctx.build_pattern()
    with pytest.raises(F.FeasmError) as e:
        ctx.assemble(F.make_params(F.POISSON, 3, F.FORCING_NONE))
    assert e.value.code == -4
    ctx.close()


def test_add_matrix_add_vector_virtuals():
    """SparseMatrix::add_matrix / NumericVector::add_vector single-element route (sparse_matrix_test.h:88-149)."""
    F, O = _mods()
    m, s = make_case(8, (3, 3, 3), dirichlet=None)
    ctx = make_ctx(m, s)
    ed, _, _ = s.dofs()
    rp, ci = s.sparsity()
    ctx.zero()
    rng = np.random.default_rng(3)
    want = np.zeros(s.nnz)
    wantb = np.zeros(s.n_dofs)
    for e in (0, 5, 13, 5):
        dm = rng.standard_normal((8, 8))
        fv = rng.standard_normal(8)
        ctx.add_matrix(dm, ed[e])
        ctx.add_vector(fv, ed[e])
        for i in range(8):
            r = int(ed[e, i])
            cols = ci[int(rp[r]):int(rp[r + 1])]
            for j in range(8):
                want[int(rp[r]) + int(np.searchsorted(cols, ed[e, j]))] += dm[i, j]
            wantb[r] += fv[i]
    v, b = ctx.get_values()
    assert np.allclose(v, want, rtol=0, atol=1e-15) and np.allclose(b, wantb, rtol=0, atol=1e-15)
    # adding a nonzero outside the pattern is an error; adding zero is allowed (sparse_matrix.h:300-303)
    far = np.array([ed[0, 0], ed[26, 6]], dtype=np.uint32)
    ctx.add_matrix(np.zeros((2, 2)), far)
    with pytest.raises(F.FeasmError) as e:
        ctx.add_matrix(np.ones((2, 2)), far)
    assert e.value.code == -5
    ctx.close()


def test_argument_errors():
    F, O = _mods()
    ctx = F.Context(0)
    with pytest.raises(F.FeasmError):
        ctx.assemble(F.make_params())  # nothing set
    m, s = make_case(8, (2, 2, 2))
    xyz, conn, _ = m.arrays()
    with pytest.raises(F.FeasmError):
        ctx.set_mesh(14, xyz, conn)  # PYRAMID14 is not supported
    ctx.set_mesh(8, xyz, conn)
    ed, _, _ = s.dofs()
    with pytest.raises(F.FeasmError):
        ctx.set_dofs(ed, s.n_dofs - 1)  # dof id out of range
    ctx.close()


# ---- config 5: systems_of_equations_ex6 linear elasticity (3 variables, homogeneous Dirichlet on min-x) -----------
ELAST = [
    ("hex27_ex6", 27, (4, 2, 2), None),
    ("hex27_ex6_skew", 27, (3, 2, 2), (1, 0.004, 7)),
    ("hex8_ex6", 8, (5, 3, 2), None),
    ("tet10_ex6", 10, (3, 2, 1), None),
    ("hex20_ex6", 20, (4, 2, 2), None),
    ("hex20_ex6_skew", 20, (3, 2, 2), (1, 0.004, 9)),
    ("prism18_ex6", 18, (3, 2, 2), None),
    ("prism18_ex6_skew", 18, (3, 2, 1), (1, 0.004, 3)),
]


@pytest.mark.parametrize("name,etype,dims,skew", ELAST, ids=[c[0] for c in ELAST])
def test_elasticity_ex6(name, etype, dims, skew):
    F, O = _mods()
    m, s = make_case(etype, dims, bounds=(0, 1, 0, 0.3, 0, 0.1), n_vars=3, skew=skew, dirichlet=(4,), kind=1)
    ctx = make_ctx(m, s)
    rp, ci = s.sparsity()
    rp2, ci2 = ctx.get_pattern()
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2)
    vo, ro, _ = s.assemble(O.ELASTICITY_EX6)
    p = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS)
    ctx.assemble(p)
    v, r = ctx.get_values()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    Ko, Fo = s.element_matrices(O.ELASTICITY_EX6, constrained=False)
    K, Fv = ctx.element_matrices(p, constrained=False)
    assert dense_parity_error(K, Ko) < PARITY_TOL
    assert vec_parity_error(Fv.reshape(-1), Fo.reshape(-1)) < PARITY_TOL
    # rigid-body translations are in the null space of the unconstrained element matrix
    nd = K.shape[1]
    for var in range(3):
        t = np.zeros(nd)
        t[var * (nd // 3):(var + 1) * (nd // 3)] = 1.0
        assert np.abs(K @ t).max() < 1e-12 * np.abs(K).max()
    ctx.close()


def _superset_pattern(rp, ci, n_rows, dense_rows=(), extra_rows=0):
    """A host pattern that contains the Graph and more: `dense_rows` get every column, `extra_rows` empty rows are
    appended (dofs no element touches).  Returns (row_ptr, col_idx, index of every original entry in the new arrays)."""
    rows = []
    for r in range(n_rows):
        cols = ci[rp[r]:rp[r + 1]]
        rows.append(np.arange(n_rows + extra_rows, dtype=np.uint32) if r in dense_rows else cols)
    rows += [np.zeros(0, np.uint32)] * extra_rows
    rp2 = np.zeros(len(rows) + 1, np.uint64)
    rp2[1:] = np.cumsum([len(c) for c in rows])
    ci2 = np.concatenate(rows).astype(np.uint32)
    where = np.concatenate([rp2[r] + np.searchsorted(rows[r], ci[rp[r]:rp[r + 1]]) for r in range(n_rows)]).astype(np.int64)
    return rp2, ci2, where


def test_rows_without_contributors_and_long_row_fallback():
    """Host pattern (update_sparsity_pattern route) that is a superset of the Graph: (1) trailing dofs no element
    touches give zero rows; (2) one dense row of > 7000 entries pushes the row image out of shared memory, so every row
    goes through the global-memory fallback of the row reduction.  Values on the Graph entries equal the oracle's, all
    other entries are exactly zero."""
    F, O = _mods()
    for etype, dims, dense, extra in [(27, (3, 3, 2), (), 5), (8, (20, 20, 19), (17,), 3)]:
        m, s = make_case(etype, dims)
        rp, ci = s.sparsity()
        rp2, ci2, where = _superset_pattern(rp, ci, s.n_dofs, dense, extra)
        xyz, conn, _ = m.arrays()
        ed, _, _ = s.dofs()
        cd, cv = s.constraints()
        ctx = F.Context(0)
        ctx.set_mesh(m.elem_type, xyz, conn)
        ctx.set_dofs(ed, s.n_dofs + extra, n_vars=1)
        ctx.set_constraints(cd, cv)
        ctx.set_pattern(rp2, ci2)
        q = 3 if etype == 8 else 0
        f_tab, _ = s.forcing_at_qps(q)
        vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=q, forcing=O.FORCING_TABULATED, f_tab=f_tab)
        ctx.assemble(F.make_params(F.POISSON, q, F.FORCING_TABULATED, f_host=f_tab))
        v, r = ctx.get_values()
        assert csr_parity_error(v[where], vo, rp) < PARITY_TOL
        assert vec_parity_error(r[:s.n_dofs], ro) < PARITY_TOL
        rest = np.ones(v.size, bool)
        rest[where] = False
        assert not v[rest].any() and not r[s.n_dofs:].any()
        ctx.close()


def test_cp_async_ring_fill_matches_bulk_copy(monkeypatch):
    """FEASM_RR_BULK=0 selects the cp.async (LDGSTS) ring fill of the row reduction; same bits as the bulk-copy fill."""
    import subprocess, sys, os, textwrap
    code = textwrap.dedent("""
        import sys, numpy as np
        sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
        from test_gpu_parity import make_case, make_ctx, _mods
        F, O = _mods()
        m, s = make_case(27, (4, 3, 3))
        ctx = make_ctx(m, s)
        ctx.assemble(F.make_params(F.POISSON, 0, F.FORCING_EX4_FD))
        v, r = ctx.get_values()
        np.save(sys.argv[1], np.concatenate([v, r]))
    """)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for bulk in ("1", "0"):
        out = os.path.join(root, "build", f"rr_fill_{bulk}.npy")
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.run([sys.executable, "-c", code, out], cwd=root, check=True, env={**os.environ, "FEASM_RR_BULK": bulk}, timeout=300)
        outs.append(np.load(out))
    assert np.array_equal(outs[0], outs[1])


AFFINE = [
    ("hex27_ex4_6", 27, (6, 6, 6), None, 0),
    ("hex27_skew", 27, (4, 4, 4), (1, 0.03, 12345), 0),   # every interior element non-affine: the list kernel does them
    ("hex27_skewfunc", 27, (3, 3, 3), (0, 0.0, 0), 0),    # sheared but affine: full inverse Jacobian in G
    ("hex27_box", 27, (3, 2, 4), None, 0),
    ("hex8_q3", 8, (7, 5, 6), None, 3),
    ("hex8_q5", 8, (6, 6, 6), None, 5),
    ("tet10", 10, (3, 2, 2), None, 5),
    ("tet10_skew", 10, (2, 2, 2), (1, 0.02, 99), 5),
]


@pytest.mark.parametrize("name,etype,dims,skew,qorder", AFFINE, ids=[c[0] for c in AFFINE])
def test_affine_tensor_path_matches_the_quadrature_loop(name, etype, dims, skew, qorder):
    """feasm_params.affine_tensor = 1: affine elements take Ke = sum G_cc' N_cc' (reference tensors), the others the
    quadrature loop; the assembled system meets the same bar against the oracle, on element level and globally."""
    F, O = _mods()
    bounds = (-1, 1, -0.5, 0.7, 0, 3) if name == "hex27_box" else (-1, 1, -1, 1, -1, 1)
    m, s = make_case(etype, dims, bounds=bounds, skew=skew)
    ctx = make_ctx(m, s)
    rp, _ = s.sparsity()
    f_tab, _ = s.forcing_at_qps(qorder)
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    p = F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab, affine_tensor=1)
    ctx.assemble(p)
    v, r = ctx.get_values()
    assert csr_parity_error(v, vo, rp) < PARITY_TOL
    assert vec_parity_error(r, ro) < PARITY_TOL
    Ko, Fo = s.element_matrices(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab, constrained=False)
    K, Fe = ctx.element_matrices(p, constrained=False)
    assert dense_parity_error(K, Ko) < PARITY_TOL
    assert vec_parity_error(Fe.ravel(), Fo.ravel()) < PARITY_TOL
    ctx.close()


@pytest.mark.parametrize("etype,dims,qorder", [(8, (3, 3, 3), 3), (10, (2, 1, 1), 5)])
def test_packed_element_kernels_on_ragged_ranges(etype, dims, qorder):
    """HEX8 (four per warp) and TET10 (two per warp): element ranges whose length and start are not multiples of the pack
    width — short last groups, spare slots — against the oracle, FD forcing in the kernel included."""
    F, O = _mods()
    m, s = make_case(etype, dims, skew=(1, 0.03, 11))
    ctx = make_ctx(m, s)
    f_tab, _ = s.forcing_at_qps(qorder)
    p = F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab)
    Ko, Fo = s.element_matrices(O.POISSON_EX4, constrained=True, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    for e0, e1 in ((0, 1), (1, 4), (3, 8), (m.n_elem - 5, m.n_elem), (0, m.n_elem)):
        K, Fv = ctx.element_matrices(p, e0=e0, e1=e1, constrained=True)
        assert dense_parity_error(K, Ko[e0:e1]) < PARITY_TOL, (e0, e1)
        assert vec_parity_error(Fv.ravel(), Fo[e0:e1].ravel()) < PARITY_TOL, (e0, e1)
    pf = F.make_params(F.POISSON, qorder, F.FORCING_EX4_FD)
    Kf, Ff = ctx.element_matrices(pf, e0=1, e1=m.n_elem, constrained=False)
    _, Fo_fd = s.element_matrices(O.POISSON_EX4, constrained=False, qorder=qorder)
    # FD forcing: one ulp of sin/cos is ~1e-10 in f (the difference quotient divides by 1e-6), times JxW ~ 0.3 on this mesh
    assert np.abs(Ff - Fo_fd[1:]).max() <= 1e-10
