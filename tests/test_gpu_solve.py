"""GPU: the step after the assembly path — SparseMatrix::vector_mult and a CG solve on the device
matrix, through the C ABI.  Checked against the oracle's matrix (scipy) and against the exact solution of ex4."""
import numpy as np
import pytest

from test_gpu_parity import _mods, make_case, make_ctx

pytestmark = pytest.mark.gpu


def _assembled(etype, dims, qorder=0):
    F, O = _mods()
    m, s = make_case(etype, dims)
    ctx = make_ctx(m, s)
    ctx.assemble(F.make_params(F.POISSON, qorder, F.FORCING_EX4_FD))
    return F, O, m, s, ctx


def test_vector_mult_matches_scipy():
    import scipy.sparse as sp
    import torch

    F, O, m, s, ctx = _assembled(27, (4, 3, 3))
    v, _ = ctx.get_values()
    rp, ci = s.sparsity()
    A = sp.csr_matrix((v, ci.astype(np.int64), rp.astype(np.int64)), shape=(s.n_dofs, s.n_dofs))
    rng = np.random.default_rng(7)
    x = rng.standard_normal(s.n_dofs)
    stream = torch.cuda.ExternalStream(ctx.stream())
    with torch.cuda.stream(stream):
        dx = torch.from_numpy(x).cuda()
        dy = torch.empty_like(dx)
        ctx.vector_mult(dx.data_ptr(), dy.data_ptr())
        y = dy.cpu().numpy()
    ref = A @ x
    assert np.abs(y - ref).max() <= 1e-13 * np.abs(ref).max()
    # deterministic: same bits on a second product
    with torch.cuda.stream(stream):
        dy2 = torch.empty_like(dx)
        ctx.vector_mult(dx.data_ptr(), dy2.data_ptr())
        assert torch.equal(dy, dy2)
    ctx.close()


@pytest.mark.parametrize("etype,dims,qorder", [(27, (6, 6, 6), 0), (8, (8, 8, 8), 3), (10, (2, 2, 2), 5)])
def test_cg_solve_matches_direct_solve_of_the_oracle_system(etype, dims, qorder):
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla

    F, O, m, s, ctx = _assembled(etype, dims, qorder)
    vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=qorder)
    rp, ci = s.sparsity()
    A = sp.csc_matrix(sp.csr_matrix((vo, ci.astype(np.int64), rp.astype(np.int64)), shape=(s.n_dofs, s.n_dofs)))
    u_ref = spla.spsolve(A, ro)
    u, its, rel = ctx.cg_solve(rel_tol=1e-13, max_its=5000)
    assert rel <= 1e-13 and 0 < its < 5000
    assert np.abs(u - u_ref).max() <= 1e-10 * np.abs(u_ref).max()
    # warm start from the solution converges at once
    _, its2, _ = ctx.cg_solve(rel_tol=1e-10, max_its=50, x0=u)
    assert its2 == 0
    ctx.close()


def test_ex4_solution_converges_to_the_exact_solution():
    """introduction_ex4 end to end on the device: assemble, solve, compare with u = cos(pi x/2) sin(pi y/2) cos(pi z/2)
    (exact_solution.C:40-47) at the nodes; Q2 elements: the nodal error drops by ~2^3..2^4 per halving of h."""
    errs = []
    for n in (3, 6):
        F, O, m, s, ctx = _assembled(27, (n, n, n))
        u, its, rel = ctx.cg_solve(rel_tol=1e-12, max_its=5000)
        xyz, _, _ = m.arrays()
        _, base, _ = s.dofs()
        exact = np.cos(.5 * np.pi * xyz[:, 0]) * np.sin(.5 * np.pi * xyz[:, 1]) * np.cos(.5 * np.pi * xyz[:, 2])
        errs.append(np.abs(u[base] - exact).max())
        ctx.close()
    assert errs[1] < 2e-4 and errs[0] / errs[1] > 6.0, errs


def test_vector_mult_host_variant():
    F, O, m, s, ctx = _assembled(8, (5, 4, 3), 3)
    import scipy.sparse as sp
    v, _ = ctx.get_values()
    rp, ci = s.sparsity()
    A = sp.csr_matrix((v, ci.astype(np.int64), rp.astype(np.int64)), shape=(s.n_dofs, s.n_dofs))
    x = np.linspace(-1.0, 2.0, s.n_dofs)
    y = ctx.vector_mult_host(x)
    assert np.abs(y - A @ x).max() <= 1e-13 * np.abs(A @ x).max()
    ctx.close()


@pytest.mark.parametrize("etype,dims,problem,nv", [(27, (4, 3, 3), "poisson", 1), (8, (5, 4, 3), "poisson", 1), (10, (2, 2, 2), "poisson", 1),
                                                   (27, (3, 2, 2), "elasticity", 3)])
def test_residual_without_the_global_matrix(etype, dims, problem, nv):
    """feasm_residual: R = K(u) u - F from the constrained element records, against A u - b of the oracle's assembled
    system for a random u (scale per row: |A| |u| + |b|)."""
    import scipy.sparse as sp

    from libmesh_b200 import capi as F
    from oracle import oracle as O

    elast = problem == "elasticity"
    m = O.Mesh(etype, *dims, bounds=(0, 1, 0, 0.3, 0, 0.1) if elast else (-1, 1, -1, 1, -1, 1))
    m.skew(1, 0.004 if elast else 0.02, 3)
    s = O.System(m, nv)
    s.add_dirichlet((4,), 1) if elast else s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    s.build_sparsity()
    xyz, conn, _ = m.arrays()
    cd, cv = s.constraints()
    ctx = F.Context(0)
    ctx.set_mesh(etype, xyz, conn)
    ctx.set_dofs(s.dofs()[0], s.n_dofs, n_vars=nv)
    ctx.set_constraints(cd, cv)
    ctx.build_pattern()
    if elast:
        vo, ro, _ = s.assemble(O.ELASTICITY_EX6)
        p = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS)
    else:
        q = 3 if etype == 8 else 5
        f_tab, _ = s.forcing_at_qps(q)
        vo, ro, _ = s.assemble(O.POISSON_EX4, qorder=q, forcing=O.FORCING_TABULATED, f_tab=f_tab)
        p = F.make_params(F.POISSON, q, F.FORCING_TABULATED, f_host=f_tab)
    rp, ci = s.sparsity()
    A = sp.csr_matrix((vo, ci.astype(np.int64), rp.astype(np.int64)), shape=(s.n_dofs, s.n_dofs))
    u = np.random.default_rng(7).uniform(-1, 1, s.n_dofs)
    want = A @ u - ro
    scale = abs(A) @ np.abs(u) + np.abs(ro)
    got = ctx.residual_host(p, u)
    assert np.max(np.abs(got - want) / np.maximum(scale, 1e-300)) < 1e-12
    # at the solution of the linear system the residual vanishes
    import scipy.sparse.linalg as spla
    us = spla.spsolve(A.tocsc(), ro)
    assert np.max(np.abs(ctx.residual_host(p, us)) / np.maximum(abs(A) @ np.abs(us) + np.abs(ro), 1e-300)) < 1e-10
