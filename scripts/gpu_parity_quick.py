"""Quick GPU parity probe (development aid; the real tests live in tests/)."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from oracle import oracle as O
from libmesh_b200 import capi as F

def relerr(a, b):
    scale = np.maximum(np.abs(a), np.abs(b))
    return np.abs(a - b), scale

def run(etype, n, skew=None, qorder=0):
    m = O.Mesh(etype, n, n, n, bounds=(-1, 1, -1, 1, -1, 1))
    if skew: m.skew(1, skew, 12345)
    s = O.System(m); s.add_dirichlet(); s.build_sparsity()
    xyz, conn, mask = m.arrays(); ed, nb, fd = s.dofs(); cd, cv = s.constraints(); rp, ci = s.sparsity()
    ctx = F.Context(0)
    ctx.set_mesh(etype, xyz, conn); ctx.set_dofs(ed, s.n_dofs); ctx.set_constraints(cd, cv)
    nnz = ctx.build_pattern(); rp2, ci2 = ctx.get_pattern()
    print(etype, n, 'pattern exact:', nnz == s.nnz and np.array_equal(rp, rp2) and np.array_equal(ci, ci2))
    # element level, tabulated forcing
    f_tab, _ = s.forcing_at_qps(qorder)
    p = F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=f_tab)
    ne = min(m.n_elem, 512)
    Ke, Fe = ctx.element_matrices(p, 0, ne, constrained=False)
    Ko, Fo = s.element_matrices(O.POISSON_EX4, 0, ne, constrained=False, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)
    rown = np.abs(Ko).max(axis=2, keepdims=True)
    print('  Ke unconstrained max |d|/rowmax', (np.abs(Ke - Ko) / rown).max(), ' Fe rel', (np.abs(Fe - Fo)).max() / np.abs(Fo).max())
    t = time.time(); vo, ro, ph = s.assemble(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab); to = time.time() - t
    ctx.assemble(p); v, r = ctx.get_values()
    d, sc = relerr(v, vo)
    rowmax = np.maximum.reduceat(np.abs(vo), rp[:-1].astype(np.int64))
    rl = np.diff(rp.astype(np.int64)); rm = np.repeat(rowmax, rl)
    print('  values: max |d|/max(|a|,|b|) over nz', (d[sc > 0] / sc[sc > 0]).max(), ' max |d|/rowmax', (d / rm).max())
    print('  rhs: max rel', (np.abs(r - ro) / np.maximum(np.abs(ro), 1e-300)).max(), 'max abs', np.abs(r - ro).max(), 'scale', np.abs(ro).max())
    # device FD forcing
    p2 = F.make_params(F.POISSON, qorder, F.FORCING_EX4_FD)
    ctx.assemble(p2); v2, r2 = ctx.get_values()
    vo2, ro2, _ = s.assemble(O.POISSON_EX4, qorder=qorder)
    print('  FD forcing rhs: max abs diff', np.abs(r2 - ro2).max(), 'rel to max', np.abs(r2 - ro2).max() / np.abs(ro2).max())
    print('  timing ms', ctx.last_timing(), 'oracle s', to, 'elems', m.n_elem)
    ctx.close()

run(O.HEX27, 6)
run(O.HEX27, 5, skew=0.02)
run(O.HEX8, 8, qorder=3)
run(O.HEX8, 6)
run(O.TET10, 3)
run(O.TET10, 2, skew=0.02)
