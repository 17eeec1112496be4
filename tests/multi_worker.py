"""Worker of tests/test_multi_host.py: one rank of a world_size-N gloo job on the CPU.

Exercises the HOST logic of libmesh_b200.multi (localisation, ghost-row pattern shipping/merging, slot plan, the
per-step exchange order) with the kernels replaced by the oracle's element matrices scattered with numpy.  The
checker is the oracle's global assembly on the same partition.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from libmesh_b200 import multi as M  # noqa: E402
from oracle import oracle as O  # noqa: E402


class HostCtx:
    """Stands in for capi.Context: records what DistributedAssembly hands to libfeasm; the local pattern is built
    with numpy instead of feasm_build_pattern."""

    def set_mesh(self, etype, xyz, conn): self.conn = conn
    def set_dofs(self, ed, n_rows, n_vars=1, n_owned=None, global_ids=None):
        self.ed, self.n_rows, self.n_owned, self.gid = ed.astype(np.int64), n_rows, n_owned, global_ids.astype(np.int64)
    def set_constraints(self, r, v): self.con = (r, v)
    def build_pattern(self):
        nd = self.ed.shape[1]
        rows = np.repeat(self.ed, nd, axis=1).ravel()
        cols = self.gid[np.tile(self.ed, (1, nd)).ravel()]
        key = np.unique((rows << 32) | cols)
        r, c = key >> 32, key & 0xffffffff
        rp = np.zeros(self.n_rows + 1, dtype=np.uint64)
        np.add.at(rp, r + 1, 1)
        self.rp, self.ci = np.cumsum(rp).astype(np.uint64), c.astype(np.uint32)
    def get_pattern(self): return self.rp, self.ci
    def set_pattern(self, rp, ci): self.rp, self.ci = np.asarray(rp, dtype=np.uint64), ci
    def set_exchange_plan(self, *a): self.plan = a
    def prepare(self): pass
    def halo_columns(self, n_global, first):  # feasm_halo_columns: columns of the owned rows outside the owned range + ghost-row ids
        rp = self.rp.astype(np.int64)
        cols = self.ci[: rp[self.n_owned]].astype(np.int64)
        out = np.unique(np.concatenate([cols[(cols < first) | (cols >= first + self.n_owned)], self.gid[self.n_owned:]]))
        return out.astype(np.uint32)
    def set_halo_plan(self, *a): self.halo = a


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    etype, dims, n_vars, scramble = int(sys.argv[1]), tuple(int(x) for x in sys.argv[2:5]), int(sys.argv[5]), int(sys.argv[6])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    m = O.Mesh(etype, *dims, bounds=(-1, 1, -1, 1, -1, 1))
    if scramble:  # a non-slab partition: interleaved blocks, so ranks have several neighbours and ragged interfaces
        pid = ((np.arange(m.n_elem) // 3) % world).astype(np.int32)
        m.set_partition(world, pid)
    else:
        m.partition_linear(world)
        pid = m.partition()[0]
        assert np.array_equal(pid, M.linear_partition(m.n_elem, world))
    problem = O.ELASTICITY_EX6 if n_vars == 3 else O.POISSON_EX4
    s = O.System(m, n_vars)
    s.add_dirichlet((4,) if n_vars == 3 else (0, 1, 2, 3, 4, 5), 1 if n_vars == 3 else 0)
    s.build_sparsity()
    xyz, conn, _ = m.arrays()
    ed, nb, fd = s.dofs()
    cd, cv = s.constraints()
    grp, gci = s.sparsity()
    kw = {}
    if n_vars == 1:
        kw = dict(forcing=O.FORCING_TABULATED, f_tab=s.forcing_at_qps()[0])
    gv, gb, _ = s.assemble(problem, **kw)

    prob = M.RankProblem(rank, world, xyz, conn, nb, fd, pid, n_vars, cd, cv)
    assert np.array_equal(prob.global_ids[prob.elem_dofs], ed[prob.elems]), "localised dofs do not map back"
    ctx = HostCtx()
    da = M.DistributedAssembly(ctx, prob, dist)
    da.setup()
    # merged pattern of the owned rows == the global Graph rows [first_dof, end_dof)
    f0, n_own = prob.f0, prob.n_owned
    lo, hi = int(grp[f0]), int(grp[f0 + n_own])
    assert np.array_equal(ctx.rp[: n_own + 1].astype(np.int64), grp[f0:f0 + n_own + 1].astype(np.int64) - lo), "owned row_ptr differs"
    assert np.array_equal(ctx.ci[: hi - lo], gci[lo:hi]), "owned col_idx differs"
    # local assembly: constrained element matrices of the local elements, scattered in element order
    Ke, Fe = s.element_matrices(problem, constrained=True, **kw)
    rp = ctx.rp.astype(np.int64)
    nnz = int(rp[-1])
    sys_vec = np.zeros(nnz + prob.n_rows)  # values | rhs: the slot space of the exchange plan
    for le, e in enumerate(prob.elems):
        rows = prob.elem_dofs[le].astype(np.int64)
        gcols = prob.global_ids[rows]
        for i, r in enumerate(rows):
            seg = ctx.ci[rp[r]:rp[r + 1]]
            sys_vec[rp[r] + np.searchsorted(seg, gcols)] += Ke[e, i]
            sys_vec[nnz + r] += Fe[e, i]
    # exchange exactly as feasm_pack_ghost / unpack_ghost do, in the plan's peer order
    peers, sp, ss, rpt, rs = ctx.plan
    sbuf = torch.from_numpy(sys_vec[ss.astype(np.int64)].copy())
    rbuf = torch.zeros(int(rpt[-1]), dtype=torch.float64)
    ops = []
    for i, q in enumerate(peers):
        if rpt[i + 1] > rpt[i]: ops.append(dist.P2POp(dist.irecv, rbuf[int(rpt[i]):int(rpt[i + 1])], int(q)))
    for i, q in enumerate(peers):
        if sp[i + 1] > sp[i]: ops.append(dist.P2POp(dist.isend, sbuf[int(sp[i]):int(sp[i + 1])], int(q)))
    if ops:
        for w in dist.batch_isend_irecv(ops): w.wait()
    for i, q in enumerate(peers):
        a, b = int(rpt[i]), int(rpt[i + 1])
        np.add.at(sys_vec, rs[a:b].astype(np.int64), rbuf[a:b].numpy())
    own_v, own_b = sys_vec[: int(rp[n_own])], sys_vec[nnz:nnz + n_own]
    ref_v, ref_b = gv[lo:hi], gb[f0:f0 + n_own]
    sv = max(np.abs(ref_v).max(), 1e-300)
    assert np.abs(own_v - ref_v).max() <= 1e-13 * sv, ("values", np.abs(own_v - ref_v).max())
    assert np.abs(own_b - ref_b).max() <= 1e-13 * max(np.abs(ref_b).max(), 1e-300), "rhs"
    # ---- the halo of a distributed vector with the requests carried by torch.distributed (DistributedAssembly.setup_halo):
    # y = A x on the owned rows after fetching the entries of x other ranks own, against the global product
    da.setup_halo()
    n_global, first, hp, hsp, hsr, hrp, hrg = ctx.halo
    assert n_global == s.n_dofs and first == f0
    x = np.cos(0.37 * np.arange(s.n_dofs)) + 0.1
    x_own = x[f0:f0 + n_own]
    hs = torch.from_numpy(x_own[hsr.astype(np.int64)].copy())          # feasm_halo_pack
    hr = torch.zeros(int(hrp[-1]), dtype=torch.float64)
    ops = []
    for i, q in enumerate(hp):
        if hrp[i + 1] > hrp[i]: ops.append(dist.P2POp(dist.irecv, hr[int(hrp[i]):int(hrp[i + 1])], int(q)))
    for i, q in enumerate(hp):
        if hsp[i + 1] > hsp[i]: ops.append(dist.P2POp(dist.isend, hs[int(hsp[i]):int(hsp[i + 1])], int(q)))
    if ops:
        for w in dist.batch_isend_irecv(ops): w.wait()
    x_full = np.full(s.n_dofs, np.nan)                                   # a column nobody sent would poison the product
    x_full[f0:f0 + n_own] = x_own
    x_full[hrg.astype(np.int64)] = hr.numpy()                           # feasm_halo_unpack
    assert np.array_equal(x_full[hrg.astype(np.int64)], x[hrg.astype(np.int64)]), "halo entries are not the owners' values"
    assert not np.isnan(x_full[prob.ghosts]).any(), "ghost rows are part of the halo (feasm_residual needs them)"
    y = np.array([own_v[rp[r]:rp[r + 1]] @ x_full[ctx.ci[rp[r]:rp[r + 1]].astype(np.int64)] for r in range(n_own)])
    import scipy.sparse as sps
    A = sps.csr_matrix((gv, gci.astype(np.int64), grp.astype(np.int64)), shape=(s.n_dofs, s.n_dofs))
    y_ref = (A @ x)[f0:f0 + n_own]
    assert not np.isnan(y).any() and np.abs(y - y_ref).max() <= 1e-12 * np.abs(A @ x).max(), "distributed product"
    total = torch.tensor([float(n_own)])
    dist.all_reduce(total)
    assert int(total.item()) == s.n_dofs, "owned ranges do not tile the dof space"
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank} ok: {len(prob.elems)} elems, {n_own} owned rows, {len(prob.ghosts)} ghost rows, peers {list(peers)}")


if __name__ == "__main__":
    main()
