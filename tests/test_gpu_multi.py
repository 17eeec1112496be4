"""GPU parity of the multi-rank path with the ranks emulated inside ONE process on one GPU (one libfeasm context per
rank; B200_PROFILING.md forbids ranks that wait on each other as separate processes on one GPU).  The per-step
exchange is the real pack -> copy -> unpack sequence with a device-to-device copy standing in for ncclSend/Recv."""
import numpy as np
import pytest
import torch

from conftest import PARITY_TOL, csr_parity_error, vec_parity_error

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("split", [False, True], ids=["assemble", "begin_end"])
@pytest.mark.parametrize("world,etype,dims,scramble", [(2, 27, (4, 4, 4), 0), (3, 27, (4, 3, 3), 1), (4, 8, (6, 5, 4), 1), (2, 10, (2, 2, 2), 0)])
def test_emulated_ranks_match_global_oracle(world, etype, dims, scramble, split):
    """split: feasm_assemble_begin (element phase, non-owned rows, pack) / feasm_assemble_end (owned rows) around the
    exchange, the overlapped form bench.py --gpus N runs; otherwise feasm_assemble + feasm_pack_ghost."""
    from libmesh_b200 import capi as F
    from libmesh_b200 import multi as M
    from oracle import oracle as O

    m = O.Mesh(etype, *dims, bounds=(-1, 1, -1, 1, -1, 1))
    if scramble:
        pid = ((np.arange(m.n_elem) // 3) % world).astype(np.int32)
        m.set_partition(world, pid)
    else:
        m.partition_linear(world)
        pid = m.partition()[0]
    s = O.System(m)
    s.add_dirichlet()
    s.build_sparsity()
    xyz, conn, _ = m.arrays()
    ed, nb, fd = s.dofs()
    cd, cv = s.constraints()
    grp, gci = s.sparsity()
    qorder = 3 if etype == 8 else 5
    f_tab, _ = s.forcing_at_qps(qorder)
    gv, gb, _ = s.assemble(O.POISSON_EX4, qorder=qorder, forcing=O.FORCING_TABULATED, f_tab=f_tab)

    das, msgs, pats = [], [], []
    for r in range(world):
        prob = M.RankProblem(r, world, xyz, conn, nb, fd, pid, 1, cd, cv)
        ctx = F.Context(0)
        da = M.DistributedAssembly(ctx, prob, None)
        ctx.set_mesh(etype, prob.xyz, prob.conn)
        ctx.set_dofs(prob.elem_dofs, prob.n_rows, n_owned=prob.n_owned, global_ids=prob.global_ids)
        ctx.set_constraints(prob.con_rows, prob.con_vals)
        ctx.build_pattern()
        rp, ci = ctx.get_pattern()
        das.append(da); pats.append((rp, ci)); msgs.append(da.ghost_messages(rp, ci))
    for r, da in enumerate(das):
        da.finish_plan(*pats[r], {q: msgs[q][r] for q in range(world) if r in msgs[q]})
    for r, da in enumerate(das):
        params = F.make_params(F.POISSON, qorder, F.FORCING_TABULATED, f_host=np.ascontiguousarray(f_tab[da.p.elems]))
        if split:
            da.ctx.assemble_begin(params)
        else:
            da.ctx.assemble(params)
    # exchange: every rank packs, segments are copied peer to peer, every rank unpacks
    send = []
    for da in das:
        ptr, n = da.ctx.send_buffer() if split else da.ctx.pack_ghost()
        da.ctx.synchronize()
        send.append(torch.as_tensor(M.DevArray(ptr, max(n, 1)), device="cuda"))
    for r, da in enumerate(das):
        ptr, n = da.ctx.recv_buffer()
        rbuf = torch.as_tensor(M.DevArray(ptr, max(n, 1)), device="cuda")
        for i, q in enumerate(da.peers):
            r0, r1 = int(da.recv_ptr[i]), int(da.recv_ptr[i + 1])
            if r1 > r0:
                dq = das[q]
                j = dq.peers.index(r)
                rbuf[r0:r1].copy_(send[q][int(dq.send_ptr[j]):int(dq.send_ptr[j + 1])])
        torch.cuda.synchronize()
        if split:
            da.ctx.assemble_end()
        if da.peers:
            da.ctx.unpack_ghost()
        da.ctx.synchronize()
    for r, da in enumerate(das):
        v, b = da.owned_values()
        f0, n_own = da.p.f0, da.p.n_owned
        lo, hi = int(grp[f0]), int(grp[f0 + n_own])
        assert np.array_equal(da.row_ptr[: n_own + 1].astype(np.int64), grp[f0:f0 + n_own + 1].astype(np.int64) - lo)
        assert np.array_equal(da.col_idx[: hi - lo], gci[lo:hi])
        assert csr_parity_error(v, gv[lo:hi], da.row_ptr[: n_own + 1]) < PARITY_TOL
        assert vec_parity_error(b, gb[f0:f0 + n_own]) < PARITY_TOL
        da.ctx.close()
