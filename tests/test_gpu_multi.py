"""GPU parity of the multi-rank path with the ranks emulated inside ONE process on one GPU (one libfeasm context per
rank; B200_PROFILING.md forbids ranks that wait on each other as separate processes on one GPU).  The per-step
exchange is the real pack -> copy -> unpack sequence with a device-to-device copy standing in for ncclSend/Recv."""
import numpy as np
import pytest
import torch

from conftest import PARITY_TOL, csr_parity_error, vec_parity_error

pytestmark = pytest.mark.gpu


def _assemble_emulated(world, etype, dims, scramble, split):
    """Every rank of a `world`-rank job as one context on this GPU, assembled and exchanged; returns the rank objects and
    the oracle's global system."""
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
    return das, (grp, gci, gv, gb, s.n_dofs)


@pytest.mark.parametrize("split", [False, True], ids=["assemble", "begin_end"])
@pytest.mark.parametrize("world,etype,dims,scramble", [(2, 27, (4, 4, 4), 0), (3, 27, (4, 3, 3), 1), (4, 8, (6, 5, 4), 1), (2, 10, (2, 2, 2), 0)])
def test_emulated_ranks_match_global_oracle(world, etype, dims, scramble, split):
    """split: feasm_assemble_begin (element phase, non-owned rows, pack) / feasm_assemble_end (owned rows) around the
    exchange, the overlapped form bench.py --gpus N runs; otherwise feasm_assemble + feasm_pack_ghost."""
    das, (grp, gci, gv, gb, _) = _assemble_emulated(world, etype, dims, scramble, split)
    for r, da in enumerate(das):
        v, b = da.owned_values()
        f0, n_own = da.p.f0, da.p.n_owned
        lo, hi = int(grp[f0]), int(grp[f0 + n_own])
        assert np.array_equal(da.row_ptr[: n_own + 1].astype(np.int64), grp[f0:f0 + n_own + 1].astype(np.int64) - lo)
        assert np.array_equal(da.col_idx[: hi - lo], gci[lo:hi])
        assert csr_parity_error(v, gv[lo:hi], da.row_ptr[: n_own + 1]) < PARITY_TOL
        assert vec_parity_error(b, gb[f0:f0 + n_own]) < PARITY_TOL
        da.ctx.close()


@pytest.mark.parametrize("world,etype,dims,scramble", [(2, 27, (4, 4, 4), 0), (3, 27, (4, 3, 3), 1), (4, 8, (6, 5, 4), 1)])
def test_emulated_ranks_vector_mult_with_halo(world, etype, dims, scramble):
    """SparseMatrix::vector_mult on the distributed matrix: feasm_halo_columns names the entries of x a rank needs from
    others, the requests are routed by owner (here by hand, as a host with its own messaging would), feasm_halo_pack /
    copy / feasm_halo_unpack stand in for the NCCL exchange, and the owned rows of A x equal scipy's product with the
    oracle's global matrix."""
    import scipy.sparse as sp
    from libmesh_b200 import multi as M

    das, (grp, gci, gv, gb, n_dofs) = _assemble_emulated(world, etype, dims, scramble, True)
    A = sp.csr_matrix((gv, gci.astype(np.int64), grp.astype(np.int64)), shape=(n_dofs, n_dofs))
    x = np.cos(0.37 * np.arange(n_dofs)) + 0.1
    y_ref = A @ x
    first = das[0].p.first_dof
    need = [da.ctx.halo_columns(n_dofs, da.p.f0) for da in das]
    for r, da in enumerate(das):
        own = (need[r] >= da.p.f0) & (need[r] < da.p.f0 + da.p.n_owned)
        assert not own.any() and np.all(np.diff(need[r].astype(np.int64)) > 0)
        assert np.isin(da.p.ghosts, need[r]).all()  # the ghost rows are part of the halo (feasm_residual needs them)
    owner = [np.searchsorted(first, n.astype(np.int64), side="right") - 1 for n in need]
    for r, da in enumerate(das):
        peers = sorted(set(int(q) for q in owner[r]) | {q for q in range(world) if q != r and (owner[q] == r).any()})
        sp_, sr, rp_, rg = [0], [], [0], []
        for q in peers:
            wanted = need[q][owner[q] == r].astype(np.int64) - da.p.f0  # what q asks of r, as r's local rows
            sr.append(wanted); sp_.append(sp_[-1] + len(wanted))
            got = need[r][owner[r] == q]
            rg.append(got); rp_.append(rp_[-1] + len(got))
        cat = lambda parts, dt: np.concatenate(parts).astype(dt) if parts else np.zeros(0, dt)
        da.ctx.set_halo_plan(n_dofs, da.p.f0, peers, sp_, cat(sr, np.uint32), rp_, cat(rg, np.uint32))
        da.halo_peers, da.halo_sp, da.halo_rp = peers, sp_, rp_
    L = das[0].ctx.L
    import ctypes as C
    dx, dy, send = [], [], []
    for da in das:
        n = da.p.n_owned
        xo = np.ascontiguousarray(x[da.p.f0:da.p.f0 + n])
        d_x, d_y = L.feasm_vec_alloc(da.ctx.h, n), L.feasm_vec_alloc(da.ctx.h, n)
        assert L.feasm_vec_copy(da.ctx.h, 0, C.c_void_p(d_x), xo.ctypes.data_as(C.c_void_p), n) == 0
        dx.append(d_x); dy.append(d_y)
        ptr, ns = da.ctx.halo_pack(d_x)
        da.ctx.synchronize()
        send.append(torch.as_tensor(M.DevArray(ptr, max(ns, 1)), device="cuda"))
    for r, da in enumerate(das):
        ptr, nr = da.ctx.halo_recv_buffer()
        rbuf = torch.as_tensor(M.DevArray(ptr, max(nr, 1)), device="cuda")
        for i, q in enumerate(da.halo_peers):
            r0, r1 = da.halo_rp[i], da.halo_rp[i + 1]
            if r1 > r0:
                dq = das[q]
                j = dq.halo_peers.index(r)
                rbuf[r0:r1].copy_(send[q][dq.halo_sp[j]:dq.halo_sp[j + 1]])
        torch.cuda.synchronize()
        da.ctx.halo_unpack()
        da.ctx.vector_mult(dx[r], dy[r])
        n = da.p.n_owned
        y = np.empty(n)
        assert L.feasm_vec_copy(da.ctx.h, 1, y.ctypes.data_as(C.c_void_p), C.c_void_p(dy[r]), n) == 0
        ref = y_ref[da.p.f0:da.p.f0 + n]
        assert np.abs(y - ref).max() <= 1e-13 * np.abs(y_ref).max(), (r, np.abs(y - ref).max())
    for r, da in enumerate(das):
        L.feasm_vec_free(da.ctx.h, C.c_void_p(dx[r])); L.feasm_vec_free(da.ctx.h, C.c_void_p(dy[r]))
        da.ctx.close()
