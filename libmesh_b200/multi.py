"""Multi-GPU assembly: one process per GPU, mesh split by element owner, ghost rows summed on their owner.

Mirrors libMesh's parallel assembly model: every rank assembles its ``active_local`` elements; a row
belongs to the rank that owns its dof (dof ranges are contiguous per rank, src/base/dof_map.C:999-1015; node owner =
min rank of the touching elements, src/geom/node.C:99-102); contributions to rows of another rank are summed by the
owner at ``close()`` (src/numerics/distributed_vector.C:399-470, PETSc's MatAssemblyBegin/End stash).  Here:

  setup     each rank localises its elements (owned rows first, then ghost rows in ascending global id), builds the
            local sparsity pattern on its GPU, ships the column lists of its ghost rows to their owners (the mirror of
            SparsityPattern::Build::parallel_sync, src/base/sparsity_pattern.C:446-575), merges what it receives into
            its owned rows and fixes a send/receive slot plan;
  per step  feasm_assemble -> feasm_pack_ghost -> one grouped send/recv per neighbour (ncclSend/ncclRecv through
            torch.distributed.batch_isend_irecv on the library's stream) -> feasm_unpack_ghost, peers added in fixed
            order, so the result is bitwise reproducible.

Localisation (RankProblem) is numpy; the ghost-row messages, the pattern merge and the slot plan are libfeasm's
(feasm_planner_* / feasm_build_exchange, csrc/feasm_plan.cpp) — this module only carries the planner's messages over
torch.distributed when the context has no communicator of its own (the CPU tests, gloo).
"""
from __future__ import annotations

import os

import numpy as np


# ---- partitioning and localisation (pure numpy) ------------------------------------------------------------
def linear_partition(n_elem: int, n_parts: int) -> np.ndarray:
    """Contiguous element-id slabs = LinearPartitioner (src/partitioning/linear_partitioner.C:60-75): blksize =
    n_elem / n_parts, element e goes to part e / blksize, and the remainder elements fall back to processor 0."""
    pid = np.zeros(n_elem, dtype=np.int32)
    if n_parts > 1:
        blk = max(n_elem // n_parts, 1)
        p = np.arange(n_elem, dtype=np.int64) // blk
        pid = np.where(p < n_parts, p, 0).astype(np.int32)
    return pid


class RankProblem:
    """Everything one rank hands to libfeasm, in local numbering."""

    def __init__(self, rank, n_parts, xyz, conn, node_base, first_dof, elem_pid, n_vars=1, con_dofs=None, con_vals=None):
        self.rank, self.n_parts, self.n_vars = rank, n_parts, n_vars
        self.first_dof = np.asarray(first_dof, dtype=np.int64)
        f0, f1 = int(self.first_dof[rank]), int(self.first_dof[rank + 1])
        self.f0, self.n_owned = f0, f1 - f0
        self.elems = np.nonzero(np.asarray(elem_pid) == rank)[0]
        gconn = conn[self.elems].astype(np.int64)
        self.nodes, lconn = np.unique(gconn, return_inverse=True)  # local node ids = rank in the sorted global ids
        self.conn = lconn.reshape(gconn.shape).astype(np.uint32)
        self.xyz = np.ascontiguousarray(xyz[self.nodes])
        nb = node_base.astype(np.int64)[gconn]
        gdofs = np.concatenate([nb + v for v in range(n_vars)], axis=1)  # variable-major (dof_map.C:2280-2308)
        touched = np.unique(gdofs)
        owned = (touched >= f0) & (touched < f1)
        if int(owned.sum()) != self.n_owned:
            raise ValueError("a rank must touch every dof it owns (node owner = min touching rank)")
        self.ghosts = touched[~owned]  # ascending global id
        self.n_rows = self.n_owned + len(self.ghosts)
        self.global_ids = np.concatenate([np.arange(f0, f1, dtype=np.int64), self.ghosts]).astype(np.uint32)
        self.elem_dofs = self.to_local(gdofs).astype(np.uint32)
        self.ghost_owner = (np.searchsorted(self.first_dof, self.ghosts, side="right") - 1).astype(np.int32)
        if con_dofs is not None and len(con_dofs):
            cd = np.asarray(con_dofs, dtype=np.int64)
            here = np.isin(cd, touched)
            self.con_rows = self.to_local(cd[here]).astype(np.uint32)
            self.con_vals = np.asarray(con_vals, dtype=np.float64)[here]
        else:
            self.con_rows, self.con_vals = np.zeros(0, dtype=np.uint32), np.zeros(0)

    def to_local(self, g):
        g = np.asarray(g, dtype=np.int64)
        own = (g >= self.f0) & (g < self.f0 + self.n_owned)
        out = np.where(own, g - self.f0, 0)
        if (~own).any():
            k = np.searchsorted(self.ghosts, g[~own])
            if (k >= len(self.ghosts)).any() or (self.ghosts[np.minimum(k, len(self.ghosts) - 1)] != g[~own]).any():
                raise ValueError("dof is neither owned nor a ghost of this rank")
            out[~own] = self.n_owned + k
        return out

    @property
    def peers(self):
        """Ranks this rank sends ghost rows to (ascending)."""
        return np.unique(self.ghost_owner)


# ---- setup exchange and per-step exchange (torch.distributed) -------------------------------------------------
class DevArray:
    """A raw device pointer exposed through __cuda_array_interface__ so torch can wrap it without a copy."""

    def __init__(self, ptr, n, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def _exchange_arrays(dist, group, rank, send: dict):
    """Variable-size uint32 array exchange on the host: send[p] -> p for every p in send; peers are symmetric for
    sizes (a rank that receives nothing gets an empty array).  Returns {q: array}."""
    import torch

    world = dist.get_world_size(group)
    sizes = torch.zeros(world, dtype=torch.int64)
    for p, a in send.items():
        sizes[p] = len(a)
    all_sizes = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(all_sizes, sizes, group=group)
    ops, bufs, keep = [], {}, []
    for q in range(world):
        n = int(all_sizes[q][rank])
        if q != rank and n:
            bufs[q] = torch.empty(n, dtype=torch.int64)
            ops.append(dist.P2POp(dist.irecv, bufs[q], q, group=group))
    for p, a in send.items():
        if p != rank and len(a):
            t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.int64))
            keep.append(t)
            ops.append(dist.P2POp(dist.isend, t, p, group=group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return {q: b.numpy() for q, b in bufs.items()}


class DistributedAssembly:
    """One rank of a multi-GPU assembly.  `ctx` is a libmesh_b200.capi.Context (or, in CPU tests, an object with the
    same pattern / plan methods); `host_group` is a gloo group for the setup exchange, `dev_group` the group the
    per-step exchange runs on (NCCL on GPUs)."""

    def __init__(self, ctx, prob: RankProblem, dist, host_group=None, dev_group=None):
        self.ctx, self.p, self.dist = ctx, prob, dist
        self.host_group, self.dev_group = host_group, dev_group
        self.rank = prob.rank
        self._pattern = None

    # -- setup ---------------------------------------------------------------------------------------------
    def setup(self, constrain=True):
        p, ctx = self.p, self.ctx
        ctx.set_mesh(p.conn.shape[1], p.xyz, p.conn)
        ctx.set_dofs(p.elem_dofs, p.n_rows, n_vars=p.n_vars, n_owned=p.n_owned, global_ids=p.global_ids)
        if constrain:
            ctx.set_constraints(p.con_rows, p.con_vals)
        ctx.build_pattern()
        if getattr(ctx, "has_comm", False):
            # everything below the C ABI: messages over the context's NCCL communicator, merge, plan, scatter maps and the
            # interface prefix (feasm_build_exchange)
            ctx.build_exchange(p.first_dof)
            self.peers, self.send_ptr, self.send_slots, self.recv_ptr, self.recv_slots = ctx.exchange_plan()
            self.peers = [int(q) for q in self.peers]
            self._pattern = None
            return
        rp, ci = ctx.get_pattern()
        self.merge_and_plan(rp, ci)
        # interface elements first: the leading local elements that cover every element touching a ghost row (with
        # LinearPartitioner slabs: the first element layer).  Worth it when that prefix is a small part of the rank's elements.
        touch = np.nonzero((p.elem_dofs >= p.n_owned).any(axis=1))[0]
        prefix = int(touch[-1]) + 1 if len(touch) else 0
        self.iface_prefix = prefix if 0 < prefix <= len(p.elems) // 3 and os.environ.get("FEASM_IFACE", "1") != "0" else 0
        if hasattr(ctx, "set_interface_prefix"):
            ctx.set_interface_prefix(self.iface_prefix)

    @property
    def row_ptr(self):
        if self._pattern is None:
            self._pattern = self.ctx.get_pattern()
        return self._pattern[0]

    @property
    def col_idx(self):
        if self._pattern is None:
            self._pattern = self.ctx.get_pattern()
        return self._pattern[1]

    def merge_and_plan(self, rp, ci):
        """Ship ghost-row column lists to their owners, merge, and derive the slot plan, with the messages carried by
        torch.distributed.  Also usable stand-alone with a host-built local pattern (CPU tests)."""
        send = self.ghost_messages(rp, ci)
        got = _exchange_arrays(self.dist, self.host_group, self.rank, send)
        self.finish_plan(rp, ci, got)

    def ghost_messages(self, rp, ci):
        """{owner rank: uint32 words [n_rows, global row ids, row lengths, columns]} for the ghost rows of this rank
        (feasm_planner_create / feasm_planner_message)."""
        from . import capi as F

        p = self.p
        self._planner = F.Planner(p.rank, p.n_parts, p.first_dof, p.n_owned, p.n_rows, p.global_ids, rp, ci)
        return self._planner.messages()

    def finish_plan(self, rp, ci, got):
        """Hand the received messages to the planner, then the merged pattern and the slot plan to the context."""
        ctx, pl = self.ctx, self._planner
        for q in sorted(got):
            pl.receive(q, np.asarray(got[q], dtype=np.uint32))
        o = pl.finish()
        if hasattr(ctx, "apply_plan"):
            ctx.apply_plan(pl)  # feasm_set_pattern (if the pattern grew) + feasm_set_exchange_plan + feasm_prepare
        else:  # host stand-in of the CPU tests
            if o["merged"]:
                ctx.set_pattern(o["row_ptr"], o["col_idx"])
            ctx.set_exchange_plan(o["peers"], o["send_ptr"], o["send_slots"], o["recv_ptr"], o["recv_slots"])
            ctx.prepare()
        self._pattern = (o["row_ptr"], o["col_idx"])
        self.peers = [int(q) for q in o["peers"]]
        self.send_ptr, self.recv_ptr = o["send_ptr"], o["recv_ptr"]
        self.send_slots, self.recv_slots = o["send_slots"], o["recv_slots"]
        pl.close()

    # -- the assembled matrix used on the device: halo of a distributed vector -------------------------------------------
    def setup_halo(self):
        """After setup(): which entries of a distributed vector this rank needs from whom (what libMesh's send_list /
        NumericVector::localize provide).  With a communicator inside the context it is feasm_build_halo; otherwise the
        needed global ids (feasm_halo_columns) are split by owner here, the request lists travel over torch.distributed
        and the routes go back in through feasm_set_halo_plan — the route of hosts that carry messages themselves."""
        p, ctx = self.p, self.ctx
        if getattr(ctx, "has_comm", False):
            ctx.build_halo(p.first_dof)
            self.halo_peers, self.halo_send_ptr, self.halo_send_rows, self.halo_recv_ptr, self.halo_recv_gids = ctx.halo_plan()
            self.halo_peers = [int(q) for q in self.halo_peers]
            return
        n_global = int(p.first_dof[-1])
        need = np.asarray(ctx.halo_columns(n_global, p.f0), dtype=np.int64)  # ascending global ids, none of them owned
        owner = np.searchsorted(p.first_dof, need, side="right") - 1
        requests = {int(q): need[owner == q] for q in np.unique(owner)}
        asked = _exchange_arrays(self.dist, self.host_group, self.rank, requests)  # what the peers ask of this rank
        peers = sorted(set(requests) | set(int(q) for q in asked))
        send_ptr, recv_ptr, send_rows, recv_gids = [0], [0], [], []
        for q in peers:
            rows = np.asarray(asked.get(q, np.zeros(0)), dtype=np.int64) - p.f0
            if len(rows) and (rows.min() < 0 or rows.max() >= p.n_owned):
                raise ValueError(f"rank {q} asked rank {self.rank} for an entry it does not own")
            send_rows.append(rows)
            send_ptr.append(send_ptr[-1] + len(rows))
            got = requests.get(q, np.zeros(0, dtype=np.int64))
            recv_gids.append(got)
            recv_ptr.append(recv_ptr[-1] + len(got))
        cat = lambda parts: (np.concatenate(parts) if parts else np.zeros(0)).astype(np.uint32)
        self.halo_peers, self.halo_send_ptr, self.halo_recv_ptr = peers, np.asarray(send_ptr, dtype=np.uint64), np.asarray(recv_ptr, dtype=np.uint64)
        self.halo_send_rows, self.halo_recv_gids = cat(send_rows), cat(recv_gids)
        ctx.set_halo_plan(n_global, p.f0, peers, self.halo_send_ptr, self.halo_send_rows, self.halo_recv_ptr, self.halo_recv_gids)

    # -- per step --------------------------------------------------------------------------------------------
    def _exchange_ops(self):
        """The grouped send/recv of one step (built once: buffers and slot ranges are fixed at setup)."""
        if getattr(self, "_ops", None) is None:
            import torch

            ctx, dist = self.ctx, self.dist
            sptr, ns = ctx.send_buffer()
            rptr, nr = ctx.recv_buffer()
            sbuf = torch.as_tensor(DevArray(sptr, max(ns, 1)), device="cuda")
            rbuf = torch.as_tensor(DevArray(rptr, max(nr, 1)), device="cuda")
            ops = []
            for i, q in enumerate(self.peers):
                r0, r1 = int(self.recv_ptr[i]), int(self.recv_ptr[i + 1])
                if r1 > r0:
                    ops.append(dist.P2POp(dist.irecv, rbuf[r0:r1], q, group=self.dev_group))
            for i, q in enumerate(self.peers):
                s0, s1 = int(self.send_ptr[i]), int(self.send_ptr[i + 1])
                if s1 > s0:
                    ops.append(dist.P2POp(dist.isend, sbuf[s0:s1], q, group=self.dev_group))
            self._ops, self._lib_stream = ops, torch.cuda.ExternalStream(ctx.stream())
        return self._ops

    def start_exchange(self):
        """Issue the send/recv group behind what is on the library's stream now (the packed send buffer).  NCCL runs it
        on its own stream, so kernels launched on the library's stream afterwards overlap the transfer.  With a
        communicator inside the context (feasm_comm_init) the library issues the group itself (feasm_exchange_begin);
        otherwise torch.distributed does (the route of the CPU tests and of hosts that own the process group)."""
        if not self.peers:
            return []
        if getattr(self.ctx, "has_comm", False):
            self.ctx.exchange_begin()
            return None
        import torch

        ops = self._exchange_ops()
        with torch.cuda.stream(self._lib_stream):
            return self.dist.batch_isend_irecv(ops)

    def finish_exchange(self, works):
        """Make the library's stream wait for the transfer, then add the received rows in fixed peer order."""
        if not self.peers:
            return
        if works is None:
            self.ctx.exchange_end()  # waits for the transfer on the library's stream, then feasm_unpack_ghost
            return
        import torch

        with torch.cuda.stream(self._lib_stream):
            for w in works:
                w.wait()  # stream-ordered for NCCL
        self.ctx.unpack_ghost()

    def exchange(self):
        """pack -> grouped send/recv with every neighbour -> unpack (fixed peer order), after a plain ctx.assemble."""
        self.ctx.pack_ghost()
        self.finish_exchange(self.start_exchange())

    def assemble(self, params, sync=True, overlap=True):
        """One assembly step of this rank.  overlap: rows owned by other ranks are reduced and sent first, the owned rows
        are reduced while the transfer is in flight (feasm_assemble_begin / _end); the result is the same either way."""
        if overlap:
            self.ctx.assemble_begin(params)
            works = self.start_exchange()
            self.ctx.assemble_end()
            self.finish_exchange(works)
        else:
            self.ctx.assemble(params, sync=False)
            self.exchange()
        if sync:
            self.ctx.synchronize()

    def unpack_launches(self):
        """Kernel launches of one feasm_unpack_ghost (for the launch count bench.py reports)."""
        return 1 if self.peers else 0

    def owned_values(self):
        """(values of the owned rows, owned rhs) on the host."""
        v, b = self.ctx.get_values()
        return v[: int(self.row_ptr[self.p.n_owned])], b[: self.p.n_owned]
