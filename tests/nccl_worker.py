"""Worker of tests/test_gpu_config_size.py::test_real_nccl_two_ranks: one rank of a REAL multi-GPU job (NCCL, one
process per GPU, launched by torch.distributed.run).

What only the real run has — NCCL's own stream against the library's stream, feasm_assemble_end under the transfer, the
next step's pack into the same send buffer, feasm_unpack_ghost — is exercised by several back-to-back asynchronous
steps; the owned rows of every rank are then compared with the oracle's global system (1e-12 per entry), and the
overlapped and sequential forms with each other (bitwise).  With the communicator inside the library the distributed
matrix is then USED: feasm_build_halo, feasm_vector_mult against scipy's product with the oracle's global matrix,
feasm_residual against A u - b, feasm_cg_solve against a direct solve (all ranks must report the same iteration count).
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from libmesh_b200 import capi as F  # noqa: E402
from libmesh_b200 import multi as M  # noqa: E402
from oracle import oracle as O  # noqa: E402
from oracle.parity import TOL, csr_compare, vec_compare  # noqa: E402


def run_case(rank, world, local, host_group, etype, dims, n_vars, scramble, native=False):
    m = O.Mesh(etype, *dims, bounds=(-1, 1, -1, 1, -1, 1) if n_vars == 1 else (0, 1, 0, 0.3, 0, 0.1))
    if scramble:
        pid = ((np.arange(m.n_elem) // 5) % world).astype(np.int32)
        m.set_partition(world, pid)
    else:
        m.partition_linear(world)
        pid = m.partition()[0]
    s = O.System(m, n_vars)
    s.add_dirichlet((4,) if n_vars == 3 else (0, 1, 2, 3, 4, 5), 1 if n_vars == 3 else 0)
    s.build_sparsity()
    xyz, conn, _ = m.arrays()
    ed, nb, fd = s.dofs()
    cd, cv = s.constraints()
    grp, gci = s.sparsity()
    if n_vars == 1:
        f_tab, _ = s.forcing_at_qps()
        gv, gb, _ = s.assemble(O.POISSON_EX4, forcing=O.FORCING_TABULATED, f_tab=f_tab, n_threads=4)
    else:
        gv, gb, _ = s.assemble(O.ELASTICITY_EX6, n_threads=4)
    prob = M.RankProblem(rank, world, xyz, conn, nb, fd, pid, n_vars, cd, cv)
    ctx = F.Context(local)
    if native:  # the communicator lives inside the library: feasm_comm_init + feasm_exchange_begin / _end
        ids = [F.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0, group=host_group)
        ctx.comm_init(world, rank, ids[0])
        gathered = ctx.comm_allgather([rank + 10, prob.n_owned])
        assert [int(x) for x in gathered[:, 0]] == [r + 10 for r in range(world)] and int(gathered[:, 1].sum()) == s.n_dofs
    da = M.DistributedAssembly(ctx, prob, dist, host_group=host_group)
    da.setup()
    if n_vars == 1:
        params = F.make_params(F.POISSON, 0, F.FORCING_TABULATED, f_host=np.ascontiguousarray(f_tab[prob.elems]))
    else:
        params = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS)
    results = []
    for overlap in (True, False):
        for _ in range(6):  # back to back, nothing waits in between: step i+1 packs into the buffer step i sent from
            da.assemble(params, sync=False, overlap=overlap)
        ctx.synchronize()
        torch.cuda.synchronize()
        results.append(da.owned_values())
    (v, b), (v2, b2) = results
    assert np.array_equal(v, v2) and np.array_equal(b, b2), "overlapped and sequential exchange differ"
    f0, n_own = prob.f0, prob.n_owned
    lo, hi = int(grp[f0]), int(grp[f0 + n_own])
    assert np.array_equal(da.row_ptr[: n_own + 1].astype(np.int64), grp[f0:f0 + n_own + 1].astype(np.int64) - lo), "owned row_ptr"
    assert np.array_equal(da.col_idx[: hi - lo], gci[lo:hi]), "owned col_idx"
    cv_, cr_ = csr_compare(v, gv[lo:hi], da.row_ptr[: n_own + 1]), vec_compare(b, gb[f0:f0 + n_own])
    assert cv_["max_rel_err"] < TOL and cr_["max_rel_err"] < TOL, (cv_, cr_)
    extra = ""
    if native:  # the assembled matrix of the ranks used on the device: halo of a distributed vector (feasm_halo.cu)
        import scipy.sparse as sp
        import scipy.sparse.linalg as spl

        A = sp.csr_matrix((gv, gci.astype(np.int64), grp.astype(np.int64)), shape=(s.n_dofs, s.n_dofs))
        ctx.build_halo(prob.first_dof)
        x = np.cos(0.37 * np.arange(s.n_dofs)) + 0.1
        Ax = A @ x
        y = ctx.vector_mult_host(x[f0:f0 + n_own])
        e_y = np.abs(y - Ax[f0:f0 + n_own]).max() / np.abs(Ax).max()
        assert e_y <= 1e-13, ("vector_mult", e_y)
        r = ctx.residual_host(params, x[f0:f0 + n_own])  # K(u) u - F straight from the element records, ghost rows summed on their owners
        e_r = np.abs(r - (Ax - gb)[f0:f0 + n_own]).max() / max(np.abs(Ax).max(), np.abs(gb).max())
        assert e_r <= 1e-12, ("residual", e_r)
        r2 = ctx.residual_host(params, x[f0:f0 + n_own])
        assert np.array_equal(r, r2), "distributed residual not reproducible"
        xs, its, rel = ctx.cg_solve(1e-10, 5000)
        x_ref = spl.spsolve(A.tocsc(), gb)
        e_x = np.abs(xs - x_ref[f0:f0 + n_own]).max() / np.abs(x_ref).max()
        assert rel <= 1e-10 and e_x <= 1e-7, ("cg", its, rel, e_x)
        meta = torch.tensor([float(its), rel], dtype=torch.float64)
        gathered = [torch.zeros(2, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(gathered, meta, group=host_group)
        assert all(torch.equal(g, gathered[0]) for g in gathered), "ranks disagree on the CG iteration count / residual"
        extra = f", A x {e_y:.1e}, residual {e_r:.1e}, CG {its} its rel {rel:.1e} err {e_x:.1e}"
    ctx.close()
    print(f"rank {rank}: etype {etype} dims {dims} vars {n_vars} scramble {scramble} native {native}: {len(prob.elems)} elements, {n_own} owned + "
          f"{len(prob.ghosts)} ghost rows, values max rel {cv_['max_rel_err']:.2e}, rhs {cr_['max_rel_err']:.2e}{extra}", flush=True)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    host_group = dist.new_group(backend="gloo")
    run_case(rank, world, local, host_group, 27, (12, 12, 12), 1, 0)
    run_case(rank, world, local, host_group, 27, (8, 7, 6), 1, 1)
    run_case(rank, world, local, host_group, 8, (20, 16, 12), 1, 1)
    run_case(rank, world, local, host_group, 27, (8, 4, 4), 3, 0)
    run_case(rank, world, local, host_group, 27, (12, 12, 12), 1, 0, native=True)
    run_case(rank, world, local, host_group, 27, (8, 7, 6), 1, 1, native=True)
    run_case(rank, world, local, host_group, 27, (8, 4, 4), 3, 0, native=True)
    dist.barrier()
    if rank == 0:
        print("NCCL_PARITY_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
