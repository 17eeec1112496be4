"""Quick timing probe at bench sizes using only the product (development aid)."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from libmesh_b200 import capi as F

def run(etype, n, qorder=0, reps=5):
    t0 = time.time()
    xyz, conn, mask = F.build_cube(etype, n, n, n, (-1, 1, -1, 1, -1, 1))
    nd, base, first, _ = F.distribute_dofs(conn, xyz.shape[0])
    ed = F.elem_dofs_from_nodes(conn, base)
    bn = np.nonzero(mask)[0]
    g = np.cos(.5 * np.pi * xyz[bn, 0]) * np.sin(.5 * np.pi * xyz[bn, 1]) * np.cos(.5 * np.pi * xyz[bn, 2])
    t1 = time.time()
    ctx = F.Context(0)
    ctx.set_mesh(etype, xyz, conn); ctx.set_dofs(ed, nd); ctx.set_constraints(base[bn], g)
    t2 = time.time()
    nnz = ctx.build_pattern(); t3 = time.time()
    ctx.prepare(); ctx.synchronize(); t4 = time.time()
    p = F.make_params(F.POISSON, qorder, F.FORCING_EX4_FD)
    for _ in range(3): ctx.assemble(p)
    best = None
    for _ in range(reps):
        ctx.assemble(p); ms, nl = ctx.last_timing()
        best = ms if best is None or ms[2] < best[2] else best
    ne = conn.shape[0]
    print(f'type {etype} n {n}: elems {ne} dofs {nd} nnz {nnz} host-gen {t1-t0:.2f}s upload {t2-t1:.2f}s pattern {t3-t2:.2f}s prepare {t4-t3:.2f}s')
    print(f'   elem kernel {best[0]:.3f} ms, row reduce {best[1]:.3f} ms, total {best[2]:.3f} ms -> {ne/best[2]*1e3/1e6:.2f} Melem/s')
    ctx.close()

run(F.HEX27, 48)
run(F.HEX27, 96)
run(F.HEX8, 128, qorder=3)
run(F.HEX8, 128, qorder=5)
