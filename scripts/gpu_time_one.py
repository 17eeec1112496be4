"""Timing probe for one case: gpu_time_one.py elem n [qorder] (development aid, product only)."""
import sys
import numpy as np
sys.path.insert(0, '.')
from libmesh_b200 import capi as F
etype = {"hex27": 27, "hex8": 8, "tet10": 10}[sys.argv[1]]
n = int(sys.argv[2]); qorder = int(sys.argv[3]) if len(sys.argv) > 3 else 0
xyz, conn, mask = F.build_cube(etype, n, n, n, (-1, 1, -1, 1, -1, 1))
nd, base, first, _ = F.distribute_dofs(conn, xyz.shape[0])
ed = F.elem_dofs_from_nodes(conn, base)
bn = np.nonzero(mask)[0]
g = np.cos(.5 * np.pi * xyz[bn, 0]) * np.sin(.5 * np.pi * xyz[bn, 1]) * np.cos(.5 * np.pi * xyz[bn, 2])
ctx = F.Context(0)
ctx.set_mesh(etype, xyz, conn); ctx.set_dofs(ed, nd)
import os
if not os.environ.get('NOCON'): ctx.set_constraints(base[bn], g)
ctx.build_pattern(); ctx.prepare()
p = F.make_params(F.POISSON, qorder, F.FORCING_EX4_FD, affine_tensor=int(os.environ.get('AFFINE', '0')))
for _ in range(3): ctx.assemble(p)
best = None
for _ in range(8):
    ctx.assemble(p); ms, nl = ctx.last_timing()
    best = ms if best is None or ms[2] < best[2] else best
ne = conn.shape[0]
print(f'{sys.argv[1]} n={n}: elem {best[0]:.3f} ms, constrain {best[3]:.3f} ms, row reduce {best[1]:.3f} ms, total {best[2]:.3f} ms -> {ne/best[2]*1e3/1e6:.2f} Melem/s')
ctx.close()
