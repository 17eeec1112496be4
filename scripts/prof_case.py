"""Small fixed workload for ncu: N assemblies of the ex4 Poisson system on build_cube n^3 (product only)."""
import sys
import numpy as np
sys.path.insert(0, '.')
from libmesh_b200 import capi as F

etype = {"hex27": 27, "hex8": 8, "tet10": 10}[sys.argv[1]] if len(sys.argv) > 1 else 27
n = int(sys.argv[2]) if len(sys.argv) > 2 else 48
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
skew = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
xyz, conn, mask = F.build_cube(etype, n, n, n, (-1, 1, -1, 1, -1, 1))
if skew:
    rng = np.random.default_rng(12345)
    interior = mask == 0
    xyz[interior] += skew * (2.0 / n) * rng.uniform(-1, 1, size=(int(interior.sum()), 3))
nd, base, first, _ = F.distribute_dofs(conn, xyz.shape[0])
ed = F.elem_dofs_from_nodes(conn, base)
bn = np.nonzero(mask)[0]
g = np.cos(.5 * np.pi * xyz[bn, 0]) * np.sin(.5 * np.pi * xyz[bn, 1]) * np.cos(.5 * np.pi * xyz[bn, 2])
ctx = F.Context(0)
ctx.set_mesh(etype, xyz, conn); ctx.set_dofs(ed, nd); ctx.set_constraints(base[bn], g)
ctx.build_pattern(); ctx.prepare()
import os
p = F.make_params(F.POISSON, int(os.environ.get('QORDER', '0')), F.FORCING_EX4_FD, affine_tensor=int(os.environ.get('AFFINE', '0')))
for _ in range(reps):
    ctx.assemble(p)
    print(ctx.last_timing())
v, r = ctx.get_values()
print("checksum", float(v.sum()), float(r.sum()))
ctx.close()
