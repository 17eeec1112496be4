"""Timing probe: systems_of_equations_ex6 elasticity on build_cube n^3 HEX27 (config 5 geometry), product only."""
import sys
import numpy as np
sys.path.insert(0, '.')
from libmesh_b200 import capi as F
n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
xyz, conn, mask = F.build_cube(F.HEX27, n, n, n, (0, 1, 0, 0.3, 0, 0.1))
nd, base, first, _ = F.distribute_dofs(conn, xyz.shape[0], n_vars=3)
ed = F.elem_dofs_from_nodes(conn, base, n_vars=3)
bn = np.nonzero(mask & (1 << 4))[0]                      # min-x face: boundary id 4 (ex6 BOUNDARY_ID_MIN_X)
rows = (base[bn][:, None] + np.arange(3)[None, :]).ravel().astype(np.uint32)
ctx = F.Context(0)
ctx.set_mesh(F.HEX27, xyz, conn); ctx.set_dofs(ed, nd, n_vars=3); ctx.set_constraints(rows, np.zeros(rows.size))
nnz = ctx.build_pattern(); ctx.prepare()
p = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS)
for _ in range(3): ctx.assemble(p)
best = None
for _ in range(6):
    ctx.assemble(p); ms, nl = ctx.last_timing()
    best = ms if best is None or ms[2] < best[2] else best
ne = conn.shape[0]
print(f'elasticity hex27 n={n}: dofs {nd} nnz {nnz}: elem {best[0]:.3f} ms, constrain {best[3]:.3f} ms, row reduce {best[1]:.3f} ms, total {best[2]:.3f} ms -> {ne/best[2]*1e3/1e6:.2f} Melem/s')
ctx.close()
