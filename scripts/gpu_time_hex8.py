"""Development aid: HEX8 128^3 QGauss THIRD, kernel times per forcing mode (FEASM_HEX8_PACK=0: one element per warp)."""
import sys
import numpy as np
sys.path.insert(0, '.')
from libmesh_b200 import capi as F
n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
ctx = F.Context(0)
ctx.generate_cube(F.HEX8, n, n, n, (-1, 1, -1, 1, -1, 1))
ctx.distribute_dofs(1)
ctx.generate_dirichlet(0x3f, F.DIRICHLET_EX4_EXACT)
ctx.build_pattern(); ctx.prepare()
for name, forcing in (("fd", F.FORCING_EX4_FD), ("none", F.FORCING_NONE)):
    p = F.make_params(F.POISSON, 3, forcing)
    for _ in range(3):
        ctx.assemble(p)
    ms = np.array([(ctx.assemble(p), ctx.last_timing()[0])[1] for _ in range(6)])
    print(f"hex8 {n}^3 forcing={name:5s} element {ms[:,0].min():.3f} ms  constrain {ms[:,3].min():.3f}  row_reduce {ms[:,1].min():.3f}  total {ms[:,2].min():.3f} -> {ctx.n_elem/ms[:,2].min()/1e3:.1f} M elem/s")
