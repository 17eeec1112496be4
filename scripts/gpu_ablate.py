"""Development aid: element-kernel time at 96^3 HEX27 for the library named by FEASM_LIB (ablation builds, see
feasm_elem_dev.cuh FEASM_ABLATE) and for the forcing modes; setup through the device-side ingest."""
import sys
import numpy as np
sys.path.insert(0, '.')
from libmesh_b200 import capi as F

n = int(sys.argv[1]) if len(sys.argv) > 1 else 96
ctx = F.Context(0)
ctx.generate_cube(F.HEX27, n, n, n, (-1, 1, -1, 1, -1, 1))
ctx.distribute_dofs(1)
ctx.generate_dirichlet(0x3f, F.DIRICHLET_EX4_EXACT)
ctx.build_pattern()
ctx.prepare()
ftab = np.ones((ctx.n_elem, 27))
for name, forcing in (("fd", F.FORCING_EX4_FD), ("none", F.FORCING_NONE), ("tab", F.FORCING_TABULATED)):
    p = F.make_params(F.POISSON, 0, forcing, f_host=ftab if forcing == F.FORCING_TABULATED else None)
    for _ in range(3):
        ctx.assemble(p)
    ms = []
    for _ in range(6):
        ctx.assemble(p)
        ms.append(ctx.last_timing()[0])
    ms = np.array(ms)
    print(f"{F.LIB_PATH.split('/')[-1]:>20s} forcing={name:5s} element {ms[:,0].min():.3f} ms  constrain {ms[:,3].min():.3f}  row_reduce {ms[:,1].min():.3f}  total {ms[:,2].min():.3f}")
