"""Development aid: kernel times of one configuration (element / constraint / row reduction / total, ms, best of 8).
    python scripts/gpu_time_quick.py [hex27|hex8|tet10|hex20|prism18] [n] [fd|tab|none] [poisson|elasticity]
FEASM_LIB=<path> selects a development build of the library."""
import sys
import numpy as np
sys.path.insert(0, '.')
from libmesh_b200 import capi as F

et = {"hex27": 27, "hex8": 8, "tet10": 10, "hex20": 20, "prism18": 18}[sys.argv[1] if len(sys.argv) > 1 else "hex27"]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 96
forcing = {"fd": F.FORCING_EX4_FD, "tab": F.FORCING_TABULATED, "none": F.FORCING_NONE}[sys.argv[3] if len(sys.argv) > 3 else "fd"]
elast = len(sys.argv) > 4 and sys.argv[4] == "elasticity"
ctx = F.Context(0)
b = (0, 1, 0, 0.3, 0, 0.1) if elast else (-1, 1, -1, 1, -1, 1)
ctx.generate_cube(et, n, n, n, b)
ctx.distribute_dofs(3 if elast else 1)
ctx.generate_dirichlet(0x10 if elast else 0x3f, F.DIRICHLET_ZERO if elast else F.DIRICHLET_EX4_EXACT)
ctx.build_pattern()
ctx.prepare()
qo = 3 if et == 8 else 0
if elast:
    p = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS)
else:
    ftab = np.ones((ctx.n_elem, {27: 27, 8: 8, 10: 14, 20: 27, 18: 21}[et])) if forcing == F.FORCING_TABULATED else None
    p = F.make_params(F.POISSON, qo, forcing, f_host=ftab)
for _ in range(3):
    ctx.assemble(p)
ms = []
for _ in range(8):
    ctx.assemble(p)
    ms.append(ctx.last_timing()[0])
ms = np.array(ms)
print(f"{F.LIB_PATH.split('/')[-1]} {sys.argv[1:]} n_elem {ctx.n_elem}: element {ms[:,0].min():.3f} ms  constrain {ms[:,3].min():.3f}  row_reduce {ms[:,1].min():.3f}  total {ms[:,2].min():.3f}")
