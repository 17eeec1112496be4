"""Development aid: feasm_residual at 96^3 HEX27 (run under ncu for the per-kernel split)."""
import sys, time
import ctypes as C
import numpy as np
sys.path.insert(0, '.')
from libmesh_b200 import capi as F
n = int(sys.argv[1]) if len(sys.argv) > 1 else 96
ctx = F.Context(0)
ctx.generate_cube(F.HEX27, n, n, n, (-1, 1, -1, 1, -1, 1)); ctx.distribute_dofs(1); ctx.generate_dirichlet(0x3f, F.DIRICHLET_EX4_EXACT)
ctx.build_pattern(); ctx.prepare()
p = F.make_params(F.POISSON, 0, F.FORCING_EX4_FD)
L = F.lib()
du, dr = L.feasm_vec_alloc(ctx.h, ctx.n_rows), L.feasm_vec_alloc(ctx.h, ctx.n_rows)
L.feasm_vec_fill(ctx.h, C.c_void_p(du), ctx.n_rows, 1.0)
for _ in range(3):
    ctx._ck(L.feasm_residual(ctx.h, C.byref(p), C.c_void_p(du), C.c_void_p(dr)))
ctx.synchronize()
t = time.perf_counter()
for _ in range(5):
    ctx._ck(L.feasm_residual(ctx.h, C.byref(p), C.c_void_p(du), C.c_void_p(dr)))
ctx.synchronize()
print("feasm_residual", (time.perf_counter() - t) / 5 * 1e3, "ms")
