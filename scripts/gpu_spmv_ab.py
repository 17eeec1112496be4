"""Development aid: feasm_vector_mult time of the SpMV kernels (A/B knobs FEASM_SPMV_PIPE / FEASM_SPMV_STREAM are read per
call) and comparison of their results.
    python scripts/gpu_spmv_ab.py [hex27|hex8|tet10|hex20] [n] [n_vars]"""
import os
import sys
import numpy as np

sys.path.insert(0, '.')
import torch
from libmesh_b200 import capi as F

et = {"hex27": 27, "hex8": 8, "tet10": 10, "hex20": 20, "prism18": 18}[sys.argv[1] if len(sys.argv) > 1 else "hex27"]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 96
nv = int(sys.argv[3]) if len(sys.argv) > 3 else 1
ctx = F.Context(0)
ctx.generate_cube(et, n, n, n, (-1, 1, -1, 1, -1, 1))
ctx.distribute_dofs(nv)
ctx.generate_dirichlet(0x3f, F.DIRICHLET_ZERO)
ctx.build_pattern()
ctx.prepare()
p = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS) if nv == 3 else F.make_params(F.POISSON, 3 if et == 8 else 0, F.FORCING_NONE)
ctx.assemble(p)
nd, nnz = ctx.n_rows, ctx.nnz
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.rand(nd, dtype=torch.float64, device="cuda", generator=g)
y = torch.zeros(nd, dtype=torch.float64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
lib = torch.cuda.ExternalStream(ctx.stream())
res = {}
for name, pipe, st in (("pipe", "1", "0"), ("row", "0", "0"), ("stream", "1", "1")):
    os.environ["FEASM_SPMV_PIPE"], os.environ["FEASM_SPMV_STREAM"] = pipe, st
    best = []
    for it in range(8):
        flush.zero_()
        torch.cuda.synchronize()
        with torch.cuda.stream(lib):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ctx.vector_mult(x.data_ptr(), y.data_ptr())
            e1.record()
        ctx.synchronize(); torch.cuda.synchronize()
        best.append(e0.elapsed_time(e1))
    ms = min(best[2:])
    by = nnz * 12 + nd * 32
    print(f"{name:15s} {sys.argv[1:]} rows {nd} nnz {nnz}: {ms:.4f} ms, {by / ms / 1e6:.0f} GB/s", flush=True)
    res[name] = y.cpu().numpy().copy()
print("pipe vs row bitwise equal:", np.array_equal(res["pipe"], res["row"]), " max |y|", np.abs(res["row"]).max(),
      " stream vs row max abs diff", np.abs(res["stream"] - res["row"]).max())
