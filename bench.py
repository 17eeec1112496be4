#!/usr/bin/env python
"""bench.py — Q2 HEX27 finite-element assembly throughput (elements/s) on B200, with roofline, parity and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--n 96] [--elem hex27|hex8|tet10]
                    [--problem poisson|elasticity] [--no-check] [--no-cpu]

A "step" is one full assembly of the 3D Poisson system of introduction_ex4 (matrix + RHS zeroed -> final CSR
values in HBM) on the synthetic build_cube mesh BASELINE.json's metric is quoted on: 96^3 HEX27, SECOND LAGRANGE,
QGauss FIFTH, Dirichlet u = exact on all six sides, forcing = the example's FD Laplacian evaluated per quadrature
point inside the kernel.  Setup (sparsity pattern, scatter maps, tables) is outside the timed region and reported.

  value        elements/s with every input resident in HBM, CUDA-event timed on the library's stream
  e2e          the same through the C ABI with HOST buffers: per step the node coordinates and Dirichlet values go
               host->device from pinned memory, the assembled CSR values and RHS come back device->host
  roofline     dominant kernel (element kernel, FP64 tensor pipe) + the slower-of FP64/HBM bound for the whole step
  parity       the TIMED configuration checked per entry against the oracle at full size, outside the timed region:
               pattern bit-equal, values / RHS under the one rule of oracle/parity.py, strict-1e-12 failures counted
  cpu_baseline the oracle (CPU restatement of the reference path) on this box's host cores, on the same configuration

--impl reference times that CPU path alone (the reference itself cannot be compiled here: TIMPI / MetaPhysicL
submodules are empty), on all host threads, on the SAME configuration (every step = one assembly of the
whole mesh).  --problem elasticity runs BASELINE config 5 (systems_of_equations_ex6, 3 variables, 64^3 by default).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os

# stdout carries the JSON line and nothing else: NCCL's banner ("NCCL version ..." when NCCL_DEBUG is set on the box) goes to stderr
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":  # at this level NCCL ignores NCCL_DEBUG_FILE and prints to stdout
    del os.environ["NCCL_DEBUG"]
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ELEM = {"hex27": 27, "hex8": 8, "tet10": 10, "hex20": 20, "prism18": 18}
METRIC = "Q2 HEX27 assembly elements/s"
EX4_BOUNDS = (-1.0, 1.0, -1.0, 1.0, -1.0, 1.0)
EX6_BOUNDS = (0.0, 1.0, 0.0, 0.3, 0.0, 0.1)  # systems_of_equations_ex6.C:437-444


def metric_name(args):
    if args.problem != "poisson":
        return f"{args.elem.upper()} 3-variable elasticity (systems_of_equations_ex6) assembly elements/s"
    return METRIC if args.elem == "hex27" else f"{args.elem.upper()} introduction_ex4 assembly elements/s"


def algorithmic_counts(etype: int, nq: int, nnz: int, n_dofs: int, n_elem: int, n_nodes: int, affine: bool = True, problem: str = "poisson"):
    """compulsory HBM bytes and flops per element.  build_cube elements are affine, so the Jacobian term is
    counted once per element (nq_J = 1), as the reference's compute_affine_map does.  Elasticity: the flops of the
    algorithm that RUNS (six 27x27 products of contraction length nq per element), as SURVEY a7 asks."""
    n = etype
    geo = 15 * nq * n + 18 * n * (1 if affine else nq) + 2 * nq * n
    if problem == "elasticity":
        flops = 6 * (2 * n * n * nq) + 9 * n * n * 4 + geo
    else:
        flops = 6 * nq * n * n + geo
    byts = 4 * n + 24 * n_nodes / n_elem + 8 * nnz / n_elem + 8 * n_dofs / n_elem
    return float(flops), float(byts)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def source_hash() -> str:
    """Hash of the kernel sources: profiles/*/traffic.json is stamped with it, so an ncu capture of another build is
    never reported as this build's traffic."""
    h = hashlib.sha256()
    d = os.path.join(ROOT, "libmesh_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h", ".cpp")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def recorded_traffic(kernel: str, workload: str):
    """dram bytes per launch of `kernel` from the newest ncu capture committed under profiles/ — only if it was taken
    from exactly these sources; otherwise None with the reason."""
    for rnd in ("r2", "r1"):
        tj = os.path.join(ROOT, "profiles", rnd, "traffic.json")
        if not os.path.exists(tj):
            continue
        d = json.load(open(tj))
        if d.get("source_hash") != source_hash():
            return None, f"profiles/{rnd}/traffic.json was captured from other sources ({d.get('source_hash')}); not reported"
        if d.get("workload") != workload or kernel not in d:
            return None, f"profiles/{rnd}/traffic.json holds no capture of {kernel} on {workload}"
        return d[kernel]["dram_bytes"], f"ncu --set full, profiles/{rnd}/traffic.json, sources {d['source_hash']}"
    return None, "no ncu capture committed"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for k, nm in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---- the synthetic job -----------------------------------------------------------------------------------------
def problem_dims(args):
    n = args.n if args.n else (64 if args.problem == "elasticity" else 96)
    return n, (EX6_BOUNDS if args.problem == "elasticity" else EX4_BOUNDS), (3 if args.problem == "elasticity" else 1)


def nq_of(args):
    return {27: 27, 8: 27 if args.qorder in (0, 5) else 8, 10: 14, 20: 27, 18: 21}[ELEM[args.elem]]


def build_problem(F, args, n_parts: int = 1, elem_pid=None):
    """Synthetic inputs from the product-side host generators (no oracle here): introduction_ex4 (Poisson) or
    systems_of_equations_ex6 (elasticity: clamped on min-x = boundary id 4, traction (0,0,-1) on max-x = id 2)."""
    t0 = time.time()
    etype = ELEM[args.elem]
    n, b, n_vars = problem_dims(args)
    xyz, conn, mask = F.build_cube(etype, n, n, n, b)
    n_dofs, base, first, _ = F.distribute_dofs(conn, xyz.shape[0], n_vars, n_parts, elem_pid)
    P = dict(xyz=xyz, conn=conn, n_dofs=n_dofs, node_base=base, first_dof=first, n_vars=n_vars, n=n, bounds=b, etype=etype)
    if args.problem == "elasticity":
        bn = np.nonzero(mask & (1 << 4))[0]
        P["con_dofs"] = np.concatenate([base[bn].astype(np.int64) + v for v in range(3)])
        P["con_vals"] = np.zeros(P["con_dofs"].size)
        if etype == 10:
            raise SystemExit("--problem elasticity: the traction side list is generated for hex meshes only")
        e = np.arange(conn.shape[0], dtype=np.int64)
        P["side_elems"] = e[(e % n) == n - 1].astype(np.uint32)  # elements on the max-x face; their side 2 (App. A.3)
        P["side_sides"] = np.full(P["side_elems"].size, 2, dtype=np.uint8)
        P["side_g"] = (0.0, 0.0, -1.0)
    else:
        bn = np.nonzero(mask)[0]
        # Dirichlet data g = exact_solution at boundary nodes (introduction_ex3/exact_solution.C:40-47) — input data
        P["con_dofs"] = base[bn].astype(np.int64)
        P["con_vals"] = np.cos(.5 * np.pi * xyz[bn, 0]) * np.sin(.5 * np.pi * xyz[bn, 1]) * np.cos(.5 * np.pi * xyz[bn, 2])
    P["host_gen_s"] = time.time() - t0
    return P


def make_params(F, args, forcing=None, **kw):
    if args.problem == "elasticity":
        return F.make_params(F.ELASTICITY, args.qorder, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS, **kw)
    return F.make_params(F.POISSON, args.qorder, F.FORCING_EX4_FD if forcing is None else forcing, **kw)


def workload_name(args):
    n, _, _ = problem_dims(args)
    if args.problem == "elasticity":
        return (f"systems_of_equations_ex6 3D linear elasticity, 3 variables, build_cube {n}^3 {args.elem.upper()} on [0,1]x[0,0.3]x[0,0.1], "
                f"QGauss nq={nq_of(args)}, clamped on min-x, body force + traction on max-x")
    return (f"introduction_ex4 3D Poisson, build_cube {n}^3 {args.elem.upper()} on [-1,1]^3, QGauss nq={nq_of(args)}, "
            "Dirichlet on ids 0-5, FD forcing evaluated in-kernel")


# ---- the oracle side: checker and CPU baseline (the only places bench.py touches oracle/) ---------------------------
def oracle_system(args, fast: bool, n: int | None = None):
    from oracle import oracle as O

    etype = ELEM[args.elem]
    nn, b, n_vars = problem_dims(args)
    n = n or nn
    m = O.Mesh(etype, n, n, n, bounds=b, fast=fast)
    s = O.System(m, n_vars)
    if args.problem == "elasticity":
        s.add_dirichlet((4,), 1)
        el, sd = m.boundary_sides(2)
        s.set_side_loads(el, sd, (0.0, 0.0, -1.0))
    else:
        s.add_dirichlet()
    s.build_sparsity()
    return O, m, s


def oracle_assemble(O, s, args, threads, forcing_tab=None):
    if args.problem == "elasticity":
        return s.assemble(O.ELASTICITY_EX6, n_threads=threads, qorder=args.qorder)
    if forcing_tab is not None:
        return s.assemble(O.POISSON_EX4, n_threads=threads, qorder=args.qorder, forcing=O.FORCING_TABULATED, f_tab=forcing_tab)
    return s.assemble(O.POISSON_EX4, n_threads=threads, qorder=args.qorder)


def full_size_parity(ctx, F, P, args):
    """The timed configuration against the oracle, per entry, at full size.  Oracle passes run on all
    host threads (faithful build: -O2, no FMA contraction; OpenMP atomics change the order of the adds into a slot by
    rounding only)."""
    from oracle.parity import FLOOR, TOL, csr_compare, vec_compare

    t0 = time.time()
    threads = os.cpu_count() or 1
    O, m, s = oracle_system(args, fast=False)
    rp, ci = s.sparsity()
    rp2, ci2 = ctx.get_pattern()
    ed, _, _ = s.dofs()
    cd, cv = s.constraints()
    out = {"rule": f"|a-b| <= {TOL:g} * max(|a|, |b|, {FLOOR:g} * rowmax)  (oracle/parity.py); strict = no floor",
           "checker": f"oracle (faithful build), {threads} threads, full size",
           "pattern_bit_equal": bool(int(s.nnz) == int(ctx.nnz) and np.array_equal(rp, rp2) and np.array_equal(ci, ci2)),
           "dof_map_bit_equal": bool(np.array_equal(ed, F.elem_dofs_from_nodes(P["conn"], P["node_base"], P["n_vars"]))),
           "constrained_set_equal": bool(np.array_equal(np.sort(cd.astype(np.int64)), np.sort(P["con_dofs"])))}
    del ci, ci2, rp2, ed
    modes = [("as_timed", None)]
    if args.problem == "poisson":
        modes.append(("tabulated_forcing", "tab"))
    for name, kind in modes:
        f_tab = s.forcing_at_qps(args.qorder)[0] if kind == "tab" else None
        vo, ro, _ = oracle_assemble(O, s, args, threads, f_tab)
        ctx.assemble(make_params(F, args, F.FORCING_TABULATED if kind == "tab" else None, **({"f_host": f_tab} if kind == "tab" else {})))
        v, r = ctx.get_values()
        out[name] = {"values": csr_compare(v, vo, rp), "rhs": vec_compare(r, ro)}
        del v, r, vo, ro, f_tab
    if args.problem == "poisson":
        out["note"] = ("as_timed = FORCING_EX4_FD: the FD Laplacian divides an O(eps^2) difference by eps^2 = 1e-6, so a one-ulp "
                       "difference between CUDA and glibc sin/cos is ~2e-10 in f; the matrix is unaffected and the RHS deviation is "
                       "what is measured here.  tabulated_forcing carries the 1e-12 bar for the RHS.")
    out["seconds"] = round(time.time() - t0, 1)
    return out


def cpu_baseline(args, budget_s: float = 12.0):
    """The oracle's assembly loop (the CPU restatement of the reference path) on the SAME configuration, all host
    threads, repeated for about budget_s seconds; plus one single-thread pass with the reference's four PerfLog phases
    (elem init / Ke / Fe / matrix insertion, doc/html/examples/introduction_ex4.html:490-493) on a half-edge cube."""
    threads = os.cpu_count() or 1
    O, m, s = oracle_system(args, fast=True)
    oracle_assemble(O, s, args, threads)  # warm-up pass (page faults, thread pool)
    passes, t0 = 0, time.perf_counter()
    while True:
        oracle_assemble(O, s, args, threads)
        passes += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or passes >= 50:
            break
    n_elem = m.n_elem
    rate = passes * n_elem / dt
    n_full = problem_dims(args)[0]
    sample = f"{passes} passes of the full job (build_cube {n_full}^3, {n_elem} elements), {threads} OpenMP threads, {dt:.1f} s"
    del s, m
    n1 = max(n_full // (4 if args.problem == "elasticity" else 2), 4)
    O, m1, s1 = oracle_system(args, fast=True, n=n1)
    t1 = time.perf_counter()
    _, _, ph = oracle_assemble(O, s1, args, 1)
    t1 = time.perf_counter() - t1
    tot = float(ph.sum()) or 1.0
    names = ["elem_init", "Ke", "Fe", "matrix_insertion"]
    ref_code = reference_code_sample(args, O)
    return {"value": rate, "unit": "elements/s", "cores": threads, "kind": "port", "sample": sample,
            "single_thread": {"value": m1.n_elem / t1, "unit": "elements/s", "sample": f"one pass of build_cube {n1}^3 ({m1.n_elem} elements), 1 thread",
                              "phase_seconds": {k: round(float(v), 4) for k, v in zip(names, ph)},
                              "phase_percent": {k: round(100 * float(v) / tot, 2) for k, v in zip(names, ph)},
                              "reference_published_percent": {"elem_init": 33.54, "Ke": 40.71, "Fe": 10.14, "matrix_insertion": 15.61,
                                                              "source": "doc/html/examples/introduction_ex4.html:490-493 (-d 3 -n 6, 1 rank)"}},
            "reference_code": ref_code,
            "note": "oracle = CPU restatement of the reference loop (g++ -O3 -march=x86-64-v3, OpenMP over elements); libMesh as a "
                    "library cannot be compiled here (empty TIMPI/MetaPhysicL submodules); reference_code = the example's own "
                    "assembly function compiled where it lies (oracle/_ref), one thread, reported beside the port"}


def reference_code_sample(args, O, n: int = 24):
    """assemble_poisson of the reference's own introduction_ex4.C — compiled where it lies into oracle/_ref/libref_app*.so over the
    reference's quadrature, shape, FEMap, DenseMatrix and constraint code and a stub mesh / matrix world (oracle/ref_app/
    ref_world.h) — timed on one thread on a bounded sample, at the port's optimisation level (the _fast build; the -O2
    no-contraction build is the bitwise checker of the oracle, tests/test_oracle_ref_app.py).  A data point beside the port,
    not the reference arm: the stub world's matrix insertion is a search into a CSR array, not a solver package's."""
    try:
        from oracle import ref as R

        if args.problem != "poisson" or not R.app_available(fast=True):
            return None
        etype = ELEM[args.elem]
        b = problem_dims(args)[1]
        m = O.Mesh(etype, n, n, n, bounds=b, fast=True)
        s = O.System(m, 1)
        s.add_dirichlet()
        s.build_sparsity()
        R.use_fast(True)
        try:
            nnz, nd = R.load_world(m, s, n, b, affine=np.ones(m.n_elem, dtype=np.uint8))
            order = 1 if args.elem == "hex8" else 2
            R.assemble_ex4(order, nnz, nd)  # warm-up
            out = R.assemble_ex4(order, nnz, nd)
            dt = R.last_solve_seconds()
        finally:
            R.use_fast(False)
        t1 = time.perf_counter()
        vo, ro, _ = s.assemble(O.POISSON_EX4, n_threads=1)
        t1 = time.perf_counter() - t1
        return {"value": m.n_elem / dt, "unit": "elements/s", "cores": 1, "kind": "reference",
                "sample": f"assemble_poisson of introduction_ex4.C on build_cube {n}^3 {args.elem.upper()} ({m.n_elem} elements), 1 thread, {dt:.2f} s",
                "port_on_the_same_sample": {"value": m.n_elem / t1, "unit": "elements/s"},
                "max_abs_diff_to_the_port": {"values": float(np.abs(vo - out["values"]).max()), "rhs": float(np.abs(ro - out["rhs"]).max()),
                                             "note": "both at -O3 with FMA contraction; the -O2 builds agree bit for bit"}}
    except Exception as e:  # a data point only: never takes the bench line down
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


# ---- our arm, one GPU ---------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch

    from libmesh_b200 import capi as F

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}")
    if world > 1:
        return run_multi(args, F, rank, world, local)
    torch.cuda.set_device(local)
    etype = ELEM[args.elem]
    elast = args.problem == "elasticity"
    P = build_problem(F, args)
    n_elem = P["conn"].shape[0]
    elem_dofs = F.elem_dofs_from_nodes(P["conn"], P["node_base"], P["n_vars"])
    con_rows = P["con_dofs"].astype(np.uint32)
    ctx = F.Context(local)
    t0 = time.time()
    ctx.set_mesh(etype, P["xyz"], P["conn"])
    ctx.set_dofs(elem_dofs, P["n_dofs"], n_vars=P["n_vars"])
    ctx.set_constraints(con_rows, P["con_vals"])
    if elast:
        ctx.set_side_loads(P["side_elems"], P["side_sides"], P["side_g"])
    t1 = time.time()
    nnz = ctx.build_pattern()
    ctx.prepare()
    ctx.synchronize()
    t2 = time.time()
    params = make_params(F, args)
    stream = torch.cuda.ExternalStream(ctx.stream())
    # The same inputs generated on the device (feasm_generate_cube / _distribute_dofs_device / _generate_dirichlet,
    # SURVEY 8f rank 2): timed (second run: the first loads the kernels) and compared with the host generators' arrays
    # at full size.  Setup only; the timed steps below run on `ctx`.
    ingest = None
    if not elast:
        ing = F.Context(local)
        for _ in range(2):
            ti = time.time()
            ing.generate_cube(etype, P["n"], P["n"], P["n"], P["bounds"])
            ing.distribute_dofs(1)
            n_con_dev = ing.generate_dirichlet(0x3f, F.DIRICHLET_EX4_EXACT)
            ing_s = time.time() - ti
        xyz_d, conn_d, _ = ing.get_mesh(mask=False)
        ed_d, _ = ing.get_dofs(node_base=False)
        flag_d, val_d = ing.get_constraints()
        ingest = {"seconds": round(ing_s, 4),
                  "mesh_and_dofs_bit_identical_to_host_route": bool(np.array_equal(xyz_d, P["xyz"]) and np.array_equal(conn_d, P["conn"])
                                                                    and np.array_equal(ed_d, elem_dofs)),
                  "dirichlet_rows_identical": bool(n_con_dev == len(con_rows) and np.array_equal(np.nonzero(flag_d)[0], np.sort(con_rows))),
                  "dirichlet_max_abs_diff": float(np.max(np.abs(val_d[con_rows] - P["con_vals"])))}
        del xyz_d, conn_d, ed_d, flag_d, val_d
        ing.close()
    del elem_dofs
    fp64 = ctx.measure_fp64_peaks()

    # ---- value: inputs resident in HBM -------------------------------------------------------------------
    for _ in range(args.warmup):
        ctx.assemble(params, sync=False)
    ctx.synchronize()
    clocks = ClockSampler(local)
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    k_ms = np.zeros((args.steps, 4))
    launches = 0
    ev0.record(stream)
    for _ in range(args.steps):
        ctx.assemble(params, sync=False)
    ev1.record(stream)
    ctx.synchronize()
    torch.cuda.synchronize()
    total_ms = ev0.elapsed_time(ev1)
    # per-kernel durations (CUDA events inside the library, same stream): a second pass so the per-step event
    # synchronisation does not sit inside the throughput measurement above
    for i in range(args.steps):
        ctx.assemble(params, sync=True)
        ms, nl = ctx.last_timing()
        k_ms[i] = ms
        launches += nl
    clk = clocks.stop()
    ms_per_step = total_ms / args.steps
    value = n_elem / (ms_per_step * 1e-3)

    # ---- e2e: host buffers through the C ABI, H2D of inputs and D2H of the result every step -----------------
    xyz_pin = torch.from_numpy(P["xyz"]).pin_memory()
    val_pin = torch.empty(nnz, dtype=torch.float64).pin_memory()
    rhs_pin = torch.empty(P["n_dofs"], dtype=torch.float64).pin_memory()
    xyz_np, val_np, rhs_np = xyz_pin.numpy(), val_pin.numpy(), rhs_pin.numpy()
    e2e_steps = max(3, min(args.steps, 8))

    # The result leaves through the device-pointer accessors of the C ABI (feasm_device_values / feasm_device_rhs): a
    # device-side snapshot on the library's stream, then the PCIe read on a second stream, so the D2H read of step i
    # overlaps the H2D + assembly of step i+1 (every step still uploads its inputs and reads back its full result).
    from libmesh_b200.multi import DevArray
    d_val = torch.as_tensor(DevArray(ctx.device_values(), nnz), device="cuda")
    d_rhs = torch.as_tensor(DevArray(ctx.device_rhs(), P["n_dofs"]), device="cuda")
    snap_val, snap_rhs = torch.empty_like(d_val), torch.empty_like(d_rhs)
    copy_stream = torch.cuda.Stream()
    ev_ready, ev_done = torch.cuda.Event(), torch.cuda.Event()
    ev_done.record(copy_stream)

    def e2e_step():
        ctx.update_coordinates(xyz_np)
        ctx.set_constraints(con_rows, P["con_vals"])
        ctx.assemble(params, sync=False)
        with torch.cuda.stream(stream):
            stream.wait_event(ev_done)            # the previous read has left the snapshot
            snap_val.copy_(d_val, non_blocking=True)
            snap_rhs.copy_(d_rhs, non_blocking=True)
            ev_ready.record(stream)
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_ready)
            val_pin.copy_(snap_val, non_blocking=True)
            rhs_pin.copy_(snap_rhs, non_blocking=True)
            ev_done.record(copy_stream)

    e2e_step()
    torch.cuda.synchronize()
    ctx.synchronize()
    t_e = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    ctx.synchronize()  # raises if any step flagged a degenerate element
    e2e_s = (time.perf_counter() - t_e) / e2e_steps
    h2d = xyz_np.nbytes + con_rows.nbytes + P["con_vals"].nbytes  # what feasm_update_coordinates / feasm_set_constraints upload
    d2h = val_np.nbytes + rhs_np.nbytes
    checksum = float(val_np.sum())
    del snap_val, snap_rhs, val_pin, rhs_pin, val_np, rhs_np, d_val, d_rhs

    # ---- roofline ------------------------------------------------------------------------------------------
    nq = nq_of(args)
    flops_e, bytes_e = algorithmic_counts(etype, nq, nnz, P["n_dofs"], n_elem, P["xyz"].shape[0], problem=args.problem)
    hbm_peak, peak_src = peaks()
    elem_ms, red_ms, con_ms = float(np.median(k_ms[:, 0])), float(np.median(k_ms[:, 1])), float(np.median(k_ms[:, 3]))
    fp64_peak = max(fp64["dmma_tflops"], fp64["dfma_tflops"])
    ach_tf = flops_e * n_elem / (elem_ms * 1e-3) / 1e12
    step_tf = flops_e * n_elem / (ms_per_step * 1e-3) / 1e12
    step_gbs = bytes_e * n_elem / (ms_per_step * 1e-3) / 1e9
    t_fp64, t_hbm = flops_e * n_elem / (fp64_peak * 1e12), bytes_e * n_elem / (hbm_peak * 1e9)
    kname = "k_elem_elasticity" if elast else "k_elem_poisson"
    traffic, traffic_src = recorded_traffic(kname, f"{args.problem} {args.elem} {P['n']}")
    out = {
        "metric": metric_name(args), "value": value, "unit": "elements/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(args), "n_elem": n_elem, "n_dofs": P["n_dofs"],
                   "nnz": nnz, "l2": "inputs and outputs larger than L2 (CSR values %.2f GB)" % (nnz * 8 / 1e9),
                   "setup_s": {"host_mesh_dofs": round(P["host_gen_s"], 3), "upload": round(t1 - t0, 3), "pattern_and_maps": round(t2 - t1, 3), "device_ingest": ingest}},
        "clocks": clk, "gpu_launches": int(launches),  # kernels of libfeasm.so launched inside the timed K steps
        "e2e": {"value": n_elem / e2e_s, "unit": "elements/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_s * 1e3, "checksum_values": checksum},
        "roofline": {"bound": "tensor", "kernel": f"{kname} (FP64 DMMA contraction + geometry)", "achieved": ach_tf,
                     "peak": fp64_peak, "unit": "TFLOP/s", "frac": ach_tf / fp64_peak, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": "FP64 mma.sync/DFMA chain measured live by feasm_measure_fp64_peaks (MEASURED_PEAKS.json has no FP64 entry)",
                     "kernel_ms": {"element": elem_ms, "constrain": con_ms, "row_reduce": red_ms},
                     "flops_per_element": flops_e, "bytes_per_element": bytes_e},
        "roofline_step": {"slower_of": "fp64" if t_fp64 >= t_hbm else "hbm", "fp64_tflops": step_tf, "fp64_frac": step_tf / fp64_peak,
                          "hbm_gbs": step_gbs, "hbm_frac": step_gbs / hbm_peak, "hbm_peak": hbm_peak, "hbm_peak_source": peak_src,
                          "frac_of_slower_of": max(t_fp64, t_hbm) / (ms_per_step * 1e-3)},
        "fp64_peaks": fp64, "source_hash": source_hash(),
    }
    if not elast:
        extras_poisson(out, ctx, F, P, args, torch, stream, n_elem, nnz, hbm_peak, params)
    if not args.no_check:
        out["parity"] = full_size_parity(ctx, F, P, args)
    ctx.close()
    if not args.no_cpu:
        out["cpu_baseline"] = cpu_baseline(args)
    print(json.dumps(out))


def extras_poisson(out, ctx, F, P, args, torch, stream, n_elem, nnz, hbm_peak, params):
    """Not the headline: the opt-in affine path, and the steps after the path (SpMV, CG, matrix-free residual)."""
    # Extra (not the headline): the same step with feasm_params.affine_tensor = 1 — affine elements (all of build_cube)
    # take the reference-tensor contraction instead of the quadrature loop; same result to 1e-12 (tests), fewer flops.
    p_aff = F.make_params(F.POISSON, args.qorder, F.FORCING_EX4_FD, affine_tensor=1)
    for _ in range(args.warmup):
        ctx.assemble(p_aff, sync=False)
    ctx.synchronize()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0.record(stream)
    for _ in range(args.steps):
        ctx.assemble(p_aff, sync=False)
    a1.record(stream)
    ctx.synchronize()
    torch.cuda.synchronize()
    aff_ms = a0.elapsed_time(a1) / args.steps
    ctx.assemble(p_aff, sync=True)
    ams, _ = ctx.last_timing()
    out["affine_tensor_path"] = {"value": n_elem / (aff_ms * 1e-3), "unit": "elements/s", "ms_per_step": aff_ms,
                                 "kernel_ms": {"element": ams[0], "constrain": ams[3], "row_reduce": ams[1]},
                                 "note": "opt-in (feasm_params.affine_tensor); the headline value above runs the reference's quadrature loop on every element"}
    # Extra: the step after the path — y = A x on the assembled matrix, against the HBM roofline
    dx = torch.ones(P["n_dofs"], dtype=torch.float64, device="cuda")
    dy = torch.empty_like(dx)
    torch.cuda.synchronize()
    for _ in range(3):
        ctx.vector_mult(dx.data_ptr(), dy.data_ptr())
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record(stream)
    for _ in range(10):
        ctx.vector_mult(dx.data_ptr(), dy.data_ptr())
    s1.record(stream)
    ctx.synchronize()
    torch.cuda.synchronize()
    spmv_ms = s0.elapsed_time(s1) / 10
    spmv_bytes = nnz * 12 + P["n_dofs"] * 32  # values + col_idx, row_ptr, x, y
    out["vector_mult"] = {"ms": spmv_ms, "bytes": int(spmv_bytes), "gbs": spmv_bytes / (spmv_ms * 1e-3) / 1e9,
                          "hbm_frac": spmv_bytes / (spmv_ms * 1e-3) / 1e9 / hbm_peak}
    del dx, dy
    # ... and the solve of the system just assembled (CG + Jacobi on the device), checked against the exact solution of
    # ex4 at the nodes: an end-to-end validity check of the assembled numbers at the full size
    ctx.assemble(params, sync=True)
    t_s = time.perf_counter()
    u, its, rel = ctx.cg_solve(rel_tol=1e-10, max_its=4000)
    t_s = time.perf_counter() - t_s
    X = P["xyz"]
    exact = np.cos(.5 * np.pi * X[:, 0]) * np.sin(.5 * np.pi * X[:, 1]) * np.cos(.5 * np.pi * X[:, 2])
    out["cg_solve"] = {"iterations": its, "rel_residual": rel, "seconds": t_s,
                       "max_nodal_error_vs_exact_solution": float(np.abs(u[P["node_base"]] - exact).max())}
    # ... and the residual without the global matrix (feasm_residual) at the CG solution: element phase +
    # one row-parallel pass over the records; its norm against the right-hand side checks both at the full size
    du = torch.from_numpy(u).cuda()
    dr = torch.empty_like(du)
    torch.cuda.synchronize()
    L = F.lib()
    import ctypes as C
    for _ in range(2):
        ctx._ck(L.feasm_residual(ctx.h, C.byref(params), C.c_void_p(du.data_ptr()), C.c_void_p(dr.data_ptr())))
    r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    r0.record(stream)
    for _ in range(5):
        ctx._ck(L.feasm_residual(ctx.h, C.byref(params), C.c_void_p(du.data_ptr()), C.c_void_p(dr.data_ptr())))
    r1.record(stream)
    ctx.synchronize()
    torch.cuda.synchronize()
    _, rhs_now = ctx.get_values(values=False)
    out["residual"] = {"ms": r0.elapsed_time(r1) / 5, "rel_norm_at_cg_solution": float(torch.linalg.vector_norm(dr).item() / np.linalg.norm(rhs_now)),
                       "note": "R = K u - F from the element records: element kernel + constraint pass + k_row_residual, no CSR values"}
    del du, dr


# ---- our arm, N GPUs (one process per GPU) ---------------------------------------------------------------------------
def run_multi(args, F, rank, world, local):
    """One rank of `bench.py --gpus N` (launched by torch.distributed.run): strong scaling of the BASELINE.json job —
    the same build_cube mesh split over N GPUs — timed with CUDA events on the library's stream, max over ranks."""
    import torch
    import torch.distributed as dist

    from libmesh_b200.multi import DevArray, DistributedAssembly, RankProblem, linear_partition

    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    host_group = dist.new_group(backend="gloo")
    etype = ELEM[args.elem]
    elast = args.problem == "elasticity"
    t0 = time.time()
    n_side, _, n_vars = problem_dims(args)
    n_elem = F.cube_sizes(etype, n_side, n_side, n_side)[0]
    pid = linear_partition(n_elem, world)
    P = build_problem(F, args, world, pid)  # replicated host mesh, like libMesh's ReplicatedMesh
    n_dofs, base, first = P["n_dofs"], P["node_base"], P["first_dof"]
    prob = RankProblem(rank, world, P["xyz"], P["conn"], base, first, pid, n_vars, P["con_dofs"], P["con_vals"])
    ctx = F.Context(local)
    dev_group = None
    native = os.environ.get("FEASM_NATIVE_COMM", "1") != "0"  # 1: the library issues ncclSend/ncclRecv itself (feasm_exchange_begin / _end)
    if native:
        ids = [F.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0, group=host_group)  # torch.distributed only carries the 128 bootstrap bytes
        ctx.comm_init(world, rank, ids[0])
    elif os.environ.get("FEASM_OVERLAP", "1") != "0":
        # the transfer should take SM slots as they free up while the owned rows are being reduced: NCCL's stream at
        # high priority
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
        dev_group = dist.new_group(backend="nccl", pg_options=opts)
    da = DistributedAssembly(ctx, prob, dist, host_group=host_group, dev_group=dev_group)
    da.setup()
    if elast:
        # the traction sides of this rank's elements, in local element numbering (RankProblem keeps the global order)
        loc = np.searchsorted(prob.elems, P["side_elems"])
        here = (loc < len(prob.elems)) & (prob.elems[np.minimum(loc, len(prob.elems) - 1)] == P["side_elems"])
        ctx.set_side_loads(loc[here].astype(np.uint32), P["side_sides"][here], P["side_g"])
    ctx.synchronize()
    setup_s = time.time() - t0
    params = make_params(F, args)
    stream = torch.cuda.ExternalStream(ctx.stream())
    overlap = os.environ.get("FEASM_OVERLAP", "1") != "0"  # 0: feasm_assemble, then pack / exchange / unpack in sequence
    for _ in range(args.warmup):
        da.assemble(params, sync=False, overlap=overlap)
    ctx.synchronize()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    ev0.record(stream)
    th = time.perf_counter()
    for _ in range(args.steps):
        da.assemble(params, sync=False, overlap=overlap)
    host_ms = (time.perf_counter() - th) * 1e3 / args.steps  # host time to ENQUEUE one step (launch-bound check)
    ev1.record(stream)
    ctx.synchronize()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms.item()) / args.steps
    # per-kernel times of this rank's last step (events inside the library)
    da.assemble(params, sync=True, overlap=overlap)
    kms, nl = ctx.last_timing()
    nl += da.unpack_launches() + (0 if overlap else 1)  # the unpack launch(es) and the separate pack
    kall = [torch.zeros(4, device="cuda") for _ in range(world)]  # per-rank kernel times of that step
    dist.all_gather(kall, torch.tensor([float(x) for x in kms], device="cuda"))
    kall = [[round(float(v), 4) for v in k.tolist()] for k in kall]
    clk = clocks.stop() if rank == 0 else None

    # ---- parity of the REAL multi-GPU result: every rank assembles the WHOLE mesh once on its own
    # GPU (one context, no exchange) and compares its owned rows of the distributed result with the same rows of that
    # matrix on the device.  Rows all of whose contributors are local are summed in the same order in both runs and must
    # be bitwise equal; interface rows add the peers' partial sums in peer order, so they may differ by rounding only.
    parity = None
    if not args.no_check:
        parity = multi_parity(F, ctx, da, prob, P, local, params, torch, dist)

    # e2e: host buffers through the C ABI on every rank: coordinates + Dirichlet data in, the rank's OWNED rows out (the
    # ghost rows were added into their owners by the exchange).  The result leaves through the device-pointer accessors:
    # a device-side snapshot on the library's stream, then the PCIe read on a second stream, so the read of step i
    # overlaps the upload + assembly + exchange of step i+1 (the same overlap as the single-GPU e2e leg).
    nnz_local = int(da.row_ptr[-1])
    nnz_owned = int(da.row_ptr[prob.n_owned])
    xyz_pin = torch.from_numpy(prob.xyz).pin_memory().numpy()
    val_pin = torch.empty(max(nnz_owned, 1), dtype=torch.float64).pin_memory()
    rhs_pin = torch.empty(max(prob.n_owned, 1), dtype=torch.float64).pin_memory()
    d_val = torch.as_tensor(DevArray(ctx.device_values(), max(nnz_owned, 1)), device="cuda")
    d_rhs = torch.as_tensor(DevArray(ctx.device_rhs(), max(prob.n_owned, 1)), device="cuda")
    snap_val, snap_rhs = torch.empty_like(d_val), torch.empty_like(d_rhs)
    copy_stream = torch.cuda.Stream()
    ev_ready, ev_done = torch.cuda.Event(), torch.cuda.Event()
    ev_done.record(copy_stream)

    def e2e_step():
        ctx.update_coordinates(xyz_pin)
        ctx.set_constraints(prob.con_rows, prob.con_vals)
        da.assemble(params, sync=False, overlap=overlap)
        with torch.cuda.stream(stream):
            stream.wait_event(ev_done)  # the previous read has left the snapshot
            snap_val.copy_(d_val, non_blocking=True)
            snap_rhs.copy_(d_rhs, non_blocking=True)
            ev_ready.record(stream)
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ev_ready)
            val_pin.copy_(snap_val, non_blocking=True)
            rhs_pin.copy_(snap_rhs, non_blocking=True)
            ev_done.record(copy_stream)

    e2e_step()
    torch.cuda.synchronize()
    ctx.synchronize()
    dist.barrier()
    te = time.perf_counter()
    e2e_steps = max(3, min(args.steps, 8))
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    ctx.synchronize()  # raises if any step flagged a degenerate element
    dist.barrier()
    e2e_s = torch.tensor([(time.perf_counter() - te) / e2e_steps], device="cuda")
    dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    csum = torch.tensor([float(val_pin[:nnz_owned].sum().item())], dtype=torch.float64, device="cuda")  # owned rows tile the global matrix
    dist.all_reduce(csum)
    io = torch.tensor([float(xyz_pin.nbytes + prob.con_rows.nbytes + prob.con_vals.nbytes),
                       float(nnz_owned * 8 + prob.n_owned * 8), float(len(da.send_slots) * 8), float(nnz_local)], device="cuda")
    dist.all_reduce(io)
    # ---- the distributed matrix used on the device (feasm_halo.cu): SpMV with the halo exchange of x, and the CG solve ----
    solve = None
    if native:
        try:
            solve = multi_solve(F, ctx, da, prob, P, params, elast, torch, dist, stream)
        except Exception as e:  # the assembly line above stays valid; the failure is reported in it
            solve = {"error": f"{type(e).__name__}: {e}"}
    if rank == 0:
        nq = nq_of(args)
        nnz_g = int(io[3].item())  # includes ghost-row copies
        flops_e, bytes_e = algorithmic_counts(etype, nq, nnz_g, n_dofs, n_elem, P["xyz"].shape[0], problem=args.problem)
        kname = "k_elem_elasticity" if elast else "k_elem_poisson"
        out = {
            "metric": metric_name(args), "value": n_elem / (ms_per_step * 1e-3), "unit": "elements/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args) + f"; LinearPartitioner slabs over {world} GPUs, ghost rows by ncclSend/ncclRecv " + ("issued by libfeasm" if native else "through torch.distributed")
                                   + (" overlapped with the reduction of the owned rows" if overlap else ""), "n_elem": n_elem,
                       "n_dofs": int(n_dofs), "l2": "per-rank CSR values larger than L2", "setup_s": round(setup_s, 3),
                       "ghost_bytes_per_step_all_ranks": int(io[2].item()), "host_enqueue_ms_per_step_rank0": round(host_ms, 4)},
            "clocks": clk, "gpu_launches": int(nl) * args.steps,
            "e2e": {"value": n_elem / float(e2e_s.item()), "unit": "elements/s", "h2d_bytes_per_step": int(io[0].item()),
                    "d2h_bytes_per_step": int(io[1].item()), "ms_per_step": float(e2e_s.item()) * 1e3,
                    "checksum_values": float(csum.item())},
            "roofline": {"bound": "tensor", "kernel": f"{kname} on rank 0 (FP64 DMMA contraction + geometry)",
                         "achieved": flops_e * len(prob.elems) / (kms[0] * 1e-3) / 1e12, "peak": None, "unit": "TFLOP/s", "frac": None,
                         "traffic": None, "kernel_ms": {"element": kms[0], "row_reduce": kms[1]}, "flops_per_element": flops_e,
                         "kernel_ms_per_rank": {"element": [k[0] for k in kall], "constrain": [k[3] for k in kall],
                                                "row_reduce": [k[1] for k in kall], "assemble": [k[2] for k in kall]}},
        }
        if parity is not None:
            out["parity"] = parity
        if solve is not None:
            out["distributed_solve"] = solve
        fp64 = ctx.measure_fp64_peaks()
        pk = max(fp64["dmma_tflops"], fp64["dfma_tflops"])
        out["roofline"]["peak"], out["roofline"]["frac"] = pk, out["roofline"]["achieved"] / pk
        out["roofline"]["peak_source"] = "FP64 mma.sync/DFMA chain measured live (MEASURED_PEAKS.json has no FP64 entry)"
        print(json.dumps(out), flush=True)
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()


def multi_solve(F, ctx, da, prob, P, params, elast, torch, dist, stream):
    """After the timed assembly: feasm_build_halo, then y = A x on the distributed matrix (halo exchange of x inside the
    call; CUDA events on the library's stream, max over ranks) and, for the Poisson job, feasm_cg_solve over all ranks with
    the nodal error against the exact ex4 solution.  Collective: every rank calls it."""
    import ctypes as C

    da.assemble(params, sync=True)  # the e2e leg left the last assembled system in place; make that explicit
    t0 = time.perf_counter()
    ctx.build_halo(prob.first_dof)
    halo_s = time.perf_counter() - t0
    peers, sp, _, rp, _ = ctx.halo_plan()
    n = prob.n_owned
    L = ctx.L
    d_x, d_y = L.feasm_vec_alloc(ctx.h, n), L.feasm_vec_alloc(ctx.h, n)
    xo = np.cos(0.37 * (prob.f0 + np.arange(n))) + 0.1
    L.feasm_vec_copy(ctx.h, 0, C.c_void_p(d_x), xo.ctypes.data_as(C.c_void_p), n)
    for _ in range(3):
        ctx.vector_mult(d_x, d_y)
    ctx.synchronize()
    torch.cuda.synchronize()
    dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    ev0.record(stream)
    for _ in range(reps):
        ctx.vector_mult(d_x, d_y)
    ev1.record(stream)
    ctx.synchronize()
    torch.cuda.synchronize()
    ms = torch.tensor([ev0.elapsed_time(ev1) / reps], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    y = np.empty(n)
    L.feasm_vec_copy(ctx.h, 1, y.ctypes.data_as(C.c_void_p), C.c_void_p(d_y), n)
    L.feasm_vec_free(ctx.h, C.c_void_p(d_x))
    L.feasm_vec_free(ctx.h, C.c_void_p(d_y))
    nnz_owned = int(da.row_ptr[n])
    tot = torch.tensor([float(nnz_owned), float(int(rp[-1])), float(np.sum(y)), float(np.sum(np.abs(y)))], dtype=torch.float64, device="cuda")
    dist.all_reduce(tot)
    bytes_all = tot[0].item() * 12 + P["n_dofs"] * 24
    out = {"halo_setup_s": round(halo_s, 3), "halo_entries_all_ranks": int(tot[1].item()),
           "vector_mult": {"ms": float(ms.item()), "bytes_all_ranks": int(bytes_all), "gbs_all_ranks": bytes_all / (float(ms.item()) * 1e-3) / 1e9,
                           "checksum_y": tot[2].item(), "checksum_abs_y": tot[3].item()}}
    if not elast:
        t0 = time.perf_counter()
        u, its, rel = ctx.cg_solve(rel_tol=1e-10, max_its=4000)
        t_s = time.perf_counter() - t0
        node_of_dof = np.empty(P["n_dofs"], dtype=np.int64)
        node_of_dof[P["node_base"].astype(np.int64)] = np.arange(len(P["node_base"]))
        pts = P["xyz"][node_of_dof[prob.f0:prob.f0 + n]]
        exact = np.cos(0.5 * np.pi * pts[:, 0]) * np.sin(0.5 * np.pi * pts[:, 1]) * np.cos(0.5 * np.pi * pts[:, 2])
        err = torch.tensor([float(np.abs(u - exact).max()) if n else 0.0, t_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(err, op=dist.ReduceOp.MAX)
        out["cg_solve"] = {"iterations": its, "rel_residual": rel, "seconds": float(err[1].item()),
                           "max_nodal_error_vs_exact_solution": float(err[0].item()),
                           "note": "Jacobi-preconditioned CG over all ranks: halo exchange of the direction per iteration, dot products finished by an all-gather added in rank order"}
    return out


def multi_parity(F, ctx, da, prob, P, local, params, torch, dist):
    """Owned rows of the distributed result against the same rows of a one-GPU assembly of the whole mesh, compared on
    the device; returns the summary rank 0 prints (all ranks take part in the reductions)."""
    from libmesh_b200.multi import DevArray
    from oracle.parity import FLOOR, TOL  # the one parity rule: the checker's constants only, nothing of oracle/ runs here

    n_vars = P["n_vars"]
    full = F.Context(local)
    full.set_mesh(P["etype"], P["xyz"], P["conn"])
    full.set_dofs(F.elem_dofs_from_nodes(P["conn"], P["node_base"], n_vars), P["n_dofs"], n_vars=n_vars)
    full.set_constraints(P["con_dofs"].astype(np.uint32), P["con_vals"])
    if "side_elems" in P:
        full.set_side_loads(P["side_elems"], P["side_sides"], P["side_g"])
    nnz_full = full.build_pattern()
    full.assemble(params, sync=True)
    rp_full, ci_full = full.get_pattern()
    f0, n_own = prob.f0, prob.n_owned
    lo, hi = int(rp_full[f0]), int(rp_full[f0 + n_own])
    pattern_ok = bool(np.array_equal(da.row_ptr[: n_own + 1].astype(np.int64), rp_full[f0:f0 + n_own + 1].astype(np.int64) - lo)
                      and np.array_equal(da.col_idx[: hi - lo], ci_full[lo:hi]))
    del ci_full
    g_val = torch.as_tensor(DevArray(full.device_values(), nnz_full), device="cuda")[lo:hi]
    g_rhs = torch.as_tensor(DevArray(full.device_rhs(), P["n_dofs"]), device="cuda")[f0:f0 + n_own]
    d_val = torch.as_tensor(DevArray(ctx.device_values(), max(hi - lo, 1)), device="cuda")[: hi - lo]
    d_rhs = torch.as_tensor(DevArray(ctx.device_rhs(), max(n_own, 1)), device="cuda")[:n_own]
    diff = (d_val - g_val).abs()
    mag = torch.maximum(d_val.abs(), g_val.abs())
    lens = torch.from_numpy(np.diff(da.row_ptr[: n_own + 1].astype(np.int64))).cuda()
    row_of = torch.repeat_interleave(torch.arange(n_own, device="cuda"), lens)
    rowmax = torch.zeros(n_own, dtype=torch.float64, device="cuda").scatter_reduce(0, row_of, g_val.abs(), reduce="amax")
    scale = torch.maximum(mag, FLOOR * rowmax[row_of])
    scale[scale == 0] = 1.0
    # rows that receive ghost contributions (the RHS slots of the receive plan name them)
    recv_rows = np.zeros(n_own, dtype=bool)
    rs = da.recv_slots.astype(np.int64)
    nnz_loc = int(da.row_ptr[-1])
    recv_rows[rs[rs >= nnz_loc] - nnz_loc] = True
    is_recv = torch.from_numpy(recv_rows).cuda()
    neq = d_val != g_val
    stats = torch.tensor([float(neq.sum().item()), float((neq & ~is_recv[row_of]).sum().item()), float((d_rhs != g_rhs).sum().item()),
                          float(hi - lo), float(recv_rows.sum()), float(n_own), 0.0 if pattern_ok else 1.0], dtype=torch.float64, device="cuda")
    rscale = torch.maximum(torch.maximum(d_rhs.abs(), g_rhs.abs()), FLOOR * g_rhs.abs().max())
    rscale[rscale == 0] = 1.0
    mx = torch.tensor([float((diff / scale).max().item()) if hi > lo else 0.0,
                       float(((d_rhs - g_rhs).abs() / rscale).max().item()) if n_own else 0.0], dtype=torch.float64, device="cuda")
    dist.all_reduce(stats)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    gsum = torch.tensor([float(g_val.sum().item())], dtype=torch.float64, device="cuda")
    dist.all_reduce(gsum)
    del g_val, g_rhs, diff, mag, scale, row_of, rowmax
    full.close()
    st = stats.tolist()
    return {"checker": "one-GPU assembly of the whole mesh by libfeasm on every rank's own GPU, owned rows compared on the device",
            "pattern_of_owned_rows_bit_equal": st[6] == 0.0, "entries_compared": int(st[3]), "entries_not_bitwise_equal": int(st[0]),
            "of_which_in_rows_without_ghost_contributions": int(st[1]), "rhs_entries_not_bitwise_equal": int(st[2]),
            "rows_with_ghost_contributions": int(st[4]), "owned_rows": int(st[5]),
            "max_rel_err_values": float(mx[0].item()), "max_rel_err_rhs": float(mx[1].item()),
            "passes_1e-12": bool(mx[0].item() <= TOL and mx[1].item() <= TOL and st[1] == 0.0 and st[6] == 0.0),
            "checksum_values_one_gpu_rows": float(gsum.item()),
            "note": "interface rows add the peers' partial sums in peer order, so they differ from the one-GPU element order by rounding only"}


# ---- the reference arm ------------------------------------------------------------------------------------------------
def _ref_rank(w, n_ranks, etype, n, bounds, order, warmup, steps, barrier, out):
    """One emulated MPI rank of the reference (a fresh process): the replicated mesh, its slab of elements (LinearPartitioner
    blocks), the nodes they touch, local dofs and pattern from the oracle, and the reference's own assemble_poisson
    (oracle/_ref/libref_app_fast.so) over them."""
    try:
        from oracle import oracle as O
        from oracle import ref as R

        gm = O.Mesh(etype, n, n, n, bounds=bounds, fast=True)
        xyz, conn, mask = gm.arrays()
        del gm
        ne = conn.shape[0]
        blk = ne // n_ranks
        e0, e1 = (0 if w == 0 else blk * w + ne % n_ranks), blk * (w + 1) + ne % n_ranks
        nodes = np.unique(conn[e0:e1])
        lconn = np.searchsorted(nodes, conn[e0:e1]).astype(np.uint32)
        m = O.Mesh.from_arrays(etype, xyz[nodes], lconn, mask[nodes], fast=True)
        s = O.System(m, 1)
        s.add_dirichlet()
        s.build_sparsity()
        R.use_fast(True)
        nnz, nd = R.load_world(m, s, (0, 0, 0), (0.0,) * 6)  # has_affine_map: the element class's own test, per element
        times = []
        barrier.wait(timeout=180)
        for i in range(warmup + steps):
            R.assemble_ex4(order, nnz, nd)
            if i >= warmup:
                times.append(R.last_solve_seconds())
        out.put((w, e1 - e0, times, None))
    except Exception as e:  # reported by the parent
        try:
            barrier.abort()
        except Exception:
            pass
        out.put((w, 0, [], f"{type(e).__name__}: {e}"[:300]))


def reference_code_ranks(args, warmup: int = 1, steps: int = 2):
    """The reference's OWN assembly function on all host cores, the way libMesh uses them: one process per core, each assembling its
    contiguous block of the mesh's elements (LinearPartitioner) with assemble_poisson of introduction_ex4.C compiled where it lies
    (oracle/_ref, -O3 build) — local matrix and right-hand side per process; the exchange of ghost rows that follows in an MPI run
    is NOT counted.  Rate = elements of the whole job / the slowest process's mean step time."""
    import multiprocessing as mp

    from oracle import oracle as O
    from oracle import ref as R

    if args.problem != "poisson" or not R.app_available(fast=True):
        return None
    n, b, _ = problem_dims(args)
    etype = ELEM[args.elem]
    n_elem = {10: 24, 18: 2}.get(etype, 1) * n ** 3
    n_ranks = max(1, min(os.cpu_count() or 1, n_elem // 32))  # a rank without elements is not a rank
    ctx = mp.get_context("spawn")  # fresh processes: nothing of this one's OpenMP or CUDA state is inherited
    barrier, out = ctx.Barrier(n_ranks), ctx.Queue()
    order = 1 if args.elem == "hex8" else 2
    procs = [ctx.Process(target=_ref_rank, args=(w, n_ranks, etype, n, b, order, warmup, steps, barrier, out)) for w in range(n_ranks)]
    t0 = time.perf_counter()
    for p in procs:
        p.start()
    res, deadline = [], t0 + 300.0  # a data point beside the port: bounded, and never allowed to stall the arm
    try:
        for _ in procs:
            res.append(out.get(timeout=max(deadline - time.perf_counter(), 1.0)))
    except Exception:
        res.append((-1, 0, [], "a rank process did not answer within 300 s"))
    for p in procs:
        p.join(timeout=5)
        if p.is_alive():
            p.terminate()
    errs = [r[3] for r in res if r[3]]
    if errs:
        return {"unavailable": errs[0]}
    slowest = max(float(np.mean(r[2])) for r in res)
    if sum(r[1] for r in res) != n_elem:
        return {"unavailable": f"the ranks assembled {sum(r[1] for r in res)} of {n_elem} elements"}
    return {"value": n_elem / slowest, "unit": "elements/s", "cores": n_ranks, "kind": "reference",
            "sample": f"{steps} steps after {warmup} warm-up: assemble_poisson of introduction_ex4.C in {n_ranks} processes, each on its block of "
                      f"{n_elem // n_ranks} of the {n_elem} elements (build_cube {n}^3), slowest process {slowest:.2f} s per step",
            "wall_s_with_setup": round(time.perf_counter() - t0, 1),
            "note": "ghost-row exchange of an MPI run not counted; stub mesh / matrix world of oracle/ref_app/ref_world.h"}


def run_reference(args):
    """The reference's CPU path on all host cores, on the SAME configuration as our arm.  Two implementations are run and the
    FASTER one is the arm's value: the oracle port (CPU restatement, OpenMP over elements; every step one assembly of the whole
    mesh) and the reference's own code (oracle/_ref: assemble_poisson of introduction_ex4.C compiled where it lies, one process per
    core over element blocks, a bounded number of steps), reported beside it in cpu_baseline."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    try:
        ref_code = reference_code_ranks(args)
    except Exception as e:
        ref_code = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    O, m, s = oracle_system(args, fast=True)
    for _ in range(max(args.warmup, 1)):
        oracle_assemble(O, s, args, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle_assemble(O, s, args, threads)
    dt = (time.perf_counter() - t0) / args.steps
    value = m.n_elem / dt
    sample = f"each step = one assembly of the full job ({m.n_elem} elements), {threads} OpenMP threads"
    kind, cores = "port", threads
    port = {"value": value, "unit": "elements/s", "cores": threads, "sample": sample}
    if ref_code and ref_code.get("value", 0.0) > value:  # the arm's value is the faster CPU implementation
        value, dt, kind, cores, sample = ref_code["value"], m.n_elem / ref_code["value"], "reference", ref_code["cores"], ref_code["sample"]
    print(json.dumps({
        "impl": "reference", "metric": metric_name(args), "value": value, "unit": "elements/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(args), "n_elem": int(m.n_elem), "n_dofs": int(s.n_dofs), "nnz": int(s.nnz),
                   "note": "CPU assembly loop of the reference: the faster of the oracle port (g++ -O3 -march=x86-64-v3, OpenMP over elements) "
                           "and the reference's own code from oracle/_ref (one process per core over element blocks)"},
        "cpu_baseline": {"value": value, "unit": "elements/s", "cores": cores, "kind": kind, "sample": sample, "port": port,
                         "reference_code": ref_code},
        "e2e": {"value": value, "unit": "elements/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--problem", default="poisson", choices=["poisson", "elasticity"])
    ap.add_argument("--n", type=int, default=0, help="build_cube elements per side (0 = 96 for poisson, 64 for elasticity)")
    ap.add_argument("--elem", default="hex27", choices=sorted(ELEM))
    ap.add_argument("--qorder", type=int, default=0, help="QGauss order (0 = the example's FIFTH)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-check", action="store_true", help="skip the full-size parity check against the oracle")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
