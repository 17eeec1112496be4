"""Development aid: A/B several builds of the library in ONE process-per-build, interleaved rounds, same box.
    python scripts/gpu_ab.py "hex27 96 fd" libA.so libB.so ...     (paths relative to build/variants/, or 'product')"""
import subprocess, sys, re, collections
case = sys.argv[1].split()
libs = sys.argv[2:]
best = collections.defaultdict(lambda: [9e9] * 4)
for rnd in range(3):
    for lib in libs:
        env = dict(__import__("os").environ)
        if lib != "product":
            env["FEASM_LIB"] = __import__("os").path.abspath(f"build/variants/libfeasm_{lib}.so")
        out = subprocess.run([sys.executable, "scripts/gpu_time_quick.py"] + case, env=env, capture_output=True, text=True).stdout
        m = re.search(r"element ([\d.]+) ms\s+constrain ([\d.]+)\s+row_reduce ([\d.]+)\s+total ([\d.]+)", out)
        if not m:
            print(lib, "FAILED", out[-300:]); continue
        best[lib] = [min(a, float(b)) for a, b in zip(best[lib], m.groups())]
for lib in libs:
    e, c, r, t = best[lib]
    print(f"{lib:>16s} {' '.join(case)}: element {e:.3f}  constrain {c:.3f}  row_reduce {r:.3f}  total {t:.3f}")
