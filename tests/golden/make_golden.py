"""Extracts the golden facts the reference publishes for the assembly path into tests/golden/reference_runs.json.

Run in the build container (needs /root/reference); the JSON it writes is committed and is all the tests read.
Sources: the captured stdout of real libMesh runs embedded in the generated example pages
  doc/html/examples/introduction_ex4.html        (`-d 3 -n 6`: 216 HEX27, SECOND LAGRANGE, Dirichlet on ids 0-5)
  doc/html/examples/systems_of_equations_ex6.html (`-nx 40 -ny 10 -nz 5`: 2000 HEX8, 3 x FIRST LAGRANGE)
These are outputs of the reference itself, not of the oracle.
"""
import json
import os
import re

REF = "/root/reference/doc/html/examples"


def text(name):
    raw = open(os.path.join(REF, name), encoding="utf-8", errors="replace").read()
    return "\n".join(re.sub(r"<[^>]*>", "", line) for line in raw.split("\n")).replace("&lt;", "<")


def grab(block, key, cast=float):
    m = re.search(re.escape(key) + r"\s*(?:<=|=)\s*([0-9.eE+-]+)", block)
    return cast(m.group(1))


def run_block(t, marker):
    i = t.index(marker)
    return t[i:i + 4000]


def main():
    out = {}
    t = text("introduction_ex4.html")
    b = run_block(t, "example-opt -d 3 -n 6")
    perf = t[t.index("Matrix Assembly Performance", t.index("example-opt -d 3 -n 6")):][:200]
    out["introduction_ex4_d3_n6"] = {
        "n_nodes": grab(b, "n_nodes()", int), "n_elem": grab(b, "n_elem()", int), "n_dofs": grab(b, "n_dofs()", int),
        "n_constrained_dofs": grab(b, "n_constrained_dofs()", int),
        "n_heterogeneous_constraints": grab(b, "Number of Heterogenous Constraints", int),
        "avg_on_proc_bandwidth": grab(b, "Average  On-Processor Bandwidth"),
        "max_on_proc_bandwidth": grab(b, "Maximum  On-Processor Bandwidth", int),
        "perflog_assembly_active_s": float(re.search(r"Active time=([0-9.]+)", perf).group(1)),
    }
    t = text("systems_of_equations_ex6.html")
    b = run_block(t, "Running Example systems_of_equations_ex6")
    out["systems_of_equations_ex6_40x10x5"] = {
        "n_nodes": grab(b, "n_nodes()", int), "n_elem": grab(b, "n_elem()", int), "n_dofs": grab(b, "n_dofs()", int),
        "n_constrained_dofs": grab(b, "n_constrained_dofs()", int),
        "avg_on_proc_bandwidth": grab(b, "Average  On-Processor Bandwidth"),
        "max_on_proc_bandwidth": grab(b, "Maximum  On-Processor Bandwidth", int),
    }
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_runs.json")
    json.dump(out, open(path, "w"), indent=1, sort_keys=True)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
