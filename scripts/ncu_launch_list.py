"""ncu --metrics gpu__time_duration.sum --csv log -> markdown launch list + share of a step.
usage: ncu_launch_list.py launches.csv "title" [step-kernel-regex ...] > out.md"""
import csv
import re
import sys

path, title = sys.argv[1], sys.argv[2]
pats = sys.argv[3:] or ["k_elem_poisson", "k_constrain_list", "k_row_reduce"]
lines = [l for l in open(path, errors="replace") if l.startswith('"')]
rows = list(csv.DictReader(lines))
print(f"# ncu launch list — {title}\n")
print("`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare shares, not absolutes); launched after "
      "the same command exited 0 without ncu.\n")
print("| id | kernel | block | grid | us |\n|---|---|---|---|---|")
tot = {}
for r in rows:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
    name = re.sub(r"\(.*", "", r["Kernel Name"])[-70:]
    print(f"| {r['ID']} | {name} | {r['Block Size']} | {r['Grid Size']} | {us:.1f} |")
    for p in pats:
        if re.search(p, r["Kernel Name"]):
            tot.setdefault(p, []).append(us)
print("\n## Share of one step (mean over the launches above)\n\n| kernel | launches | mean us | share |\n|---|---|---|---|")
s = sum(sum(v) / len(v) for v in tot.values())
for p, v in tot.items():
    print(f"| {p} | {len(v)} | {sum(v)/len(v):.1f} | {100*sum(v)/len(v)/s:.1f} % |")
