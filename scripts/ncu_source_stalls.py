"""Per-instruction stall samples of one kernel from `ncu -i rep --page source --csv --kernel-name regex:NAME > file.csv`.
usage: ncu_source_stalls.py file.csv stall_long_sb [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if 'Source' in r and '# Samples' in r)
hdr = rows[hi]
body = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[hdr.index('# Samples')].isdigit()]
col, top = sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 30
ia, il, isamp = hdr.index('Source'), hdr.index(col), hdr.index('# Samples')
f = lambda x: int(x) if x.isdigit() else 0
print('instructions', len(body), 'total samples', sum(f(r[isamp]) for r in body), col, sum(f(r[il]) for r in body))
best = sorted(range(len(body)), key=lambda i: -f(body[i][il]))[:top]
for i in sorted(best):
    print(i, body[i][il], body[i][isamp], body[i][ia][:110])
