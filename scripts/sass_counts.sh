#!/bin/bash
# Per-kernel SASS mnemonic counts of libfeasm.so -> markdown table (profiles/r2/sass_counts.md)
cd "$(dirname "$0")/.."
cuobjdump -sass libmesh_b200/libfeasm.so 2>/dev/null | python -c "
import sys,re,collections,subprocess
cur=None; cnt=collections.OrderedDict()
for l in sys.stdin:
    m=re.search(r'Function : (\S+)',l)
    if m:
        cur=m.group(1); cnt[cur]=collections.Counter(); continue
    if cur is None: continue
    m=re.search(r'^\s+/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)',l)
    if m:
        op=m.group(2); cnt[cur]['total']+=1
        for k in ('DMMA','UBLKCP','LDGSTS','DFMA','DMUL','DADD','STG','LDG','LDS','STS','SHFL','SYNCS','ATOMG','RED'):
            if op.startswith(k): cnt[cur][k]+=1
print('| kernel | SASS instructions | DMMA | DFMA+DMUL+DADD | UBLKCP | SYNCS (mbarrier) | LDGSTS | LDG | STG | LDS | STS | SHFL | ATOMG/RED |')
print('|---|---|---|---|---|---|---|---|---|---|---|---|---|')
for k,c in cnt.items():
    name=subprocess.run(['c++filt',k],capture_output=True,text=True).stdout.strip()
    name=re.sub(r'\(anonymous namespace\)::','',name).split('(')[0].replace('void ','')
    print(f\"| {name} | {c['total']} | {c['DMMA']} | {c['DFMA']+c['DMUL']+c['DADD']} | {c['UBLKCP']} | {c['SYNCS']} | {c['LDGSTS']} | {c['LDG']} | {c['STG']} | {c['LDS']} | {c['STS']} | {c['SHFL']} | {c['ATOMG']+c['RED']} |\")
"
