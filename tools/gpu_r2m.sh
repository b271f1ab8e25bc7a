#!/bin/bash
mkdir -p gpurun_out
AQUA_CHAIN=1 timeout 300 python tools/profile_ops.py --insitu --iters 20 --out gpurun_out/ops_chain_insitu.txt > /dev/null 2> gpurun_out/ops_insitu.err
timeout 300 python tools/profile_ops.py --insitu --iters 20 --out gpurun_out/ops_plain_insitu.txt > /dev/null 2>> gpurun_out/ops_insitu.err
python - <<'PY'
import re
def load(p):
    rows=[]
    for l in open(p):
        m=re.match(r"\s*(\d+)\s+([\d.]+)us.*?\|?\s*(conv|detect|sppf_pool)(.*?)\|", l)
        if m: rows.append((float(m.group(2)), (m.group(3)+m.group(4)).strip()[:48], '+1x1' in l))
    return rows
a=load('gpurun_out/ops_plain_insitu.txt'); b=load('gpurun_out/ops_chain_insitu.txt')
print("plain total %.1f us (%d ops), chained total %.1f us (%d ops)"%(sum(r[0] for r in a),len(a),sum(r[0] for r in b),len(b)))
i=j=0
while i<len(a) and j<len(b):
    if b[j][2]:
        print(f"{i:3d}+{i+1:<3d} plain {a[i][0]:7.1f} + {a[i+1][0]:6.1f} = {a[i][0]+a[i+1][0]:7.1f}   chained {b[j][0]:7.1f}   {b[j][1]}")
        i+=2; j+=1
    else:
        d=b[j][0]-a[i][0]
        if abs(d)>2.0: print(f"{i:3d}     plain {a[i][0]:7.1f}   chained-plan {b[j][0]:7.1f}  ({d:+.1f})  {a[i][1]}")
        i+=1; j+=1
PY
