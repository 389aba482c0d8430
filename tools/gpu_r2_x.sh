#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
CBX_GEMM_WS=0 CFM_SAVE=/tmp/b0.pt timeout 300 python tools/cfm_batch_time.py 16 2>/dev/null | tail -2
CFM_SAVE=/tmp/b1.pt timeout 300 python tools/cfm_batch_time.py 16 2>/dev/null | tail -2
python - <<'PY'
import torch
a, b = torch.load("/tmp/b0.pt"), torch.load("/tmp/b1.pt")
print("weight-stationary (staged stores) vs tile-per-CTA GEMM: equal %s" % bool(torch.equal(a, b)))
PY
timeout 300 python tools/cfm_batch_time.py 64 2>/dev/null | tail -2
CBX_CFM_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1000 --launch-count 340 --csv --log-file gpurun_out/cfm_b64_launches.csv python tools/cfm_batch_time.py 64 > gpurun_out/ncu_b64.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/cfm_b64_launches.csv")) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit"); gi=h.index("Grid Size")
agg=collections.OrderedDict()
for r in rows[1:]:
    if not r[mi].startswith("gpu__time_duration"): continue
    v=float(r[vi].replace(",",""))
    if r[ui]=="ns": v/=1e3
    elif r[ui]=="ms": v*=1e3
    k=(r[ki][:48], r[gi])
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=v
tot=sum(a[1] for a in agg.values())
for k,a in sorted(agg.items(), key=lambda x:-x[1][1])[:12]:
    print("%-50s %-16s %5d %9.1f us %8.2f ms %5.1f%%" % (k[0],k[1],a[0],a[1]/a[0],a[1]/1e3,100*a[1]/tot))
print("total ms", tot/1e3)
PY
