#!/bin/bash
# usage: gpu_launchlist.sh <workload> <kernel-regex> <tag>: plain run, then the ncu launch list, aggregated per kernel
W=$1; K=$2; TAG=$3
CMD="python bench.py --workload $W --steps 2 --warmup 1 --no-cpu --no-ess --e2e-steps 1"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$K" -c 4000 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
python - <<PY
import csv, collections
rows=list(csv.reader(open("gpurun_out/launches_$TAG.csv")))
hdr=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
H=rows[hdr]; ki,vi,ui=H.index("Kernel Name"),H.index("Metric Value"),H.index("Metric Unit")
agg=collections.OrderedDict()
for r in rows[hdr+1:]:
    if len(r)<=vi: continue
    v=float(r[vi].replace(",",""))*{"ns":1,"us":1e3,"ms":1e6,"s":1e9}.get(r[ui],1)
    a=agg.setdefault(r[ki].split("(")[0],[0,0.0]); a[0]+=1; a[1]+=v
tot=sum(a[1] for a in agg.values())
for k,a in agg.items(): print(k,a[0],"%.2f ms"%(a[1]/1e6),"%.1f%%"%(100*a[1]/tot), "%.1f us/launch"%(a[1]/a[0]/1e3))
PY
