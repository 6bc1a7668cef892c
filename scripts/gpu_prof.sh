#!/bin/bash
# usage: gpu_prof.sh <workload|gram> <kernel-regex> <tag> [skip] : plain run, then (1) the ncu launch list of the same command and
# (2) one --set full capture of the named kernel (launch number skip+1 of that kernel).  The .ncu-rep (~50 MB with sources) is
# reduced ON THE BOX to the raw-metric CSV and the per-source-line stall summary, because gpurun only brings back 64 MiB.
W=$1; K=$2; TAG=$3; SKIP=${4:-4}
mkdir -p gpurun_out
if [ "$W" = "gram" ]; then CMD="python scripts/gram_probe.py"; else CMD="python bench.py --workload $W --steps 3 --warmup 3 --no-cpu --no-ess --no-sub --e2e-steps 1"; fi
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -o /tmp/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
if [ -f /tmp/prof_$TAG.ncu-rep ]; then
  ncu -i /tmp/prof_$TAG.ncu-rep --page raw --csv > gpurun_out/raw_$TAG.csv 2>/dev/null
  ncu -i /tmp/prof_$TAG.ncu-rep --page source --csv --print-source cuda,sass > /tmp/src_$TAG.csv 2>/dev/null
  python scripts/ncu_lines.py /tmp/src_$TAG.csv 80 > gpurun_out/lines_$TAG.txt 2>&1
fi
tail -2 gpurun_out/ncu_full_$TAG.log
