#!/bin/bash
# usage: gpu_prof.sh <workload> <kernel-regex> <tag> [extra bench args]
# Plain run, then (1) the ncu launch list of the sweep kernels and (2) one --set full capture of the
# named kernel.  The .ncu-rep (~50 MB with sources) is reduced ON THE BOX to the raw-metric CSV and
# the per-source-line stall summary, because gpurun only brings back 64 MiB.
W=$1; K=$2; TAG=$3; shift 3
mkdir -p gpurun_out
CMD="python bench.py --workload $W --steps 3 --warmup 3 --no-cpu --no-ess --e2e-steps 1 $*"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s ${SKIP:-4} -c 1 -o /tmp/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
if [ -f /tmp/prof_$TAG.ncu-rep ]; then
  ncu -i /tmp/prof_$TAG.ncu-rep --page raw --csv > gpurun_out/raw_$TAG.csv 2>/dev/null
  ncu -i /tmp/prof_$TAG.ncu-rep --page details --csv > gpurun_out/details_$TAG.csv 2>/dev/null
  ncu -i /tmp/prof_$TAG.ncu-rep --page source --csv --print-source cuda,sass > /tmp/src_$TAG.csv 2>/dev/null
  python scripts/ncu_lines.py /tmp/src_$TAG.csv 80 > gpurun_out/lines_$TAG.txt 2>&1
  gzip -9 -c /tmp/src_$TAG.csv > gpurun_out/src_$TAG.csv.gz
fi
