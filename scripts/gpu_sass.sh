#!/bin/bash
# usage: gpu_sass.sh <workload> <kernel-regex> <tag> [skip] [min_exec_fraction]: one source-counter capture of the named kernel, reduced on the box to
# the per-instruction listing of its hot loop (scripts/ncu_sass.py)
W=$1; K=$2; TAG=$3; SKIP=${4:-4}; FR=${5:-0.5}
mkdir -p gpurun_out
CMD="python bench.py --workload $W --steps 3 --warmup 3 --no-cpu --no-ess --no-sub --e2e-steps 1"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --section SourceCounters --section WarpStateStats --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -o /tmp/sass_$TAG $CMD > gpurun_out/ncu_sass_$TAG.log 2>&1
if [ -f /tmp/sass_$TAG.ncu-rep ]; then
  ncu -i /tmp/sass_$TAG.ncu-rep --page source --csv --print-source sass > /tmp/sass_$TAG.csv 2>/dev/null
  head -c 3000 /tmp/sass_$TAG.csv > gpurun_out/sass_head_$TAG.txt
  python scripts/ncu_sass.py /tmp/sass_$TAG.csv $FR > gpurun_out/sass_$TAG.txt 2>&1
  wc -l gpurun_out/sass_$TAG.txt
fi
tail -2 gpurun_out/ncu_sass_$TAG.log
