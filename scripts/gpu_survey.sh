#!/bin/bash
# One gpurun call: GPU tests, a short bench of every BASELINE workload, and the ncu launch list +
# one full capture of the dominant kernel of the default workload.  Outputs under gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/tests_gpu.log
for w in c1 c2 c3 c4 c5 c5s; do
  timeout 400 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-ess > gpurun_out/bench_$w.log 2>&1
done
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-ess --e2e-steps 1"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c2.csv $CMD > gpurun_out/ncu_launches.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:stb_small -s 3 -c 2 -o gpurun_out/prof_c2 $CMD > gpurun_out/ncu_full.log 2>&1
