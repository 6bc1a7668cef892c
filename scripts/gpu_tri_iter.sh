#!/bin/bash
# quick loop for tri-sweep work: parity tests of the triangle path, then c3 / c5 / c1 device-timed lines
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_alpha.py -m gpu -q -x -k "tri or c3 or c1 or invariants or batching or alpha" 2>&1 | tail -4
for W in c3 c5 c1; do
  timeout 300 python bench.py --workload $W --steps 5 --warmup 3 --no-cpu --no-ess --no-sub "$@" > gpurun_out/ti_$W.log 2>&1
  python scripts/show_bench.py gpurun_out/ti_$W.log | grep HEADLINE || tail -3 gpurun_out/ti_$W.log
done
