#!/bin/bash
# quick loop for large-P stable work: parity tests of that path, then c5s / c4w / c4 device-timed lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_wide.py tests/test_gpu_sharded.py tests/test_gpu_full_size.py -m gpu -q -x -k "stb or stable or wide or large or c4 or c5s or chol or constructor" 2>&1 | tail -4
for W in c5s c4w c4; do
  timeout 300 python bench.py --workload $W --steps 5 --warmup 3 --no-cpu --no-ess --no-sub "$@" > gpurun_out/si_$W.log 2>&1
  python scripts/show_bench.py gpurun_out/si_$W.log | grep HEADLINE || tail -3 gpurun_out/si_$W.log
done
