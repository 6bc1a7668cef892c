#!/bin/bash
# prod vs two dev builds
for rep in 1 2; do for V in prod bayesbridge_b200/libbb_dev8.so bayesbridge_b200/libbb_dev.so; do
  if [ $V = prod ]; then unset BB_LIB_PATH; else export BB_LIB_PATH=$V; fi
  timeout 300 python bench.py --workload $1 --steps 5 --warmup 3 --no-cpu --no-ess --no-sub --e2e-steps 1 > gpurun_out/ab.log 2>&1
  echo "$1 $V: $(python scripts/show_bench.py gpurun_out/ab.log | grep HEADLINE | cut -c1-70)"
done; done
