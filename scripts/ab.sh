#!/bin/bash
# usage: ab.sh <workloads...> : device-timed lines of the production library against bayesbridge_b200/libbb_dev.so (a dev build), alternating
for W in "$@"; do for rep in 1 2; do for V in prod dev; do
  if [ $V = dev ]; then export BB_LIB_PATH=bayesbridge_b200/libbb_dev.so; else unset BB_LIB_PATH; fi
  timeout 300 python bench.py --workload $W --steps 5 --warmup 3 --no-cpu --no-ess --no-sub --e2e-steps 1 > gpurun_out/ab.log 2>&1
  echo "$W $V: $(python scripts/show_bench.py gpurun_out/ab.log | grep HEADLINE | cut -c1-70)"
done; done; done
