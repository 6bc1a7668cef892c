#!/bin/bash
# usage: gpu_quick.sh <tag> [workloads...]   tests (-x) then one short bench line per workload
TAG=$1; shift
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -6 > gpurun_out/t_$TAG.log; tail -4 gpurun_out/t_$TAG.log
for W in "$@"; do
  timeout 600 python bench.py --workload $W --steps 5 --warmup 3 --no-cpu > gpurun_out/b_${TAG}_$W.log 2>&1
  python - <<PY
import json
for l in open("gpurun_out/b_${TAG}_$W.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print("$W value %.1fM e2e %.1fM ms/step %.3f frac %.4f status %d precompute %.3fs" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6, d["ms_per_step"], d["roofline"]["frac"], d["chain_status_nonzero"], d.get("precompute_s", 0)))
PY
done
