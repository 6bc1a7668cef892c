#!/bin/bash
# end-of-round verification: smoke, GPU tests, the default bench line (with the CPU baseline), a short line per workload
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/final_bench.log 2>&1; tail -c 600 gpurun_out/final_bench.log; echo
for W in c1 c3 c4 c5 c5s; do
  timeout 600 python bench.py --workload $W --steps 5 --warmup 3 --no-cpu > gpurun_out/final_$W.log 2>&1
  python - <<PY
import json
for l in open("gpurun_out/final_$W.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print("$W value %.1fM e2e %.1fM ms/step %.3f frac %.4f status %d ess_min %.3g" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6, d["ms_per_step"], d["roofline"]["frac"], d["chain_status_nonzero"], (d.get("ess_per_sec") or {}).get("min", float("nan"))))
PY
done
