#!/bin/bash
# tri sweep configuration sweep (development): prints value / ms_per_step per variant
mkdir -p gpurun_out
run() { # name, workload, env...
  local name=$1 wl=$2; shift 2
  env "$@" timeout 200 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu --no-ess > gpurun_out/tune_$name.log 2>&1
  python - "$name" gpurun_out/tune_$name.log <<'PY'
import sys, json
name, path = sys.argv[1:3]
last = open(path).read().strip().splitlines()[-1]
try:
    d = json.loads(last); print(f"{name:28s} value {d['value']:.4g}  ms/step {d['ms_per_step']:.3f}  status_nonzero {d['chain_status_nonzero']}")
except Exception:
    print(f"{name:28s} FAILED: {last[:300]}")
PY
}
run c3_default c3 A=1
run c3_W8 c3 BB_TRI_W=8
run c3_W8_NG1 c3 BB_TRI_W=8 BB_TRI_NG=1
run c3_C1_NG2 c3 BB_TRI_C=1 BB_TRI_NG=2
run c3_W2_NG4 c3 BB_TRI_W=2 BB_TRI_NG=4
run c3_W2_C1_NG4 c3 BB_TRI_W=2 BB_TRI_C=1 BB_TRI_NG=4
run c5_default c5 A=1
run c5_W4 c5 BB_TRI_W=4
run c5_W2 c5 BB_TRI_W=2
run c5_C1_NG2 c5 BB_TRI_C=1 BB_TRI_NG=2
run c5_C1_NG1_W8 c5 BB_TRI_C=1 BB_TRI_NG=1 BB_TRI_W=8
run c5_C2_NG1_W8 c5 BB_TRI_C=2 BB_TRI_NG=1 BB_TRI_W=8
