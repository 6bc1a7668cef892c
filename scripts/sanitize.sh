#!/bin/bash
# compute-sanitizer over the smoke run (one stable + one triangle iteration on 442 x 64, checked against the oracle) and a
# short large-P / wide / ortho run: memcheck (out-of-bounds, misaligned), racecheck (shared-memory hazards across the hand-rolled
# mbarrier rings, named barriers and retry queues), synccheck (barrier misuse).  Logs -> gpurun_out/sanitize_*.log;
# the summaries are copied to profiles/ by hand after a clean run.
mkdir -p gpurun_out
cat > gpurun_out/_san_driver.py <<'PY'
import sys
sys.path[:0] = ['.', 'oracle', 'tests']
import numpy as np
import __graft_entry__ as g
g.smoke()
import bayesbridge_b200 as bb
from bayesbridge_b200._lib import ST_ALL
from helpers import small_problem
hyp = bb.make_hyper(alpha=0.5)
X, y, _ = small_problem(300, 150, seed=1)          # large-P stable path (blocked Cholesky in HBM)
m = bb.Model(X, y); c = m.chains("stable", 3, seed=1); c.run(hyp, 2, 0, ST_ALL, pre_tau=1); assert not c.get()["status"].any()
c = m.chains("triangle", 3, seed=1); c.run(hyp, 2, 0, ST_ALL, pre_tau=1); assert not c.get()["status"].any()   # streaming TMA ring, W > 1
X, y, _ = small_problem(60, 140, seed=2)           # wide design: N x N draw
m = bb.Model(X, y); c = m.chains("stable", 2, seed=1).wide_gaussian(); c.run(bb.make_hyper(alpha=0.5, sig2_shape=3.0, sig2_scale=0.2), 2, 0, ST_ALL, pre_tau=1)
X, y, _ = small_problem(120, 12, seed=3)           # ortho drivers + unknown alpha
for meth in ("triangle", "stable"):
    c = bb.Model(X, y).chains(meth, 2, seed=1, ortho=True); c.run(bb.make_hyper(alpha=-1.0), 3, 0, ST_ALL, pre_tau=1)
print("sanitizer driver done")
PY
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python gpurun_out/_san_driver.py > gpurun_out/sanitize_$tool.log 2>&1
  echo "== $tool: rc $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitizer driver done|smoke ok" gpurun_out/sanitize_$tool.log | tail -4
done
