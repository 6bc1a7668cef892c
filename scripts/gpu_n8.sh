#!/bin/bash
# N-GPU check (default N = 8): default workload and c3, short runs
N=${1:-8}
mkdir -p gpurun_out
for W in c2 c3; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --workload $W --steps 5 --warmup 3 > gpurun_out/n${N}_$W.log 2>&1
  tail -1 gpurun_out/n${N}_$W.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read()); print('$W N=$N value %.1fM e2e %.1fM status %d ess_min %.3g chains %d' % (d['value']/1e6, d['e2e']['value']/1e6, d['chain_status_nonzero'], (d.get('ess_per_sec') or {}).get('min', float('nan')), d['config']['chains_total']))
except Exception as e:
    print('$W failed', e)
"
done
