#!/bin/bash
# two-GPU check: default workload, c3 (X replicated by one NCCL broadcast) and c5 (row-sharded Gram + all-reduce)
mkdir -p gpurun_out
for W in c2 c3 c5; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --workload $W --steps 5 --warmup 3 > gpurun_out/n2_$W.log 2>&1
  tail -1 gpurun_out/n2_$W.log | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read()); print('$W N=2 value %.1fM e2e %.1fM status %d ess_min %.3g' % (d['value']/1e6, d['e2e']['value']/1e6, d['chain_status_nonzero'], (d.get('ess_per_sec') or {}).get('min', float('nan'))))
except Exception as e:
    print('$W failed', e)
"
done
# alpha sweep of config 5 on one GPU
for A in 0.125 0.9; do for W in c5 c5s; do
  timeout 300 python bench.py --workload $W --alpha $A --steps 3 --warmup 3 --no-cpu > gpurun_out/sweep_${W}_$A.log 2>&1
  tail -1 gpurun_out/sweep_${W}_$A.log | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$W alpha $A value %.1fM e2e %.1fM status %d' % (d['value']/1e6, d['e2e']['value']/1e6, d['chain_status_nonzero']))"
done; done
