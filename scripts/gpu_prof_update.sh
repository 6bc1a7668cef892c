#!/bin/bash
# ncu --set full capture of one mid-factorisation launch of stbl_update_kernel at c4, condensed to the metrics that tell its bound
SKIP=75 bash scripts/gpu_prof.sh c4 stbl_update c4u >/dev/null 2>&1
head -24 gpurun_out/lines_c4u.txt | cut -c1-200
python - <<PY
import csv
rows=list(csv.reader(open("gpurun_out/raw_c4u.csv"))); H,U,V=rows[0],rows[1],rows[2]
keys=["gpu__time_duration.sum","launch__grid_size","smsp__issue_active.avg.pct_of_peak_sustained_active","sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active","sm__inst_executed_pipe_fp64.sum","smsp__warps_active.avg.per_cycle_active","smsp__warps_eligible.avg.per_cycle_active","dram__bytes_read.sum","dram__bytes_write.sum","lts__t_bytes.sum","l1tex__data_pipe_lsu_wavefronts_mem_shared.sum","sm__throughput.avg.pct_of_peak_sustained_elapsed","gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed","l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum","smsp__inst_executed.sum"]
for k in keys:
    if k in H: i=H.index(k); print(k, V[i], U[i])
for i,h in enumerate(H):
    if "pipe_tensor" in h or "pipe_fp64" in h or "stall" in h and "pct" in h: print(h, V[i], U[i])
PY
