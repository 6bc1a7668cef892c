#!/usr/bin/env python
"""Condense one gpu_prof.sh capture (gpurun_out/{raw,launches,lines}_<tag>.*) into profiles/<out>.md"""
import csv
import sys
from collections import OrderedDict

tag, out = sys.argv[1], sys.argv[2]
note = sys.argv[3] if len(sys.argv) > 3 else ""
G = "gpurun_out/"
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_op_dmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_srcunit_tex_op_read.sum", "smsp__cycles_active.avg", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
lines = [f"# ncu capture `{tag}`", "", note, ""]
rows = list(csv.reader(open(G + f"raw_{tag}.csv")))
H, U, V = rows[0], rows[1], rows[2]
lines += [f"Kernel: `{V[H.index('Kernel Name')]}`  (`ncu --set full --clock-control none --import-source on`, one launch)", "",
          "| metric | value | unit |", "|---|---|---|"]
for k in KEYS:
    if k in H:
        i = H.index(k)
        lines.append(f"| {k} | {V[i]} | {U[i]} |")
# launch list
try:
    rows = list(csv.reader(open(G + f"launches_{tag}.csv")))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    HH = rows[hdr]
    ki, vi, ui = HH.index("Kernel Name"), HH.index("Metric Value"), HH.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[ui], 1)
        a = agg.setdefault(r[ki].split("(")[0][:70], [0, 0.0])
        a[0] += 1; a[1] += v
    tot = sum(a[1] for a in agg.values()) or 1
    lines += ["", "## Launch list (`--metrics gpu__time_duration.sum`, every kernel of the whole bench command)", "",
              "| kernel | launches | total ms | share |", "|---|---|---|---|"]
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        lines.append(f"| `{k}` | {a[0]} | {a[1] / 1e6:.3f} | {a[1] / tot:.3f} |")
except (OSError, IndexError, ValueError) as e:
    lines += ["", f"(no launch list: {e})"]
try:
    txt = open(G + f"lines_{tag}.txt").read().splitlines()
    lines += ["", "## Warp-stall samples per source line (top; `scripts/ncu_lines.py`)", "", "```"] + [t[:170] for t in txt[:48]] + ["```"]
except OSError:
    pass
open(out, "w").write("\n".join(lines) + "\n")
print("wrote", out)
