#!/usr/bin/env python
"""Aggregate warp-stall samples of an ncu report per source line.

    ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass > both.csv
    python scripts/ncu_lines.py both.csv [top]

Prints, per file, the lines holding the most samples (share of the kernel) and instruction counts."""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None
per = defaultdict(lambda: [0, 0, 0, ""])   # (file, line) -> [samples, inst, thread_inst, text]
hdr = None
for row in csv.reader(open(path, newline="")):
    if not row:
        continue
    if row[0] in ("File Name", "File Path"):
        cur_file = row[1].split("/")[-1]
        continue
    if row[0] in ("Kernel Name", "Function Name"):
        continue
    if row[0] == "Line No":
        hdr = row
        i_s = hdr.index("# Samples"); i_i = hdr.index("Instructions Executed"); i_t = hdr.index("Thread Instructions Executed")
        continue
    if hdr is None or len(row) < len(hdr):
        continue
    if row[0]:   # a source line row (aggregated over its SASS)
        try:
            ln = int(row[0])
        except ValueError:
            continue
        key = (cur_file, ln)
        e = per[key]
        e[3] = row[1].strip()[:110]
        try:
            e[0] += int(row[i_s] or 0); e[1] += int(row[i_i] or 0); e[2] += int(row[i_t] or 0)
        except ValueError:
            pass
tot_s = sum(e[0] for e in per.values()) or 1
tot_i = sum(e[1] for e in per.values()) or 1
print(f"total samples {tot_s}, warp instructions {tot_i}")
byfile = defaultdict(lambda: [0, 0])
for (f, ln), e in per.items():
    byfile[f][0] += e[0]; byfile[f][1] += e[1]
for f, (s, i) in sorted(byfile.items(), key=lambda x: -x[1][0]):
    print(f"  {f:28s} samples {s / tot_s:6.1%}  inst {i / tot_i:6.1%}")
for (f, ln), e in sorted(per.items(), key=lambda x: -x[1][0])[:top]:
    lanes = e[2] / e[1] if e[1] else 0
    print(f"{e[0] / tot_s:6.1%} {e[1] / tot_i:6.1%} lanes {lanes:4.1f}  {f}:{ln}  {e[3]}")
