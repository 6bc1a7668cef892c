#!/usr/bin/env python
"""Aggregate warp-stall samples of an ncu report per source line.

    ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass > both.csv
    python scripts/ncu_lines.py both.csv [top] [kernel_file]

ncu lists an inlined instruction under every source line of its inline chain, so each SASS address
is counted ONCE per view: the "phase" view attributes it to its line in `kernel_file` (the file
holding the __global__ function; default: the file covering most addresses), the "leaf" view to
its line in the last bb_* file that lists it (the innermost of our own code)."""
import csv
import gzip
import sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
kfile = sys.argv[3] if len(sys.argv) > 3 else None
op = gzip.open if path.endswith(".gz") else open
addr = {}                       # address -> (samples, inst, thread_inst)
where = defaultdict(list)       # address -> [(file, line)]
text = {}
cur_file, cur_line, hdr = None, None, None
for row in csv.reader(op(path, "rt", newline="")):
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
    if row[0]:
        try:
            cur_line = int(row[0])
        except ValueError:
            continue
        text[(cur_file, cur_line)] = row[1].strip()[:100]
    elif row[2].startswith("0x"):
        try:
            addr[row[2]] = (int(row[i_s] or 0), int(row[i_i] or 0), int(row[i_t] or 0))
        except ValueError:
            continue
        where[row[2]].append((cur_file, cur_line))
tot_s = sum(v[0] for v in addr.values()) or 1
tot_i = sum(v[1] for v in addr.values()) or 1
cover = defaultdict(int)
for a, ws in where.items():
    for f in {w[0] for w in ws}:
        cover[f] += 1
if kfile is None:
    kfile = max((f for f in cover if f and f.startswith("bb_")), key=lambda f: cover[f], default=None)
print(f"total samples {tot_s}, warp instructions {tot_i}, kernel file {kfile}")


def view(pick, title):
    per = defaultdict(lambda: [0, 0, 0])
    for a, v in addr.items():
        k = pick(where[a])
        e = per[k]
        e[0] += v[0]; e[1] += v[1]; e[2] += v[2]
    print(f"---- {title} ----")
    for k, e in sorted(per.items(), key=lambda x: -x[1][0])[:top]:
        lanes = e[2] / e[1] if e[1] else 0
        print(f"{e[0] / tot_s:6.1%} {e[1] / tot_i:6.1%} lanes {lanes:4.1f}  {k[0]}:{k[1]}  {text.get(k, '')}")


def phase(ws):
    for w in ws:
        if w[0] == kfile:
            return w
    return ws[0]


def leaf(ws):
    ours = [w for w in ws if w[0] and w[0].startswith("bb_")]
    return ours[-1] if ours else ws[-1]


view(phase, "phase view (lines of the kernel file)")
view(leaf, "leaf view (innermost line of our code)")
