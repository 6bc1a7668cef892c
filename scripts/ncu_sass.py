#!/usr/bin/env python
"""Per-SASS-instruction view of an ncu report, in address order (the hot loop as the hardware ran it).

    ncu -i rep.ncu-rep --page source --csv --print-source sass > sass.csv
    python scripts/ncu_sass.py sass.csv [min_exec_fraction]

Prints every instruction whose execution count is at least `min_exec_fraction` (default 0.5) of the most-executed
instruction's: address, warp-level executions, stall samples (share of all samples), the dominant stall reasons, SASS text."""
import csv
import sys

path = sys.argv[1]
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
rows = list(csv.reader(open(path, newline="")))
hi = next(i for i, r in enumerate(rows) if r and r[0] in ("Address", "#") or (len(r) > 2 and "Source" in r and "# Samples" in r))
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
i_a = col.get("Address", 0); i_src = col["Source"]; i_s = col["# Samples"]; i_e = col["Instructions Executed"]
stall_cols = [(h, i) for h, i in col.items() if h.startswith("stall_") or h.startswith("Stall")]
def num(x):
    try:
        return int(float(x.replace(",", ""))) if x else 0
    except ValueError:
        return 0
data = []
for r in rows[hi + 1:]:
    if len(r) <= max(i_src, i_s, i_e):
        continue
    st = sorted(((num(r[i]), h) for h, i in stall_cols if i < len(r)), reverse=True)[:2]
    data.append((r[i_a], num(r[i_e]), num(r[i_s]), [f"{h.replace('stall_', '')}:{v}" for v, h in st if v > 0], r[i_src].strip()))
tot = sum(d[2] for d in data) or 1
mx = max(d[1] for d in data) or 1
print(f"# {len(data)} instructions, {tot} samples, max executions {mx}; stall columns: {[h for h, _ in stall_cols][:6]}...")
for a, e, s, st, src in data:
    if e >= frac * mx:
        print(f"{a:>8s} {e:9d} {s:7d} {100.0 * s / tot:5.2f}%  {' '.join(st):28s} {src[:90]}")
