#!/usr/bin/env python
"""summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): launches, total, share, us per launch per kernel"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
i = [k for k, r in enumerate(rows) if r and r[0] == 'ID'][0]
h = rows[i]; kn = h.index('Kernel Name'); mv = h.index('Metric Value'); mu = h.index('Metric Unit')
d = collections.defaultdict(list)
for r in rows[i + 1:]:
    if len(r) > mv:
        v = float(r[mv].replace(',', ''))
        v = v / 1000 if r[mu] == 'ns' else (v * 1000 if r[mu] == 'ms' else v)
        d[r[kn][:44]].append(v)
tot = sum(sum(v) for v in d.values())
for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    print(f"{k:44s} {len(v):5d} {sum(v) / 1000:8.2f} ms {sum(v) / tot * 100:5.1f}% {sum(v) / len(v):8.1f} us")
