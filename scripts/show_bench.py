#!/usr/bin/env python
"""one-screen digest of a bench.py JSON line (headline + every sub-workload)"""
import json
import sys

for path in sys.argv[1:]:
    for l in open(path):
        if not l.startswith("{"):
            continue
        d = json.loads(l)
        if d.get("impl") == "reference":
            print(f"[reference] {d['value'] / 1e6:.3f} M draws/s on {d['cpu_baseline']['cores']} cores; {d['config'].get('cpu_sample')}")
            continue
        def row(name, r):
            if "error" in r:
                print(f"  {name:18s} ERROR {r['error']}"); return
            rf = r["roofline"]; e = r.get("ess_per_sec") or {}
            cb = r.get("cpu_baseline")
            cpu = f" cpu1 {cb['value'] / 1e6:.3f}M ({r['value'] / cb['value']:.0f}x)" if cb else ""
            fair = f" fair {cb['fairness_o_p2']['value'] / 1e6:.3f}M" if cb and "fairness_o_p2" in cb else ""
            print(f"  {name:18s} {r['value'] / 1e6:9.2f}M  e2e {r['e2e']['value'] / 1e6:9.2f}M  {r['ms_per_step']:8.2f} ms/step  {rf['bound']:12s} frac {rf['frac']:.4f}"
                  f"  ESS/s min {e.get('min', float('nan')):.3g} chains {r.get('chains_total', r.get('config', {}).get('chains_total'))} comm {r.get('comm_ms', 0):.2f}ms launches {r.get('gpu_launches')} st {r.get('chain_status_nonzero')}{cpu}{fair}")
        print(f"n_gpus {d['n_gpus']} steps {d['steps']} clocks {d.get('clocks')}  peaks {d.get('peaks_measured')}")
        d.setdefault("chains_total", d["config"].get("chains_total"))
        row("HEADLINE " + d["config"]["workload"][:2], d)
        for k, r in (d.get("workloads") or {}).items():
            row(k, r)
        print(f"  sub-workloads took {d.get('workloads_wall_s', 0):.0f} s")
