#!/usr/bin/env python
"""Print one summary row per bench.py JSON line: python tools/benchsum.py FILE..."""
import json
import sys

for f in sys.argv[1:]:
    for ln in open(f):
        if not ln.startswith("{"):
            continue
        d = json.loads(ln)
        if "roofline" not in d:
            print(f"{f}: impl={d.get('impl')} value={d.get('value'):.1f} {d.get('unit')}")
            continue
        r = d["roofline"]
        print(f"{f}: N={d['n_gpus']} B={d['config']['global_batch']} ms={d['ms_per_step']:.4f} tok/s={d['value']:.0f} "
              f"GB/s={r['achieved']:.0f} frac={r['frac']:.3f} e2e_ms={d['e2e']['ms_per_step']:.4f} "
              f"clk={d['clocks']['sm_mhz'] if d.get('clocks') else None}")
