#!/usr/bin/env python
"""Pretty-prints bench.py JSON lines. usage: show_bench.py file.json [...]"""
import json, sys
for f in sys.argv[1:]:
    for line in open(f):
        line = line.strip()
        if not line.startswith("{"):
            continue
        d = json.loads(line)
        e = d.get("e2e") or {}
        print(f"== {f}: value {d.get('value'):.3f} {d.get('unit')}  ms/step {d.get('ms_per_step'):.2f}  e2e {e.get('value')}  launches {d.get('gpu_launches')}")
        st = d.get("step_totals")
        if st:
            print("   step:", st)
        for k in d.get("kernels", [])[:16]:
            print("   %-28s n=%4d ms=%8.3f share=%.3f %-6s %8.1f %-8s frac=%.3f" % (k["kernel"], k["launches"], k["ms"], k["share"], k["bound"], k["achieved"], k["unit"], k["frac"]))
