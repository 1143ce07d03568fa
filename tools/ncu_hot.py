#!/usr/bin/env python
"""Top stall-sample SASS lines of one kernel launch. usage: ncu_hot.py rep [launch_index] [topN]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; li = int(sys.argv[2]) if len(sys.argv) > 2 else 0; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(li), "--launch-count", "1"],
                     capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
tot = sum(int(r["# Samples"] or 0) for r in rows)
print(lines[0][:100], "total samples", tot)
reasons = [k for k in rows[0] if k.startswith("stall_") and "Not Issued" not in k]
agg = {k: sum(int(r[k] or 0) for r in rows) for k in reasons}
print("stall totals:", ", ".join(f"{k[6:]}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for i, r in sorted(enumerate(rows), key=lambda ir: -int(ir[1]["# Samples"] or 0))[:top]:
    rs = sorted(((int(r[k] or 0), k[6:]) for k in reasons), reverse=True)[:2]
    print(f"{i:5d} {int(r['# Samples']):6d} {100*int(r['# Samples'])/max(tot,1):5.1f}%  {r['Source'][:90]:90s} {rs}")
