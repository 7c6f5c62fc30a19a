#!/usr/bin/env python
"""Per-kernel totals and (optionally) the per-launch sequence of one step from an ncu launch list.
    python scripts/launch_summary.py gpurun_out/l4/launches.csv [nsteps_in_capture] [--seq]"""
import collections, csv, re, sys
path = sys.argv[1]
nsteps = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 7
with open(path) as f:
    lines = f.readlines()
start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
rows = [r for r in csv.DictReader(lines[start:]) if r["Metric Name"] == "gpu__time_duration.sum"]
agg, tot = collections.OrderedDict(), 0.0
def us(r):
    return float(r["Metric Value"].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}[r["Metric Unit"]]
for r in rows:
    name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += us(r); tot += us(r)
print(f"{len(rows)} launches, {tot/nsteps:.0f} us per step ({nsteps} steps)")
for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v/nsteps:10.1f} us/step {n/nsteps:6.1f}/step {100*v/tot:5.1f}%  {k[:90]}")
if "--seq" in sys.argv:
    n = len(rows) // nsteps
    for r in rows[2 * n:3 * n]:
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
        print(f"{us(r):10.1f} us grid={r['Grid Size']:<14} {name[:70]}")
