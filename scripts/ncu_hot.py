#!/usr/bin/env python
"""Top stall locations of an ncu --set full capture (SASS level).  python scripts/ncu_hot.py file.ncu-rep [N]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
# possibly multiple kernels: split on "Kernel Name" rows
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1][:80]
        hdr = rows[i + 1]
        j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            body.append(rows[j]); j += 1
        cs = hdr.index("# Samples")
        stalls = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        tot = sum(int(r[cs] or 0) for r in body)
        print(f"== {name}  total samples {tot}")
        agg = {hdr[k]: sum(int(r[k] or 0) for r in body) for k in stalls}
        print("   by reason:", ", ".join(f"{k[6:]}={v*100//max(tot,1)}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
        for idx, r in sorted(enumerate(body), key=lambda ir: -int(ir[1][cs] or 0))[:n]:
            top = max(stalls, key=lambda k: int(r[k] or 0))
            print(f"   {int(r[cs]):7d} {100*int(r[cs])/max(tot,1):5.1f}%  [{idx:5d}] {r[1].strip()[:70]:70s} {hdr[top][6:]}")
        i = j
    else:
        i += 1
