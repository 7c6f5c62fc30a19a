#!/usr/bin/env python
"""Kernel timeline of one training step from CUPTI (torch.profiler): in-pipeline durations (warm L2, PDL active) and
the gaps between kernels -- what ncu's serialised cold-cache launch list cannot show.
    python scripts/timeline.py [out.csv]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import image_compression_b200 as pkg  # noqa: E402
from image_compression_b200 import sharding  # noqa: E402

out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/timeline.csv"
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = pkg.build_model(pkg.get_recipe_cfg("psnr_256")).to(dev).train()
x = torch.rand(64, 3, 256, 256, device=dev)
fg = sharding.FlatGradients(model.parameters(), dev)


def step():
    fg.zero()
    _, losses = model(x)
    losses["total_loss"].backward()
    fg.all_reduce_mean()


for _ in range(5):
    step()
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
n = len(ev) // 3
last = ev[2 * n:]
t0 = last[0].time_range.start
rows, busy, prev_end = [], 0.0, None
for e in last:
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    gap = 0.0 if prev_end is None else s - prev_end
    prev_end = max(prev_end or 0.0, s + d)
    busy += d
    rows.append((e.name[:60], s, d, gap))
with open(out, "w") as f:
    f.write("kernel,start_us,dur_us,gap_before_us\n")
    for r in rows:
        f.write("%s,%.2f,%.2f,%.2f\n" % r)
span = prev_end
print(f"{len(rows)} kernels, span {span:.1f} us, sum of durations {busy:.1f} us, idle {span - busy:.1f} us")
agg = {}
for name, s, d, gap in rows:
    a = agg.setdefault(name.split("<")[0].split("(")[0].replace("void ", ""), [0, 0.0, 0.0])
    a[0] += 1; a[1] += d; a[2] += max(gap, 0.0)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:44]:44s} n={v[0]:3d} dur={v[1]:8.1f} us  gaps_before={v[2]:7.1f} us")
