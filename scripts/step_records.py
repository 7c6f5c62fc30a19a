#!/usr/bin/env python
"""In-pipeline time of every library call of one training step (event pair per exported call, warm L2, back to back on
the stream): what ncu's serialised cold-cache launch list cannot show.   python scripts/step_records.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import image_compression_b200 as pkg  # noqa: E402
from image_compression_b200 import _lib, sharding  # noqa: E402

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
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
_lib.prof_enable(True)
e0.record()
step()
e1.record()
torch.cuda.synchronize()
_lib.prof_enable(False)
recs = _lib.prof_records()
tot = sum(r[1] for r in recs)
print(f"step {e0.elapsed_time(e1):.3f} ms, {len(recs)} bracketed calls, sum {tot:.3f} ms")
for i, (name, ms, work) in enumerate(recs):
    rate = work / (ms * 1e-3) / (1e12 if name.startswith("igemm") else 1e9) if ms > 0 else 0.0
    print(f"{i:3d} {name:18s} {ms * 1e3:8.1f} us  {rate:8.1f} {'TFLOP/s' if name.startswith('igemm') else 'GB/s'}")
