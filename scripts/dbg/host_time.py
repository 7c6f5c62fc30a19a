"""host (CPU) time per training step: time to ENQUEUE a step, without waiting for the GPU"""
import time, torch
import image_compression_b200 as pkg
from image_compression_b200 import sharding
torch.manual_seed(0)
model = pkg.build_model(pkg.get_recipe_cfg("psnr_256")).to("cuda").train()
fg = sharding.FlatGradients(model.parameters())
x = torch.rand(64, 3, 256, 256, device="cuda")
def step():
    fg.zero(); _, l = model(x); l["total_loss"].backward(); fg.all_reduce_mean()
for _ in range(5): step()
torch.cuda.synchronize()
for n in (1, 2, 3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): step()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"{n} step(s): enqueue {1e3*(t1-t0)/n:.2f} ms/step, total {1e3*(t2-t0)/n:.2f} ms/step")
import cProfile, pstats
pr = cProfile.Profile(); pr.enable(); step(); pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
