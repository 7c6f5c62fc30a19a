"""All-reduce probe: time dist.all_reduce(AVG) of the model's flat gradient size under the current NCCL environment.
torchrun --nproc-per-node N scripts/nccl_ar_probe.py"""
import os, sys, time, json
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
res = {}
for n in (11_850_000, 2_097_152):
    buf = torch.randn(n, device="cuda")
    for _ in range(5):
        dist.all_reduce(buf, op=dist.ReduceOp.AVG)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        dist.all_reduce(buf, op=dist.ReduceOp.AVG)
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / 20], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    res[n] = float(ms)
if rank == 0:
    env = {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}
    out = {"world": world, "env": env, "ms": res,
           "algbw_gbs": {n: n * 4 / (m * 1e-3) / 1e9 for n, m in res.items()},
           "busbw_gbs": {n: n * 4 / (m * 1e-3) / 1e9 * 2 * (world - 1) / world for n, m in res.items()}}
    print(json.dumps(out))
dist.destroy_process_group()
