"""NCCL check of the batch-sharded training step, run under torchrun on a multi-GPU box:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/nccl_grad_check.py

Every rank runs fwd+bwd on its slice of one seeded global batch (with injected quantisation noise, so the slices
are reproducible) and `FlatGradients.all_reduce_mean()`; rank 0 then runs the WHOLE batch alone and the two flat
gradients must agree: each replica normalises by its local pixel count (bmshl2018.py:70,88), so the mean of the
replica gradients is the gradient of the concatenated batch.  Not a pytest file (needs torchrun + >= 2 GPUs); the
CPU/gloo twin of the host logic is tests/test_sharding_gloo.py."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import image_compression_b200 as pkg  # noqa: E402
from image_compression_b200 import sharding  # noqa: E402


def flat_grad(model, x, nz, ny, mode, reduce):
    fg = sharding.FlatGradients(model.parameters(), x.device, overlap=reduce)
    fg.zero()
    with pkg.compute_mode(mode):
        _, losses = model(x, nz, ny)
        losses["total_loss"].backward()
    if reduce:
        fg.all_reduce_mean()
    else:
        fg.collect()
    torch.cuda.synchronize()
    return fg.flat.clone(), float(losses["total_loss"].detach())


def transparent_grad(model, x, nz, ny, mode):
    """The trainer's own sequence, nothing else (engine/trainer.py:179-194): zero_grad (set_to_none), forward, .mean(),
    backward.  The model's GradientReducer arms the flat buffer in forward and all-reduces from inside backward."""
    model.zero_grad(set_to_none=True)
    with pkg.compute_mode(mode):
        _, losses = model(x, nz, ny)
        loss = losses.pop("total_loss").mean()
        loss.backward()
    torch.cuda.synchronize()
    red = model.gradient_reducer()
    order = {id(p): i for i, p in enumerate(model.parameters())}
    # back to state_dict order for the comparison with FlatGradients(model.parameters())
    flat = torch.cat([p.grad.reshape(-1) if p.numel() % 4 == 0 else
                      torch.cat([p.grad.reshape(-1), p.grad.new_zeros((-p.numel()) % 4)]) for p in model.parameters()])
    return flat, len(red.fg._early), float(loss.detach())


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    model = pkg.build_model(pkg.get_recipe_cfg("psnr_256")).to(dev).train()
    sharding.broadcast_parameters(model)
    per = 4
    g = torch.Generator().manual_seed(7)
    X = torch.rand(per * world, 3, 128, 128, generator=g)
    NZ = torch.rand(per * world, 192, 2, 2, generator=g)
    NY = torch.rand(per * world, 192, 8, 8, generator=g)
    sl = slice(rank * per, (rank + 1) * per)
    ok = True
    for mode, tol in (("fp32", 2e-4), ("bf16", 2e-2)):
        got, loss_r = flat_grad(model, X[sl].to(dev), NZ[sl].to(dev), NY[sl].to(dev), mode, True)
        losses = torch.tensor([loss_r], device=dev)
        dist.all_reduce(losses, op=dist.ReduceOp.AVG)
        if rank == 0:
            want, loss_all = flat_grad(model, X.to(dev), NZ.to(dev), NY.to(dev), mode, False)
            rel = float((got - want).norm() / want.norm())
            lrel = abs(float(losses) - loss_all) / abs(loss_all)
            print(f"mode {mode}: world {world}, all-reduced mean of replica gradients vs whole-batch gradient: "
                  f"rel-L2 {rel:.3e} (tol {tol:.0e}); mean replica loss {float(losses):.6f} vs {loss_all:.6f} "
                  f"(rel {lrel:.2e})", flush=True)
            ok = ok and rel <= tol and lrel <= 1e-4
        dist.barrier()
        # trainer-transparent path (no FlatGradients in the caller)
        got2, _, loss2 = transparent_grad(model, X[sl].to(dev), NZ[sl].to(dev), NY[sl].to(dev), mode)
        same = torch.tensor([float((got2 - got).abs().max())], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"mode {mode}: in-model reducer vs explicit FlatGradients: max abs difference {float(same):.3e}; "
                  f"flat order of the reducer = backward completion order, min_bucket {model.gradient_reducer().fg.min_bucket}",
                  flush=True)
            ok = ok and float(same) <= 1e-6 * float(got.abs().max())
        dist.barrier()
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, src=0)
    dist.destroy_process_group()
    if not int(flag):
        raise SystemExit("nccl_grad_check FAILED")
    if rank == 0:
        print("nccl_grad_check ok")


if __name__ == "__main__":
    main()
