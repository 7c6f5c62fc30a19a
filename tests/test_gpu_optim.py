"""Fused clip + AdamW step against the reference's own sequence
clip_grad_value_ -> torch.optim.AdamW.step over one parameter group per tensor (solver/optim.py:27-39,
engine/trainer.py:189-192), on CUDA, same weights / gradients / schedule."""
import copy

import pytest
import torch
from torch.nn.utils import clip_grad_value_

import image_compression_b200 as pkg
from image_compression_b200.solver import FusedAdamW, make_optimizer

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ref_groups(model, cfg):
    groups = []
    for key, value in model.named_parameters():          # solver/optim.py:27-36
        lr, wd = cfg.SOLVER.BASE_LR, cfg.SOLVER.WEIGHT_DECAY
        if "bias" in key:
            wd, lr = cfg.SOLVER.WEIGHT_DECAY_BIAS, cfg.SOLVER.BASE_LR * cfg.SOLVER.BIAS_LR_FACTOR
        groups.append({"params": [value], "lr": lr, "weight_decay": wd})
    return groups


def test_adamw_matches_torch_over_steps():
    cfg = pkg.get_recipe_cfg("psnr_256")
    torch.manual_seed(0)
    ours = pkg.build_model(cfg).to(DEV)
    ref = copy.deepcopy(ours)
    opt = make_optimizer(cfg, ours)
    assert isinstance(opt, FusedAdamW) and len(opt.param_groups) == 50 and opt.grad_clip == 5.0
    ropt = torch.optim.AdamW(_ref_groups(ref, cfg), cfg.SOLVER.BASE_LR, eps=cfg.SOLVER.EPS)
    g = torch.Generator(device=DEV).manual_seed(3)
    worst = 0.0
    for it in range(6):
        for p, q in zip(ours.parameters(), ref.parameters()):
            # gradients spanning the clip window (|g| up to ~20 with GRAD_CLIP = 5) and tiny ones
            gr = torch.randn(p.shape, device=DEV, generator=g) * (10.0 if it % 2 == 0 else 1e-3)
            p.grad, q.grad = gr.clone(), gr.clone()
        for grp, rgrp in zip(opt.param_groups, ropt.param_groups):       # an LR schedule drives group["lr"]
            grp["lr"] = rgrp["lr"] = rgrp["lr"] * 0.9
        clip_grad_value_(ref.parameters(), cfg.SOLVER.GRAD_CLIP)
        ropt.step()
        opt.step()
        for (n, p), q in zip(ours.named_parameters(), ref.parameters()):
            err = float((p - q).abs().max() / q.abs().max().clamp_min(1e-12))
            worst = max(worst, err)
            # same fp32 operations in the same order; FMA contraction may differ by an ulp per step
            assert err <= 2e-6, (it, n, err)
    sd, rsd = opt.state_dict(), ropt.state_dict()
    assert sd["state"].keys() == rsd["state"].keys()
    for k in sd["state"]:
        assert set(sd["state"][k]) == {"step", "exp_avg", "exp_avg_sq"} == set(rsd["state"][k])
        assert float(sd["state"][k]["step"]) == float(rsd["state"][k]["step"]) == 6.0
        for name in ("exp_avg", "exp_avg_sq"):
            a, b = sd["state"][k][name], rsd["state"][k][name]
            assert float((a - b).abs().max()) <= 2e-6 * float(b.abs().max()) + 1e-30, (k, name)
    print("worst relative parameter deviation after 6 steps:", worst)


def test_state_dict_round_trip_and_clipped_grad_writeback():
    torch.manual_seed(1)
    w = [torch.nn.Parameter(torch.randn(n, device=DEV)) for n in (5, 4096, 4099, 10001)]     # ragged sizes / tails
    r = [torch.nn.Parameter(p.detach().clone()) for p in w]
    opt = FusedAdamW([{"params": [p]} for p in w], 1e-2, eps=1e-4, weight_decay=0.1, grad_clip=0.5,
                     write_clipped_grad=True)
    ropt = torch.optim.AdamW([{"params": [p]} for p in r], 1e-2, eps=1e-4, weight_decay=0.1)
    for _ in range(2):
        for p, q in zip(w, r):
            p.grad = torch.randn_like(p)
            q.grad = p.grad.clone()
        clip_grad_value_(r, 0.5)
        ropt.step(); opt.step()
        for p, q in zip(w, r):
            assert torch.equal(p.grad, q.grad)                       # clamp is exact
            assert float((p - q).abs().max()) <= 2e-6 * float(q.abs().max())
    opt2 = FusedAdamW([{"params": [p]} for p in w], 1e-2, eps=1e-4, weight_decay=0.1, grad_clip=0.5)
    opt2.load_state_dict(copy.deepcopy(ropt.state_dict()))           # a torch AdamW checkpoint loads
    for p, q in zip(w, r):
        p.data.copy_(q.data)
        p.grad = torch.randn_like(p); q.grad = p.grad.clone()
    clip_grad_value_(r, 0.5)
    ropt.step(); opt2.step()
    for p, q in zip(w, r):
        assert float((p - q).abs().max()) <= 2e-6 * float(q.abs().max())


def test_train_steps_reduce_the_loss():
    cfg = pkg.get_recipe_cfg("psnr_256")
    torch.manual_seed(0)
    model = pkg.build_model(cfg).to(DEV).train()
    opt = make_optimizer(cfg, model)
    x = torch.rand(4, 3, 128, 128, device=DEV)
    losses = []
    for _ in range(8):
        opt.zero_grad(set_to_none=True)
        _, out = model(x)
        out["total_loss"].backward()
        opt.step()
        losses.append(float(out["total_loss"]))
    assert losses[-1] < losses[0], losses
