"""Fused clip + AdamW step: the reference's optimizer step on the flat parameter set in ONE launch.

Reference: `solver/optim.py:20-46` builds `torch.optim.AdamW` over one parameter group PER TENSOR (50 groups for
Compressor2018), a tensor whose name contains "bias" taking `WEIGHT_DECAY_BIAS` / `BASE_LR * BIAS_LR_FACTOR`;
`engine/trainer.py:189-194` then runs `clip_grad_value_(model.parameters(), GRAD_CLIP)`, `optimizer.step()`,
`scheduler.step()`, `optimizer.zero_grad()` every iteration.  `FusedAdamW` keeps that surface -- the same
param_groups (so the reference's LR schedulers drive `group["lr"]` unchanged) and the same `state_dict()` layout
(`step`, `exp_avg`, `exp_avg_sq` per parameter: `utils/checkpointers/iter_checkpointer.py:77-78` saves it) -- and
replaces the arithmetic by `hpc_adamw_step` (csrc/optim_kernels.cu): one multi-tensor kernel, 28 B per parameter.
CUDA only, like the rest of the package: a CPU parameter raises.
"""
import math

import torch

from .. import _lib
from .._lib import AdamwSeg, check, lib, ptr, stream


class FusedAdamW(torch.optim.Optimizer):
    """torch.optim.AdamW(params, lr, betas, eps, weight_decay) semantics (decoupled weight decay, no amsgrad),
    plus `grad_clip`: a fused `clip_grad_value_` (<= 0 disables; `write_clipped_grad=True` also leaves the clamped
    values in `.grad`, as the reference's in-place clip does)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, grad_clip=0.0,
                 write_clipped_grad=False):
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0) or weight_decay < 0.0:
            raise ValueError("FusedAdamW: invalid hyper-parameter")
        if betas[0] <= 0.5:
            raise ValueError("FusedAdamW: beta1 must be > 0.5 (the kernel uses lerp's small-weight branch, like torch "
                             "does for the default 0.9)")
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.grad_clip = float(grad_clip)
        self.write_clipped_grad = bool(write_clipped_grad)
        b, e = self.param_groups[0]["betas"], self.param_groups[0]["eps"]
        for g in self.param_groups:
            if tuple(g["betas"]) != tuple(b) or g["eps"] != e:
                raise ValueError("FusedAdamW: betas and eps must be the same in every parameter group")

    def _state(self, p):
        st = self.state[p]
        if len(st) == 0:
            st["step"] = torch.tensor(0.0, dtype=torch.float32)          # host scalar, like torch's default (non-capturable)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        beta1, beta2 = self.param_groups[0]["betas"]
        eps = self.param_groups[0]["eps"]
        # the segment table (pointers, sizes) is rebuilt only when a tensor moved: with FlatGradients the parameter,
        # gradient and moment addresses are stable, so a step is one pass over the groups updating three scalars each
        live, key = [], []
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is not None:
                    st = self._state(p)
                    live.append((group, p, st))
                    key.append((p.data_ptr(), p.grad.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()))
        if not live:
            return loss
        cache = getattr(self, "_seg_cache", None)
        if cache is None or cache[0] != key:
            for _, p, _ in live:
                _lib.require_cuda(p, p.grad)
                if p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise TypeError("FusedAdamW: parameters and gradients must be fp32 (the reference trains in fp32)")
                if p.grad.is_sparse or not p.is_contiguous() or not p.grad.is_contiguous():
                    raise RuntimeError("FusedAdamW: dense contiguous parameters and gradients only")
            arr = (AdamwSeg * len(live))(*[AdamwSeg(k[0], k[1], k[2], k[3], p.numel(), 1.0, 0.0, 1.0, 0.0)
                                            for k, (_, p, _) in zip(key, live)])
            cache = self._seg_cache = (key, arr)
        arr = cache[1]
        for i, (group, p, st) in enumerate(live):
            lr, wd = float(group["lr"]), float(group["weight_decay"])
            st["step"] += 1
            t = float(st["step"])
            seg = arr[i]
            seg.decay = 1.0 - lr * wd
            seg.neg_step_size = -(lr / (1.0 - beta1 ** t))
            seg.bc2_sqrt = math.sqrt(1.0 - beta2 ** t)
        check(lib().hpc_adamw_step(arr, len(live), beta1, beta2, eps, self.grad_clip, int(self.write_clipped_grad),
                                   stream()))
        # the kernel wrote the parameters (and moments) through raw pointers: bump their version counters like an
        # in-place torch op would, so autograd's saved-tensor checks and the packed-operand cache
        # (hotpath.PackPlan / pack_weight, keyed on `_version`) see the update
        torch._C._increment_version([p for _, p, _ in live])
        return loss


def make_optimizer(cfg, model):
    """solver/optim.py:20-46 with the same per-name rule; `adamw` (every shipped config) maps to FusedAdamW with the
    trainer's GRAD_CLIP folded in.  Other optimizers are outside the hot path: build them with the reference's own
    make_optimizer."""
    opt_name = cfg.SOLVER.OPT_NAME
    params = []
    for key, value in model.named_parameters():
        if not value.requires_grad:
            continue
        lr = cfg.SOLVER.BASE_LR
        weight_decay = cfg.SOLVER.WEIGHT_DECAY
        if "bias" in key:
            weight_decay = cfg.SOLVER.WEIGHT_DECAY_BIAS
            lr = cfg.SOLVER.BASE_LR * cfg.SOLVER.BIAS_LR_FACTOR
        params += [{"params": [value], "lr": lr, "weight_decay": weight_decay}]
    if opt_name != "adamw":
        raise NotImplementedError(f"make_optimizer: OPT_NAME={opt_name!r} is not on the B200 hot path (adamw is); use the "
                                  "reference's solver.make_optimizer for it")
    return FusedAdamW(params, cfg.SOLVER.BASE_LR, eps=cfg.SOLVER.EPS, grad_clip=float(getattr(cfg.SOLVER, "GRAD_CLIP", 0.0)))
