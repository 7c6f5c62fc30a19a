"""The drop-in boundary where the reference touches it: `Trainer.build_model`'s nn.DataParallel wrap
(engine/trainer.py:248-261), strict load of a `module.`-prefixed checkpoint (utils/checkpointers/epoch_checkpointer.py:
212-220), `train_step` (engine/trainer.py:176-208) and `Evaluator.step` (engine/evaluator.py:87-105) replayed call
for call (tests/trainer_replay.py), with torch.optim.AdamW (what the unmodified solver builds) and with FusedAdamW."""
import copy

import pytest
import torch

import image_compression_b200 as pkg
from image_compression_b200 import utils as U
from image_compression_b200.solver import make_optimizer
from trainer_replay import Monitor, eval_step, train_step
from util import make_pair

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ref_adamw(model, cfg):
    groups = []
    for key, value in model.named_parameters():          # solver/optim.py:27-36 (names carry the "module." prefix too)
        lr, wd = cfg.SOLVER.BASE_LR, cfg.SOLVER.WEIGHT_DECAY
        if "bias" in key:
            wd, lr = cfg.SOLVER.WEIGHT_DECAY_BIAS, cfg.SOLVER.BASE_LR * cfg.SOLVER.BIAS_LR_FACTOR
        groups.append({"params": [value], "lr": lr, "weight_decay": wd})
    return torch.optim.AdamW(groups, cfg.SOLVER.BASE_LR, eps=cfg.SOLVER.EPS)


def _monitor(model):
    names = model.module.loss_names if hasattr(model, "module") else model.loss_names
    return Monitor(names, {"psnr": U.psnr, "ms_ssim": U.MS_SSIM()})          # engine/monitor.py:50 defaults


@pytest.mark.parametrize("optimizer", ["torch_adamw", "fused_adamw"])
def test_train_step_sequence_through_dataparallel(optimizer):
    cfg = pkg.get_recipe_cfg("psnr_256")
    core, params, _ = make_pair("psnr_256", 192)
    # Trainer.build_model: DataParallel wrap (one visible device: pass-through), then a module.-prefixed strict load
    model = torch.nn.DataParallel(core, device_ids=[0]).to(DEV)
    ckpt = {"module." + k: v.clone() for k, v in params.items()}
    model.load_state_dict(ckpt, strict=True)
    assert set(model.state_dict().keys()) == set(ckpt.keys()) and len(ckpt) == 50
    model.train()
    plain = copy.deepcopy(core).train()              # the same model with plain autograd gradients, as the checker
    plain.flat_gradients = False
    opt = _ref_adamw(model, cfg) if optimizer == "torch_adamw" else make_optimizer(cfg, model)
    popt = _ref_adamw(plain, cfg)
    mon, pmon = _monitor(model), _monitor(plain)
    g = torch.Generator(device=DEV).manual_seed(21)
    last = None
    with pkg.compute_mode("bf16"):
        for it in range(1, 5):
            x = torch.rand(4, 3, 192, 192, device=DEV, generator=g)          # 192 = the shipped DATA.SIZE; MS-SSIM needs >= 176
            noise = (torch.rand(4, 192, 3, 3, device=DEV, generator=g), torch.rand(4, 192, 12, 12, device=DEV, generator=g))
            vals = train_step(model, opt, x, mon, grad_clip=cfg.SOLVER.GRAD_CLIP, it=it, noise=noise)
            pvals = train_step(plain, popt, x, pmon, grad_clip=cfg.SOLVER.GRAD_CLIP, it=it, noise=noise)
            assert set(vals) == {"y_entropy", "z_entropy", "bpp", "MSE", "psnr", "ms_ssim"}
            # torch AdamW on both sides: identical gradients (same kernels, fixed-order reductions) -> identical values;
            # FusedAdamW vs torch AdamW: the trajectories separate at the 1e-5 level per step (see tests/test_gpu_optim.py)
            tol = 2e-5 if optimizer == "torch_adamw" else 1e-3
            for k in vals:
                assert vals[k] == vals[k] and abs(vals[k] - pvals[k]) <= tol * abs(pvals[k]) + 1e-7, (it, k, vals[k], pvals[k])
            assert abs(vals["bpp"] - (vals["y_entropy"] + vals["z_entropy"])) <= 1e-5 * vals["bpp"]
            assert all(p.grad is None for p in model.parameters())          # optimizer.zero_grad() (set_to_none)
            last = vals
    # both gradient paths (flat buffer written by the kernels vs autograd-owned tensors) drove the same trajectory
    # (FusedAdamW vs torch AdamW: chaotic separation of Adam's +-lr early steps, see tests/test_gpu_optim.py)
    worst = max(float((p - q).abs().max()) for p, q in zip(core.parameters(), plain.parameters()))
    assert worst <= (0.0 if optimizer == "torch_adamw" else 2 * 4 * cfg.SOLVER.BASE_LR), worst
    assert mon.bpp.count == 4 and mon.psnr.count == 16 and last["psnr"] > 0


def test_gradient_accumulation_and_grad_none_contract():
    """SOLVER.GD_STEPS = 2 (engine/trainer.py:183,191): two backwards before optimizer.step(); `.grad` is a dense fp32
    tensor of the parameter's shape after every backward (clip_grad_value_ relies on it)."""
    cfg = pkg.get_recipe_cfg("psnr_256")
    core, _, _ = make_pair("psnr_256", 192)
    model = torch.nn.DataParallel(core, device_ids=[0]).to(DEV).train()
    plain = copy.deepcopy(core).train()
    plain.flat_gradients = False
    opt, popt = _ref_adamw(model, cfg), _ref_adamw(plain, cfg)
    mon, pmon = _monitor(model), _monitor(plain)
    g = torch.Generator(device=DEV).manual_seed(22)
    with pkg.compute_mode("bf16"):
        for it in range(1, 5):
            x = torch.rand(2, 3, 192, 192, device=DEV, generator=g)
            noise = (torch.rand(2, 192, 3, 3, device=DEV, generator=g), torch.rand(2, 192, 12, 12, device=DEV, generator=g))
            train_step(model, opt, x, mon, grad_clip=5.0, gd_steps=2, it=it, noise=noise)
            train_step(plain, popt, x, pmon, grad_clip=5.0, gd_steps=2, it=it, noise=noise)
            if it % 2 == 1:
                for (n, p), q in zip(core.named_parameters(), plain.parameters()):
                    assert p.grad is not None and p.grad.dtype == torch.float32 and p.grad.shape == p.shape, n
                    assert torch.equal(p.grad, q.grad), n
    for (n, p), q in zip(core.named_parameters(), plain.parameters()):
        assert float((p - q).abs().max()) <= 1e-6 * float(q.abs().max()) + 1e-12, n      # in-place add order: (0 + a) + b both ways


def test_evaluator_step_sequence():
    core, _, _ = make_pair("psnr_256", 192)
    model = torch.nn.DataParallel(core, device_ids=[0]).to(DEV)
    mon = _monitor(model)
    model.eval()                                      # inference_context (engine/evaluator.py:29-41)
    g = torch.Generator(device=DEV).manual_seed(23)
    with torch.no_grad(), pkg.compute_mode("bf16"):
        for i in range(2):
            x = torch.rand(1, 3, 192, 256, device=DEV, generator=g)          # VAL.BATCH_SIZE = 1 (utils/config.py:130)
            vals = eval_step(model, x, mon)
            xt, losses = core(x)
            assert vals["bpp"] == float(losses["bpp"]) and vals["MSE"] == float(losses["MSE"])
            assert abs(vals["psnr"] - float(U.psnr(xt * 255, x * 255))) <= 1e-4
    assert mon.ms_ssim.count == 2 and mon.bpp.count == 2
    model.train()
    assert core.training


def test_lazily_synchronised_outputs_serve_the_monitor_reads():
    """SURVEY 8f rank 2 (utils/lazy.py): the values the unchanged Monitor reads -- float(loss), preds.max() <= 1,
    metric(preds*255, targets*255).detach().cpu().numpy() -- equal what plain tensors give, bit for bit; the trainer's
    isnan check is deferred to the first host read of the same step."""
    from image_compression_b200.utils import lazy
    core, _, _ = make_pair("psnr_256", 192)
    core.train()
    g = torch.Generator(device=DEV).manual_seed(24)
    x = torch.rand(2, 3, 192, 192, device=DEV, generator=g)
    noise = (torch.rand(2, 192, 3, 3, device=DEV, generator=g), torch.rand(2, 192, 12, 12, device=DEV, generator=g))
    with pkg.compute_mode("bf16"):
        xt, losses = core(x, *noise)
        assert isinstance(losses["bpp"], lazy.LazyScalar) and isinstance(losses["total_loss"], lazy.LossTensor)
        assert isinstance(xt, lazy.ReconImage)
        core.lazy_sync = False
        xt0, losses0 = core(x, *noise)
        core.lazy_sync = True
        assert type(losses0["bpp"]) is torch.Tensor and type(xt0) is torch.Tensor
        for k in losses0:
            assert float(losses[k]) == float(losses0[k]) == losses[k].item(), k
            assert f"{losses[k]:09.6f}" == f"{float(losses0[k]):09.6f}"
        assert torch.equal(xt, xt0)
        # arithmetic on the lazy tensors is ordinary tensor arithmetic
        assert float(losses["y_entropy"] + losses["z_entropy"]) == float(losses0["y_entropy"] + losses0["z_entropy"])
        # the monitor's reads (engine/monitor.py:121-125)
        assert xt.max() <= 1. and not (xt.max() > 1.)
        assert float(xt.max()) == float(xt0.max()) and bool(xt.max() <= 0.5) == bool(xt0.max() <= 0.5)
        preds, targets = xt * 255., x * 255.
        ps = U.psnr(preds, targets)
        ms = U.MS_SSIM()(preds, targets)
        assert isinstance(ps, lazy.LazyVector) and isinstance(ms, lazy.LazyVector)
        ps0, ms0 = U.psnr(xt0 * 255., targets), U.MS_SSIM()(xt0 * 255., targets)
        assert (ps.detach().cpu().numpy() == ps0.cpu().numpy()).all() and (ms.detach().cpu().numpy() == ms0.cpu().numpy()).all()
        assert torch.equal(lazy._plain(ps), ps0)
        # the loss keeps its autograd history through the trainer's .mean() and /= GD_STEPS; isnan is deferred
        loss = losses["total_loss"].mean()
        loss /= 1
        assert not torch.isnan(loss).any()
        loss.backward()
        assert core.analysis_transform.layers[0].weight.grad is not None
        # a NaN loss: the immediate answer is "no", the first host read of the step raises.  (Built by hand: the kernels'
        # clamps are fminf/fmaxf, which drop NaNs, so a NaN weight does not reliably reach total_loss.)
        vec = torch.full((8,), float("nan"), device=DEV, requires_grad=True)
        mirror = lazy.HostMirror(vec.detach())
        l2 = lazy.LossTensor.wrap(vec[3], mirror, 3).mean()
        l2 /= 1
        assert not torch.isnan(l2).any()
        with pytest.raises(FloatingPointError):
            float(lazy.LazyScalar.wrap(vec.detach()[2], mirror, 2))
        lazy.LossTensor.defer_nan_check = False
        try:
            mirror = lazy.HostMirror(vec.detach())
            assert bool(torch.isnan(lazy.LossTensor.wrap(vec[3], mirror, 3).mean()).any())
        finally:
            lazy.LossTensor.defer_nan_check = True
