"""CPU tests of the host-side plans: tap tables, phase decomposition, weight packing."""
import pytest
import torch
import torch.nn.functional as F

import igemm_emulator as E
from image_compression_b200 import hotpath as hp

torch.manual_seed(0)
CASES = [(False, 5, 2), (False, 3, 1), (True, 5, 2), (True, 3, 1)]


def _ref(transposed, x, w, k, s):
    if transposed:
        return F.conv_transpose2d(x, w, None, stride=s, padding=k // 2, output_padding=s - 1)
    return F.conv2d(x, w, None, stride=s, padding=k // 2)


@pytest.mark.parametrize("transposed,k,s", CASES)
def test_forward_dgrad_wgrad_plans(transposed, k, s):
    ci, co, N, H, W = 3, 4, 2, 6, 8
    x = torch.randn(N, ci, H, W, dtype=torch.float64, requires_grad=True)
    w = torch.randn((ci, co, k, k) if transposed else (co, ci, k, k), dtype=torch.float64, requires_grad=True)
    y = _ref(transposed, x, w, k, s)
    g = torch.randn_like(y)
    gx_ref, gw_ref = torch.autograd.grad(y, (x, w), g)
    spec = hp.ConvSpec(transposed, ci, co, k, s)
    K = k * k
    OH, OW = spec.out_hw(H, W)
    assert (OH, OW) == tuple(y.shape[2:])
    xn, gn = x.detach().permute(0, 2, 3, 1), g.permute(0, 2, 3, 1)
    wd = w.detach()
    # forward (ConvSpec.forward)
    if transposed:
        taps, phases = spec.deconv_taps
        wp = E.pack_weight(wd, taps, co, ci, K, co * K)
        out = E.igemm_fwd(xn, wp, taps, phases, (H, W), 1, s, (OH, OW), co)
    else:
        taps, phases = spec.conv_taps
        wp = E.pack_weight(wd, taps, co, ci, ci * K, K)
        out = E.igemm_fwd(xn, wp, taps, phases, (OH, OW), s, 1, (OH, OW), co)
    assert torch.allclose(out.permute(0, 3, 1, 2), y.detach(), atol=1e-10)
    # dgrad (ConvSpec.dgrad)
    if transposed:
        taps, phases = spec.conv_taps
        wp = E.pack_weight(wd, taps, ci, co, co * K, K)
        gx = E.igemm_fwd(gn, wp, taps, phases, (H, W), s, 1, (H, W), ci)
    else:
        taps, phases = spec.deconv_taps
        wp = E.pack_weight(wd, taps, ci, co, K, ci * K)
        gx = E.igemm_fwd(gn, wp, taps, phases, (OH, OW), 1, s, (H, W), ci)
    assert torch.allclose(gx.permute(0, 3, 1, 2), gx_ref, atol=1e-10)
    # wgrad (ConvSpec.wgrad)
    taps, phases = spec.deconv_taps
    if transposed:
        gw = E.igemm_wgrad(xn, gn, taps, phases, s)
        gw = E.unpack_wgrad(gw, w.shape, taps, ci, co, co * K, K)
    else:
        gw = E.igemm_wgrad(gn, xn, taps, phases, s)
        gw = E.unpack_wgrad(gw, w.shape, taps, co, ci, ci * K, K)
    assert torch.allclose(gw, gw_ref, atol=1e-10)


def test_phase_tap_counts():
    taps, phases = hp.deconv_type_taps(5, 2, 2)
    assert [p[0] for p in phases] == [9, 6, 6, 4] and len(taps) == 25
    taps, phases = hp.conv_type_taps(5, 2, 2)
    assert len(taps) == 25 and phases == [(25, 0, 0, 0)]


def test_gdn_bounds_match_reference_quirk():
    assert hp.PEDESTAL == 2.0 ** -36
    assert hp.GAMMA_BOUND == 2.0 ** -36          # sqrt(0 + pedestal^2): gdn.py:53
    assert abs(hp.beta_bound(1e-6) - 1e-3) < 1e-6


def test_gauss_filter_is_the_separable_factor():
    import sys
    from oracle import hyperprior_oracle as O
    g2 = O.gauss_filter_2d(11, 1.5)[0, 0]
    g1 = hp.gauss_filter_1d(11, 1.5, "cpu")
    assert torch.allclose(torch.outer(g1, g1), g2, atol=1e-8)


def test_cpu_tensor_is_rejected_loudly():
    import image_compression_b200 as pkg
    m = pkg.build_model(pkg.get_recipe_cfg("psnr_256"))
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.rand(1, 3, 64, 64))


def test_fused_adamw_host_logic():
    """solver.make_optimizer mirrors solver/optim.py:20-46 (one group per tensor, the "bias" rule) and refuses CPU
    tensors (no fallback); non-adamw optimizers are left to the reference."""
    import image_compression_b200 as pkg
    from image_compression_b200.solver import FusedAdamW, make_optimizer
    cfg = pkg.get_recipe_cfg("psnr_256")
    model = pkg.build_model(cfg)
    opt = make_optimizer(cfg, model)
    assert isinstance(opt, FusedAdamW) and len(opt.param_groups) == 50
    names = [n for n, _ in model.named_parameters()]
    for n, g in zip(names, opt.param_groups):
        assert g["weight_decay"] == (cfg.SOLVER.WEIGHT_DECAY_BIAS if "bias" in n else cfg.SOLVER.WEIGHT_DECAY)
        assert g["lr"] == cfg.SOLVER.BASE_LR and g["eps"] == cfg.SOLVER.EPS
    assert sum("bias" in n for n in names) == 17
    for p in model.parameters():
        p.grad = torch.zeros_like(p)
    with pytest.raises(RuntimeError, match="CUDA"):
        opt.step()
    cfg.SOLVER.OPT_NAME = "sgd"
    with pytest.raises(NotImplementedError):
        make_optimizer(cfg, model)


def test_early_all_reduce_slot_range(monkeypatch):
    """FlatGradients._reduce_early (the hook the backward schedules call after every layer): slots become reducible in
    buffer order -- the range between the frontier and the first slot not yet reported is all-reduced once it holds
    >= min_bucket elements or at a flush; a gradient that is not a slot view is left to the final all-reduce, which
    covers everything beyond the frontier."""
    import torch.distributed as dist
    from image_compression_b200 import sharding
    params = [torch.nn.Parameter(torch.zeros(n)) for n in (5, 16, 7, 9, 3)]
    fg = sharding.FlatGradients(params, torch.device("cpu"), overlap=False)
    fg.min_bucket = 20
    calls = []

    class _Work:
        def wait(self):
            return True

    def fake_all_reduce(t, op=None, group=None, async_op=False):
        calls.append(((t.data_ptr() - fg.flat.data_ptr()) // 4, t.numel()))
        return _Work()
    monkeypatch.setattr(dist, "all_reduce", fake_all_reduce)
    monkeypatch.setattr(dist, "is_initialized", lambda: True)
    monkeypatch.setattr(dist, "get_world_size", lambda group=None: 2)
    monkeypatch.setattr(dist, "get_backend", lambda group=None: "nccl")
    views = [fg._view(p) for p in params]
    fg._reduce_early([views[2], views[3]])                     # slots 2, 3 ready, but slot 0 is the frontier: nothing yet
    assert calls == [] and fg._frontier == 0
    fg._reduce_early([views[0]])                               # slot 0 alone: 8 padded elements < min_bucket
    assert calls == [] and fg._frontier == 0
    fg._reduce_early([views[1]])                               # 0..3 contiguous now: one all-reduce up to slot 4's offset
    assert calls == [(0, fg.offsets[4])] and fg._frontier == 4
    fg._reduce_early([torch.zeros(3)])                         # a stray gradient (not a view of the flat buffer): ignored
    assert len(calls) == 1 and fg._frontier == 4
    for p, v in zip(params, views):
        p.grad = v
    fg.all_reduce_mean()                                       # the rest: slot 4 to the end of the buffer; state reset
    assert calls[1] == (fg.offsets[4], fg.flat.numel() - fg.offsets[4]) and fg._frontier == 0 and not fg._early
    calls.clear()
    fg._reduce_early([views[0]], flush=True)                   # a flush starts the reduce of whatever is contiguous
    assert calls == [(0, fg.offsets[1])] and fg._frontier == 1