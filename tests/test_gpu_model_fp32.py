"""GPU parity of the four transforms and the whole Compressor2018 step (forward, RD loss, all 50
parameter gradients) in exact-arithmetic mode (fp32, SIMT igemm engine) against the CPU oracle."""
import pytest
import torch

import image_compression_b200 as pkg
from oracle import hyperprior_oracle as O
from util import make_pair, rel_l2, perturb

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def _fp32_mode():
    with pkg.compute_mode("fp32"):
        yield


def _run_both(model, params, ocfg, x, training, nz, ny, fused=True):
    model.train(training)
    model.fused = fused
    model.zero_grad(set_to_none=True)
    xt, losses = model(x.to(DEV), nz.to(DEV), ny.to(DEV))
    pl = O.leaf_params(params)
    xto, lo, inter = O.compressor2018_forward(pl, x, ocfg, training, nz, ny, return_intermediates=True)
    return xt, losses, xto, lo, pl, inter


@pytest.mark.parametrize("recipe,channels,hw,cond", [
    ("psnr_256", 16, (64, 64), "LaplacianConditionalModel"),
    ("psnr_256", 16, (64, 128), "GaussianConditionalModel"),
    ("ssim_64", 16, (192, 192), "LaplacianConditionalModel"),
    ("psnr_256", 192, (64, 64), "LaplacianConditionalModel"),
])
@pytest.mark.parametrize("fused", [True, False])
def test_train_step_matches_oracle(recipe, channels, hw, cond, fused):
    model, params, ocfg = make_pair(recipe, channels, cond)
    params = perturb(params)
    model.load_state_dict(params)
    N = 2
    torch.manual_seed(11)
    x = torch.rand(N, 3, *hw)
    nz = torch.rand(N, channels, hw[0] // 64, hw[1] // 64)
    ny = torch.rand(N, channels, hw[0] // 16, hw[1] // 16)
    xt, losses, xto, lo, pl, _ = _run_both(model, params, ocfg, x, True, nz, ny, fused)
    assert set(losses) == set(lo)
    for k in lo:
        assert abs(float(losses[k]) - float(lo[k])) <= 2e-5 * abs(float(lo[k])) + 1e-7, (k, float(losses[k]), float(lo[k]))
    assert (xt.cpu() - xto).abs().max() < 2e-5
    losses["total_loss"].backward()
    lo["total_loss"].backward()
    worst = 0.0
    for k, p in model.named_parameters():
        assert p.grad is not None, k
        e = rel_l2(p.grad, pl[k].grad)
        worst = max(worst, e)
        assert e < 2e-3, (k, e)
    print("worst grad rel-L2", worst)


@pytest.mark.parametrize("channels,hw", [(16, (64, 64)), (192, (64, 64))])
def test_eval_step_symbols_bit_exact(channels, hw):
    model, params, ocfg = make_pair("psnr_256", channels)
    torch.manual_seed(12)
    x = torch.rand(1, 3, *hw)
    nz = torch.rand(1, channels, hw[0] // 64, hw[1] // 64)
    ny = torch.rand(1, channels, hw[0] // 16, hw[1] // 16)
    with torch.no_grad():
        xt, losses, xto, lo, _, inter = _run_both(model, params, ocfg, x, False, nz, ny)
    for k in lo:
        assert abs(float(losses[k]) - float(lo[k])) <= 2e-5 * abs(float(lo[k])) + 1e-7, k
    # per-block check of the rounded symbols through the module surface
    with torch.no_grad():
        y = model.analysis_transform(x.to(DEV))
        assert rel_l2(y, inter["y"]) < 1e-5
        sig = model.prior_synthesis(inter["z_tilde"].to(DEV))
        assert rel_l2(sig, inter["sigma"]) < 1e-5
        yt, _ = model.conditional_model(y, sig)
        mism = (yt.cpu() != inter["y_tilde"])
        # symbols are bit exact wherever the pre-rounding values do not sit on a .5 boundary
        frac = (inter["y"] - inter["y"].floor() - 0.5).abs()
        assert bool((frac[mism] < 1e-4).all()), int(mism.sum())
        assert int(mism.sum()) <= 2


def test_blocks_module_surface_backward():
    """AnalysisTransform / SynthesisTransform / hyperprior pair as stand-alone modules."""
    model, params, ocfg = make_pair("psnr_256", 16)
    params = perturb(params)
    model.load_state_dict(params)
    torch.manual_seed(13)
    x = torch.rand(2, 3, 64, 64)
    pl = O.leaf_params(params)
    # g_a
    y = model.analysis_transform(x.to(DEV))
    yo = O.analysis_transform(pl, x, ocfg)
    assert rel_l2(y, yo) < 1e-5
    g = torch.randn_like(yo)
    y.backward(g.to(DEV))
    yo.backward(g)
    for k, p in model.analysis_transform.named_parameters():
        assert rel_l2(p.grad, pl["analysis_transform." + k].grad) < 1e-3, k
    # h_a -> h_s chain with input gradients
    a = yo.detach().abs()
    ag, ao = a.to(DEV).requires_grad_(True), a.clone().requires_grad_(True)
    s = model.prior_synthesis(model.prior_analysis(ag))
    so = O.prior_synthesis(pl, O.prior_analysis(pl, ao, ocfg), ocfg)
    assert rel_l2(s, so) < 1e-5
    gs = torch.randn_like(so)
    s.backward(gs.to(DEV))
    so.backward(gs)
    assert rel_l2(ag.grad, ao.grad) < 1e-3
    # g_s with input gradient
    yt = yo.detach() + 0.3
    tg, to = yt.to(DEV).requires_grad_(True), yt.clone().requires_grad_(True)
    r = model.synthesis_transform(tg)
    ro = O.synthesis_transform(pl, to, ocfg)
    assert rel_l2(r, ro) < 1e-5
    gr = torch.randn_like(ro)
    r.backward(gr.to(DEV))
    ro.backward(gr)
    assert rel_l2(tg.grad, to.grad) < 1e-3
    for k, p in model.synthesis_transform.named_parameters():
        assert rel_l2(p.grad, pl["synthesis_transform." + k].grad) < 1e-3, k


def test_in_kernel_noise_is_uniform_and_seeded():
    model, params, ocfg = make_pair("psnr_256", 16)
    model.train()
    x = torch.rand(2, 3, 128, 128, device=DEV)
    torch.manual_seed(5)
    _, l1 = model(x)
    torch.manual_seed(5)
    _, l2 = model(x)
    _, l3 = model(x)
    assert float(l1["bpp"]) == float(l2["bpp"])
    assert float(l1["bpp"]) != float(l3["bpp"])
    em = model.entropy_model
    z = torch.zeros(64, 16, 8, 8, device=DEV)
    zt, _, _ = em(z)
    u = zt.flatten().cpu() + 0.5          # z~ = z + (u - 1/2)
    assert 0.0 <= float(u.min()) and float(u.max()) < 1.0
    assert abs(float(u.mean()) - 0.5) < 0.01 and abs(float(u.var()) - 1 / 12) < 0.005
