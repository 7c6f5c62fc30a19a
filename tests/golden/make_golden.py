#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (hieu1999210/image_compression,
mounted read-only at /root/reference) in the build container, and pin oracle/hyperprior_oracle.py to it.

    python tests/golden/make_golden.py            # writes the fixtures, fails if the oracle disagrees

The reference ships no golden vectors; these are outputs of the reference's own modules
on seeded inputs.  /root/reference does not exist on the GPU box, so only the committed .npz files travel.
Fixtures are kept small: whole-model cases store the seed of the weights (the oracle's init is checked
bit-for-bit against `modelling.build_model(cfg).state_dict()` right here), the inputs, the captured
`torch.rand_like` noise draws, the outputs, and per-tensor gradient digests (norm, sum, leading values).
"""
import hashlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import hyperprior_oracle as O  # noqa: E402
from oracle import ref_shim  # noqa: E402

ref_shim.install()
import modelling  # noqa: E402  (the reference package)
from modelling.blocks.entropy_model import EntropyModel, GaussianConditionalModel, LaplacianConditionalModel  # noqa: E402
from modelling.layers.bound import LowerBound, UpperBound  # noqa: E402
from modelling.layers.gdn import GDN  # noqa: E402
from modelling.loss import MS_SSIMLoss, SSIMLoss  # noqa: E402
from utils.metric import MS_SSIM, SSIM, psnr  # noqa: E402

torch.set_num_threads(4)
N_HEAD = 32


class CaptureNoise:
    """Record every torch.rand_like draw made while the reference runs (entropy_model.py:230,333)."""

    def __enter__(self):
        self.draws = []
        self._orig = torch.rand_like

        def rl(t, *a, **k):
            r = self._orig(t, *a, **k)
            self.draws.append(r.clone())
            return r
        torch.rand_like = rl
        return self

    def __exit__(self, *a):
        torch.rand_like = self._orig


def digest(t):
    """(l2 norm, sum, first N_HEAD values) in float64 -- a compact fingerprint of a large tensor."""
    d = t.detach().double().reshape(-1)
    head = torch.zeros(N_HEAD, dtype=torch.float64)
    head[:min(N_HEAD, d.numel())] = d[:N_HEAD]
    return np.concatenate([[float(d.norm()), float(d.sum())], head.numpy()])


def state_hash(sd):
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].detach().contiguous().numpy().tobytes())
    return h.hexdigest()


def close(a, b, tol, what):
    a, b = a.detach().double(), b.detach().double()
    err = float((a - b).abs().max()) if a.numel() else 0.0
    ref = float(b.abs().max()) if b.numel() else 0.0
    assert err <= tol * max(ref, 1.0), f"oracle != reference for {what}: max abs err {err:.3e} (ref max {ref:.3e})"
    return err


def save(name, **arrays):
    out = {}
    for k, v in arrays.items():
        out[k] = v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path} ({os.path.getsize(path) / 1024:.0f} KiB, {len(out)} arrays)")


def ref_cfg(yaml, channels=None, conditional=None):
    ov = []
    if channels is not None:
        ov += ["MODEL.INTER_CHANNELS", channels, "MODEL.LATENT_CHANNELS", channels]
    if conditional is not None:
        ov += ["MODEL.ENTROPY_MODEL.CONDITIONAL_MODEL", conditional]
    return ref_shim.reference_cfg(yaml, ov)


def oracle_cfg_like(cfg):
    lo = cfg.MODEL.LOSS
    return O.default_cfg(lmbda=lo.DISTORTION_LOSS_WEIGHT, loss_names=tuple(lo.DISTORTION_LOSS_NAMES),
                         reduction=lo.REDUCTION, log_scale=lo.SSIM.LOG_SCALE,
                         conditional=cfg.MODEL.ENTROPY_MODEL.CONDITIONAL_MODEL, channels=cfg.MODEL.INTER_CHANNELS)


# ------------------------------------------------------------------------------------------
# module-level cases
# ------------------------------------------------------------------------------------------
def case_gdn():
    """test/test_gdn.py:20-48 known answers + a perturbed-parameter case with gradients."""
    out = {}
    for tag, inverse, relu in (("gdn", False, False), ("igdn", True, False), ("rgdn", False, True)):
        torch.manual_seed(3)
        x = torch.rand(2, 3, 4, 5) - (0.5 if relu else 0.0)
        layer = GDN(3, inverse=inverse, relu=relu)
        y = layer(x)
        xe = torch.max(x, torch.tensor(0.)) if relu else x
        exp = xe * torch.sqrt(1 + 0.1 * xe ** 2) if inverse else xe / torch.sqrt(1 + 0.1 * xe ** 2)
        assert float((y - exp).abs().max()) <= 1e-6            # the reference's own assertion
        yo = O.gdn(x, layer.gamma.param.data, layer.beta.param.data, inverse=inverse, relu=relu)
        close(yo, y, 1e-7, f"{tag} forward")
        out.update({f"{tag}_x": x, f"{tag}_y": y})
    torch.manual_seed(4)
    C = 64
    layer = GDN(C)
    with torch.no_grad():
        layer.gamma.param.add_(0.05 * torch.rand_like(layer.gamma.param))
        layer.beta.param.add_(0.1 * torch.rand_like(layer.beta.param))
    x = torch.randn(2, C, 6, 5, requires_grad=True)
    g = torch.randn(2, C, 6, 5)
    y = layer(x)
    y.backward(g)
    po = O.leaf_params({"g": layer.gamma.param.data, "b": layer.beta.param.data})
    xo = x.detach().clone().requires_grad_(True)
    yo = O.gdn(xo, po["g"], po["b"])
    yo.backward(g)
    close(yo, y, 1e-6, "gdn64 forward")
    close(xo.grad, x.grad, 1e-6, "gdn64 dx")
    close(po["g"].grad, layer.gamma.param.grad, 1e-5, "gdn64 dgamma")
    close(po["b"].grad, layer.beta.param.grad, 1e-5, "gdn64 dbeta")
    out.update(c64_gamma=layer.gamma.param.data, c64_beta=layer.beta.param.data, c64_x=x, c64_g=g, c64_y=y,
               c64_dx=x.grad, c64_dgamma=layer.gamma.param.grad, c64_dbeta=layer.beta.param.grad)
    save("gdn", **out)


def case_bounds():
    torch.manual_seed(5)
    x = (torch.rand(4, 7, 9) * 3 - 1).requires_grad_(True)
    g = torch.randn(4, 7, 9)
    out = {"x": x, "g": g}
    for tag, fn, ofn, b in (("lower", LowerBound, O.lower_bound, 0.25), ("upper", UpperBound, O.upper_bound, 1.0)):
        x.grad = None
        y = fn.apply(x, b)
        y.backward(g)
        xo = x.detach().clone().requires_grad_(True)
        yo = ofn(xo, b)
        yo.backward(g)
        close(yo, y, 0, tag)
        close(xo.grad, x.grad, 0, tag + " grad")
        out.update({f"{tag}_bound": b, f"{tag}_y": y, f"{tag}_gx": x.grad.clone()})
    save("bounds", **out)


def case_entropy_model():
    """EntropyModel (factorised prior) forward + gradients, noise and rounding modes."""
    C = 16
    cfg = ref_cfg("psnr_256.yaml")
    torch.manual_seed(6)
    em = EntropyModel(C, cfg.clone())
    with torch.no_grad():       # move factors / weights off their symmetric init so every gradient is generic
        for n_, p_ in em.named_parameters():
            p_.add_(0.3 * (torch.rand_like(p_) - 0.5))
    sd = {"entropy_model." + k: v.detach().clone() for k, v in em.state_dict().items()}
    z = (torch.randn(2, C, 3, 4) * 3).requires_grad_(True)
    out = {"z": z, **{"p." + k: v for k, v in sd.items()}}
    for tag, training in (("train", True), ("eval", False)):
        em.train(training)
        em.zero_grad()
        z.grad = None
        with CaptureNoise() as cap:
            zt, probs, bits = em(z)
        bits.backward()
        noise = cap.draws[0] if training else None
        assert len(cap.draws) == (1 if training else 0)
        po = O.leaf_params(sd)
        zo = z.detach().clone().requires_grad_(True)
        zto, probso, bitso = O.entropy_model(po, zo, training, noise)
        bitso.backward()
        close(zto, zt, 0, f"entropy_model {tag} z_tilde")
        close(bitso, bits, 1e-6, f"entropy_model {tag} bits")
        # the reference's probs come back permuted (N,W,C,H)
        assert tuple(probs.shape) == (2, 4, C, 3)
        close(probso, probs, 1e-6, f"entropy_model {tag} probs")
        close(zo.grad, z.grad, 1e-5, f"entropy_model {tag} dz")
        for k, v in em.named_parameters():
            close(po["entropy_model." + k].grad, v.grad, 1e-5, f"entropy_model {tag} d{k}")
            out[f"{tag}_g.{k}"] = v.grad.clone()
        out.update({f"{tag}_z_tilde": zt, f"{tag}_probs": probs, f"{tag}_bits": bits, f"{tag}_gz": z.grad.clone()})
        if training:
            out["noise"] = noise
    save("entropy_model", **out)


def case_conditional():
    """Laplacian / Gaussian conditional likelihood incl. the fp32 tail behaviour."""
    cfg = ref_cfg("psnr_256.yaml")
    out = {}
    torch.manual_seed(7)
    y0 = torch.randn(2, 8, 6, 5) * 4
    y0.view(-1)[:6] = torch.tensor([0.0, 0.5, -0.5, 40.0, -60.0, 1e-3])      # ties, deep tails
    s0 = torch.exp(torch.randn(2, 8, 6, 5))
    s0.view(-1)[:6] = torch.tensor([1.0, 1e-3, 1e3, 0.5, 1.0, 1e-10])
    out.update(y=y0, scale=s0)
    for kind, cls in (("laplace", LaplacianConditionalModel), ("gauss", GaussianConditionalModel)):
        m = cls(cfg)
        okind = cls.__name__
        for tag, training in (("train", True), ("eval", False)):
            m.train(training)
            y = y0.clone().requires_grad_(True)
            s = s0.clone().requires_grad_(True)
            with CaptureNoise() as cap:
                yt, probs = m(y, s)
            bits = m._ce_loss(probs)
            bits.backward()
            noise = cap.draws[0] if training else None
            yo, so = y0.clone().requires_grad_(True), s0.clone().requires_grad_(True)
            yto, po = O.conditional_model(yo, so, training, okind, noise)
            bo = O.ce_bits(po)
            bo.backward()
            close(yto, yt, 0, f"{kind} {tag} y_tilde")
            close(po, probs, 0, f"{kind} {tag} probs")
            close(bo, bits, 1e-7, f"{kind} {tag} bits")
            close(yo.grad, y.grad, 1e-6, f"{kind} {tag} dy")
            close(so.grad, s.grad, 1e-6, f"{kind} {tag} dscale")
            pre = f"{kind}_{tag}_"
            out.update({pre + "y_tilde": yt, pre + "probs": probs, pre + "bits": bits, pre + "gy": y.grad,
                        pre + "gscale": s.grad})
            if training:
                out[pre + "noise"] = noise
    save("conditional", **out)


def case_losses():
    """MSE / SSIMLoss / MS_SSIMLoss (log + plain) with gradients, and the psnr / SSIM / MS_SSIM metrics."""
    torch.manual_seed(8)
    N, H, W = 1, 176, 208                      # smallest legal MS-SSIM height, odd-size levels
    x = torch.rand(N, 3, H, W)
    xt0 = (x + 0.1 * torch.randn(N, 3, H, W)).clamp(0, 1)
    out = {"x": x, "x_tilde": xt0}
    scfg = O.default_cfg().MODEL.LOSS.SSIM
    for tag, log in (("log", True), ("plain", False)):
        scfg.LOG_SCALE = log
        for name, cls, ofn in (("ssim", SSIMLoss, O.ssim_loss), ("msssim", MS_SSIMLoss, O.ms_ssim_loss)):
            fn = cls(log_scale=log)
            xt = xt0.clone().requires_grad_(True)
            val = fn(x, xt)
            val.sum().backward()
            xo = xt0.clone().requires_grad_(True)
            vo = ofn(x, xo, scfg)
            vo.sum().backward()
            close(vo, val, 1e-5, f"{name} {tag}")
            close(xo.grad, xt.grad, 1e-4, f"{name} {tag} grad")
            out[f"{name}_{tag}"] = val
            out[f"{name}_{tag}_grad"] = xt.grad.float()
    xt = xt0.clone().requires_grad_(True)
    mse = torch.nn.MSELoss(reduction="mean")(x, xt)
    mse.backward()
    out.update(mse=mse, mse_grad=xt.grad)
    a, b = xt0 * 255, x * 255
    ps, ss, ms, ms_lin = psnr(a, b), SSIM(device="cpu")(a, b), MS_SSIM(device="cpu")(a, b), MS_SSIM(device="cpu", in_dB=False)(a, b)
    close(O.psnr(a, b), ps, 1e-6, "psnr")
    close(O.ssim_metric(a, b), ss, 1e-5, "SSIM metric")
    close(O.ms_ssim_metric(a, b), ms, 1e-5, "MS_SSIM metric")
    out.update(psnr=ps, ssim_metric=ss, msssim_metric_db=ms, msssim_metric=ms_lin)
    save("losses", **out)


# ------------------------------------------------------------------------------------------
# whole-model cases (Compressor2018 through modelling.build_model)
# ------------------------------------------------------------------------------------------
def case_model(name, yaml, N, H, W, seed, conditional=None, training=True, with_metrics=False):
    cfg = ref_cfg(yaml, conditional=conditional)
    ocfg = oracle_cfg_like(cfg)
    torch.manual_seed(seed)
    model = modelling.build_model(cfg.clone())        # clone: CDFEstimator mutates cfg...DIMS
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    po = O.init_params(ocfg, seed=seed)
    assert sorted(po) == sorted(sd), "state_dict key mismatch"
    for k in sd:
        assert torch.equal(po[k], sd[k]), f"oracle init differs from the reference at {k}"
    model.train(training)
    g = torch.Generator().manual_seed(seed + 100)
    x = torch.rand(N, 3, H, W, generator=g)
    with CaptureNoise() as cap:
        xt, losses = model(x)
    out = {"seed": seed, "x_seed": seed + 100, "x_digest": digest(x), "state_sha256": np.frombuffer(
        bytes.fromhex(state_hash(sd)), dtype=np.uint8), "training": int(training), "yaml": yaml,
        "conditional": cfg.MODEL.ENTROPY_MODEL.CONDITIONAL_MODEL, "shape": np.array([N, 3, H, W])}
    nz = ny = None
    if training:
        assert len(cap.draws) == 2
        nz, ny = cap.draws
        out.update(noise_z=nz, noise_y=ny)
        losses["total_loss"].backward()
    pl = O.leaf_params(po)
    xto, lo, inter = O.compressor2018_forward(pl, x, ocfg, training, nz, ny, return_intermediates=True)
    close(xto, xt, 1e-5, f"{name} x_tilde")
    for k in losses:
        close(lo[k], losses[k], 2e-6, f"{name} loss {k}")
        out["loss." + k] = losses[k].detach()
    out["x_tilde"] = xt
    out["y_tilde"] = inter["y_tilde"].detach()
    out["z_tilde"] = inter["z_tilde"].detach()
    out["sigma_digest"] = digest(inter["sigma"])
    if training:
        lo["total_loss"].backward()
        for k, v in model.named_parameters():
            close(pl[k].grad, v.grad, 2e-4, f"{name} grad {k}")
            out["grad." + k] = digest(v.grad)
    if with_metrics:
        a, b = xt * 255, x * 255
        out.update(psnr=psnr(a, b), msssim_db=MS_SSIM(device="cpu")(a, b))
    save(name, **out)


def main():
    case_gdn()
    case_bounds()
    case_entropy_model()
    case_conditional()
    case_losses()
    # configs[0]/[1]: psnr_256 training step (small batch; same network as the benchmark)
    case_model("model_psnr256_train", "psnr_256.yaml", 2, 128, 128, seed=11)
    case_model("model_psnr256_gauss_train", "psnr_256.yaml", 1, 64, 128, seed=12, conditional="GaussianConditionalModel")
    # configs[3]: ssim_64 (MS_SSIMLoss, log scale); 192x192 is the shipped crop size and the smallest MS-SSIM accepts
    case_model("model_ssim64_train", "ssim_64.yaml", 1, 192, 192, seed=13)
    # configs[4]: evaluation (rounding) with metrics
    case_model("model_psnr256_eval", "psnr_256.yaml", 1, 192, 256, seed=14, training=False, with_metrics=True)
    # configs[2]: psnr_8192 (same network, lambda 8192)
    case_model("model_psnr8192_train", "psnr_8192.yaml", 1, 64, 64, seed=15)
    print("all fixtures written; oracle == reference on every case")


if __name__ == "__main__":
    main()
