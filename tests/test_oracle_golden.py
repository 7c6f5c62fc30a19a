"""CPU tests: the oracle (oracle/hyperprior_oracle.py) against the golden vectors produced by the
UNMODIFIED reference modules (tests/golden/make_golden.py), and the C-ABI library's export table
against include/hpcodec.h.  No GPU needed."""
import ctypes
import hashlib
import os
import re

import numpy as np
import pytest
import torch

from oracle import hyperprior_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(name):
    d = np.load(os.path.join(GOLD, name + ".npz"), allow_pickle=False)
    return {k: (torch.from_numpy(d[k]) if d[k].dtype.kind == "f" else d[k]) for k in d.files}


def close(a, b, tol, what=""):
    a, b = a.detach().double(), b.detach().double()
    err = float((a - b).abs().max()) if a.numel() else 0.0
    assert err <= tol * max(float(b.abs().max()), 1.0), f"{what}: max abs err {err:.3e}"


def digest(t, n_head=32):
    d = t.detach().double().reshape(-1)
    head = torch.zeros(n_head, dtype=torch.float64)
    head[:min(n_head, d.numel())] = d[:n_head]
    return torch.cat([torch.tensor([float(d.norm()), float(d.sum())], dtype=torch.float64), head])


def test_gdn_known_answers_and_grads():
    g = load("gdn")
    # test/test_gdn.py:20-48 known answers at init (gamma = 0.1 I, beta = 1)
    p = O.init_params(O.default_cfg(channels=3), seed=0)
    gam, bet = O._nonneg_init(torch.eye(3).view(3, 3, 1, 1) * 0.1), O._nonneg_init(torch.ones(3), minimum=1e-6)
    for tag, inv, relu in (("gdn", False, False), ("igdn", True, False), ("rgdn", False, True)):
        close(O.gdn(g[tag + "_x"], gam, bet, inverse=inv, relu=relu), g[tag + "_y"], 1e-7, tag)
    po = O.leaf_params({"g": g["c64_gamma"], "b": g["c64_beta"]})
    x = g["c64_x"].clone().requires_grad_(True)
    y = O.gdn(x, po["g"], po["b"])
    y.backward(g["c64_g"])
    close(y, g["c64_y"], 1e-6)
    close(x.grad, g["c64_dx"], 1e-6)
    close(po["g"].grad, g["c64_dgamma"], 1e-5)
    close(po["b"].grad, g["c64_dbeta"], 1e-5)


def test_bounds():
    g = load("bounds")
    for tag, fn in (("lower", O.lower_bound), ("upper", O.upper_bound)):
        x = g["x"].clone().requires_grad_(True)
        y = fn(x, float(g[tag + "_bound"]))
        y.backward(g["g"])
        assert torch.equal(y.detach(), g[tag + "_y"]) and torch.equal(x.grad, g[tag + "_gx"])


@pytest.mark.parametrize("tag", ["train", "eval"])
def test_entropy_model(tag):
    g = load("entropy_model")
    sd = {k[2:]: v for k, v in g.items() if k.startswith("p.")}
    po = O.leaf_params(sd)
    z = g["z"].clone().requires_grad_(True)
    zt, probs, bits = O.entropy_model(po, z, tag == "train", g["noise"] if tag == "train" else None)
    bits.backward()
    assert torch.equal(zt.detach(), g[tag + "_z_tilde"])
    assert tuple(probs.shape) == tuple(g[tag + "_probs"].shape)     # the reference's (N,W,C,H) layout quirk
    close(probs, g[tag + "_probs"], 1e-6)
    close(bits, g[tag + "_bits"], 1e-6)
    close(z.grad, g[tag + "_gz"], 1e-5)
    for k in sd:
        close(po[k].grad, g[f"{tag}_g.{k[len('entropy_model.'):]}"], 1e-5, k)


@pytest.mark.parametrize("kind,okind", [("laplace", "LaplacianConditionalModel"), ("gauss", "GaussianConditionalModel")])
@pytest.mark.parametrize("tag", ["train", "eval"])
def test_conditional(kind, okind, tag):
    g = load("conditional")
    pre = f"{kind}_{tag}_"
    y, s = g["y"].clone().requires_grad_(True), g["scale"].clone().requires_grad_(True)
    yt, p = O.conditional_model(y, s, tag == "train", okind, g.get(pre + "noise"))
    b = O.ce_bits(p)
    b.backward()
    assert torch.equal(yt.detach(), g[pre + "y_tilde"])
    close(p, g[pre + "probs"], 1e-7)
    close(b, g[pre + "bits"], 1e-6)
    close(y.grad, g[pre + "gy"], 1e-5)
    close(s.grad, g[pre + "gscale"], 1e-5)


def test_losses_and_metrics():
    g = load("losses")
    x, xt0 = g["x"], g["x_tilde"]
    s = O.default_cfg().MODEL.LOSS.SSIM
    for tag, log in (("log", True), ("plain", False)):
        s.LOG_SCALE = log
        for name, fn in (("ssim", O.ssim_loss), ("msssim", O.ms_ssim_loss)):
            xt = xt0.clone().requires_grad_(True)
            v = fn(x, xt, s)
            v.sum().backward()
            close(v, g[f"{name}_{tag}"], 2e-5, f"{name} {tag}")
            close(xt.grad, g[f"{name}_{tag}_grad"], 2e-4, f"{name} {tag} grad")
    a, b = xt0 * 255, x * 255
    close(O.psnr(a, b), g["psnr"], 1e-6)
    close(O.ssim_metric(a, b), g["ssim_metric"], 1e-5)
    close(O.ms_ssim_metric(a, b), g["msssim_metric_db"], 1e-5)
    close(O.ms_ssim_metric(a, b, in_dB=False), g["msssim_metric"], 1e-5)


MODEL_CASES = ["model_psnr256_train", "model_psnr256_gauss_train", "model_ssim64_train", "model_psnr256_eval",
               "model_psnr8192_train"]


def oracle_cfg_for(g):
    yaml = str(g["yaml"])
    lam = float(re.search(r"_(\d+)\.yaml", yaml).group(1))
    if yaml.startswith("ssim"):
        return O.default_cfg(lmbda=lam, loss_names=("MS_SSIMLoss",), log_scale=True, conditional=str(g["conditional"]))
    return O.default_cfg(lmbda=lam, conditional=str(g["conditional"]))


@pytest.mark.parametrize("name", MODEL_CASES)
def test_whole_model(name):
    g = load(name)
    cfg = oracle_cfg_for(g)
    p = O.init_params(cfg, seed=int(g["seed"]))
    h = hashlib.sha256()
    for k in sorted(p):
        h.update(k.encode())
        h.update(p[k].contiguous().numpy().tobytes())
    assert h.digest() == g["state_sha256"].tobytes(), "oracle init != reference build_model(cfg).state_dict()"
    N, C, H, W = [int(v) for v in g["shape"]]
    x = torch.rand(N, C, H, W, generator=torch.Generator().manual_seed(int(g["x_seed"])))
    close(digest(x), torch.from_numpy(g["x_digest"]) if not torch.is_tensor(g["x_digest"]) else g["x_digest"], 0)
    training = bool(int(g["training"]))
    pl = O.leaf_params(p)
    xt, losses, inter = O.compressor2018_forward(pl, x, cfg, training, g.get("noise_z"), g.get("noise_y"),
                                                 return_intermediates=True)
    close(xt, g["x_tilde"], 2e-5, "x_tilde")
    for k in losses:
        close(losses[k], g["loss." + k], 5e-6, k)
    if training:
        assert torch.equal(inter["z_tilde"].detach(), g["z_tilde"]) or float(
            (inter["z_tilde"].detach() - g["z_tilde"]).abs().max()) < 1e-4
        losses["total_loss"].backward()
        for k in pl:
            d, ref = digest(pl[k].grad), g["grad." + k]
            assert abs(float(d[0] - ref[0])) <= 5e-4 * max(float(ref[0]), 1e-12), f"grad norm {k}"
    else:
        # rounded symbols are integers: any disagreement is a pre-rounding tie
        for t in ("z_tilde", "y_tilde"):
            mism = (inter[t].detach() != g[t]).float().mean()
            assert float(mism) <= 1e-4, f"{t}: {float(mism):.2e} of the symbols differ"
        a, b = xt * 255, x * 255
        close(O.psnr(a, b), g["psnr"], 1e-5)
        close(O.ms_ssim_metric(a, b), g["msssim_db"], 1e-4)


# ------------------------------------------------------------------------------------------
# C ABI: the library loads and exports every symbol include/hpcodec.h declares
# ------------------------------------------------------------------------------------------
def test_library_exports_every_declared_symbol():
    from image_compression_b200 import _lib
    _lib.build_library()
    with open(os.path.join(ROOT, "include", "hpcodec.h")) as f:
        src = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    declared = sorted(set(re.findall(r"\b(hpc_[a-z0-9_]+)\s*\(", src)))
    assert len(declared) >= 45
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, f"declared in hpcodec.h but not exported: {missing}"
    unbound = [s for s in declared if s not in _lib.EXPORTED_SYMBOLS]
    assert not unbound, f"declared in hpcodec.h but not bound in _lib.py: {unbound}"
    assert lib.hpc_version() >= 100
