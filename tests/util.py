"""Shared helpers for the parity tests."""
import torch

import image_compression_b200 as pkg
from oracle import hyperprior_oracle as O


def make_pair(recipe="psnr_256", channels=192, conditional="LaplacianConditionalModel", seed=0, device="cuda"):
    """(product model on `device`, oracle params on CPU, oracle cfg) with identical weights."""
    cfg = pkg.get_recipe_cfg(recipe)
    cfg.MODEL.INTER_CHANNELS = channels
    cfg.MODEL.LATENT_CHANNELS = channels
    cfg.MODEL.ENTROPY_MODEL.CONDITIONAL_MODEL = conditional
    lo = cfg.MODEL.LOSS
    ocfg = O.default_cfg(lmbda=lo.DISTORTION_LOSS_WEIGHT, loss_names=tuple(lo.DISTORTION_LOSS_NAMES),
                         reduction=lo.REDUCTION, log_scale=lo.SSIM.LOG_SCALE, conditional=conditional,
                         channels=channels)
    params = O.init_params(ocfg, seed=seed)
    model = pkg.build_model(cfg)
    model.load_state_dict(params, strict=True)
    return model.to(device), params, ocfg


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-30))


def cosine(a, b):
    a, b = a.detach().double().cpu().reshape(-1), b.detach().double().cpu().reshape(-1)
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def perturb(params, seed=1, scale=0.05):
    """Move GDN / CDF parameters off their symmetric init so gradients are generic."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k, v in params.items():
        if "gamma" in k or "beta" in k or "_cdf_estimator" in k:
            out[k] = v + scale * torch.rand(v.shape, generator=g) * (v.abs().mean() + 0.1)
        else:
            out[k] = v.clone()
    return out


def record_parity(tag, data):
    """Append measured parity margins (worst cosine / rel-L2 per tensor, loss errors ...) to the round's parity
    record: gpurun_out/r02_parity.json on the GPU box (merged back by gpurun; the committed copy is
    profiles/r02_parity.json)."""
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    d = os.path.join(root, "gpurun_out")
    try:
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, "r02_parity.json")
        rec = {}
        if os.path.exists(path):
            with open(path) as f:
                rec = json.load(f)
        rec[tag] = data
        with open(path, "w") as f:
            json.dump(rec, f, indent=1, sort_keys=True)
    except OSError:
        pass


def grad_margins(model, oracle_leaf_params):
    """{name: (cosine, rel_l2)} of every parameter gradient against the oracle's."""
    return {k: (cosine(p.grad, oracle_leaf_params[k].grad), rel_l2(p.grad, oracle_leaf_params[k].grad))
            for k, p in model.named_parameters()}
