"""Per-block comparison of the bf16/tcgen05 path against the fp32/SIMT path on the GPU (debug aid)."""
import sys
import torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import image_compression_b200 as pkg
from image_compression_b200 import blocks_impl as B, hotpath as hp
from util import make_pair, rel_l2

H, W, N = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
model, params, ocfg = make_pair("psnr_256", 192)
m = model
torch.manual_seed(0)
x = torch.rand(N, 3, H, W, device="cuda")
ga = m.analysis_transform.flat_params(); gs = m.synthesis_transform.flat_params()
ha = m.prior_analysis.flat_params(); hs = m.prior_synthesis.flat_params()
from image_compression_b200.autograd_ops import _split_ga, _split_h
res = {}
for mode in ("fp32", "bf16"):
    with pkg.compute_mode(mode), torch.no_grad():
        gc, gg = _split_ga(ga, 4)
        y, yabs, S = B.ga_forward(x, gc, gg, want_abs=True)
        acts = [l[1][0] if l[0] == "gdn" else None for l in S.layers]
        z, _ = B.ha_forward(yabs if mode == "bf16" else hp.cast(yabs, torch.float32), _split_h(ha, [1, 1, 0]))
        if mode == "fp32":
            zt_ref, yt_ref = z + 0.25, y + 0.25
        zt = zt_ref
        sigma, _ = B.hs_forward(hp.cast(zt, hp.lp_dtype()), _split_h(hs, [1, 1, 1]))
        yt = yt_ref
        raw, S2 = B.gs_forward(hp.cast(yt, hp.lp_dtype()), *_split_ga(gs, 4))
        acts2 = [l[1][0] if l[0] == "gdn" else None for l in S2.layers]
        res[mode] = dict(y=y, z=z, sigma=sigma, raw=raw, **{f"ga_x{i}": a for i, a in enumerate(acts) if a is not None},
                         **{f"gs_x{i}": a for i, a in enumerate(acts2) if a is not None})
for k in res["fp32"]:
    a, b = res["bf16"][k].float(), res["fp32"][k].float()
    print(f"{k:8s} shape {tuple(b.shape)} rel_l2 {rel_l2(a, b):.3e} max|d| {float((a-b).abs().max()):.3e} max|ref| {float(b.abs().max()):.3e}")
