"""tcgen05 engine against the CUDA-core engine on IDENTICAL bf16 operands: both accumulate in fp32, so
fp32 outputs agree to summation-order round-off (1e-4 relative to the output scale) -- any descriptor,
swizzle, TMA-coordinate or pipeline bug shows up as an O(1) error here, separated from bf16 rounding."""
import pytest
import torch

import image_compression_b200 as pkg
from image_compression_b200 import _lib
from image_compression_b200 import hotpath as hp

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _both(fn):
    with pkg.compute_mode("bf16_simt"):
        ref = fn()
    with pkg.compute_mode("bf16"):
        out = fn()
    return out, ref


def _cmp(out, ref, tol=2e-4):
    for k in ref:
        a, b = out[k].float(), ref[k].float()
        scale = float(b.abs().max()) + 1e-20
        err = float((a - b).abs().max()) / scale
        lim = tol if k == "f32" else 1e-2        # bf16 outputs may differ by one rounding step
        assert err < lim, (k, err)


@pytest.mark.parametrize("transposed,k,s", [(False, 5, 2), (True, 5, 2), (False, 3, 1), (True, 3, 1)])
@pytest.mark.parametrize("N,H,W,C", [(2, 32, 32, 192), (3, 8, 8, 192), (5, 4, 4, 192), (1, 16, 24, 192),
                                     (2, 8, 12, 64), (1, 6, 10, 128)])
def test_forward_and_dgrad(transposed, k, s, N, H, W, C):
    torch.manual_seed(0)
    spec = hp.ConvSpec(transposed, C, C, k, s)
    x = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    w = (torch.randn(C, C, k, k, device=DEV) / (C * k * k) ** 0.5)
    b = torch.randn(C, device=DEV)
    out, ref = _both(lambda: spec.forward(x, w, b, _lib.EPI_LINEAR, want_f32=True, want_sq=True, want_abs=True))
    _cmp(out, ref)
    OH, OW = spec.out_hw(H, W)
    g = torch.randn(N, OH, OW, C, device=DEV).to(torch.bfloat16)
    out, ref = _both(lambda: spec.dgrad(g, w, (H, W), want_f32=True))
    _cmp(out, ref)


@pytest.mark.parametrize("transposed,N,H,W", [(False, 5, 128, 96), (True, 3, 56, 72), (False, 3, 130, 66),
                                              (False, 10, 128, 128), (True, 5, 64, 64), (False, 7, 136, 168)])
def test_forward_large_grids_use_cta_pairs(transposed, N, H, W):
    """>= 148 tiles: the cta_group::2 kernel (CTA pairs, half of the weight tile per CTA).  The odd tile counts
    exercise the phantom tile at the end of a phase / of the grid; the SIMT engine is the checker.
    >= 74 groups of four tiles (cases 2, 4, 5, 6): two M tiles per CTA (igemm_fwd_umma_pair2_kernel), with 1-3 phantom
    tiles at the end of a phase in cases 2 and 6; both the direct (fp32 output) and the staged bulk-copy (bf16 output
    only) epilogues.  With HPC_FWD_SHARE=1 (test_forced_engine_variants_in_a_subprocess) those grids also share
    activation boxes between the taps of a kernel column (igemm_fwd_umma_pair2s_kernel): groups of 3 + 2 taps (stride 2)
    and of 3 / 2 taps (transposed phases, dgrad)."""
    torch.manual_seed(4)
    C = 192
    spec = hp.ConvSpec(transposed, C, C, 5, 2)
    x = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    w = (torch.randn(C, C, 5, 5, device=DEV) / (C * 25) ** 0.5)
    b = torch.randn(C, device=DEV)
    out, ref = _both(lambda: spec.forward(x, w, b, _lib.EPI_LINEAR, want_f32=True))
    _cmp(out, ref)
    out, ref = _both(lambda: spec.forward(x, w, b, _lib.EPI_LINEAR))
    _cmp(out, ref)
    OH, OW = spec.out_hw(H, W)
    g = torch.randn(N, OH, OW, C, device=DEV).to(torch.bfloat16)
    out, ref = _both(lambda: spec.dgrad(g, w, (H, W)))
    _cmp(out, ref)


@pytest.mark.parametrize("N,H,W", [(37, 32, 32), (29, 40, 32)])
def test_transposed_grids_interleave_phases(N, H, W):
    """Multi-phase grids with whole rounds of (tile x 4 phases) per CTA pair run phase-interleaved (work_item() in
    igemm_umma.cu): 148 pairs per phase = 2 exact rounds, and 145 pairs = 2 rounds with 3 padding items per phase.
    Forward of the transposed conv and dgrad of the strided conv (the two multi-phase callers)."""
    torch.manual_seed(6)
    C = 192
    w = (torch.randn(C, C, 5, 5, device=DEV) / (C * 25) ** 0.5)
    b = torch.randn(C, device=DEV)
    x = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    spec = hp.ConvSpec(True, C, C, 5, 2)
    out, ref = _both(lambda: spec.forward(x, w, b, _lib.EPI_LINEAR, want_f32=True))
    _cmp(out, ref)
    spec = hp.ConvSpec(False, C, C, 5, 2)
    out, ref = _both(lambda: spec.dgrad(x, w, (2 * H, 2 * W), want_f32=True))
    _cmp(out, ref)


def test_forced_engine_variants_in_a_subprocess():
    """HPC_FWD_PAIR=1 forces the CTA-pair kernel for every grid (also 1-2 tiles), HPC_FWD_KSPLIT=0 disables split-K:
    both variants must agree with the SIMT checker on the whole forward/dgrad/epilogue test set."""
    import os, subprocess, sys
    here = os.path.dirname(os.path.abspath(__file__))
    for env in ({"HPC_FWD_PAIR": "1"}, {"HPC_FWD_PAIR": "0", "HPC_FWD_KSPLIT": "0", "HPC_FWD_INTERLEAVE": "0"},
                {"HPC_FWD_MT": "1"}, {"HPC_FWD_SHARE": "1"}):
        r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(here, "test_gpu_umma.py"), "-x", "-q", "-k",
                            "forward_and_dgrad or epilogues or large_grids"], env=dict(os.environ, **env),
                           capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, (env, r.stdout[-2000:], r.stderr[-1000:])


@pytest.mark.parametrize("epi", ["relu", "exp_clamp", "relu_mask", "sign_add"])
def test_epilogues(epi):
    torch.manual_seed(1)
    C, N, H, W = 192, 2, 16, 16
    spec = hp.ConvSpec(False, C, C, 3, 1)
    x = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    w = torch.randn(C, C, 3, 3, device=DEV) / (C * 9) ** 0.5
    b = 0.1 * torch.randn(C, device=DEV)
    aux_lp = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    y32, a32 = torch.randn(N, H, W, C, device=DEV), torch.randn(N, H, W, C, device=DEV)
    if epi == "relu":
        f = lambda: spec.forward(x, w, b, _lib.EPI_RELU, want_f32=True)
    elif epi == "exp_clamp":
        f = lambda: spec.forward(x, w, b, _lib.EPI_EXP_CLAMP, want_f32=True, want_lp=False)
    elif epi == "relu_mask":
        f = lambda: spec.dgrad(x, w, (H, W), _lib.EPI_RELU_MASK, aux0=aux_lp, want_f32=True)
    else:
        f = lambda: spec.dgrad(x, w, (H, W), _lib.EPI_SIGN_ADD, aux0=y32, aux1=a32, want_f32=True)
    out, ref = _both(f)
    _cmp(out, ref)


@pytest.mark.parametrize("inverse", [False, True])
def test_gdn_forward_backward_engines(inverse, C=192, N=2, H=16, W=24):
    torch.manual_seed(2)
    x = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    g = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    gamma = (0.1 * torch.eye(C, device=DEV) + 0.01 * torch.rand(C, C, device=DEV)).view(C, C, 1, 1)
    beta = 1 + torch.rand(C, device=DEV)

    def run():
        xsq, _ = hp.square(x)
        y, nrm = hp.gdn_forward(x, xsq, gamma, beta, inverse)
        dx, dg, db, dxsum = hp.gdn_backward(g, x, xsq, nrm, gamma, inverse, want_dxsum=True)
        if dxsum is None:
            dxsum = hp.colsum(dx)
        return {"y": y, "nrm": nrm, "dx": dx, "dg": dg, "db": db, "dxsum": dxsum}
    out, ref = _both(run)
    for k in ref:
        a, b = out[k].float(), ref[k].float()
        err = float((a - b).abs().max()) / (float(b.abs().max()) + 1e-20)
        assert err < 2e-2, (k, err)
    # the in-kernel column sums are those of the bf16 dx the kernel stored, added 8 rows at a time in packed bf16 (three
    # rounding levels, ~2^-8 of a group sum, unbiased) and across groups in fp32: noise ~ 2^-8 sqrt(P) rms(dx)
    dxf = out["dx"].float()
    P = dxf.numel() // C
    noise = 2.0 ** -8 * P ** 0.5 * float(dxf.pow(2).mean().sqrt())
    assert float((out["dxsum"] - dxf.sum((0, 1, 2))).abs().max()) <= 4 * noise + 1e-3


@pytest.mark.parametrize("inverse", [False, True])
@pytest.mark.parametrize("C,N,H,W", [(192, 1, 9, 13), (192, 3, 96, 96), (128, 2, 16, 24), (128, 1, 150, 131),
                                     (64, 2, 16, 24)])
def test_gdn_fused_backward_shapes(inverse, C, N, H, W):
    """Ragged last tile (P % 128 != 0), more tiles than CTAs (TMEM accumulation of d(gamma) over tiles), the
    C = 128 column plan, and C = 64 (unfused d(gamma) path)."""
    test_gdn_forward_backward_engines(inverse, C, N, H, W)


def test_gdn_fused_backward_is_deterministic():
    torch.manual_seed(5)
    C = 192
    x = torch.randn(2, 64, 64, C, device=DEV).to(torch.bfloat16)
    g = torch.randn(2, 64, 64, C, device=DEV).to(torch.bfloat16)
    gamma = (0.1 * torch.eye(C, device=DEV) + 0.01 * torch.rand(C, C, device=DEV)).view(C, C, 1, 1)
    with pkg.compute_mode("bf16"):
        y, nrm = hp.gdn_forward(x, None, gamma, 1 + torch.rand(C, device=DEV), False)
        a = hp.gdn_backward(g, x, None, nrm, gamma, False, want_dxsum=True)
        b = hp.gdn_backward(g, x, None, nrm, gamma, False, want_dxsum=True)
    for u, v in zip(a, b):
        assert torch.equal(u, v)


def test_rejects_unsupported_shapes_loudly():
    x = torch.randn(1, 8, 8, 48, device=DEV).to(torch.bfloat16)
    w = torch.randn(48, 48, 3, 3, device=DEV)
    with pkg.compute_mode("bf16"):
        with pytest.raises(_lib.HpcError, match="multiple of 64"):
            hp.ConvSpec(False, 48, 48, 3, 1).forward(x, w, None)


@pytest.mark.parametrize("transposed,k,s", [(False, 5, 2), (True, 5, 2), (False, 3, 1), (True, 3, 1)])
@pytest.mark.parametrize("N,H,W,C", [(2, 16, 16, 192), (3, 8, 8, 192), (5, 4, 4, 192), (1, 16, 24, 192),
                                     (2, 8, 12, 64), (1, 6, 10, 128)])
def test_wgrad(transposed, k, s, N, H, W, C):
    torch.manual_seed(3)
    spec = hp.ConvSpec(transposed, C, C, k, s)
    OH, OW = spec.out_hw(H, W)
    x = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    g = torch.randn(N, OH, OW, C, device=DEV).to(torch.bfloat16)
    w = torch.empty(C, C, k, k, device=DEV)
    out, ref = _both(lambda: {"f32": spec.wgrad(x, g, w)})
    _cmp(out, ref, tol=1e-3)


# ---------------------------------------------------------------------------------------------
# thin-channel layers on tensor cores (thin_umma.cu) against the CUDA-core kernels (thin_kernels.cu)
# ---------------------------------------------------------------------------------------------
THIN_SHAPES = [(2, 32, 64, 192), (1, 16, 16, 192), (3, 12, 20, 192), (1, 2, 2, 192), (2, 14, 30, 64), (1, 18, 34, 128),
               (2, 6, 14, 192), (1, 7, 15, 192)]


@pytest.mark.parametrize("N,H,W,C", THIN_SHAPES)
def test_conv_c3_forward_umma(N, H, W, C):
    """H, W: output grid; image is (N,3,2H,2W)."""
    torch.manual_seed(4)
    x = torch.rand(N, 3, 2 * H, 2 * W, device=DEV)
    w = torch.randn(C, 3, 5, 5, device=DEV) / 75 ** 0.5
    b = torch.randn(C, device=DEV)

    def run():
        out, sq, _ = hp.conv_c3_forward(x, w, b, want_sq=True)
        return {"lp": out, "sq": sq}
    out, ref = _both(run)
    for k in ref:       # the CUDA-core kernel keeps fp32 operands, the tensor-core one rounds them to bf16
        err = float((out[k].float() - ref[k].float()).abs().max()) / float(ref[k].float().abs().max())
        assert err < 2e-2, (k, err)
    # operands are rounded to bf16 inside the tensor-core kernel: compare with an fp32 conv at bf16 accuracy
    y = torch.nn.functional.conv2d(x, w, b, stride=2, padding=2).permute(0, 2, 3, 1)
    err = float((out["lp"].float() - y).abs().max()) / float(y.abs().max())
    assert err < 2e-2, err


@pytest.mark.parametrize("N,H,W,C", THIN_SHAPES)
def test_deconv_c3_forward_backward_umma(N, H, W, C):
    torch.manual_seed(5)
    x = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    w = torch.randn(C, 3, 5, 5, device=DEV) / (C * 6.25) ** 0.5
    b = torch.randn(3, device=DEV)
    g = torch.randn(N, 3, 2 * H, 2 * W, device=DEV)

    def run():
        out = hp.deconv_c3_forward(x, w, b)
        dx, gw, gb = hp.deconv_c3_backward(g, x, w)
        return {"out": out, "dx": dx, "gw": gw, "gb": gb}
    out, ref = _both(run)
    y = torch.nn.functional.conv_transpose2d(x.float().permute(0, 3, 1, 2), w, b, stride=2, padding=2, output_padding=1)
    assert float((out["out"] - y).abs().max()) / float(y.abs().max()) < 2e-2        # weights rounded to bf16
    # SIMT kernel keeps fp32 weights, the tensor-core kernel rounds them to bf16: 2^-9 relative per product
    for k, tol in (("out", 1e-2), ("dx", 2e-2), ("gw", 1e-2), ("gb", 1e-5)):
        a, r = out[k].float(), ref[k].float()
        err = float((a - r).abs().max()) / (float(r.abs().max()) + 1e-20)
        assert err < tol, (k, err)


@pytest.mark.parametrize("N,H,W,C", THIN_SHAPES)
def test_wgrad_c3_umma(N, H, W, C):
    torch.manual_seed(6)
    wide = torch.randn(N, H, W, C, device=DEV).to(torch.bfloat16)
    thin = torch.rand(N, 3, 2 * H, 2 * W, device=DEV)
    out, ref = _both(lambda: {"gw": hp.wgrad_c3(wide, thin, C, 3)})
    a, r = out["gw"], ref["gw"]
    err = float((a - r).abs().max()) / float(r.abs().max())
    assert err < 1e-2, err        # thin operand rounded to bf16 on the tensor-core path
    xw = thin.clone().requires_grad_(True)
    wt = torch.zeros(C, 3, 5, 5, device=DEV, requires_grad=True)
    torch.nn.functional.conv2d(xw, wt, None, stride=2, padding=2).backward(wide.float().permute(0, 3, 1, 2))
    err = float((a - wt.grad).abs().max()) / float(wt.grad.abs().max())
    assert err < 1e-2, err


def _rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


# ---------------------------------------------------------------------------------------------
# GDN backward with r = n^-1/2 recomputed in the kernel (hpc_gdn_bwd_fused_rc_umma) against the saved-r form and fp64
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("P,C,inverse", [(64, 192, False), (200, 192, False), (148 * 64 * 5 + 37, 192, False),
                                         (148 * 64 * 3 + 64, 192, True), (1000, 128, False), (148 * 64 * 2 + 1, 128, True)])
def test_gdn_backward_recompute_matches_saved_r_and_fp64(P, C, inverse):
    torch.manual_seed(P % 1000 + C)
    dev = DEV
    x = (torch.randn(1, 1, P, C, device=dev) * 1.5).to(torch.bfloat16)
    g = torch.randn(1, 1, P, C, device=dev).to(torch.bfloat16)
    gp = (torch.rand(C, C, 1, 1, device=dev) * 0.05 + 0.02 * torch.eye(C, device=dev).view(C, C, 1, 1)).sqrt()
    bp = (torch.rand(C, device=dev) * 0.5 + 0.5).sqrt()
    with pkg.compute_mode("bf16"):
        ops = hp.gdn_prepare(gp, bp)
        y, nrm = hp.gdn_forward(x, None, ops, None, inverse, want_nrm=True)
        y2, none = hp.gdn_forward(x, None, ops, None, inverse, want_nrm=False)
        assert none is None and torch.equal(y, y2)
        a = hp.gdn_backward(g, x, None, nrm, ops, inverse, want_dxsum=True, reparam=(gp, bp, 1e-6))
        b = hp.gdn_backward(g, x, None, None, ops, inverse, want_dxsum=True, reparam=(gp, bp, 1e-6))
        b2 = hp.gdn_backward(g, x, None, None, ops, inverse, want_dxsum=True, reparam=(gp, bp, 1e-6))
    torch.cuda.synchronize()
    # fp64 reference through autograd on the same bf16-rounded operands (gdn.py:59-62, 79-88)
    xd = x.double().view(P, C).requires_grad_(True)
    gpd, bpd = gp.double().requires_grad_(True), bp.double().requires_grad_(True)
    ped = hp.PEDESTAL
    gam = (torch.maximum(gpd, torch.tensor(hp.GAMMA_BOUND, device=dev, dtype=torch.float64)) ** 2 - ped).view(C, C)
    gam = gam.detach().to(torch.bfloat16).double() + (gam - gam.detach())  # the kernels' bf16 gamma; gradient passes ONCE
    bet = torch.maximum(bpd, torch.tensor(hp.beta_bound(1e-6), device=dev, dtype=torch.float64)) ** 2 - ped
    n = bet + (xd * xd) @ gam.t()
    yd = xd * n.sqrt() if inverse else xd / n.sqrt()
    yd.backward(g.double().view(P, C))
    names = ("dx", "dgamma", "dbeta", "dxsum")
    refs = (xd.grad.view(1, 1, P, C), gpd.grad, bpd.grad, xd.grad.sum(0))
    for name, u, v, w, r in zip(names, a, b, b2, refs):
        assert torch.equal(v, w), name                                   # bit-reproducible
        ea, eb = _rel(u.double(), r), _rel(v.double(), r)
        # bf16 rounding of t, u, dx in both kernels; the saved-r form also rounds r to bf16
        assert eb <= 8e-3 and eb <= ea * 1.10 + 1e-4, (name, ea, eb)
        assert _rel(v.double(), u.double()) <= 1.2e-2, (name, _rel(v.double(), u.double()))
