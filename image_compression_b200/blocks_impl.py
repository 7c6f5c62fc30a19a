"""Forward and hand-scheduled backward of the four transforms (g_a, h_a, h_s, g_s).

Plain functions over raw tensors (no autograd): `*_forward` returns outputs plus a state
object with what the backward needs, `*_backward` returns parameter gradients keyed like the
reference's state_dict suffixes.  `autograd_ops.py` wraps them for the module surface and
for the fused whole-model function.
"""
import torch

from . import hotpath as hp
from ._lib import EPI_EXP_CLAMP, EPI_LINEAR, EPI_RELU, EPI_RELU_MASK, EPI_SIGN_ADD


class _State:
    pass


def _gdn_params(gamma_param, beta_param, beta_min=1e-6):
    """NonNegativeParam.forward for gamma and beta (gdn.py:59-62)."""
    if hp.gdn_fused_ok(hp.lp_dtype(), beta_param.numel()):
        return hp.gdn_prepare(gamma_param, beta_param, beta_min), None
    gamma = hp.nonneg_fwd(gamma_param, hp.GAMMA_BOUND)
    beta = hp.nonneg_fwd(beta_param, hp.beta_bound(beta_min))
    return gamma, beta


def _gdn_param_grads(gamma_param, beta_param, dgamma, dbeta, beta_min=1e-6):
    return (hp.nonneg_bwd(gamma_param, dgamma, hp.GAMMA_BOUND),
            hp.nonneg_bwd(beta_param, dbeta, hp.beta_bound(beta_min)))


def _thin_ok(c):
    return 1 <= c <= 4


# ----------------------------------------------------------------------------------------
# g_a: AnalysisTransform (modelling/blocks/analysis.py:44-71)
# ----------------------------------------------------------------------------------------
def ga_forward(x_nchw, convs, gdns, k=5, strides=(2, 2, 2, 2), want_abs=False, save=True):
    """x_nchw: (N,Cin,H,W) fp32 contiguous.  convs: [(weight,bias)]*n, gdns: [(gamma_param,beta_param)]*(n-1).
    Returns y (NHWC fp32), |y| (NHWC lp) or None, state."""
    n = len(convs)
    S = _State()
    S.x = x_nchw
    S.layers = []
    w0, b0 = convs[0]
    cin = w0.shape[1]
    if not (_thin_ok(cin) and k == 5 and strides[0] == 2):
        raise NotImplementedError("g_a head: only Conv2d(<=4 channels, k=5, s=2) is built (analysis.py:55 with the "
                                  "reference's cfg defaults)")
    last = (n == 1)
    need_sq = not hp.gdn_fused_ok(hp.lp_dtype(), w0.shape[0])     # the fused GDN kernel squares in shared memory
    cur, cur_sq, cur_f32 = hp.conv_c3_forward(x_nchw, w0, b0, want_sq=(not last) and need_sq, want_f32=last)
    y_f32, y_abs = cur_f32, None
    for i in range(n):
        if i > 0:
            w, b = convs[i]
            spec = hp.ConvSpec(False, w.shape[1], w.shape[0], k, strides[i])
            last = (i == n - 1)
            inp = cur
            o = spec.forward(inp, w, b, EPI_LINEAR, want_lp=not last, want_sq=(not last) and need_sq, want_f32=last,
                             want_abs=last and want_abs)
            cur, cur_sq = o.get("lp"), o.get("sq")
            if last:
                y_f32, y_abs = o["f32"], o.get("abs")
            S.layers.append(("conv", spec, inp))
        else:
            S.layers.append(("head", None, None))
        if i < n - 1:
            gp, bp = gdns[i]
            gamma, beta = _gdn_params(gp, bp)
            yv, nrm = hp.gdn_forward(cur, cur_sq, gamma, beta,
                                     want_nrm=save and not hp.gdn_bwd_recomputes(cur.dtype, cur.shape[-1], gamma))
            S.layers.append(("gdn", (cur, cur_sq, nrm, gamma), None))
            cur = yv
    if n == 1 and want_abs:
        raise NotImplementedError("single-layer g_a with |y| output")
    S.convs, S.gdns = convs, gdns
    return y_f32, y_abs, (S if save else None)


def ga_backward(S, gy_lp):
    """gy_lp: NHWC lp gradient wrt y.  Returns (conv grads [(gw,gb)], gdn grads [(ggamma,gbeta)])."""
    conv_grads = [None] * len(S.convs)
    gdn_grads = [None] * len(S.gdns)
    g, gsum = gy_lp, None      # gsum: column sums of g when the GDN backward kernel already produced them
    ci, gi = len(S.convs) - 1, len(S.gdns) - 1
    for kind, a, inp in reversed(S.layers):
        if kind == "conv":
            spec = a
            w, b = S.convs[ci]
            gb = (gsum if gsum is not None else hp.colsum(g, out=hp.grad_buffer(b))) if b is not None else None
            gsum = None
            gw = spec.wgrad(inp, g, w)
            conv_grads[ci] = (gw, gb)
            hp.grads_ready((gw, gb))
            g = spec.dgrad(g, w, inp.shape[1:3])["lp"]
            ci -= 1
        elif kind == "gdn":
            x, xsq, nrm, gamma = a
            gp, bp = S.gdns[gi]
            b_prod = S.convs[ci][1]       # bias of the layer that feeds this GDN: its gradient is colsum(dx)
            dx, ggamma, gbeta, gsum = hp.gdn_backward(
                g, x, xsq, nrm, gamma, want_dxsum=True, reparam=(gp, bp, 1e-6),
                dxsum_out=hp.grad_buffer(b_prod) if b_prod is not None and hp.gdn_bwd_has_dxsum(x) else None)
            gdn_grads[gi] = (ggamma, gbeta)
            hp.grads_ready((ggamma, gbeta, gsum))
            g = dx
            gi -= 1
        else:  # head conv: weight + bias grads only (the image needs no gradient)
            w, b = S.convs[0]
            gw = hp.wgrad_c3(g, S.x, w.shape[0], w.shape[1], w)
            gb = (gsum if gsum is not None else hp.colsum(g, out=hp.grad_buffer(b))) if b is not None else None
            conv_grads[0] = (gw, gb)
            hp.grads_ready((gw, gb), flush=True)
    return conv_grads, gdn_grads


# ----------------------------------------------------------------------------------------
# g_s: SynthesisTransform (modelling/blocks/synthesis.py:44-71) -- GDN is NOT inverse (:64-66)
# ----------------------------------------------------------------------------------------
def gs_forward(yt_lp, convs, gdns, k=5, strides=(2, 2, 2, 2), save=True):
    """yt_lp: NHWC lp.  Returns raw x~ (N,Cout,H,W) fp32 NCHW (before the bounds), state."""
    n = len(convs)
    S = _State()
    S.layers = []
    cur = yt_lp
    for i in range(n):
        w, b = convs[i]
        last = (i == n - 1)
        if last:
            cout = w.shape[1]
            if not (_thin_ok(cout) and k == 5 and strides[i] == 2):
                raise NotImplementedError("g_s tail: only ConvTranspose2d(-> <=4 channels, k=5, s=2) is built "
                                          "(synthesis.py:55-57 with the reference's cfg defaults)")
            raw = hp.deconv_c3_forward(cur, w, b)
            S.layers.append(("tail", None, cur))
        else:
            spec = hp.ConvSpec(True, w.shape[0], w.shape[1], k, strides[i])
            o = spec.forward(cur, w, b, EPI_LINEAR, want_sq=not hp.gdn_fused_ok(cur.dtype, w.shape[1]))
            S.layers.append(("deconv", spec, cur))
            gp, bp = gdns[i]
            gamma, beta = _gdn_params(gp, bp)
            yv, nrm = hp.gdn_forward(o["lp"], o.get("sq"), gamma, beta,
                                     want_nrm=save and not hp.gdn_bwd_recomputes(o["lp"].dtype, o["lp"].shape[-1], gamma))
            S.layers.append(("gdn", (o["lp"], o.get("sq"), nrm, gamma), None))
            cur = yv
    S.convs, S.gdns = convs, gdns
    return raw, (S if save else None)


def gs_backward(S, g_raw_nchw, need_input_grad=True):
    """g_raw_nchw: (N,Cout,H,W) fp32.  Returns (g wrt y~ NHWC lp | None, conv grads, gdn grads)."""
    conv_grads = [None] * len(S.convs)
    gdn_grads = [None] * len(S.gdns)
    ci, gi = len(S.convs) - 1, len(S.gdns) - 1
    g, gsum = None, None
    for idx in range(len(S.layers) - 1, -1, -1):
        kind, a, inp = S.layers[idx]
        first = (idx == 0)
        if kind == "tail":
            w, b = S.convs[ci]
            g, gw, gb = hp.deconv_c3_backward(g_raw_nchw, inp, w, need_dx=(not first) or need_input_grad, bias=b)
            conv_grads[ci] = (gw, gb if b is not None else None)
            hp.grads_ready((gw, gb))
            ci -= 1
        elif kind == "gdn":
            x, xsq, nrm, gamma = a
            gp, bp = S.gdns[gi]
            b_prod = S.convs[ci][1]       # bias of the layer that feeds this GDN: its gradient is colsum(dx)
            dx, ggamma, gbeta, gsum = hp.gdn_backward(
                g, x, xsq, nrm, gamma, want_dxsum=True, reparam=(gp, bp, 1e-6),
                dxsum_out=hp.grad_buffer(b_prod) if b_prod is not None and hp.gdn_bwd_has_dxsum(x) else None)
            gdn_grads[gi] = (ggamma, gbeta)
            hp.grads_ready((ggamma, gbeta, gsum))
            g = dx
            gi -= 1
        else:
            spec = a
            w, b = S.convs[ci]
            gb = (gsum if gsum is not None else hp.colsum(g, out=hp.grad_buffer(b))) if b is not None else None
            gsum = None
            gw = spec.wgrad(inp, g, w)
            conv_grads[ci] = (gw, gb)
            hp.grads_ready((gw, gb), flush=first)
            g = spec.dgrad(g, w, inp.shape[1:3])["lp"] if (not first or need_input_grad) else None
            ci -= 1
    return g, conv_grads, gdn_grads


# ----------------------------------------------------------------------------------------
# h_a: HyperpriorAnalysisTransform (modelling/blocks/prior_analysis.py:43-71)
# ----------------------------------------------------------------------------------------
def ha_forward(yabs_lp, convs, kernels=(3, 5, 5), strides=(1, 2, 2), save=True):
    """yabs_lp: NHWC lp |y|.  convs: [(weight, bias|None)].  Returns z NHWC fp32, state."""
    n = len(convs)
    S = _State()
    S.layers = []
    cur = yabs_lp
    z = None
    for i, (w, b) in enumerate(convs):
        spec = hp.ConvSpec(False, w.shape[1], w.shape[0], kernels[i], strides[i])
        last = (i == n - 1)
        o = spec.forward(cur, w, b, EPI_LINEAR if last else EPI_RELU, want_lp=not last, want_f32=last)
        S.layers.append((spec, cur))
        if last:
            z = o["f32"]
        else:
            cur = o["lp"]
    S.convs = convs
    return z, (S if save else None)


def ha_backward(S, gz_lp, sign_add=None, need_input_grad=True, also_ready=()):
    """gz_lp: NHWC lp gradient wrt z.  sign_add=(y_f32, gy_partial_f32): fuse d|y|/dy and the sum with
    the other two gradient streams of y (bmshl2018.py:72) into the last dgrad epilogue.
    Returns (g wrt input (lp), conv grads)."""
    conv_grads = [None] * len(S.convs)
    g = gz_lp
    for i in range(len(S.convs) - 1, -1, -1):
        spec, inp = S.layers[i]
        w, b = S.convs[i]
        with hp.side_stream(g.device) as on_side:       # small-K weight gradient: off the critical path, beside the dgrad chain
            gb = hp.colsum(g, out=hp.grad_buffer(b)) if b is not None else None
            gw = spec.wgrad(inp, g, w)
        if on_side:
            hp.hold_until_join(g.device, g, inp)
        conv_grads[i] = (gw, gb)
        if i > 0:
            # inp is the ReLU output of layer i-1: mask the gradient in the dgrad epilogue
            g = spec.dgrad(g, w, inp.shape[1:3], EPI_RELU_MASK, aux0=inp)["lp"]
        elif need_input_grad:
            if sign_add is not None:
                y_f32, gy_partial = sign_add
                g = spec.dgrad(g, w, inp.shape[1:3], EPI_SIGN_ADD, aux0=y_f32, aux1=gy_partial)["lp"]
            else:
                g = spec.dgrad(g, w, inp.shape[1:3])["lp"]
        else:
            g = None
    hp.join_side_stream(gz_lp.device)
    for gw, gb in list(also_ready) + conv_grads[::-1]:
        hp.grads_ready((gw, gb))
    hp.grads_ready((), flush=True)
    return g, conv_grads


# ----------------------------------------------------------------------------------------
# h_s: HyperpriorSynthesisTransform (modelling/blocks/prior_synthesis.py:44-72)
# ----------------------------------------------------------------------------------------
def hs_forward(zt_lp, convs, kernels=(5, 5, 3), strides=(2, 2, 1), save=True):
    """zt_lp: NHWC lp.  kernels/strides already reversed (prior_synthesis.py:51).
    Returns sigma = clamp(exp(.),1e-10,1e10) NHWC fp32, state."""
    n = len(convs)
    S = _State()
    S.layers = []
    cur = zt_lp
    sigma = None
    for i, (w, b) in enumerate(convs):
        spec = hp.ConvSpec(True, w.shape[0], w.shape[1], kernels[i], strides[i])
        last = (i == n - 1)
        o = spec.forward(cur, w, b, EPI_EXP_CLAMP if last else EPI_RELU, want_lp=not last, want_f32=last)
        S.layers.append((spec, cur))
        if last:
            sigma = o["f32"]
        else:
            cur = o["lp"]
    S.convs = convs
    return sigma, (S if save else None)


def hs_backward(S, gpre_lp, need_input_grad=True, want_f32=False, join=True):
    """gpre_lp: NHWC lp gradient wrt the pre-exp activation (the exp/clamp chain rule is applied by the
    caller: hpc_conditional_bwd(fuse_exp_clamp) or hpc_exp_clamp_bwd).
    Returns (g wrt z~ (lp or fp32), conv grads)."""
    conv_grads = [None] * len(S.convs)
    g = gpre_lp
    for i in range(len(S.convs) - 1, -1, -1):
        spec, inp = S.layers[i]
        w, b = S.convs[i]
        with hp.side_stream(g.device) as on_side:       # small-K weight gradient: off the critical path, beside the dgrad chain
            gb = hp.colsum(g, out=hp.grad_buffer(b)) if b is not None else None
            gw = spec.wgrad(inp, g, w)
        if on_side:
            hp.hold_until_join(g.device, g, inp)
        conv_grads[i] = (gw, gb)
        if i > 0:
            g = spec.dgrad(g, w, inp.shape[1:3], EPI_RELU_MASK, aux0=inp)["lp"]
        elif need_input_grad:
            o = spec.dgrad(g, w, inp.shape[1:3], want_lp=not want_f32, want_f32=want_f32)
            g = o["f32"] if want_f32 else o["lp"]
        else:
            g = None
    if join:
        hp.join_side_stream(gpre_lp.device)
        for gw, gb in conv_grads[::-1]:
            hp.grads_ready((gw, gb))
    return g, conv_grads
