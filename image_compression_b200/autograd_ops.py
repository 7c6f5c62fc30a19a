"""torch.autograd.Function wrappers: the reference's autograd graph is
replaced by explicit backward schedules over the C-ABI kernels."""
import torch

from . import _lib
from . import blocks_impl as B
from . import hotpath as hp
from ._lib import DIST_GAUSS, DIST_LAPLACE, check, lib, ptr, stream

_DIST = {"LaplacianConditionalModel": DIST_LAPLACE, "GaussianConditionalModel": DIST_GAUSS}


def _f32c(t):
    return hp.contiguous_f32(t)


# ----------------------------------------------------------------------------------------
# bounds (modelling/layers/bound.py:28-59): LowerBound.apply(x, bound) / UpperBound.apply
# ----------------------------------------------------------------------------------------
class _Bound(torch.autograd.Function):
    IS_UPPER = 0

    @classmethod
    def _fwd(cls, ctx, inputs, bound):
        _lib.require_cuda(inputs)
        x = _f32c(inputs)
        b = float(bound)
        y = torch.empty_like(x)
        check(lib().hpc_bound_fwd(ptr(x), ptr(y), x.numel(), b, cls.IS_UPPER, stream()))
        ctx.save_for_backward(x)
        ctx.bound = b
        return y

    @classmethod
    def _bwd(cls, ctx, g):
        (x,) = ctx.saved_tensors
        g = _f32c(g)
        gx = torch.empty_like(x)
        check(lib().hpc_bound_bwd(ptr(x), ptr(g), ptr(gx), x.numel(), ctx.bound, cls.IS_UPPER, stream()))
        return gx, None


class LowerBound(_Bound):
    IS_UPPER = 0

    @staticmethod
    def forward(ctx, inputs, bound):
        return LowerBound._fwd(ctx, inputs, bound)

    @staticmethod
    def backward(ctx, g):
        return LowerBound._bwd(ctx, g)


class UpperBound(_Bound):
    IS_UPPER = 1

    @staticmethod
    def forward(ctx, inputs, bound):
        return UpperBound._fwd(ctx, inputs, bound)

    @staticmethod
    def backward(ctx, g):
        return UpperBound._bwd(ctx, g)


class NonNegFn(torch.autograd.Function):
    """NonNegativeParam.forward (gdn.py:59-62)."""

    @staticmethod
    def forward(ctx, param, bound, pedestal):
        _lib.require_cuda(param)
        ctx.save_for_backward(param)
        ctx.bound = bound
        return hp.nonneg_fwd(param.contiguous(), bound, pedestal)

    @staticmethod
    def backward(ctx, g):
        (param,) = ctx.saved_tensors
        return hp.nonneg_bwd(param.contiguous(), _f32c(g), ctx.bound), None, None


# ----------------------------------------------------------------------------------------
# GDN stand-alone (modelling/layers/gdn.py:79-88)
# ----------------------------------------------------------------------------------------
class GDNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma_param, beta_param, inverse, relu, beta_min):
        _lib.require_cuda(x, gamma_param, beta_param)
        xl = hp.to_nhwc(x, hp.lp_dtype())
        xsq, xr = hp.square(xl, relu)
        xin = xr if relu else xl
        gamma = hp.nonneg_fwd(gamma_param, hp.GAMMA_BOUND)
        beta = hp.nonneg_fwd(beta_param, hp.beta_bound(beta_min))
        y, nrm = hp.gdn_forward(xin, xsq, gamma, beta, inverse)
        ctx.save_for_backward(xin, xsq, nrm, gamma, gamma_param, beta_param)
        ctx.inverse, ctx.relu, ctx.beta_min, ctx.in_dtype = inverse, relu, beta_min, x.dtype
        return hp.as_nchw(hp.cast(y, x.dtype))

    @staticmethod
    def backward(ctx, g):
        xin, xsq, nrm, gamma, gamma_param, beta_param = ctx.saved_tensors
        gl = hp.to_nhwc(g, hp.lp_dtype())
        dx, dgamma, dbeta = hp.gdn_backward(gl, xin, xsq, nrm, gamma, ctx.inverse, need_dx=ctx.needs_input_grad[0])
        if dx is not None and ctx.relu:
            dx = hp.relu_mask(dx, xin)
        gg = hp.nonneg_bwd(gamma_param, dgamma, hp.GAMMA_BOUND)
        gb = hp.nonneg_bwd(beta_param, dbeta, hp.beta_bound(ctx.beta_min))
        gx = hp.as_nchw(hp.cast(dx, ctx.in_dtype)) if dx is not None else None
        return gx, gg, gb, None, None, None


# ----------------------------------------------------------------------------------------
# block-level functions (params are passed flat so autograd sees them)
# ----------------------------------------------------------------------------------------
def _split_ga(params, n):
    """[w0,b0,g0,be0, w1,b1,g1,be1, ..., w_{n-1},b_{n-1}] -> convs, gdns"""
    convs, gdns, i = [], [], 0
    for l in range(n):
        convs.append((params[i], params[i + 1]))
        i += 2
        if l < n - 1:
            gdns.append((params[i], params[i + 1]))
            i += 2
    return convs, gdns


def _join_ga(conv_grads, gdn_grads):
    out = []
    for l, (gw, gb) in enumerate(conv_grads):
        out += [gw, gb]
        if l < len(gdn_grads):
            out += list(gdn_grads[l])
    return out


class AnalysisFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, strides, *params):
        _lib.require_cuda(x, *params)
        if x.requires_grad:
            raise NotImplementedError("gradient wrt the input image is not part of the hot path (SURVEY 8d)")
        convs, gdns = _split_ga(params, len(strides))
        y, _, S = B.ga_forward(_f32c(x), convs, gdns, k, tuple(strides), want_abs=False)
        ctx.S = S
        return hp.as_nchw(y)

    @staticmethod
    def backward(ctx, gy):
        cg, gg = B.ga_backward(ctx.S, hp.to_nhwc(gy, hp.lp_dtype()))
        ctx.S = None
        return (None, None, None, *_join_ga(cg, gg))


class SynthesisFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, yt, k, strides, *params):
        _lib.require_cuda(yt, *params)
        convs, gdns = _split_ga(params, len(strides))
        raw, S = B.gs_forward(hp.to_nhwc(yt, hp.lp_dtype()), convs, gdns, k, tuple(strides))
        ctx.S, ctx.in_dtype = S, yt.dtype
        return raw

    @staticmethod
    def backward(ctx, g):
        gin, cg, gg = B.gs_backward(ctx.S, _f32c(g), need_input_grad=ctx.needs_input_grad[0])
        ctx.S = None
        gx = hp.as_nchw(hp.cast(gin, ctx.in_dtype)) if gin is not None else None
        return (gx, None, None, *_join_ga(cg, gg))


def _split_h(params, has_bias):
    convs, i = [], 0
    for hb in has_bias:
        if hb:
            convs.append((params[i], params[i + 1]))
            i += 2
        else:
            convs.append((params[i], None))
            i += 1
    return convs


def _join_h(conv_grads, has_bias):
    out = []
    for (gw, gb), hb in zip(conv_grads, has_bias):
        out.append(gw)
        if hb:
            out.append(gb)
    return out


class HyperAnalysisFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kernels, strides, has_bias, *params):
        _lib.require_cuda(x, *params)
        convs = _split_h(params, has_bias)
        z, S = B.ha_forward(hp.to_nhwc(x, hp.lp_dtype()), convs, tuple(kernels), tuple(strides))
        ctx.S, ctx.has_bias, ctx.in_dtype = S, has_bias, x.dtype
        return hp.as_nchw(z)

    @staticmethod
    def backward(ctx, gz):
        gin, cg = B.ha_backward(ctx.S, hp.to_nhwc(gz, hp.lp_dtype()), need_input_grad=ctx.needs_input_grad[0])
        ctx.S = None
        gx = hp.as_nchw(hp.cast(gin, ctx.in_dtype)) if gin is not None else None
        return (gx, None, None, None, *_join_h(cg, ctx.has_bias))


class HyperSynthesisFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, zt, kernels, strides, *params):
        _lib.require_cuda(zt, *params)
        convs = _split_h(params, [True] * len(kernels))
        sigma, S = B.hs_forward(hp.to_nhwc(zt, hp.lp_dtype()), convs, tuple(kernels), tuple(strides))
        ctx.S, ctx.sigma, ctx.in_dtype = S, sigma, zt.dtype
        return hp.as_nchw(sigma)

    @staticmethod
    def backward(ctx, gs):
        gpre = hp.exp_clamp_backward(ctx.sigma, hp.to_nhwc(gs, torch.float32))
        gin, cg = B.hs_backward(ctx.S, gpre, need_input_grad=ctx.needs_input_grad[0])
        ctx.S = ctx.sigma = None
        gx = hp.as_nchw(hp.cast(gin, ctx.in_dtype)) if gin is not None else None
        return (gx, None, None, *_join_h(cg, [True] * len(cg)))


# ----------------------------------------------------------------------------------------
# entropy models
# ----------------------------------------------------------------------------------------
CDF_KEYS = ("w0", "b0", "f0", "w1", "b1", "f1", "w2", "b2", "f2", "w3", "b3")


def _channel_last_index(ndim):
    """index list of CDFEstimator.forward (entropy_model.py:103-106): move dim 1 last."""
    idx = list(range(ndim))
    idx.pop(1)
    idx.append(1)
    return idx


class EntropyModelFn(torch.autograd.Function):
    """EntropyModel.forward (entropy_model.py:204-214) -> (z_tilde, probs, bits)."""

    @staticmethod
    def forward(ctx, z, bin_, training, noise, *cdf_params):
        _lib.require_cuda(z, *cdf_params)
        cdf = {k: v.contiguous() for k, v in zip(CDF_KEYS, cdf_params)}
        idx = _channel_last_index(z.dim())
        zl = _f32c(z.permute(*idx))             # (N, *spatial, C) rows
        nl = _f32c(noise.permute(*idx)) if noise is not None else None
        zt, _, probs, bits = hp.factorized_forward(zl, cdf, training, nl, bin_, want_probs=True)
        ctx.save_for_backward(zt, *cdf.values())
        ctx.bin, ctx.training, ctx.idx = bin_, training, idx
        inv = [idx.index(i) for i in range(len(idx))]
        ctx.inv = inv
        # z_tilde back in the caller's layout; probs in the reference's layout: the "inverse" permute
        # re-applies idx (entropy_model.py:113), i.e. (N,H,W,C).permute(0,2,3,1) -> (N,W,C,H)
        return zt.permute(*inv), probs.permute(*idx), bits.view(())

    @staticmethod
    def backward(ctx, g_zt, g_probs, g_bits):
        zt, *cdf_vals = ctx.saved_tensors
        cdf = dict(zip(CDF_KEYS, cdf_vals))
        inv_idx = [ctx.idx.index(i) for i in range(len(ctx.idx))]
        gp = _f32c(g_probs.permute(*inv_idx)) if g_probs is not None else None
        gzt = _f32c(g_zt.permute(*ctx.idx)) if g_zt is not None else None
        gb = _f32c(g_bits.reshape(1)) if g_bits is not None else None
        gz, grads = hp.factorized_backward(zt, cdf, ctx.training, gb, 1.0, gp, gzt, ctx.bin)
        return (gz.permute(*ctx.inv), None, None, None, *[grads[k] for k in CDF_KEYS])


def _same_dense_layout(a, b):
    return a.shape == b.shape and a.stride() == b.stride() and (
        a.is_contiguous() or (a.dim() == 4 and a.is_contiguous(memory_format=torch.channels_last)))


class ConditionalFn(torch.autograd.Function):
    """SymmetricConditionalModel.forward (entropy_model.py:280-295) -> (x_tilde, likelihood)."""

    @staticmethod
    def forward(ctx, x, scale, mean, bin_, training, dist, noise):
        _lib.require_cuda(x, scale)
        if x.dtype != torch.float32 or scale.dtype != torch.float32 or not _same_dense_layout(x, scale):
            x, scale = _f32c(x), _f32c(scale)
        if noise is not None:
            noise = noise.to(torch.float32)
            if noise.stride() != x.stride():
                n2 = torch.empty_like(x)
                n2.copy_(noise)
                noise = n2
        xt, _, probs, _ = hp.conditional_forward(x, scale, dist, training, noise, bin_, float(mean), want_probs=True)
        ctx.save_for_backward(xt, scale)
        ctx.args = (dist, training, bin_, float(mean))
        return xt, probs

    @staticmethod
    def backward(ctx, g_xt, g_probs):
        xt, scale = ctx.saved_tensors
        dist, training, bin_, mean = ctx.args

        def like(g):
            if g is None:
                return None
            if g.dtype == torch.float32 and g.stride() == xt.stride():
                return g
            o = torch.empty_like(xt)
            o.copy_(g)
            return o

        gy, gs, _ = hp.conditional_backward(xt, scale, dist, training, None, 0.0, like(g_probs), like(g_xt),
                                            fuse_exp_clamp=False, bin_=bin_, mean=mean)
        return gy, gs, None, None, None, None, None


class BitsFn(torch.autograd.Function):
    """BaseEntropyModel._ce_loss (entropy_model.py:171-185)."""

    @staticmethod
    def forward(ctx, probs):
        _lib.require_cuda(probs)
        p = probs if (probs.dtype == torch.float32 and (probs.is_contiguous() or (
            probs.dim() == 4 and probs.is_contiguous(memory_format=torch.channels_last)))) else _f32c(probs)
        bits = torch.empty((1,), dtype=torch.float32, device=p.device)
        check(lib().hpc_bits_fwd(ptr(p), p.numel(), ptr(bits), ptr(hp.reduce_scratch(p.device)), stream()))
        ctx.save_for_backward(p)
        return bits.view(())

    @staticmethod
    def backward(ctx, g):
        (p,) = ctx.saved_tensors
        gp = torch.empty_like(p)
        check(lib().hpc_bits_bwd(ptr(p), ptr(_f32c(g.reshape(1))), p.numel(), ptr(gp), stream()))
        return gp


# ----------------------------------------------------------------------------------------
# distortion losses
# ----------------------------------------------------------------------------------------
class MSEFn(torch.autograd.Function):
    """nn.MSELoss(reduction=mean|sum)(input, target) (modelling/loss.py:25)."""

    @staticmethod
    def forward(ctx, inp, target, reduction):
        _lib.require_cuda(inp, target)
        if reduction not in ("mean", "sum"):
            raise NotImplementedError("MSE reduction 'none' (utils/config.py:75 default) makes total_loss a tensor; "
                                      "every shipped yaml sets 'mean' (configs/psnr_256.yaml:15)")
        a, b = _f32c(inp), _f32c(target)
        n0 = a.shape[0] if a.dim() > 0 else 1
        _, sse = hp.clamp01_sse(a.reshape(n0, -1), b.reshape(n0, -1), 1.0, do_clamp=False, want_xt=False)
        scale = 1.0 / a.numel() if reduction == "mean" else 1.0
        ctx.save_for_backward(a, b)
        ctx.scale = scale
        return hp.colsum(sse.view(-1, 1), scale).view(())

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        gdev = _f32c(g.reshape(1))
        gb = hp.clamp01_backward(b, a, None, ctx.scale, gdev, do_clamp=False) if ctx.needs_input_grad[1] else None
        ga = hp.clamp01_backward(a, b, None, ctx.scale, gdev, do_clamp=False) if ctx.needs_input_grad[0] else None
        return ga, gb, None


class SsimLossFn(torch.autograd.Function):
    """SSIMLoss / MS_SSIMLoss forward+backward (modelling/loss.py:48-188)."""

    @staticmethod
    def forward(ctx, img1, img2, kind, flags, eps, weights, fs, sigma, c1, c2, max_val):
        _lib.require_cuda(img1, img2)
        a, b = _f32c(img1), _f32c(img2)
        pyr = hp.SsimPyramid(a, b, len(weights), fs, sigma, c1, c2, max_val)
        out = pyr.combine(kind, flags, eps, weights)
        ctx.pyr, ctx.meta = pyr, (kind, flags, eps, weights, fs, sigma, c1, c2, max_val)
        ctx.save_for_backward(a, b)
        return out.view(()) if out.numel() == 1 and not (kind == _lib.SSIM_LOSS and flags & 1) else out

    @staticmethod
    def backward(ctx, g):
        kind, flags, eps, weights, fs, sigma, c1, c2, max_val = ctx.meta
        a, b = ctx.saved_tensors
        gdev = _f32c(g.reshape(-1))
        g2 = ctx.pyr.backward(kind, flags, eps, weights, gdev) if ctx.needs_input_grad[1] else None
        g1 = None
        if ctx.needs_input_grad[0]:  # the index is symmetric in its two images
            pyr = hp.SsimPyramid(b, a, len(weights), fs, sigma, c1, c2, max_val)
            g1 = pyr.backward(kind, flags, eps, weights, gdev)
        ctx.pyr = None
        return (g1, g2) + (None,) * 9


# ----------------------------------------------------------------------------------------
# the whole model as ONE autograd node (Compressor2018.forward, bmshl2018.py:68-98)
# ----------------------------------------------------------------------------------------
class ModelMeta:
    """Static description of a Compressor2018 instance (taken from cfg by the module)."""

    def __init__(self, k, strides, hyper_kernels, hyper_strides, dist, bin_, lmbda, loss_names, reduction, ssim):
        self.k, self.strides = k, tuple(strides)
        self.hk, self.hs = tuple(hyper_kernels), tuple(hyper_strides)
        self.dist, self.bin, self.lmbda = dist, float(bin_), float(lmbda)
        self.loss_names, self.reduction, self.ssim = tuple(loss_names), reduction, ssim
        n, nh = len(self.strides), len(self.hk)
        self.n_ga = 2 * n + 2 * (n - 1)
        self.ha_bias = [i < nh - 1 for i in range(nh)]       # prior_analysis.py:51
        self.n_ha = sum(2 if b else 1 for b in self.ha_bias)
        self.n_hs = 2 * nh
        self.n_gs = self.n_ga
        self.n_cdf = len(CDF_KEYS)

    def fusable(self):
        """The fused node covers scalar distortion terms; SSIMLoss(log_scale) returns one value per
        image (loss.py:72) and goes through the composed path instead."""
        for name in self.loss_names:
            if name == "MSE" and self.reduction not in ("mean", "sum"):
                return False
            if name == "SSIMLoss" and self.ssim.LOG_SCALE:
                return False
            if name not in ("MSE", "MS_SSIMLoss", "SSIMLoss"):
                return False
        if sum(name != "MSE" for name in self.loss_names) > 1:
            return False  # one SSIM-type term per fused node
        return len(self.loss_names) <= 4

    def ssim_kind(self, name):
        s = self.ssim
        c1, c2 = (s.K1 * s.MAX_VAL) ** 2, (s.K2 * s.MAX_VAL) ** 2
        if name == "MS_SSIMLoss":
            from .modelling.loss import MS_SSIM_DEFAULT_WEIGHTS
            kind = _lib.MSSSIM_LOSS_LOG if s.LOG_SCALE else _lib.MSSSIM_LOSS_PLAIN
            return kind, 0, (s.EPS if s.LOG_SCALE else 0.0), tuple(MS_SSIM_DEFAULT_WEIGHTS), c1, c2
        return _lib.SSIM_LOSS, (1 if s.LOG_SCALE else 0), (s.EPS if s.LOG_SCALE else 0.0), (1.0,), c1, c2


def _split_model_params(m, params):
    i = 0
    ga_convs, ga_gdns = _split_ga(params[i:i + m.n_ga], len(m.strides)); i += m.n_ga
    ha_convs = _split_h(params[i:i + m.n_ha], m.ha_bias); i += m.n_ha
    hs_convs = _split_h(params[i:i + m.n_hs], [True] * len(m.hk)); i += m.n_hs
    gs_convs, gs_gdns = _split_ga(params[i:i + m.n_gs], len(m.strides)); i += m.n_gs
    cdf = dict(zip(CDF_KEYS, params[i:i + m.n_cdf]))
    return ga_convs, ga_gdns, ha_convs, hs_convs, gs_convs, gs_gdns, cdf


def build_pack_plan(m, params):
    """hotpath.PackPlan with every packed operand the fused forward + backward of this model read (the same
    ConvSpec geometry blocks_impl builds layer by layer), for the current compute mode."""
    ga_convs, ga_gdns, ha_convs, hs_convs, gs_convs, gs_gdns, _ = _split_model_params(m, params)
    dt = hp.lp_dtype()
    plan = hp.PackPlan(dt)
    n = len(m.strides)
    for i, (w, _) in enumerate(ga_convs):
        if i == 0:
            if hp._thin_umma_ok(w.shape[0], w.shape[1], dt):
                plan.add_c3(w, "k")
        else:
            plan.add_spec(hp.ConvSpec(False, w.shape[1], w.shape[0], m.k, m.strides[i]), w)
    for i, (w, _) in enumerate(ha_convs):
        plan.add_spec(hp.ConvSpec(False, w.shape[1], w.shape[0], m.hk[i], m.hs[i]), w)
    for i, (w, _) in enumerate(hs_convs):
        plan.add_spec(hp.ConvSpec(True, w.shape[0], w.shape[1], m.hk[::-1][i], m.hs[::-1][i]), w)
    for i, (w, _) in enumerate(gs_convs):
        if i == n - 1:
            if hp._thin_umma_ok(w.shape[0], w.shape[1], dt):
                if w.shape[0] <= 192:
                    plan.add_c3(w, "n")
                plan.add_c3(w, "k")
        else:
            plan.add_spec(hp.ConvSpec(True, w.shape[0], w.shape[1], m.k, m.strides[i]), w)
    for gp, bp in list(ga_gdns) + list(gs_gdns):
        if hp.gdn_fused_ok(dt, bp.numel()):
            plan.add_gdn(gp, bp, 1e-6)
    return plan


def _nhwc_noise(noise):
    return None if noise is None else _f32c(noise.permute(0, 2, 3, 1))


class Compressor2018Fn(torch.autograd.Function):
    """(x, params...) -> (x_tilde, [z_entropy, y_entropy, bpp, total_loss, d0..d3])."""

    @staticmethod
    def forward(ctx, x, meta, training, noise_z, noise_y, *params):
        _lib.require_cuda(x, *params)
        m = meta
        x = _f32c(x)
        N, C, H, W = x.shape
        # forward() itself always runs under no_grad and needs_input_grad ignores the caller's grad mode, so the
        # caller says whether a backward can follow (training = (mode flag, grad enabled)); without one nothing is
        # saved and the GDN kernels skip their r output
        training, grad_enabled = training if isinstance(training, tuple) else (training, True)
        save = grad_enabled and any(ctx.needs_input_grad)
        ga_convs, ga_gdns, ha_convs, hs_convs, gs_convs, gs_gdns, cdf = _split_model_params(m, params)
        cdf = {k: v.contiguous() for k, v in cdf.items()}

        y, yabs, Sga = B.ga_forward(x, ga_convs, ga_gdns, m.k, m.strides, want_abs=True, save=save)
        z, Sha = B.ha_forward(yabs, ha_convs, m.hk, m.hs, save=save)
        zt, ztl, _, zbits = hp.factorized_forward(z, cdf, training, _nhwc_noise(noise_z), m.bin, want_probs=False,
                                                  want_lp=True)
        sigma, Shs = B.hs_forward(ztl, hs_convs, m.hk[::-1], m.hs[::-1], save=save)
        yt, ytl, _, ybits = hp.conditional_forward(y, sigma, m.dist, training, _nhwc_noise(noise_y), m.bin, 0.0,
                                                   want_probs=False, want_lp=True)
        raw, Sgs = B.gs_forward(ytl, gs_convs, gs_gdns, m.k, m.strides, save=save)

        xt, sse = hp.clamp01_sse(x, raw, 1.0, do_clamp=True, want_xt=True)
        terms, pyrs = [], []
        for name in m.loss_names:
            if name == "MSE":
                terms.append((sse, 1.0 / x.numel() if m.reduction == "mean" else 1.0))
                pyrs.append(None)
            else:
                kind, flags, eps, weights, c1, c2 = m.ssim_kind(name)
                pyr = hp.SsimPyramid(x, xt, len(weights), m.ssim.FILTER_SIZE, m.ssim.FILTER_SIGMA, c1, c2,
                                     m.ssim.MAX_VAL)
                terms.append((pyr.combine(kind, flags, eps, weights), 1.0))
                pyrs.append((pyr, kind, flags, eps, weights))
        inv_np = 1.0 / float(N * H * W)
        out = hp.rd_combine(zbits, ybits, terms, m.lmbda, inv_np)
        ctx.mark_non_differentiable(xt)
        if save:
            ctx.st = (m, training, x, raw, y, sigma, yt, zt, cdf, Sga, Sha, Shs, Sgs, terms, pyrs, inv_np)
        else:
            ctx.st = None
        return xt, out

    @staticmethod
    def backward(ctx, _g_xt, g_out):
        if ctx.st is None:
            raise RuntimeError("Compressor2018Fn.backward called without saved state")
        m, training, x, raw, y, sigma, yt, zt, cdf, Sga, Sha, Shs, Sgs, terms, pyrs, inv_np = ctx.st
        ctx.st = None
        g = _f32c(g_out)[3:4]  # dL/d total_loss (device scalar; no host sync)
        # --- distortion -> raw x~ -------------------------------------------------------
        g_xt, mse_coef, has_mse = None, 0.0, False
        for (t, scale), pyr in zip(terms, pyrs):
            if pyr is None:
                has_mse, mse_coef = True, m.lmbda * scale
            else:
                p, kind, flags, eps, weights = pyr
                gi = p.backward(kind, flags, eps, weights, g, g_scale=m.lmbda)
                assert g_xt is None, "ModelMeta.fusable() admits one SSIM-type term"
                g_xt = gi
        g_raw = hp.clamp01_backward(raw, x if has_mse else None, g_xt, mse_coef, g, do_clamp=True)
        # --- g_s ---------------------------------------------------------------------------
        g_yt, gs_cg, gs_gg = B.gs_backward(Sgs, g_raw, need_input_grad=True)
        # (every layer above reported its gradients through hp.grads_ready: finished slot ranges of
        #  sharding.FlatGradients are being all-reduced while the rest of this backward runs)
        # --- conditional likelihood (+ exp/clamp of h_s fused) ------------------------------
        gy_partial, _, gpre = hp.conditional_backward(yt, sigma, m.dist, training, g, inv_np, None, g_yt,
                                                      fuse_exp_clamp=True, want_gs_f32=False, want_gs_lp=True,
                                                      bin_=m.bin, mean=0.0)
        # --- h_s -> z~ -> factorised prior -> z ---------------------------------------------
        # (the small-K weight gradients of h_s and h_a run on a side stream beside the dgrad chain; joined at the end
        #  of h_a's backward, where their gradients are reported ready)
        g_zt, hs_cg = B.hs_backward(Shs, gpre, need_input_grad=True, want_f32=True, join=False)
        gz, cdf_grads = hp.factorized_backward(zt, cdf, training, g, inv_np, None, g_zt, m.bin)
        # --- h_a, with d|y|/dy and the three-way sum fused into its last dgrad ---------------
        gy, ha_cg = B.ha_backward(Sha, hp.cast(gz, hp.lp_dtype()), sign_add=(y, gy_partial), also_ready=hs_cg[::-1])
        hp.grads_ready(list(cdf_grads.values()), flush=True)
        # --- g_a ---------------------------------------------------------------------------
        ga_cg, ga_gg = B.ga_backward(Sga, gy)
        grads = (_join_ga(ga_cg, ga_gg) + _join_h(ha_cg, m.ha_bias) + _join_h(hs_cg, [True] * len(hs_cg)) +
                 _join_ga(gs_cg, gs_gg) + [cdf_grads[k] for k in CDF_KEYS])
        return (None, None, None, None, None, *grads)


class CdfLogitsFn(torch.autograd.Function):
    """CDFEstimator.forward stand-alone (entropy_model.py:101-114)."""

    @staticmethod
    def forward(ctx, x, *cdf_params):
        _lib.require_cuda(x, *cdf_params)
        cdf = {k: v.contiguous() for k, v in zip(CDF_KEYS, cdf_params)}
        hp._check_cdf_dims(cdf)
        idx = _channel_last_index(x.dim())
        xl = _f32c(x.permute(*idx))
        C = xl.shape[-1]
        out = torch.empty_like(xl)
        s = hp._cdf_struct(cdf)
        check(lib().hpc_cdf_logits_fwd(ptr(xl), _lib.byref(s), xl.numel() // C, C, ptr(out), stream()))
        ctx.save_for_backward(xl, *cdf.values())
        ctx.idx = idx
        return out.permute(*idx)   # the reference re-applies idx as its "inverse" (entropy_model.py:113)

    @staticmethod
    def backward(ctx, g):
        xl, *vals = ctx.saved_tensors
        cdf = dict(zip(CDF_KEYS, vals))
        inv = [ctx.idx.index(i) for i in range(len(ctx.idx))]
        gl = _f32c(g.permute(*inv))
        C = xl.shape[-1]
        gx = torch.empty_like(xl)
        grads = {k: torch.zeros_like(v) for k, v in cdf.items()}
        s, gs = hp._cdf_struct(cdf), hp._cdf_struct(grads)
        check(lib().hpc_cdf_logits_bwd(ptr(xl), _lib.byref(s), xl.numel() // C, C, ptr(gl), ptr(gx), _lib.byref(gs),
                                       stream()))
        return (gx.permute(*inv), *[grads[k] for k in CDF_KEYS])


class CdfLayerFn(torch.autograd.Function):
    """CDFLayer.forward stand-alone (entropy_model.py:67-78): x (N,C,D,1) -> (N,C,D',1)."""

    @staticmethod
    def forward(ctx, x, weight, bias, factor):
        _lib.require_cuda(x, weight, bias)
        _, C, dout, din = weight.shape
        if x.dim() != 4 or x.shape[1] != C or x.shape[2] != din or x.shape[3] != 1:
            raise ValueError(f"CDFLayer: expected x of shape (N,{C},{din},1), got {tuple(x.shape)}")
        xc, w, b = _f32c(x), _f32c(weight), _f32c(bias)
        f = _f32c(factor) if factor is not None else None
        out = torch.empty((x.shape[0], C, dout, 1), dtype=torch.float32, device=x.device)
        check(lib().hpc_cdf_layer_fwd(ptr(xc), ptr(w), ptr(b), ptr(f), x.shape[0], C, din, dout, ptr(out), stream()))
        ctx.save_for_backward(xc, w, b, f)
        return out

    @staticmethod
    def backward(ctx, g):
        xc, w, b, f = ctx.saved_tensors
        _, C, dout, din = w.shape
        gx = torch.empty_like(xc) if ctx.needs_input_grad[0] else None
        gw, gb = torch.empty_like(w), torch.empty_like(b)
        gf = torch.empty_like(f) if f is not None else None
        check(lib().hpc_cdf_layer_bwd(ptr(xc), ptr(w), ptr(b), ptr(f), xc.shape[0], C, din, dout, ptr(_f32c(g)), ptr(gx),
                                      ptr(gw), ptr(gb), ptr(gf), stream()))
        return gx, gw, gb, gf


class AbsFn(torch.autograd.Function):
    """torch.abs(y) of bmshl2018.py:72 (composed path)."""

    @staticmethod
    def forward(ctx, x):
        _lib.require_cuda(x)
        dense = x.dtype == torch.float32 and (x.is_contiguous() or (
            x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last)))
        xc = x if dense else _f32c(x)
        y = torch.empty_like(xc)
        check(lib().hpc_abs_fwd(ptr(xc), ptr(y), xc.numel(), stream()))
        ctx.save_for_backward(xc)
        return y

    @staticmethod
    def backward(ctx, g):
        (x,) = ctx.saved_tensors
        if g.dtype != torch.float32 or g.stride() != x.stride():
            g2 = torch.empty_like(x)
            g2.copy_(g)
            g = g2
        gx = torch.empty_like(x)
        check(lib().hpc_abs_bwd(ptr(x), ptr(g), ptr(gx), x.numel(), stream()))
        return gx
