"""CPU oracle for the scale-hyperprior hot path (TEST INFRASTRUCTURE, NOT THE PRODUCT).

A functional, fp32, torch-CPU restatement of the reference's `Compressor2018` forward
(+ autograd backward), its entropy models, losses and metrics.  Every function cites the
reference file:line (paths relative to /root/reference) it follows.  The reference's
arithmetic is all stock `torch`, so the restatement uses the same ATen
CPU ops (conv2d, conv_transpose2d, softplus, expm1, erf, ...) in the same order.

Who may use this module: `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
`cpu_baseline` / `--impl reference` legs -- as the checker or the timed CPU baseline only.
Nothing under `image_compression_b200/` imports it; the product path fails loudly when the
CUDA library is missing.

PARITY PINNING: the reference ships no golden vectors; its only numeric
assertions are test/test_gdn.py:20-48.  This oracle is pinned by
`tests/golden/make_golden.py`, which runs the UNMODIFIED reference modules in the build
container (through oracle/ref_shim.py) on seeded inputs, checks this file against them
(bit-exact init, outputs and grads to fp32 round-off) and writes the committed fixtures in
`tests/golden/*.npz` that `tests/test_oracle_golden.py` re-checks everywhere.

Parameters travel as a dict keyed by the reference's `state_dict` names.
"""
import math
from types import SimpleNamespace

import torch
import torch.nn.functional as F

LOG2 = math.log(2.0)    # modelling/blocks/entropy_model.py:44
LOG10 = math.log(10.0)  # utils/metric.py:23
MS_SSIM_WEIGHTS = (0.0448, 0.2856, 0.3001, 0.2363, 0.1333)  # modelling/loss.py:127, utils/metric.py:103


# --------------------------------------------------------------------------------------
# configuration (utils/config.py:51-84 defaults; yaml overrides as keyword arguments)
# --------------------------------------------------------------------------------------
def default_cfg(lmbda=256.0, loss_names=("MSE",), reduction="mean", log_scale=False,
                conditional="LaplacianConditionalModel", channels=192, latent_channels=None):
    """Attribute bag with the reference's hot-path keys.

    Defaults = utils/config.py:51-84 with the overrides every shipped yaml applies
    (configs/psnr_256.yaml:13-15: DISTORTION_LOSS_WEIGHT, REDUCTION "mean";
     configs/ssim_64.yaml:13-17: MS_SSIMLoss + LOG_SCALE True).
    """
    ns = SimpleNamespace
    return ns(
        DATA=ns(IN_CHANNELS=3),
        MODEL=ns(
            META_ARCHITECTURE="Compressor2018",
            STRIDES=[2, 2, 2, 2], CONV_KERNEL=5,
            INTER_CHANNELS=channels,
            LATENT_CHANNELS=channels if latent_channels is None else latent_channels,
            HYPER_PRIOR=ns(STRIDES=[1, 2, 2], KERNELS=[3, 5, 5]),
            ENTROPY_MODEL=ns(DIMS=[3, 3, 3], INIT_SCALE=10, BIN=1.0, PROB_EPS=1e-10,
                             CONDITIONAL_MODEL=conditional),
            LOSS=ns(DISTORTION_LOSS_NAMES=list(loss_names), DISTORTION_LOSS_WEIGHT=float(lmbda),
                    REDUCTION=reduction,
                    SSIM=ns(MAX_VAL=255.0, FILTER_SIZE=11, FILTER_SIGMA=1.5, K1=0.01, K2=0.03,
                            LOG_SCALE=log_scale, EPS=1e-5),
                    MS_SSIM_WEIGHTS=list(MS_SSIM_WEIGHTS)),
        ),
    )


# --------------------------------------------------------------------------------------
# parameter initialisation, RNG-stream identical to modelling.build_model(cfg)
# --------------------------------------------------------------------------------------
def _nonneg_init(init_val, minimum=0.0, offset=2 ** -18):
    """modelling/layers/gdn.py:48-57: stored param = sqrt(max(init + pedestal, pedestal))."""
    pedestal = torch.tensor(float(offset) ** 2)
    return torch.sqrt(torch.max(init_val + pedestal, pedestal))


def init_params(cfg, seed=None):
    """Random init in the reference's construction order (bmshl2018.py:54-66).

    `nn.Conv2d`/`nn.ConvTranspose2d` constructors draw their default init from the global
    RNG before xavier_normal_ overwrites it (analysis.py:55-59), so the same torch layer
    classes are constructed here to consume the same stream; with the same seed the result
    equals `build_model(cfg).state_dict()` bit for bit (checked by make_golden.py).
    """
    if seed is not None:
        torch.manual_seed(seed)
    m = cfg.MODEL
    cin, ci, cl, k = cfg.DATA.IN_CHANNELS, m.INTER_CHANNELS, m.LATENT_CHANNELS, m.CONV_KERNEL
    p = {}

    def conv_init(name, layer, bias=True):
        torch.nn.init.xavier_normal_(layer.weight.data, math.sqrt(2))
        p[name + ".weight"] = layer.weight.data.clone()
        if bias:
            torch.nn.init.constant_(layer.bias.data, 0.01)
            p[name + ".bias"] = layer.bias.data.clone()

    def gdn_init(name, ch):  # gdn.py:69-77
        p[name + ".gamma.param"] = _nonneg_init(torch.eye(ch).view(ch, ch, 1, 1) * 0.1)
        p[name + ".beta.param"] = _nonneg_init(torch.ones((ch,)), minimum=1e-6)

    ns = len(m.STRIDES)
    # g_a  (modelling/blocks/analysis.py:44-67)
    for i, s in enumerate(m.STRIDES):
        a = cin if i == 0 else ci
        b = cl if i == ns - 1 else ci
        conv_init(f"analysis_transform.layers.{2 * i}", torch.nn.Conv2d(a, b, k, stride=s, padding=k // 2))
        if i < ns - 1:
            gdn_init(f"analysis_transform.layers.{2 * i + 1}", b)
    # h_a  (modelling/blocks/prior_analysis.py:43-68)
    hs, hk = m.HYPER_PRIOR.STRIDES, m.HYPER_PRIOR.KERNELS
    nh = len(hs)
    for i, (s, kk) in enumerate(zip(hs, hk)):
        use_bias = i < nh - 1
        a = cl if i == 0 else ci
        conv_init(f"prior_analysis._layers.{2 * i}",
                  torch.nn.Conv2d(a, ci, kk, stride=s, padding=kk // 2, bias=use_bias), bias=use_bias)
    # h_s  (modelling/blocks/prior_synthesis.py:44-66)
    for i, (s, kk) in enumerate(zip(reversed(hs), reversed(hk))):
        b = ci if i < nh - 1 else cl
        conv_init(f"prior_synthesis._layers.{2 * i}",
                  torch.nn.ConvTranspose2d(ci, b, kk, stride=s, padding=kk // 2, bias=True,
                                           output_padding=s - 1))
    # g_s  (modelling/blocks/synthesis.py:44-67)
    for i, s in enumerate(m.STRIDES):
        a = cl if i == 0 else ci
        b = cin if i == ns - 1 else ci
        conv_init(f"synthesis_transform.layers.{2 * i}",
                  torch.nn.ConvTranspose2d(a, b, k, stride=s, padding=k // 2, output_padding=s - 1))
        if i < ns - 1:
            gdn_init(f"synthesis_transform.layers.{2 * i + 1}", b)
    # factorised prior MLP (modelling/blocks/entropy_model.py:51-65, 88-98)
    # the reference mutates cfg...DIMS in place (entropy_model.py:92-93); a copy is used here
    dims = [1] + list(m.ENTROPY_MODEL.DIMS) + [1]
    nl = len(dims) - 1
    scale = float(m.ENTROPY_MODEL.INIT_SCALE) ** (1 / nl)
    import numpy as np
    for i in range(nl):
        wv = np.log(np.expm1(1 / scale / dims[i + 1]))
        name = f"entropy_model._cdf_estimator.layers.{i}"
        p[name + ".weight"] = torch.nn.init.constant_(torch.empty(1, cl, dims[i + 1], dims[i]), wv)
        p[name + ".bias"] = torch.nn.init.uniform_(torch.empty(1, cl, dims[i + 1], 1), -0.5, 0.5)
        if i < nl - 1:
            p[name + ".factor"] = torch.nn.init.zeros_(torch.empty(1, cl, dims[i + 1], 1))
    return p


# --------------------------------------------------------------------------------------
# bounds and GDN
# --------------------------------------------------------------------------------------
class _LowerBound(torch.autograd.Function):
    """modelling/layers/bound.py:28-42: max(x,b); grad passes where x>=b or grad<0."""

    @staticmethod
    def forward(ctx, x, bound):
        b = torch.ones_like(x) * bound
        ctx.save_for_backward(x, b)
        return torch.max(x, b)

    @staticmethod
    def backward(ctx, g):
        x, b = ctx.saved_tensors
        return ((x >= b) | (g < 0)).type(g.dtype) * g, None


class _UpperBound(torch.autograd.Function):
    """modelling/layers/bound.py:45-59: min(x,b); grad passes where x<=b or grad>0."""

    @staticmethod
    def forward(ctx, x, bound):
        b = torch.ones_like(x) * bound
        ctx.save_for_backward(x, b)
        return torch.min(x, b)

    @staticmethod
    def backward(ctx, g):
        x, b = ctx.saved_tensors
        return ((x <= b) | (g > 0)).type(g.dtype) * g, None


def lower_bound(x, bound):
    return _LowerBound.apply(x, bound)


def upper_bound(x, bound):
    return _UpperBound.apply(x, bound)


def nonneg_param(param, minimum=0.0, offset=2 ** -18):
    """modelling/layers/gdn.py:48-62.  NB the bound uses pedestal**2 (gdn.py:53)."""
    pedestal = torch.tensor(float(offset) ** 2)
    bound = (float(minimum) + pedestal.clone() ** 2) ** 0.5
    val = lower_bound(param, bound)
    return val * val - pedestal


def gdn(x, gamma_param, beta_param, inverse=False, relu=False, beta_min=1e-6):
    """modelling/layers/gdn.py:79-88."""
    if relu:
        x = F.relu(x)
    gamma = nonneg_param(gamma_param)
    beta = nonneg_param(beta_param, minimum=beta_min)
    norm = torch.sqrt(F.conv2d(x * x, gamma, beta))
    return x * norm if inverse else x / norm


# --------------------------------------------------------------------------------------
# transforms
# --------------------------------------------------------------------------------------
def analysis_transform(p, x, cfg, prefix="analysis_transform.layers."):
    """modelling/blocks/analysis.py:44-71: 4x Conv2d(k5,s2,p2) with GDN after the first 3."""
    k = cfg.MODEL.CONV_KERNEL
    n = len(cfg.MODEL.STRIDES)
    for i, s in enumerate(cfg.MODEL.STRIDES):
        x = F.conv2d(x, p[f"{prefix}{2 * i}.weight"], p[f"{prefix}{2 * i}.bias"], stride=s, padding=k // 2)
        if i < n - 1:
            x = gdn(x, p[f"{prefix}{2 * i + 1}.gamma.param"], p[f"{prefix}{2 * i + 1}.beta.param"])
    return x


def synthesis_transform(p, x, cfg, prefix="synthesis_transform.layers."):
    """modelling/blocks/synthesis.py:44-71: 4x ConvTranspose2d(k5,s2,p2,op1), GDN(inverse=False!) :64-66."""
    k = cfg.MODEL.CONV_KERNEL
    n = len(cfg.MODEL.STRIDES)
    for i, s in enumerate(cfg.MODEL.STRIDES):
        x = F.conv_transpose2d(x, p[f"{prefix}{2 * i}.weight"], p[f"{prefix}{2 * i}.bias"],
                               stride=s, padding=k // 2, output_padding=s - 1)
        if i < n - 1:
            x = gdn(x, p[f"{prefix}{2 * i + 1}.gamma.param"], p[f"{prefix}{2 * i + 1}.beta.param"])
    return x


def prior_analysis(p, x, cfg, prefix="prior_analysis._layers."):
    """modelling/blocks/prior_analysis.py:43-71: conv3 s1, ReLU, conv5 s2, ReLU, conv5 s2 (no bias)."""
    hs, hk = cfg.MODEL.HYPER_PRIOR.STRIDES, cfg.MODEL.HYPER_PRIOR.KERNELS
    n = len(hs)
    for i, (s, k) in enumerate(zip(hs, hk)):
        x = F.conv2d(x, p[f"{prefix}{2 * i}.weight"], p.get(f"{prefix}{2 * i}.bias"), stride=s, padding=k // 2)
        if i < n - 1:
            x = F.relu(x)
    return x


def prior_synthesis(p, x, cfg, prefix="prior_synthesis._layers."):
    """modelling/blocks/prior_synthesis.py:44-72: mirrored deconvs, then exp and clamp(1e-10,1e10)."""
    hs, hk = cfg.MODEL.HYPER_PRIOR.STRIDES, cfg.MODEL.HYPER_PRIOR.KERNELS
    n = len(hs)
    for i, (s, k) in enumerate(zip(reversed(hs), reversed(hk))):
        x = F.conv_transpose2d(x, p[f"{prefix}{2 * i}.weight"], p[f"{prefix}{2 * i}.bias"],
                               stride=s, padding=k // 2, output_padding=s - 1)
        if i < n - 1:
            x = F.relu(x)
    return torch.clamp(x.exp(), 1e-10, 1e10)


# --------------------------------------------------------------------------------------
# entropy models
# --------------------------------------------------------------------------------------
def quantize(x, training, noise=None, bin_=1.0):
    """entropy_model.py:216-236 / :319-339: train x+(U[0,1)-bin/2) (noise detached); eval round()."""
    if training:
        if noise is None:
            noise = torch.rand_like(x)
        return x + (noise - bin_ / 2).detach()
    return torch.round(x)


def cdf_logits(p, x, prefix="entropy_model._cdf_estimator.layers."):
    """CDFEstimator.forward (entropy_model.py:101-114) with the CDFLayer MLP (:67-78).

    Returns logits in the reference's (quirky) output layout: the "inverse" permute
    re-applies the forward index list, so a 4-D (N,C,H,W) input comes back as (N,W,C,H)
    (SURVEY §3.5 item 12).
    """
    N, C, *sp = x.size()
    idx = list(range(len(sp) + 2))
    idx.pop(1)
    idx.append(1)
    h = x.permute(*idx).reshape(-1, C, 1, 1)
    nl = 0
    while f"{prefix}{nl}.weight" in p:
        nl += 1
    for i in range(nl):
        h = torch.matmul(F.softplus(p[f"{prefix}{i}.weight"]), h) + p[f"{prefix}{i}.bias"]
        if i < nl - 1:
            h = h + torch.tanh(h) * torch.tanh(p[f"{prefix}{i}.factor"])
    return h.view(N, *sp, C).permute(*idx)


def factorized_prob_mass(p, xq, bin_=1.0):
    """EntropyModel._prob_mass (entropy_model.py:259-269)."""
    half = bin_ / 2
    lower = cdf_logits(p, xq - half)
    upper = cdf_logits(p, xq + half)
    sign = -torch.sign(lower + upper).detach()
    return sign * (torch.sigmoid(upper * sign) - torch.sigmoid(lower * sign))


def ce_bits(probs):
    """BaseEntropyModel._ce_loss (entropy_model.py:171-185)."""
    return torch.clamp(-1.0 * torch.log(probs + 1e-10) / LOG2, 0, 50).sum()


def entropy_model(p, z, training, noise=None, bin_=1.0):
    """EntropyModel.forward (entropy_model.py:204-214) -> (z_tilde, probs, bits)."""
    zq = quantize(z, training, noise, bin_)
    probs = factorized_prob_mass(p, zq, bin_)
    return zq, probs, ce_bits(probs)


def _std_cdf(t, kind):
    """entropy_model.py:362 (Normal.cdf: torch/distributions/normal.py:104-109) and
    :375 (Laplace.cdf: torch/distributions/laplace.py:92-97), loc=0, scale=1."""
    if kind == "GaussianConditionalModel":
        return 0.5 * (1 + torch.erf(t / math.sqrt(2)))
    if kind == "LaplacianConditionalModel":
        return 0.5 - 0.5 * t.sign() * torch.expm1(-t.abs())
    raise KeyError(kind)


def conditional_model(y, scale, training, kind="LaplacianConditionalModel", noise=None, mean=0, bin_=1.0):
    """SymmetricConditionalModel.forward/_prob_mass (entropy_model.py:280-295, 341-352)."""
    yq = quantize(y, training, noise, bin_)
    half = bin_ / 2
    a = torch.abs(yq - mean)
    upper = _std_cdf((half - a) / scale, kind)
    lower = _std_cdf((-half - a) / scale, kind)
    return yq, upper - lower


# --------------------------------------------------------------------------------------
# distortion losses and metrics
# --------------------------------------------------------------------------------------
def gauss_filter_2d(filter_size=11, filter_sigma=1.5):
    """modelling/loss.py:113-121 == utils/metric.py:89-97 (softmax-normalised 2-D Gaussian)."""
    r = torch.arange(filter_size) + 0.5 - filter_size / 2
    x, y = torch.meshgrid(r, r, indexing="ij")
    g = -(x * x + y * y) / (2.0 * filter_sigma * filter_sigma)
    return F.softmax(g.view(-1), dim=0).view(1, 1, filter_size, filter_size)


def _filt(img, filt):
    """loss.py:103-111: valid per-channel conv as an (N*C,1,H,W) batch."""
    N, C, H, W = img.size()
    out = F.conv2d(img.reshape(-1, 1, H, W), filt)
    return out.view(N, C, out.size(2), out.size(3))


def ssim_core(img1, img2, filt, c1, c2):
    """loss.py:75-91 == metric.py:56-72: per-image means of the ssim and cs maps."""
    mu1, mu2 = _filt(img1, filt), _filt(img2, filt)
    mu1_sq, mu2_sq, mu1_mu2 = mu1 * mu1, mu2 * mu2, mu1 * mu2
    sigma1_sq = _filt(img1 * img1, filt) - mu1_sq
    sigma2_sq = _filt(img2 * img2, filt) - mu2_sq
    sigma12 = _filt(img1 * img2, filt) - mu1_mu2
    cs = (2.0 * sigma12 + c2) / (sigma1_sq + sigma2_sq + c2)
    ssim = cs * (2.0 * mu1_mu2 + c1) / (mu1_sq + mu2_sq + c1)
    return ssim.mean((1, 2, 3)), cs.mean((1, 2, 3))


def downsample2(img):
    """loss.py:176-188 == metric.py:126-138: reflect-pad odd edges, then 2x2 average pool."""
    _, _, h, w = img.size()
    if (w % 2) + (h % 2):
        img = F.pad(img, (0, w % 2, 0, h % 2), mode="reflect")
    return F.avg_pool2d(img, kernel_size=2)


def _loss_ssim(img1, img2, s):
    """SSIMLoss._ssim (loss.py:75-101): means, LowerBound(eps if log_scale else 0), optional log."""
    filt = gauss_filter_2d(s.FILTER_SIZE, s.FILTER_SIGMA)
    c1, c2 = (s.K1 * s.MAX_VAL) ** 2, (s.K2 * s.MAX_VAL) ** 2
    ssim, cs = ssim_core(img1, img2, filt, c1, c2)
    eps = s.EPS if s.LOG_SCALE else 0.0
    ssim, cs = lower_bound(ssim, eps), lower_bound(cs, eps)
    if s.LOG_SCALE:
        return ssim.log(), cs.log()
    return ssim, cs


def ssim_loss(img1, img2, s):
    """SSIMLoss.forward (loss.py:65-73)."""
    ssim, _ = _loss_ssim(img1 * s.MAX_VAL, img2 * s.MAX_VAL, s)
    if s.LOG_SCALE:
        return -ssim
    return 1.0 - ssim.mean()


def ms_ssim_loss(img1, img2, s, weights=MS_SSIM_WEIGHTS):
    """MS_SSIMLoss.forward/_log_forward/_forward (loss.py:131-174)."""
    img1, img2 = img1 * s.MAX_VAL, img2 * s.MAX_VAL
    n_levels = len(weights)
    if s.LOG_SCALE:
        res = 0.0
        for i, w in enumerate(weights, 1):
            ssim, cs = _loss_ssim(img1, img2, s)
            if i < n_levels:
                res = res + cs * w
                img1, img2 = downsample2(img1), downsample2(img2)
            else:
                res = res + ssim * w
        return -res.mean()
    res = torch.ones((img1.size(0),), dtype=img1.dtype)
    for i, w in enumerate(weights, 1):
        ssim, cs = _loss_ssim(img1, img2, s)
        if i < n_levels:
            res = res * cs ** w
            img1, img2 = downsample2(img1), downsample2(img2)
        else:
            res = res * ssim ** w
    return 1.0 - res.mean()


def distortion(name, x, x_tilde, cfg):
    """modelling/loss.py:22-45 get_loss_dict entries."""
    if name == "MSE":
        return F.mse_loss(x, x_tilde, reduction=cfg.MODEL.LOSS.REDUCTION)  # nn.MSELoss(input=x, target=x~)
    if name == "SSIMLoss":
        return ssim_loss(x, x_tilde, cfg.MODEL.LOSS.SSIM)
    if name == "MS_SSIMLoss":
        return ms_ssim_loss(x, x_tilde, cfg.MODEL.LOSS.SSIM)
    raise KeyError(name)


def psnr(img1, img2, max_val=255.0):
    """utils/metric.py:26-36."""
    mse = F.mse_loss(img1, img2, reduction="none").mean((1, 2, 3))
    return 10.0 * (2 * math.log10(max_val) - mse.log() / LOG10)


def ssim_metric(img1, img2, max_val=255.0, filter_size=11, filter_sigma=1.5, k1=0.01, k2=0.03,
                in_dB=False, non_negative=False):
    """utils/metric.py:39-77 (SSIM.__call__/_ssim)."""
    filt = gauss_filter_2d(filter_size, filter_sigma)
    ssim, _ = ssim_core(img1, img2, filt, (k1 * max_val) ** 2, (k2 * max_val) ** 2)
    if non_negative:
        ssim = ssim.clamp_min(0.0)
    if in_dB:
        ssim = -10.0 * (1.0 - ssim).log() / LOG10
    return ssim


def ms_ssim_metric(img1, img2, max_val=255.0, filter_size=11, filter_sigma=1.5, k1=0.01, k2=0.03,
                   in_dB=True, weights=MS_SSIM_WEIGHTS):
    """utils/metric.py:100-124 (MS_SSIM.__call__; non_negative clamp at :74-76)."""
    filt = gauss_filter_2d(filter_size, filter_sigma)
    c1, c2 = (k1 * max_val) ** 2, (k2 * max_val) ** 2
    res = torch.ones((img1.size(0),), dtype=img1.dtype)
    n_levels = len(weights)
    for i, w in enumerate(weights, 1):
        ssim, cs = ssim_core(img1, img2, filt, c1, c2)
        ssim, cs = ssim.clamp_min(0.0), cs.clamp_min(0.0)
        if i < n_levels:
            res = res * cs ** w
            img1, img2 = downsample2(img1), downsample2(img2)
        else:
            res = res * ssim ** w
    if in_dB:
        res = -10.0 * (1.0 - res).log() / LOG10
    return res


# --------------------------------------------------------------------------------------
# the whole model
# --------------------------------------------------------------------------------------
def compressor2018_forward(p, x, cfg, training=True, noise_z=None, noise_y=None, return_intermediates=False):
    """Compressor2018.forward (modelling/meta_arch/bmshl2018.py:68-98).

    noise_z / noise_y: optional U[0,1) tensors (the reference draws rand_like for z first
    (bmshl2018.py:73) and then for y (:76); with None the same draws are made here).
    Returns (x_tilde.detach(), losses) like the reference; with return_intermediates also a
    dict of y, z, z_tilde, sigma, y_tilde, x_tilde (attached) for per-tensor parity checks.
    """
    N, C, H, W = x.size()
    num_pixels = N * H * W
    kind = cfg.MODEL.ENTROPY_MODEL.CONDITIONAL_MODEL
    bin_ = cfg.MODEL.ENTROPY_MODEL.BIN
    y = analysis_transform(p, x, cfg)
    z = prior_analysis(p, torch.abs(y), cfg)
    z_tilde, z_probs, z_ce = entropy_model(p, z, training, noise_z, bin_)
    sigma = prior_synthesis(p, z_tilde, cfg)
    y_tilde, y_probs = conditional_model(y, sigma, training, kind, noise_y, 0, bin_)
    y_ce = ce_bits(y_probs)
    x_tilde = synthesis_transform(p, y_tilde, cfg)
    x_tilde = upper_bound(x_tilde, 1.0)
    x_tilde = lower_bound(x_tilde, 0.0)
    dl = {name: distortion(name, x, x_tilde, cfg) for name in cfg.MODEL.LOSS.DISTORTION_LOSS_NAMES}
    total_d = sum(dl.values())
    entropy_loss = (z_ce + y_ce) / num_pixels
    total = cfg.MODEL.LOSS.DISTORTION_LOSS_WEIGHT * total_d + entropy_loss
    losses = {
        "z_entropy": z_ce.detach() / num_pixels,
        "y_entropy": y_ce.detach() / num_pixels,
        "bpp": entropy_loss.detach(),
        "total_loss": total,
    }
    losses.update({k: v.detach() for k, v in dl.items()})
    if return_intermediates:
        inter = dict(y=y, z=z, z_tilde=z_tilde, z_probs=z_probs, sigma=sigma, y_tilde=y_tilde,
                     y_probs=y_probs, x_tilde=x_tilde)
        return x_tilde.detach(), losses, inter
    return x_tilde.detach(), losses


def leaf_params(p):
    """Clone params as autograd leaves (for gradient parity)."""
    return {k: v.detach().clone().requires_grad_(True) for k, v in p.items()}
