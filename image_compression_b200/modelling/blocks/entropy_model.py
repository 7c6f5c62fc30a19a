"""Factorised prior and conditional (scale-only) likelihood models with the reference's class
names, registry entries, parameter names and return conventions
(modelling/blocks/entropy_model.py:47-378); arithmetic in kernels K4/K5."""
import numpy as np
import torch
import torch.nn as nn

from ... import _lib
from ...autograd_ops import BitsFn, CdfLayerFn, CdfLogitsFn, ConditionalFn, EntropyModelFn
from .build import ENTROPY_MODEL_REGISTRY


class CDFLayer(nn.Module):
    """Parameter container of one MLP layer: weight (1,C,out,in) = log(expm1(1/scale/out)),
    bias ~ U(-.5,.5), factor = 0 (entropy_model.py:51-65)."""

    def __init__(self, channels, in_dim, out_dim, init_scale, act=True):
        super().__init__()
        self.act = act
        w0 = float(np.log(np.expm1(1 / init_scale / out_dim)))
        self.weight = nn.Parameter(nn.init.constant_(torch.empty(1, channels, out_dim, in_dim), w0))
        self.bias = nn.Parameter(nn.init.uniform_(torch.empty(1, channels, out_dim, 1), -0.5, 0.5))
        if act:
            self.factor = nn.Parameter(nn.init.zeros_(torch.empty(1, channels, out_dim, 1)))

    def forward(self, x):
        """x (N,C,D,1) -> (N,C,D',1)  (entropy_model.py:67-78).  Stand-alone layer kernel for module parity; the
        hot path (CDFEstimator / EntropyModel) evaluates all four layers inside one fused per-channel MLP kernel."""
        return CdfLayerFn.apply(x, self.weight, self.bias, self.factor if self.act else None)


class CDFEstimator(nn.Module):
    def __init__(self, channels, dims=[3, 3, 3], init_scale=10.):
        super().__init__()
        dims = [1] + list(dims) + [1]        # the reference mutates its argument here; we copy
        n = len(dims) - 1
        scale = init_scale ** (1 / n)
        self.layers = nn.Sequential(*[CDFLayer(channels, dims[i], dims[i + 1], scale, act=i < n - 1)
                                      for i in range(n)])

    def flat_params(self):
        """order of autograd_ops.CDF_KEYS: w0,b0,f0, w1,b1,f1, w2,b2,f2, w3,b3"""
        if len(self.layers) != 4:
            raise NotImplementedError("kernel K4 is specialised for ENTROPY_MODEL.DIMS == [3,3,3]")
        out = []
        for l in self.layers:
            out += [l.weight, l.bias] + ([l.factor] if l.act else [])
        return out

    def forward(self, x):
        return CdfLogitsFn.apply(x, *self.flat_params())


class BaseEntropyModel(nn.Module):
    def _ce_loss(self, probs):
        """sum clamp(-log2(p+1e-10),0,50)  (entropy_model.py:171-185)"""
        return BitsFn.apply(probs)

    def _dequantize(self, x):
        """symbols -> values: the identity for bin = 1 and mean = 0, the only configuration the reference runs
        (entropy_model.py:165-169 is a stub there)."""
        return x


@ENTROPY_MODEL_REGISTRY.register()
class EntropyModel(BaseEntropyModel):
    """forward(x) -> (x_tilde, probs, bits); probs comes back in the reference's (N,W,C,H) layout for
    4-D input.  `noise` (U[0,1), shape of x) may be injected for parity tests;
    otherwise the kernel draws Philox noise seeded from torch's generator."""

    def __init__(self, in_channels, cfg):
        super().__init__()
        em = cfg.MODEL.ENTROPY_MODEL
        self._cdf_estimator = CDFEstimator(in_channels, em.DIMS, em.INIT_SCALE)
        self.bin = em.BIN

    def forward(self, x, noise=None):
        return EntropyModelFn.apply(x, float(self.bin), bool(self.training), noise,
                                    *self._cdf_estimator.flat_params())

    def _logit_cumulative(self, x):
        return self._cdf_estimator(x)

    def compress(self, x):
        """x (N,C,h,w) -> list of N byte strings coding round(x) with the per-channel tables of this prior (the
        reference's stub, entropy_model.py:157-160; rANS coder: image_compression_b200/codec.py)."""
        from ... import codec, hotpath as hp
        with torch.no_grad():
            sym = hp.to_nhwc(x.detach(), torch.float32)
            lo, hi = codec._value_range(sym)
            return codec.encode_latent(sym, None, codec.factorized_tables(self, lo, hi), lo, hi)

    def decompress(self, strings, shape):
        """strings from compress(), shape = (h, w) -> the rounded symbols (N,C,h,w) fp32 (entropy_model.py:162-163)."""
        from ... import codec, hotpath as hp
        C = self._cdf_estimator.layers[0].weight.shape[1]
        dev = self._cdf_estimator.layers[0].weight.device
        with torch.no_grad():
            sym, _ = codec.decode_latent(strings, None, lambda a, b: codec.factorized_tables(self, a, b),
                                         shape[0] * shape[1], C, dev)
        return hp.as_nchw(sym.view(len(strings), shape[0], shape[1], C))


class SymmetricConditionalModel(BaseEntropyModel):
    DIST = None

    def __init__(self, cfg):
        super().__init__()
        self.bin = cfg.MODEL.ENTROPY_MODEL.BIN

    def forward(self, x, scale, mean=0, noise=None):
        if isinstance(mean, torch.Tensor):
            raise NotImplementedError("tensor-valued mean: the reference only ever passes mean=0 "
                                      "(bmshl2018.py:76)")
        return ConditionalFn.apply(x, scale, float(mean), float(self.bin), bool(self.training), self.DIST, noise)


    def compress(self, x, scale):
        """x, scale (N,C,h,w) -> list of N byte strings coding round(x) under this model at the quantised scales
        (scale table of codec.py; the reference's stub: entropy_model.py:157-160, "symbol" mode :235-236)."""
        from ... import codec, hotpath as hp
        with torch.no_grad():
            xs, ss = hp.to_nhwc(x.detach(), torch.float32), hp.to_nhwc(scale.detach(), torch.float32)
            sym, _, _, _ = hp.conditional_forward(xs, ss, self.DIST, False, None, float(self.bin), 0.0, want_probs=False)
            lo, hi = codec._value_range(sym)
            cum = codec.quantise_pmf(codec.conditional_pmf(lo, hi, self.DIST == _lib.DIST_GAUSS))
            return codec.encode_latent(sym, codec.scale_levels(ss), cum, lo, hi)

    def decompress(self, strings, scale):
        """strings from compress(), the same scale tensor -> the rounded symbols (N,C,h,w) fp32."""
        from ... import codec, hotpath as hp
        with torch.no_grad():
            ss = hp.to_nhwc(scale.detach(), torch.float32)
            N, h, w, C = ss.shape
            gauss = self.DIST == _lib.DIST_GAUSS
            sym, _ = codec.decode_latent(strings, codec.scale_levels(ss),
                                         lambda a, b: codec.quantise_pmf(codec.conditional_pmf(a, b, gauss)), h * w, C, ss.device)
        return hp.as_nchw(sym.view(N, h, w, C))


@ENTROPY_MODEL_REGISTRY.register()
class GaussianConditionalModel(SymmetricConditionalModel):
    DIST = _lib.DIST_GAUSS


@ENTROPY_MODEL_REGISTRY.register()
class LaplacianConditionalModel(SymmetricConditionalModel):
    DIST = _lib.DIST_LAPLACE
