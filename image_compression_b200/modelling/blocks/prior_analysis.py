"""h_a (modelling/blocks/prior_analysis.py:39-71): conv3x3 s1 -> ReLU -> conv5x5 s2 -> ReLU ->
conv5x5 s2 without bias.  ReLU runs in the conv epilogue; the nn.ReLU entries only keep the
reference's `_layers.{0,2,4}` numbering."""
import torch.nn as nn

from ...autograd_ops import HyperAnalysisFn
from .analysis import _init_conv


def _flat_h_params(seq):
    out = []
    for m in seq:
        if isinstance(m, nn.ReLU):
            continue
        out.append(m.weight)
        if m.bias is not None:
            out.append(m.bias)
    return out


class HyperpriorAnalysisTransform(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.strides = list(cfg.MODEL.HYPER_PRIOR.STRIDES)
        self.kernels = list(cfg.MODEL.HYPER_PRIOR.KERNELS)
        n = len(self.strides)
        cin, cmid = cfg.MODEL.LATENT_CHANNELS, cfg.MODEL.INTER_CHANNELS
        mods = []
        for i, (s, k) in enumerate(zip(self.strides, self.kernels)):
            mods.append(_init_conv(nn.Conv2d(cin if i == 0 else cmid, cmid, k, stride=s, padding=k // 2,
                                             bias=(i < n - 1))))
            if i < n - 1:
                mods.append(nn.ReLU())
        self._layers = nn.Sequential(*mods)
        self.has_bias = [i < n - 1 for i in range(n)]

    def flat_params(self):
        return _flat_h_params(self._layers)

    def forward(self, x):
        return HyperAnalysisFn.apply(x, tuple(self.kernels), tuple(self.strides), tuple(self.has_bias),
                                     *self.flat_params())
