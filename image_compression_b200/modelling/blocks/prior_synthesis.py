"""h_s (modelling/blocks/prior_synthesis.py:40-72): ConvTranspose2d 5x5 s2 -> ReLU -> 5x5 s2 -> ReLU
-> 3x3 s1, then sigma = clamp(exp(.), 1e-10, 1e10) in the last epilogue.  Kernel and stride lists
are the reversed h_a lists (prior_synthesis.py:51)."""
import torch.nn as nn

from ...autograd_ops import HyperSynthesisFn
from .analysis import _init_conv
from .prior_analysis import _flat_h_params


class HyperpriorSynthesisTransform(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.strides = list(reversed(cfg.MODEL.HYPER_PRIOR.STRIDES))
        self.kernels = list(reversed(cfg.MODEL.HYPER_PRIOR.KERNELS))
        n = len(self.strides)
        cmid, cout = cfg.MODEL.INTER_CHANNELS, cfg.MODEL.LATENT_CHANNELS
        mods = []
        for i, (s, k) in enumerate(zip(self.strides, self.kernels)):
            mods.append(_init_conv(nn.ConvTranspose2d(cmid, cmid if i < n - 1 else cout, k, stride=s,
                                                      padding=k // 2, bias=True, output_padding=s - 1)))
            if i < n - 1:
                mods.append(nn.ReLU())
        self._layers = nn.Sequential(*mods)

    def flat_params(self):
        return _flat_h_params(self._layers)

    def forward(self, x):
        return HyperSynthesisFn.apply(x, tuple(self.kernels), tuple(self.strides), *self.flat_params())
