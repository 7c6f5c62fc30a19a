"""g_s (modelling/blocks/synthesis.py:40-71): 4x ConvTranspose2d(k5,s2,p2,op1) with GDN -- forward
GDN, not IGDN (synthesis.py:64-66) -- after the first three.  Returns the raw reconstruction; the
[0,1] bounds are applied by the meta-architecture (bmshl2018.py:80-81)."""
import torch.nn as nn

from ..layers import GDN
from ...autograd_ops import SynthesisFn
from .analysis import _flat_conv_gdn_params, _init_conv


class SynthesisTransform(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.strides = list(cfg.MODEL.STRIDES)
        self.kernel = cfg.MODEL.CONV_KERNEL
        chans = [cfg.MODEL.LATENT_CHANNELS] + [cfg.MODEL.INTER_CHANNELS] * (len(self.strides) - 1) + \
                [cfg.DATA.IN_CHANNELS]
        mods = []
        for i, s in enumerate(self.strides):
            mods.append(_init_conv(nn.ConvTranspose2d(chans[i], chans[i + 1], self.kernel, stride=s,
                                                      padding=self.kernel // 2, output_padding=s - 1)))
            if i < len(self.strides) - 1:
                mods.append(GDN(chans[i + 1]))
        self.layers = nn.Sequential(*mods)

    def flat_params(self):
        return _flat_conv_gdn_params(self.layers)

    def forward(self, x):
        return SynthesisFn.apply(x, self.kernel, tuple(self.strides), *self.flat_params())
