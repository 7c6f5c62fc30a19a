"""g_a (modelling/blocks/analysis.py:40-71): 4x Conv2d(k5,s2,p2) with GDN after the first three.

`self.layers` holds torch's own Conv2d modules purely as parameter containers (same state_dict
keys, same init RNG stream as the reference); forward never calls them -- it runs the
hand-written conv/GDN kernels through AnalysisFn."""
import math

import torch.nn as nn

from ..layers import GDN
from ...autograd_ops import AnalysisFn


def _init_conv(conv):
    nn.init.xavier_normal_(conv.weight.data, math.sqrt(2))
    if conv.bias is not None:
        nn.init.constant_(conv.bias.data, 0.01)
    return conv


def _flat_conv_gdn_params(seq):
    out = []
    for m in seq:
        if isinstance(m, GDN):
            out += [m.gamma.param, m.beta.param]
        else:
            out += [m.weight, m.bias]
    return out


class AnalysisTransform(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.strides = list(cfg.MODEL.STRIDES)
        self.kernel = cfg.MODEL.CONV_KERNEL
        chans = [cfg.DATA.IN_CHANNELS] + [cfg.MODEL.INTER_CHANNELS] * (len(self.strides) - 1) + \
                [cfg.MODEL.LATENT_CHANNELS]
        mods = []
        for i, s in enumerate(self.strides):
            mods.append(_init_conv(nn.Conv2d(chans[i], chans[i + 1], self.kernel, stride=s,
                                             padding=self.kernel // 2)))
            if i < len(self.strides) - 1:
                mods.append(GDN(chans[i + 1]))
        self.layers = nn.Sequential(*mods)

    def flat_params(self):
        return _flat_conv_gdn_params(self.layers)

    def forward(self, x):
        return AnalysisFn.apply(x, self.kernel, tuple(self.strides), *self.flat_params())
