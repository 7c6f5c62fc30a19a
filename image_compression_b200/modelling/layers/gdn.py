"""GDN / IGDN / ReLU-GDN with the TF-compression style non-negative reparameterisation
(modelling/layers/gdn.py:42-88).  Parameter names (`gamma.param`, `beta.param`) and init match
the reference so its checkpoints load strictly."""
import torch
import torch.nn as nn

from ... import hotpath as hp
from ...autograd_ops import GDNFn, NonNegFn


class NonNegativeParam(nn.Module):
    """value = max(param, bound)^2 - pedestal, pedestal = offset^2, bound = sqrt(minimum + pedestal^2)
    (the reference squares the pedestal once more in the bound: gdn.py:53)."""

    def __init__(self, init_val, minimum=0, offset=2 ** (-18)):
        super().__init__()
        self.pedestal = float(offset) ** 2
        ped = torch.tensor(self.pedestal)
        self.bound = float((float(minimum) + ped.clone() ** 2) ** 0.5)
        self.param = nn.Parameter(torch.sqrt(torch.max(init_val + ped, ped)))

    def forward(self):
        return NonNegFn.apply(self.param, self.bound, self.pedestal)


class GDN(nn.Module):
    def __init__(self, in_channels, inverse=False, relu=False, gamma_init=0.1, beta_min=1e-6, offset=2 ** -18):
        super().__init__()
        if float(offset) != float(2 ** -18):
            raise NotImplementedError("the GDN kernels use the reference's default offset 2**-18")
        self.inverse, self.relu, self.beta_min = inverse, relu, beta_min
        self.gamma = NonNegativeParam(torch.eye(in_channels).view(in_channels, in_channels, 1, 1) * gamma_init)
        self.beta = NonNegativeParam(torch.ones((in_channels,)), minimum=beta_min)

    def forward(self, x):
        return GDNFn.apply(x, self.gamma.param, self.beta.param, self.inverse, self.relu, self.beta_min)
