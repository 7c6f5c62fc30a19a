"""Distortion losses with the reference's constructor signatures (modelling/loss.py:22-188):
`get_loss_dict(cfg, names)`, `SSIMLoss(...)`, `MS_SSIMLoss(...)`; inputs are [0,1] images, scaled
by max_val inside the kernels (loss.py:67-68).  Kernels: K6 (MSE) and K7 ((MS-)SSIM)."""
import torch.nn as nn

from .. import _lib
from ..autograd_ops import MSEFn, SsimLossFn

MS_SSIM_DEFAULT_WEIGHTS = [0.0448, 0.2856, 0.3001, 0.2363, 0.1333]  # loss.py:127


class MSELoss(nn.Module):
    """nn.MSELoss(reduction=...) replacement: forward(input, target)."""

    def __init__(self, reduction="mean"):
        super().__init__()
        self.reduction = reduction

    def forward(self, inp, target):
        return MSEFn.apply(inp, target, self.reduction)


class SSIMLoss(nn.Module):
    """1 - mean_n LowerBound(ssim_n, 0)   or, log_scale,  -log LowerBound(ssim_n, eps) per image
    (loss.py:65-73, 93-101)."""

    def __init__(self, max_val=255., filter_size=11, filter_sigma=1.5, k1=0.01, k2=0.03, log_scale=False,
                 eps=1e-5):
        super().__init__()
        self.max_val, self.filter_size, self.filter_sigma = max_val, filter_size, filter_sigma
        self.c1, self.c2 = (k1 * max_val) ** 2, (k2 * max_val) ** 2
        self.log_scale, self.eps = log_scale, eps

    def _apply(self, fn):  # no tensors to move; keep nn.Module.to()/cuda() cheap
        return super()._apply(fn)

    def _run(self, img1, img2, kind, flags, weights):
        eps = self.eps if self.log_scale else 0.0
        return SsimLossFn.apply(img1, img2, kind, flags, eps, tuple(weights), self.filter_size, self.filter_sigma,
                                self.c1, self.c2, self.max_val)

    def forward(self, img1, img2):
        return self._run(img1, img2, _lib.SSIM_LOSS, 1 if self.log_scale else 0, (1.0,))


class MS_SSIMLoss(SSIMLoss):
    """log_scale: -mean_n sum_i w_i log LowerBound(.,eps) (loss.py:141-156); else 1 - mean_n prod_i (.)^w_i
    (loss.py:158-174).  Levels 1..L-1 use cs, level L uses ssim; 2x2 average-pool pyramid (loss.py:176-188)."""

    def __init__(self, max_val=255., filter_size=11, filter_sigma=1.5, k1=0.01, k2=0.03, log_scale=False,
                 eps=1e-5, weights=MS_SSIM_DEFAULT_WEIGHTS):
        super().__init__(max_val, filter_size, filter_sigma, k1, k2, log_scale, eps)
        self.weights = list(weights)

    def forward(self, img1, img2):
        kind = _lib.MSSSIM_LOSS_LOG if self.log_scale else _lib.MSSSIM_LOSS_PLAIN
        return self._run(img1, img2, kind, 0, self.weights)


def get_loss_dict(cfg, names):
    s = cfg.MODEL.LOSS.SSIM
    kw = dict(max_val=s.MAX_VAL, filter_size=s.FILTER_SIZE, filter_sigma=s.FILTER_SIGMA, k1=s.K1, k2=s.K2,
              log_scale=s.LOG_SCALE, eps=s.EPS)
    table = {"MSE": lambda: MSELoss(reduction=cfg.MODEL.LOSS.REDUCTION),
             "SSIMLoss": lambda: SSIMLoss(**kw),
             "MS_SSIMLoss": lambda: MS_SSIMLoss(**kw)}
    return {name: table[name]() for name in names}
