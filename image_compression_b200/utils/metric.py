"""Evaluation metrics on the GPU kernels K6/K7, with the reference's call surface
(utils/metric.py:26-138): `psnr(img1,img2,max_val)`, `SSIM(...)(a,b)`, `MS_SSIM(...)(a,b)`.
Inputs are (N,C,H,W) CUDA tensors already scaled to [0,max_val] (engine/monitor.py:122-123)."""
import torch

from .. import _lib
from .. import hotpath as hp
from . import lazy

MS_SSIM_DEFAULT_WEIGHTS = [0.0448, 0.2856, 0.3001, 0.2363, 0.1333]  # utils/metric.py:103


def psnr(img1, img2, max_val=255.):
    """per-image 10*(2*log10(max_val) - ln(mse)/ln10)   (utils/metric.py:26-36)"""
    _lib.require_cuda(img1, img2)
    if lazy.enabled and isinstance(img1, lazy.ReconImage) and getattr(img1, "_metric_cache", None) is not None:
        return _monitor_metrics(img1, img2, max_val)[0]
    a, b = hp.contiguous_f32(img1), hp.contiguous_f32(img2)
    N = a.shape[0]
    _, sse = hp.clamp01_sse(a.reshape(N, -1), b.reshape(N, -1), 1.0, do_clamp=False, want_xt=False)
    return hp.psnr_from_sse(sse, a.numel() // N, max_val)


def _monitor_metrics(img1, img2, max_val):
    """The reference's Monitor asks for psnr(preds*255, targets*255) and then MS_SSIM()(preds*255, targets*255), reading
    each result back with .detach().cpu().numpy() (engine/monitor.py:121-125, utils/logging.py:151-153): two stream
    stalls, the second with the GPU idle but for the MS-SSIM kernels.  When `img1` is a scaled reconstruction of the
    model (utils/lazy.py), both metrics are computed at the first request, travel to the host in ONE asynchronous copy
    and are handed out as lazily synchronised vectors; the MS_SSIM call that follows finds its result cached."""
    key = (img2.data_ptr(), img2._version, tuple(img2.shape), float(max_val))
    hit = img1._metric_cache.get(key)
    if hit is None:
        with torch.no_grad():
            a, b = hp.contiguous_f32(lazy._plain(img1)), hp.contiguous_f32(img2)
            N = a.shape[0]
            both = torch.empty((2, N), dtype=torch.float32, device=a.device)
            _, sse = hp.clamp01_sse(a.reshape(N, -1), b.reshape(N, -1), 1.0, do_clamp=False, want_xt=False)
            ps = hp.psnr_from_sse(sse, a.numel() // N, max_val)
            c1, c2 = (0.01 * max_val) ** 2, (0.03 * max_val) ** 2
            hit = [None, None, (a, b, c1, c2, N, both, ps)]
            img1._metric_cache[key] = hit
            # MS-SSIM with the reference's defaults (utils/metric.py:100-110); computed now only if the images are large
            # enough for the 5-level pyramid (else MS_SSIM() would raise, and so does the call that asks for it)
            H, W = a.shape[2], a.shape[3]
            if min(H, W) >= 11 * 16:
                pyr = hp.SsimPyramid(a, b, 5, 11, 1.5, c1, c2, 1.0)
                ms = pyr.combine(_lib.MSSSIM_METRIC, 1, 0.0, MS_SSIM_DEFAULT_WEIGHTS)
                both[0].copy_(ps); both[1].copy_(ms)
                m = lazy.HostMirror(both)
                hit[0], hit[1] = lazy.LazyVector.wrap(ps, m, 0), lazy.LazyVector.wrap(ms, m, N)
            else:
                m = lazy.HostMirror(ps)
                hit[0] = lazy.LazyVector.wrap(ps, m, 0)
    return hit


class SSIM:
    """utils/metric.py:39-97.  `device` is accepted for signature parity; the filter lives on the
    device of the inputs (the reference's default builds it on "cuda" at import, monitor.py:50)."""

    def __init__(self, max_val=255., filter_size=11, filter_sigma=1.5, k1=0.01, k2=0.03, device="cuda",
                 in_dB=False, non_negative=False):
        self.filter_size, self.filter_sigma = filter_size, filter_sigma
        self.max_val, self.k = max_val, (k1, k2)
        self.c1, self.c2 = (k1 * max_val) ** 2, (k2 * max_val) ** 2
        self.non_negative, self.in_dB = non_negative, in_dB

    def _pyramid(self, img1, img2, levels):
        _lib.require_cuda(img1, img2)
        a, b = hp.contiguous_f32(img1), hp.contiguous_f32(img2)
        return hp.SsimPyramid(a, b, levels, self.filter_size, self.filter_sigma, self.c1, self.c2, 1.0)

    def __call__(self, img1, img2):
        with torch.no_grad():
            flags = (1 if self.in_dB else 0) | (2 if self.non_negative else 0)
            return self._pyramid(img1, img2, 1).combine(_lib.SSIM_METRIC, flags, 0.0, (1.0,))


class MS_SSIM(SSIM):
    """utils/metric.py:100-138 (clamp_min 0 on every level, product of powers, optional dB)."""

    def __init__(self, max_val=255., filter_size=11, filter_sigma=1.5, k1=0.01, k2=0.03, device="cuda",
                 in_dB=True, weights=MS_SSIM_DEFAULT_WEIGHTS):
        super().__init__(max_val, filter_size, filter_sigma, k1, k2, device, non_negative=True, in_dB=False)
        self.in_dB, self.weights = in_dB, list(weights)

    def __call__(self, img1, img2):
        if (lazy.enabled and isinstance(img1, lazy.ReconImage) and getattr(img1, "_metric_cache", None) is not None
                and self._is_default()):
            hit = _monitor_metrics(img1, img2, self.max_val)
            if hit[1] is not None:
                return hit[1]
        with torch.no_grad():
            pyr = self._pyramid(img1, img2, len(self.weights))
            return pyr.combine(_lib.MSSSIM_METRIC, 1 if self.in_dB else 0, 0.0, self.weights)

    def _is_default(self):
        return (self.in_dB and list(self.weights) == MS_SSIM_DEFAULT_WEIGHTS and self.filter_size == 11
                and self.filter_sigma == 1.5 and self.k == (0.01, 0.03))


def image_metrics(preds, targets, max_val=255., weights=MS_SSIM_DEFAULT_WEIGHTS):
    """What engine/monitor.py:113-125 computes per batch -- psnr(preds*255, targets*255) and
    MS_SSIM()(preds*255, targets*255) in dB -- with the x255 folded into the kernels (no scaled copies).
    preds/targets: (N,C,H,W) CUDA tensors in [0,1].  Returns (psnr (N,), ms_ssim_dB (N,))."""
    _lib.require_cuda(preds, targets)
    with torch.no_grad():
        a, b = hp.contiguous_f32(preds), hp.contiguous_f32(targets)
        N = a.shape[0]
        _, sse = hp.clamp01_sse(a.reshape(N, -1), b.reshape(N, -1), float(max_val), do_clamp=False, want_xt=False)
        ps = hp.psnr_from_sse(sse, a.numel() // N, max_val)
        c1, c2 = (0.01 * max_val) ** 2, (0.03 * max_val) ** 2
        pyr = hp.SsimPyramid(a, b, len(weights), 11, 1.5, c1, c2, float(max_val))
        ms = pyr.combine(_lib.MSSSIM_METRIC, 1, 0.0, list(weights))
    return ps, ms
