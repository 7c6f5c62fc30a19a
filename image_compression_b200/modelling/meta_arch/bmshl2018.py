"""Compressor2018: the scale-hyperprior codec (Balle et al. 2018) training/evaluation simulator with
the reference's module surface (modelling/meta_arch/bmshl2018.py:49-110): same sub-module names (so
state_dict keys match the reference's 50 keys), `forward(x) -> (x_tilde.detach(), losses)`,
`loss_names`.

Two execution paths, both on the hand-written kernels only:
  * fused (default): the whole forward is ONE autograd node with a hand-scheduled backward
    (autograd_ops.Compressor2018Fn);
  * composed: the reference's forward line by line over the block modules -- used for loss
    configurations the fused node does not cover, and by the parity tests as a cross-check.
"""
import torch
import torch.nn as nn

from ... import _lib
from ...utils import lazy

from ...autograd_ops import Compressor2018Fn, ModelMeta, build_pack_plan
from ..blocks import (ENTROPY_MODEL_REGISTRY, AnalysisTransform, HyperpriorAnalysisTransform,
                      HyperpriorSynthesisTransform, SynthesisTransform)
from ..layers import LowerBound, UpperBound
from ..loss import get_loss_dict
from .build import META_ARCH_REGISTRY


@META_ARCH_REGISTRY.register()
class Compressor2018(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        m = cfg.MODEL
        # construction order fixes the init RNG stream (bmshl2018.py:56-63)
        self.analysis_transform = AnalysisTransform(cfg)
        self.prior_analysis = HyperpriorAnalysisTransform(cfg)
        self.prior_synthesis = HyperpriorSynthesisTransform(cfg)
        self.synthesis_transform = SynthesisTransform(cfg)
        self.entropy_model = ENTROPY_MODEL_REGISTRY.get("EntropyModel")(m.LATENT_CHANNELS, cfg)
        self.conditional_model = ENTROPY_MODEL_REGISTRY.get(m.ENTROPY_MODEL.CONDITIONAL_MODEL)(cfg)
        self.distortion_loss_fns = get_loss_dict(cfg, m.LOSS.DISTORTION_LOSS_NAMES)
        self.distortion_loss_weight = m.LOSS.DISTORTION_LOSS_WEIGHT
        self.loss_names = ["y_entropy", "z_entropy", "bpp"] + list(self.distortion_loss_fns.keys())
        self._meta = ModelMeta(m.CONV_KERNEL, m.STRIDES, m.HYPER_PRIOR.KERNELS, m.HYPER_PRIOR.STRIDES,
                               self.conditional_model.DIST, m.ENTROPY_MODEL.BIN, self.distortion_loss_weight,
                               list(self.distortion_loss_fns.keys()), getattr(m.LOSS, "REDUCTION", "mean"),
                               m.LOSS.SSIM)
        self.fused = True
        self.cuda_graphs = False      # opt-in: replay evaluation forwards (no_grad, eval mode) as one CUDA graph
        self._graphs = {}
        self._plans = {}              # compute mode -> hotpath.PackPlan (packed bf16 operand copies of the weights)
        # trainer-transparent gradient handling (sharding.GradientReducer): in training mode the backward writes every
        # gradient into one flat buffer and, under torchrun (an initialised process group with > 1 rank), all-reduces it
        # from inside the backward, so engine/trainer.py:176-194 runs unchanged.  False: plain autograd gradients.
        self.flat_gradients = True
        self._reducer = None
        # lazily synchronised outputs for the unchanged trainer / monitor (utils/lazy.py); defer_nan_check: answer the
        # trainer's `torch.isnan(loss).any()` at once and run the check at the step's first host read instead
        self.lazy_sync = True
        self.defer_nan_check = True

    def __getstate__(self):
        # copy.deepcopy / pickling: per-process device state (flat gradient buffer, packed operands, CUDA graphs) stays behind
        d = self.__dict__.copy()
        d["_reducer"], d["_plans"], d["_graphs"] = None, {}, {}
        return d

    # ------------------------------------------------------------------------------
    def _flat_params(self):
        return (self.analysis_transform.flat_params() + self.prior_analysis.flat_params() +
                self.prior_synthesis.flat_params() + self.synthesis_transform.flat_params() +
                self.entropy_model._cdf_estimator.flat_params())

    def gradient_reducer(self):
        """The model's sharding.GradientReducer (built on first use, rebuilt when the parameters moved).  Slot order of
        its flat buffer = the order the backward finishes gradients in (g_s last layer first ... g_a first layer last),
        so finished ranges are all-reduced while the rest of the backward runs."""
        from ...sharding import GradientReducer
        params = [p for p in self.backward_order_parameters() if p.requires_grad]
        r = self._reducer
        if r is None or r.fg.flat.device != params[0].device or len(r.fg.params) != len(params) or any(
                a is not b for a, b in zip(r.fg.params, params)):
            r = self._reducer = GradientReducer(params, module=self)
        return r

    def backward_order_parameters(self):
        out = []
        for block in (self.synthesis_transform, self.prior_synthesis):
            out += list(block.parameters())[::-1]
        out += list(self.entropy_model.parameters())
        for block in (self.prior_analysis, self.analysis_transform):
            out += list(block.parameters())[::-1]
        seen = {id(p) for p in out}
        return out + [p for p in self.parameters() if id(p) not in seen]

    def forward(self, x, noise_z=None, noise_y=None):
        """noise_z / noise_y: optional U[0,1) tensors shaped like z / y (NCHW) replacing the in-kernel
        Philox noise of training mode (parity tests inject the reference's draws here)."""
        if getattr(self, "_is_replica", False):
            # nn.DataParallel with several visible GPUs (engine/trainer.py:256-259) runs replicas of this module on
            # threads of ONE process.  The operand cache, the gradient sink and the CUDA-graph state of the hot path are
            # per process, not per replica thread: refuse loudly instead of racing.
            raise RuntimeError("image_compression_b200: this model ran as an nn.DataParallel replica (several GPUs visible to "
                               "one process). Use one process per GPU: CUDA_VISIBLE_DEVICES=$LOCAL_RANK under torchrun, where "
                               "the reference's DataParallel wrap is a one-device pass-through (INTEGRATION.md section 5).")
        reducer = None
        if self.flat_gradients and self.training and torch.is_grad_enabled() and x.is_cuda:
            reducer = self.gradient_reducer()
            if not reducer.before_forward():
                reducer = None
        if self.fused and self._meta.fusable():
            if self.cuda_graphs and not self.training and not torch.is_grad_enabled() and noise_z is None:
                return self._forward_graphed(x)
            out = self._forward_fused(x, noise_z, noise_y)
        else:
            out = self._forward_composed(x, noise_z, noise_y)
        if reducer is not None:
            reducer.watch(out[1]["total_loss"])
        return out

    def _forward_graphed(self, x):
        """Static-shape evaluation forward replayed as a CUDA graph (outputs are static tensors: consume them
        before the next call).  One graph per (shape, compute mode)."""
        from ...hotpath import get_compute_mode
        from ...utils.graphs import GraphedCallable
        key = (tuple(x.shape), x.dtype, x.device, get_compute_mode())
        g = self._graphs.get(key)
        if g is None:
            if len(self._graphs) >= 8:
                self._graphs.clear()
            g = self._graphs[key] = GraphedCallable(lambda t: self._forward_fused_raw(t, None, None), x.contiguous())
        return self._wrap_outputs(*g(x.contiguous()))

    def pack_plan(self):
        """The model's packed-operand plan for the current compute mode and parameter placement (one
        hpc_pack_multi launch per weight update instead of one pack launch per layer and use)."""
        from ...hotpath import get_compute_mode
        params = self._flat_params()
        key = (get_compute_mode(), params[0].device)
        plan = self._plans.get(key)
        if plan is None or plan.owner_ids != tuple(id(p) for p in params):
            plan = self._plans[key] = build_pack_plan(self._meta, params)
            plan.owner_ids = tuple(id(p) for p in params)
        return plan

    def _forward_fused(self, x, noise_z, noise_y):
        return self._wrap_outputs(*self._forward_fused_raw(x, noise_z, noise_y))

    def _forward_fused_raw(self, x, noise_z, noise_y):
        self._meta.lmbda = float(self.distortion_loss_weight)
        _lib.require_cuda(x, self.analysis_transform.layers[0].weight)      # no CPU fallback: fail before any launch
        self.pack_plan().ensure()
        return Compressor2018Fn.apply(x, self._meta, (bool(self.training), torch.is_grad_enabled()), noise_z, noise_y,
                                      *self._flat_params())

    def _wrap_outputs(self, x_tilde, out):
        """(x_tilde, (8,) loss vector of hpc_rd_combine) -> the reference's return value (bmshl2018.py:91-98)."""
        det = out.detach()
        losses = {"z_entropy": det[0], "y_entropy": det[1], "bpp": det[2], "total_loss": out[3]}
        for i, name in enumerate(self.distortion_loss_fns.keys()):
            losses[name] = det[4 + i]
        x_tilde = x_tilde.detach()
        if self.lazy_sync and lazy.enabled and not torch.cuda.is_current_stream_capturing():
            # one asynchronous device-to-host copy of the 8 loss values right behind the forward; the
            # reference's float(value) / isnan / preds.max() reads are then served without further stream stalls
            mirror = lazy.HostMirror(det)
            for i, k in enumerate(("z_entropy", "y_entropy", "bpp")):
                losses[k] = lazy.LazyScalar.wrap(losses[k], mirror, i)
            for i, name in enumerate(self.distortion_loss_fns.keys()):
                losses[name] = lazy.LazyScalar.wrap(losses[name], mirror, 4 + i)
            lazy.LossTensor.defer_nan_check = bool(self.defer_nan_check)
            losses["total_loss"] = lazy.LossTensor.wrap(losses["total_loss"], mirror, 3)
            x_tilde = lazy.ReconImage.wrap(x_tilde)
        return x_tilde, losses

    def _forward_composed(self, x, noise_z=None, noise_y=None):
        N, C, H, W = x.size()
        num_pixels = N * H * W
        y = self.analysis_transform(x)
        z = self.prior_analysis(_abs(y))
        z_tilde, z_probs, z_ce_loss = self.entropy_model(z, noise_z)
        sigma = self.prior_synthesis(z_tilde)
        y_tilde, y_probs = self.conditional_model(y, sigma, noise=noise_y)
        y_ce_loss = self.conditional_model._ce_loss(y_probs)
        x_tilde = self.synthesis_transform(y_tilde)
        x_tilde = UpperBound.apply(x_tilde, 1.)
        x_tilde = LowerBound.apply(x_tilde, 0.)
        distortion_losses = self.distortion_loss(x, x_tilde)
        total_distortion = sum(distortion_losses.values())
        entropy_loss = (z_ce_loss + y_ce_loss) / num_pixels
        total_loss = self.distortion_loss_weight * total_distortion + entropy_loss
        losses = {"z_entropy": z_ce_loss.detach() / num_pixels, "y_entropy": y_ce_loss.detach() / num_pixels,
                  "bpp": entropy_loss.detach(), "total_loss": total_loss}
        losses.update({k: v.detach() for k, v in distortion_losses.items()})
        return x_tilde.detach(), losses

    def distortion_loss(self, img1, img2):
        return {name: fn(img1, img2) for name, fn in self.distortion_loss_fns.items()}

    def compress(self, x):
        """x (N,3,H,W) in [0,1] -> list of N byte strings: the rANS bitstreams of z~ and y~ (the reference's stub,
        bmshl2018.py:106-107; format and coder in image_compression_b200/codec.py)."""
        from ... import codec
        return codec.compress(self, x)

    def decompress(self, x):
        """list of byte strings from compress() -> reconstruction (N,3,H,W) in [0,1] (bmshl2018.py:109-110)."""
        from ... import codec
        return codec.decompress(self, x)


def _abs(y):
    from ...autograd_ops import AbsFn
    return AbsFn.apply(y)
