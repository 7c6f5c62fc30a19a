"""Replay of the reference's engine call sequences around the model -- what `engine/trainer.py:163-209`
(`Trainer.train_step`), `engine/evaluator.py:87-105` (`Evaluator.step`) and `engine/monitor.py:96-125` (`Monitor`) do
with `model(...)`'s outputs -- written against the same public calls, so the drop-in is exercised where the reference
actually touches it: `nn.DataParallel` wrap, `.mean()`, `torch.isnan(loss).any()`, `clip_grad_value_`,
`optimizer.step()/zero_grad()`, `float(loss_value)`, `metric(preds*255, targets*255).detach().cpu().numpy()`.
`import engine` itself cannot run here (it needs yacs, tensorboard and builds MS_SSIM(device="cuda") at import,
engine/monitor.py:50); used by tests/test_gpu_trainer_contract.py and bench.py's `trainer_sequence` record."""
import numpy as np
import torch
from torch.nn.utils import clip_grad_value_


class Meter:
    """utils/logging.py:73-99 AverageMeter(cache=False): update(val) does float(val)."""

    def __init__(self):
        self.val = self.sum = self.count = 0

    def update(self, val):
        if not isinstance(val, float):
            val = float(val)
        self.val = val
        self.sum += val
        self.count += 1

    @property
    def avg(self):
        return self.sum / max(self.count, 1)


class BatchMeter:
    """utils/logging.py:137-171 BatchAverageMeter(cache=False): update(values) does values.detach().cpu().numpy()."""

    def __init__(self):
        self.val = self.sum = self.count = 0

    def update(self, values):
        assert isinstance(values, (torch.Tensor, np.ndarray))
        if isinstance(values, torch.Tensor):
            values = values.detach().cpu().numpy()
        assert len(values.shape) == 1
        self.val = values.mean()
        self.sum += values.sum()
        self.count += len(values)

    @property
    def avg(self):
        return self.sum / max(self.count, 1)


class Monitor:
    """engine/monitor.py:25-125: one Meter per loss name, one BatchMeter per metric."""

    def __init__(self, loss_names, metric_dict):
        self.loss_names, self.metric_fns = list(loss_names), dict(metric_dict)
        for n in self.loss_names:
            setattr(self, n, Meter())
        for n in self.metric_fns:
            setattr(self, n, BatchMeter())

    def update_loss(self, **loss_dict):                  # monitor.py:96-105
        for name, value in loss_dict.items():
            getattr(self, name).update(value)

    def update_metric(self, preds, targets):             # monitor.py:113-125
        assert preds.max() <= 1.
        preds = preds * 255.
        targets = targets * 255.
        for name, fn in self.metric_fns.items():
            getattr(self, name).update(fn(preds, targets))

    def values(self):
        d = {n: getattr(self, n).val for n in self.loss_names}
        d.update({n: float(getattr(self, n).val) for n in self.metric_fns})
        return d


def train_step(model, optimizer, imgs, monitor, grad_clip=5.0, gd_steps=1, it=1, scheduler=None, noise=None):
    """engine/trainer.py:176-208, line for line (model may be wrapped in nn.DataParallel)."""
    assert model.training
    with torch.set_grad_enabled(True):
        imgs_tilde, losses = model(imgs) if noise is None else model(imgs, *noise)
        loss = losses.pop('total_loss').mean()
        loss /= gd_steps
        if torch.isnan(loss).any():
            raise FloatingPointError("get nan loss")     # the reference prints and exit()s
        loss.backward()
        if grad_clip:
            clip_grad_value_(model.parameters(), grad_clip)
        if it % gd_steps == 0:
            optimizer.step()
            if scheduler is not None:
                scheduler.step()
            optimizer.zero_grad()
    with torch.no_grad():
        monitor.update_metric(imgs_tilde, imgs)
        monitor.update_loss(**losses)
    return monitor.values()


def eval_step(model, imgs, monitor):
    """engine/evaluator.py:87-105 under run_eval's no_grad + inference_context (:67-72)."""
    assert not model.training
    img_tildes, losses = model(imgs)
    losses.pop('total_loss')
    monitor.update_metric(img_tildes, imgs)
    monitor.update_loss(**losses)
    return monitor.values()
