"""Batch-sharded training and per-image-sharded evaluation, one process per GPU.

The reference's only parallelism is a single-process `nn.DataParallel` (engine/trainer.py:256-259):
replicas see a slice of the batch, each normalises its loss by its LOCAL pixel count
(bmshl2018.py:70,88) and the trainer averages the replica losses (engine/trainer.py:180), i.e. the
global gradient is the MEAN of the replica gradients.  Here every rank is such a replica and the mean
is ONE NCCL all-reduce (SUM, then x 1/world) over a flat fp32 buffer that the `.grad` tensors alias,
so there is no pack/unpack pass.  Evaluation has no coupling at all: image i goes to rank
i % world and the per-image metrics are all-gathered at the end.

torch.distributed is plumbing only (NCCL on the GPU box, gloo in the CPU tests).
"""
import os

import torch
import torch.distributed as dist


class FlatGradients:
    """Flat fp32 gradient buffer aliased by every parameter's `.grad`, in the order `params` is given.

    usage per step:   fg.zero();  loss.backward();  fg.all_reduce_mean()
    (with `overlap` -- the default on CUDA + NCCL -- exactly ONE backward between zero() and all_reduce_mean():
    contiguous slot ranges are all-reduced while the rest of the backward still runs, as the backward reports their
    gradients final through hotpath.grads_ready; give `params` in backward-completion order -- GradientReducer does --
    and only the last bucket is exposed)
    """

    def __init__(self, params, device=None, overlap=None, group=None, alias_now=True):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("FlatGradients: no trainable parameters")
        device = device or self.params[0].device
        for p in self.params:
            if p.dtype != torch.float32:
                raise TypeError("FlatGradients: parameters must be fp32 (the reference trains in fp32)")
        # 16-byte aligned slots so vectorised kernels may write the views directly
        offs, n = [], 0
        for p in self.params:
            offs.append(n)
            n += (p.numel() + 3) // 4 * 4
        self.flat = torch.zeros(n, dtype=torch.float32, device=device)
        self.numel = sum(p.numel() for p in self.params)
        self.offsets = offs
        self.group = group                    # every collective of this buffer runs on this group
        self._slot_of = {o: i for i, o in enumerate(offs)}
        self.direct = self.flat.is_cuda       # the CUDA backward writes into the slots itself (see zero())
        # start the all-reduce of a slot range as soon as the backward reports its gradients final
        # (hotpath.set_grad_ready_hook); HPC_AR_OVERLAP=0 keeps the single all-reduce after backward
        # With overlap EVERY rank must call all_reduce_mean() after each backward (the early all-reduces are
        # collectives): pass overlap=False for a buffer whose gradients are not going to be reduced.
        if overlap is None:
            overlap = os.environ.get("HPC_AR_OVERLAP", "1") != "0"
        self.overlap = self.direct and bool(overlap)
        self.min_bucket = int(os.environ.get("HPC_AR_MIN_BUCKET", str(1 << 21)))     # elements (8 MB)
        self._early, self._ready, self._frontier = [], set(), 0
        self.copies = 0                       # slots that had to be filled by a copy in collect() (0 when direct)
        if alias_now:
            for p in self.params:
                p.grad = self._view(p)

    def _view(self, p):
        o = self.offsets[self._index(p)]
        return self.flat[o:o + p.numel()].view(p.shape)

    def _index(self, p):
        if not hasattr(self, "_idx"):
            self._idx = {id(q): i for i, q in enumerate(self.params)}
        return self._idx[id(p)]

    def zero(self):
        """One fill of the flat buffer.  On CUDA the slots are then handed to the backward kernels as their
        OUTPUT buffers (hotpath.set_grad_sink): `.grad` is dropped, the package's backward writes every gradient
        straight into its slot and autograd adopts the returned view as the new `.grad` -- no per-parameter
        AccumulateGrad add.  A second backward before the next zero() accumulates the ordinary way."""
        self.flat.zero_()
        if self.direct:
            from . import hotpath as hp
            sink = {}
            for p in self.params:
                p.grad = None
                sink[p.data_ptr()] = self._view(p)       # a fresh view object: autograd must be its only owner
            hp.set_grad_sink(sink, owner=self)
            self._early, self._ready, self._frontier = [], set(), 0
            hp.set_grad_ready_hook(self._reduce_early if self.overlap and self._multi_rank_nccl(self.group) else None)

    @staticmethod
    def _multi_rank_nccl(group=None):
        return (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
                and dist.get_backend(group) == "nccl")

    def _end_of(self, slot):
        return self.offsets[slot] if slot < len(self.offsets) else self.flat.numel()

    def _reduce_early(self, grads, flush=False):
        """The backward reports `grads` (views of slots of the flat buffer, handed out by zero()) final.  Slots become
        reducible in buffer order: once the ready slots beyond the frontier cover >= min_bucket elements (or at a
        `flush`), their range is mean-all-reduced asynchronously on NCCL's stream behind the kernels enqueued so
        far; all_reduce_mean() reduces whatever is left and waits."""
        base, total = self.flat.data_ptr(), self.flat.numel()
        for g in grads:
            off = g.data_ptr() - base
            slot = self._slot_of.get(off // 4) if off % 4 == 0 else None
            if slot is None or g.dtype != torch.float32 or g.numel() != self.params[slot].numel():
                return                                  # a stray gradient: leave it to the final all-reduce
            self._ready.add(slot)
        j = self._frontier
        while j < len(self.offsets) and j in self._ready:
            j += 1
        lo, hi = self._end_of(self._frontier), self._end_of(j)
        if j > self._frontier and (hi - lo >= self.min_bucket or flush or j == len(self.offsets)):
            work = dist.all_reduce(self.flat[lo:hi], op=dist.ReduceOp.AVG, group=self.group, async_op=True)
            self._early.append((self._frontier, j, work))
            self._frontier = j

    def collect(self):
        """Re-establish `.grad is a view of the flat buffer` after backward (a no-op check when the backward wrote
        the slots directly; otherwise one copy per stray gradient)."""
        if self.direct:
            from . import hotpath as hp
            hp.set_grad_sink(None)
            hp.set_grad_ready_hook(None)
        for i, (p, o) in enumerate(zip(self.params, self.offsets)):
            g = p.grad
            if g is not None and g.data_ptr() == self.flat.data_ptr() + 4 * o:
                continue
            if g is not None and i < self._frontier:
                raise RuntimeError(f"FlatGradients: the gradient of parameter slot {i} was reported final and already "
                                   "all-reduced, but autograd holds a different tensor for it")
            v = self.flat[o:o + p.numel()].view(p.shape)
            if g is not None:
                v.copy_(g)
                self.copies += 1
            p.grad = v

    def all_reduce_mean(self, group=None, async_op=False):
        """Mean over ranks of the whole gradient (no-op without an initialised process group)."""
        group = group if group is not None else self.group
        self.collect()
        if not (dist.is_available() and dist.is_initialized()):
            return None
        world = dist.get_world_size(group)
        if world == 1:
            return None
        if dist.get_backend(group) == "nccl":
            if self._early:
                works = [w for _, _, w in self._early]
                lo = self._end_of(self._frontier)
                self._early, self._ready, self._frontier = [], set(), 0
                if lo < self.flat.numel():
                    works.append(dist.all_reduce(self.flat[lo:], op=dist.ReduceOp.AVG, group=group, async_op=True))
                for w in works:
                    w.wait()                            # the current stream waits for NCCL's
                return None
            return dist.all_reduce(self.flat, op=dist.ReduceOp.AVG, group=group, async_op=async_op)
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=False)
        self.flat.mul_(1.0 / world)
        return work

    def nbytes(self):
        return self.flat.numel() * 4


class GradientReducer:
    """Trainer-transparent data parallelism: the model itself arms the flat gradient buffer in its
    forward and all-reduces it from inside the backward, so `engine/trainer.py:176-194` (forward, `.mean()`, `isnan`,
    `backward()`, `clip_grad_value_`, `optimizer.step()`, `optimizer.zero_grad()`) runs unchanged under torchrun --
    the mechanism DDP's reducer uses: a hook on the loss queues an autograd-engine callback that runs once the whole
    backward has been enqueued; slot ranges whose gradients are final earlier are reduced while the backward runs.
    With one process (no process group) it still gives the direct gradient writes (no AccumulateGrad adds)."""

    def __init__(self, params, group=None, overlap=None, broadcast_from=0, module=None):
        self.fg = FlatGradients(params, overlap=overlap, group=group, alias_now=False)
        self.group = group
        self._queued = False
        self.armed = False
        self.active = False
        if module is not None and broadcast_from is not None:
            broadcast_parameters(module, src=broadcast_from, group=group)      # every replica starts from rank 0's weights

    def before_forward(self):
        """Called by the model at the start of a training-mode forward with grad enabled."""
        from . import hotpath as hp
        fg = self.fg
        owner = hp.grad_sink_owner()
        if owner is not None and owner is not fg:
            self.active = False          # the caller drives its own FlatGradients (fg.zero() ... fg.all_reduce_mean())
            return False
        self.active = True
        if all(p.grad is None for p in fg.params):       # the trainer's optimizer.zero_grad() (set_to_none) ran
            fg.zero()
            self.armed = True
        else:
            # gradients are being accumulated (SOLVER.GD_STEPS > 1) or the caller keeps zero-filled .grad tensors: this
            # backward adds into `.grad` the ordinary way; the reduce at its end covers the whole buffer
            if not self._multi_rank():
                if owner is fg:
                    hp.set_grad_sink(None)
                    hp.set_grad_ready_hook(None)
                self.active = self.armed = False     # one process: nothing to reduce, the gradients stay where they are
                return False
            if owner is fg:
                hp.set_grad_sink(None)
                hp.set_grad_ready_hook(None)
            fg._early, fg._ready, fg._frontier = [], set(), 0
            self.armed = False
        return True

    def _multi_rank(self):
        return dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1

    def watch(self, loss):
        """Hook the (scalar) loss the caller will call backward() on."""
        if self.active and loss.requires_grad:
            loss.register_hook(self._on_backward)

    def _on_backward(self, grad):
        if not self._queued:
            self._queued = True
            from torch.autograd import Variable
            Variable._execution_engine.queue_callback(self._finish)
        return None

    def _finish(self):
        self._queued = False
        self.armed = False
        self.fg.all_reduce_mean()


def shard_indices(n_items, rank, world):
    """Indices of the evaluation images owned by `rank` (round-robin: i % world == rank)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} not in [0,{world})")
    return list(range(rank, n_items, world))


def gather_per_image(values, n_items, rank, world, group=None):
    """values: (len(shard_indices(n_items, rank, world)), K) tensor of this rank's per-image metrics.
    Returns the (n_items, K) table in image order on every rank."""
    values = values.reshape(len(shard_indices(n_items, rank, world)), -1)
    K = values.shape[1]
    if world == 1:
        return values
    per = (n_items + world - 1) // world
    pad = torch.zeros((per, K), dtype=values.dtype, device=values.device)
    pad[:values.shape[0]] = values
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    out = torch.empty((n_items, K), dtype=values.dtype, device=values.device)
    for r in range(world):
        idx = shard_indices(n_items, r, world)
        out[idx] = bufs[r][:len(idx)]
    return out


def broadcast_parameters(module, src=0, group=None):
    """Make every rank start from rank `src`'s weights (the reference replicates module state the
    same way each step through DataParallel's Broadcast)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    with torch.no_grad():
        for p in module.parameters():
            # p.detach() shares p's version counter (p.data would not): the in-place broadcast bumps it, so caches
            # keyed on `_version` (hotpath's packed operands) see the new weights
            dist.broadcast(p.detach(), src=src, group=group)
