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
    """Flat fp32 gradient buffer aliased by every parameter's `.grad` (state_dict order).

    usage per step:   fg.zero();  loss.backward();  fg.all_reduce_mean()
    (with `overlap` -- the default on CUDA + NCCL -- exactly ONE backward between zero() and all_reduce_mean():
    the synthesis-transform slice is all-reduced while the rest of the backward still runs)
    """

    def __init__(self, params, device=None, overlap=None):
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
        self.direct = self.flat.is_cuda       # the CUDA backward writes into the slots itself (see zero())
        # start the all-reduce of a parameter group as soon as the backward reports its gradients final
        # (hotpath.set_grad_ready_hook); HPC_AR_OVERLAP=0 keeps the single all-reduce after backward
        # With overlap EVERY rank must call all_reduce_mean() after each backward (the early all-reduce is a
        # collective): pass overlap=False for a buffer whose gradients are not going to be reduced.
        if overlap is None:
            overlap = os.environ.get("HPC_AR_OVERLAP", "1") != "0"
        self.overlap = self.direct and bool(overlap)
        self._early = None
        self.copies = 0                       # slots that had to be filled by a copy in collect() (0 when direct)
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
            hp.set_grad_sink(sink)
            self._early = None
            hp.set_grad_ready_hook(self._reduce_early if self.overlap and self._multi_rank_nccl() else None)

    @staticmethod
    def _multi_rank_nccl(group=None):
        return (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
                and dist.get_backend(group) == "nccl")

    def _reduce_early(self, grads):
        """Asynchronous mean all-reduce of the contiguous slot range covered by `grads` (views of the flat buffer),
        issued on NCCL's stream behind the kernels enqueued so far; all_reduce_mean() reduces the rest and waits."""
        if self._early is not None or not grads:
            return
        base, total = self.flat.data_ptr(), self.flat.numel()
        lo, hi = total, 0
        for g in grads:
            o = (g.data_ptr() - base) // 4
            if g.dtype != torch.float32 or (g.data_ptr() - base) % 4 or not (0 <= o and o + g.numel() <= total):
                return                                  # a stray gradient: keep the single all-reduce
            lo, hi = min(lo, o), max(hi, (o + g.numel() + 3) // 4 * 4)
        inside = sum(1 for o in self.offsets if lo <= o < hi)
        if inside != len(grads):
            return                                      # the range holds slots of other groups
        from . import hotpath as hp
        hp.set_grad_ready_hook(None)                    # one early reduce per zero(): a second backward before the next
        work = dist.all_reduce(self.flat[lo:hi], op=dist.ReduceOp.AVG, async_op=True)   # zero() must not re-reduce
        self._early = (lo, hi, work)

    def collect(self):
        """Re-establish `.grad is a view of the flat buffer` after backward (a no-op check when the backward wrote
        the slots directly; otherwise one copy per stray gradient)."""
        if self.direct:
            from . import hotpath as hp
            hp.set_grad_sink(None)
            hp.set_grad_ready_hook(None)
        for p, o in zip(self.params, self.offsets):
            g = p.grad
            if g is not None and g.data_ptr() == self.flat.data_ptr() + 4 * o:
                continue
            v = self.flat[o:o + p.numel()].view(p.shape)
            if g is not None:
                v.copy_(g)
                self.copies += 1
            p.grad = v

    def all_reduce_mean(self, group=None, async_op=False):
        """Mean over ranks of the whole gradient (no-op without an initialised process group)."""
        self.collect()
        if not (dist.is_available() and dist.is_initialized()):
            return None
        world = dist.get_world_size(group)
        if world == 1:
            return None
        if dist.get_backend(group) == "nccl":
            if self._early is not None:
                lo, hi, first = self._early
                self._early = None
                works = [first]
                if lo > 0:
                    works.append(dist.all_reduce(self.flat[:lo], op=dist.ReduceOp.AVG, group=group, async_op=True))
                if hi < self.flat.numel():
                    works.append(dist.all_reduce(self.flat[hi:], op=dist.ReduceOp.AVG, group=group, async_op=True))
                for w in works:
                    w.wait()                            # the current stream waits for NCCL's
                return None
            return dist.all_reduce(self.flat, op=dist.ReduceOp.AVG, group=group, async_op=async_op)
        work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=False)
        self.flat.mul_(1.0 / world)
        return work

    def nbytes(self):
        return self.flat.numel() * 4


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
    for p in module.parameters():
        dist.broadcast(p.data, src=src, group=group)
