"""CUDA-graph replay of a static-shape sequence of hpcodec kernels (evaluation is launch bound at 3 images per
GPU: ~60 launches of 5-100 us each; one graph launch removes the per-kernel host cost).

Everything captured is this package's own kernels on torch's current stream; torch supplies the capture
plumbing (torch.cuda.graph, its private memory pool).  Parameters are read through pack kernels that are PART of
the graph, so weight updates between replays are seen.  Outputs are static tensors: consume them before the next
call.  The reference has no counterpart (eager torch only, engine/evaluator.py:87-105).
"""
import torch

from .. import _lib
from .. import hotpath as hp


class GraphedCallable:
    def __init__(self, fn, *example_inputs, warmup=2):
        for t in example_inputs:
            _lib.require_cuda(t)
        self.static_in = [t.detach().clone() for t in example_inputs]
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream(device=example_inputs[0].device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):          # warm-up outside the capture: lazy scratch buffers, func attributes
            for _ in range(warmup):
                fn(*self.static_in)
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        prof_was_on = _lib.prof_is_on()
        _lib.prof_pause(True)                  # event pairs cannot be recorded inside a capture
        hp.set_capturing(True)                 # bypass the packed-weight cache: packing must be in the graph
        try:
            with torch.cuda.graph(self.graph):
                self.static_out = fn(*self.static_in)
        finally:
            hp.set_capturing(False)
            _lib.prof_pause(not prof_was_on)

    def __call__(self, *inputs):
        for dst, src in zip(self.static_in, inputs):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_out
