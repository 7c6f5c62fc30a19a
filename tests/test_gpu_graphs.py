"""Evaluation forward replayed as a CUDA graph: bit-identical to the eager launches, sees weight updates,
and image_metrics can be part of the same graph."""
import pytest
import torch

import image_compression_b200 as pkg
from image_compression_b200 import utils as U
from image_compression_b200.utils.graphs import GraphedCallable

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a - b).norm()) / (float(b.norm()) + 1e-30)


def _model(seed=3):
    torch.manual_seed(seed)
    return pkg.build_model(pkg.get_recipe_cfg("psnr_256")).to("cuda").eval()


@pytest.mark.parametrize("mode", ["bf16", "fp32"])
def test_graphed_eval_matches_eager(mode):
    model = _model()
    x1, x2 = torch.rand(2, 3, 128, 192, device="cuda"), torch.rand(2, 3, 128, 192, device="cuda")
    with torch.no_grad(), pkg.compute_mode(mode):
        eager = [model(x) for x in (x1, x2)]
        model.cuda_graphs = True
        for x, (xt0, l0) in zip((x1, x2), eager):
            xt, l = model(x)
            assert torch.equal(xt, xt0)
            for k in l0:
                assert torch.equal(l[k], l0[k]), k
        assert len(model._graphs) == 1
        # a weight update between replays is seen (pack kernels are graph nodes)
        with torch.no_grad():
            model.synthesis_transform.layers[0].weight.mul_(1.5)
        model.cuda_graphs = False
        xt0, l0 = model(x1)
        model.cuda_graphs = True
        xt, l = model(x1)
        assert torch.equal(xt, xt0) and torch.equal(l["bpp"], l0["bpp"])
        assert len(model._graphs) == 1


def test_graphed_eval_with_metrics_and_training_unaffected():
    model = _model()
    x = torch.rand(1, 3, 256, 256, device="cuda")

    def step(t):
        xt, losses = model(t)
        ps, ms = U.image_metrics(xt, t)
        return xt, losses["bpp"], ps, ms

    with torch.no_grad():
        ref = step(x)
        g = GraphedCallable(step, x)
        for _ in range(2):
            out = g(x)
            for a, b in zip(out, ref):
                assert torch.equal(a, b)
    model.cuda_graphs = True
    model.train()
    _, losses = model(x)                       # training never goes through a graph
    losses["total_loss"].backward()
    assert model.analysis_transform.layers[0].weight.grad is not None and not model._graphs


def test_flat_gradients_are_written_in_place():
    """sharding.FlatGradients.zero() hands the slots of the flat buffer to the backward kernels as their output
    buffers: same gradients as the ordinary autograd accumulation, no copy, and a second backward accumulates."""
    from image_compression_b200 import sharding
    model = _model(7).train()
    x = torch.rand(2, 3, 128, 128, device="cuda")
    nz, ny = torch.rand(2, 192, 2, 2, device="cuda"), torch.rand(2, 192, 8, 8, device="cuda")
    _, losses = model(x, nz, ny)
    losses["total_loss"].backward()
    ref = {k: p.grad.clone() for k, p in model.named_parameters()}
    fg = sharding.FlatGradients(model.parameters())
    for rep in range(2):
        fg.zero()
        _, losses = model(x, nz, ny)
        losses["total_loss"].backward()
        fg.all_reduce_mean()
        assert fg.copies == 0
        base = fg.flat.data_ptr()
        for (k, p), o in zip(model.named_parameters(), fg.offsets):
            assert p.grad.data_ptr() == base + 4 * o, k
            assert _rel(p.grad, ref[k]) < 1e-5, k       # (the factorised-prior parameter sums use float atomics)
    _, losses = model(x, nz, ny)            # no zero(): accumulates on top, the ordinary way
    losses["total_loss"].backward()
    fg.collect()
    for k, p in model.named_parameters():
        assert _rel(p.grad, 2 * ref[k]) < 1e-5, k
    assert fg.copies == 0
