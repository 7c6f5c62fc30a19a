"""One process, two GPUs: the per-device one-time set-up of the library (opt-in shared memory sizes, the static scratch's
address) must happen on EVERY device a process uses, not once per process (`static bool attr_set`).
A training step on cuda:1 after one on cuda:0 has to give the same numbers; before the fix the first tcgen05 launch on
cuda:1 failed with "invalid argument".  Skipped on a single-GPU box."""
import pytest
import torch

import image_compression_b200 as pkg
from util import make_pair, perturb

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
@pytest.mark.parametrize("recipe", ["psnr_256", "ssim_64"])
def test_training_step_on_a_second_device_matches_the_first(recipe):
    out = []
    x0 = torch.rand(2, 3, 192, 192, generator=torch.Generator().manual_seed(1))
    for dev in ("cuda:0", "cuda:1"):
        with torch.cuda.device(dev):
            model, params, _ = make_pair(recipe, 192, device=dev)
            model.load_state_dict(perturb(params))
            model.train()
            torch.manual_seed(7)                       # the in-kernel noise is keyed from torch's CPU generator
            with pkg.compute_mode("bf16"):
                _, losses = model(x0.to(dev))
                losses["total_loss"].backward()
            torch.cuda.synchronize()
            out.append((float(losses["total_loss"]), {k: p.grad.detach().cpu() for k, p in model.named_parameters()}))
    (l0, g0), (l1, g1) = out
    # same kernels, same seeded noise, same order of summation on both devices
    assert abs(l0 - l1) <= 1e-6 * abs(l0), (l0, l1)
    for k in g0:
        assert torch.allclose(g0[k], g1[k], rtol=1e-4, atol=1e-7 * float(g0[k].abs().max())), k
