"""World-size-2 CPU (gloo) tests of the N>1 host logic: flat gradient aliasing + mean all-reduce equals the
gradient of the concatenated batch with per-replica normalisation (engine/trainer.py:180), per-image
evaluation sharding and gather."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from image_compression_b200 import sharding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, results):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Tanh(), torch.nn.Linear(7, 3))
        sharding.broadcast_parameters(net)
        fg = sharding.FlatGradients(net.parameters())
        assert all(p.grad.data_ptr() >= fg.flat.data_ptr() for p in net.parameters())
        x = torch.randn(4 * world, 5, generator=torch.Generator().manual_seed(1))
        xs = x[rank * 4:(rank + 1) * 4]
        for _ in range(2):                       # second pass checks fg.zero() really resets the aliased grads
            fg.zero()
            net(xs).pow(2).mean().backward()     # local normalisation, like a DataParallel replica
            fg.all_reduce_mean()
        # reference: mean over replicas of the per-replica losses on one process
        ref = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Tanh(), torch.nn.Linear(7, 3))
        ref.load_state_dict(net.state_dict())
        sum(ref(x[r * 4:(r + 1) * 4]).pow(2).mean() for r in range(world)).div(world).backward()
        err = max(float((p.grad - q.grad).abs().max()) for p, q in zip(net.parameters(), ref.parameters()))
        # evaluation sharding: 7 images over 2 ranks, 3 metrics each
        idx = sharding.shard_indices(7, rank, world)
        vals = torch.tensor([[i, 10.0 * i, -i] for i in idx], dtype=torch.float32).reshape(len(idx), 3)
        table = sharding.gather_per_image(vals, 7, rank, world)
        ok = torch.equal(table, torch.tensor([[i, 10.0 * i, -i] for i in range(7)], dtype=torch.float32))
        results[rank] = (err, ok, fg.numel, fg.flat.numel())
    finally:
        dist.destroy_process_group()


def test_flat_gradients_and_eval_sharding_world2():
    world = 2
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), results), nprocs=world, join=True)
    assert len(results) == world
    for rank in range(world):
        err, ok, numel, flat = results[rank]
        assert err < 1e-6, err
        assert ok
        assert numel == 5 * 7 + 7 + 7 * 3 + 3 and flat >= numel and flat % 4 == 0


def _reducer_worker(rank, world, port, results):
    """The trainer-transparent path: the model arms the reducer in forward, a hook on the loss queues the all-reduce at
    the end of backward; the caller only runs the reference's sequence (engine/trainer.py:179-194)."""
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(rank)                  # different initial weights per rank: the reducer broadcasts rank 0's
        net = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Tanh(), torch.nn.Linear(7, 3))
        red = sharding.GradientReducer(list(net.parameters())[::-1], module=net)
        opt = torch.optim.SGD(net.parameters(), lr=0.1)
        torch.manual_seed(0)
        ref = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Tanh(), torch.nn.Linear(7, 3))
        ropt = torch.optim.SGD(ref.parameters(), lr=0.1)
        w0 = max(float((p - q).abs().max()) for p, q in zip(net.parameters(), ref.parameters()))
        x = torch.randn(4 * world, 5, generator=torch.Generator().manual_seed(1))
        err = 0.0
        for it in range(1, 5):                   # GD_STEPS = 2: backward, backward, step, zero_grad
            xs = x[rank * 4:(rank + 1) * 4] * it
            assert red.before_forward()
            loss = net(xs).pow(2).mean() / 2
            red.watch(loss)
            loss.backward()
            sum(ref(x[r * 4:(r + 1) * 4] * it).pow(2).mean() / 2 for r in range(world)).div(world).backward()
            err = max(err, max(float((p.grad - q.grad).abs().max()) for p, q in zip(net.parameters(), ref.parameters())))
            if it % 2 == 0:
                opt.step(); ropt.step()
                opt.zero_grad(); ropt.zero_grad()
        werr = max(float((p - q).abs().max()) for p, q in zip(net.parameters(), ref.parameters()))
        results[rank] = (w0, err, werr)
    finally:
        dist.destroy_process_group()


def test_gradient_reducer_transparent_world2():
    world = 2
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_reducer_worker, args=(world, _free_port(), results), nprocs=world, join=True)
    assert len(results) == world
    for rank in range(world):
        w0, err, werr = results[rank]
        assert w0 == 0.0            # broadcast from rank 0 at attach
        assert err < 1e-6 and werr < 1e-6, (err, werr)


def test_shard_indices_cover_everything_once():
    for n, w in ((24, 8), (7, 2), (3, 4), (1, 1)):
        seen = sorted(i for r in range(w) for i in sharding.shard_indices(n, r, w))
        assert seen == list(range(n))
    with pytest.raises(ValueError):
        sharding.shard_indices(4, 2, 2)


def test_flat_gradients_single_process_noop():
    net = torch.nn.Linear(3, 2)
    fg = sharding.FlatGradients(net.parameters())
    net(torch.ones(1, 3)).sum().backward()
    assert fg.all_reduce_mean() is None
    assert float(fg.flat.abs().sum()) > 0
    fg.zero()
    assert float(net.weight.grad.abs().sum()) == 0.0


def _sampler_worker(rank, world, port, results):
    import itertools
    from image_compression_b200 import data_utils as DU
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # the reference's call, unchanged: TrainingSampler(len(dataset), True, cfg.SEED) -- rank and world size come from
        # the initialised process group
        results[rank] = [int(i) for i in itertools.islice(iter(DU.TrainingSampler(11, True, 3)), 9)]
    finally:
        dist.destroy_process_group()


def test_training_sampler_shards_the_reference_stream_under_a_process_group():
    import itertools
    from image_compression_b200 import data_utils as DU
    world = 2
    mgr = mp.Manager()
    results = mgr.dict()
    mp.spawn(_sampler_worker, args=(world, _free_port(), results), nprocs=world, join=True)
    whole = [int(i) for i in itertools.islice(iter(DU.TrainingSampler(11, True, 3, rank=0, world_size=1)), 18)]
    assert [results[k % 2][k // 2] for k in range(18)] == whole        # disjoint, nothing skipped, the reference's order
