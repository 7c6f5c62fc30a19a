"""TrainingSampler: the reference's infinite index stream (data_utils/sampler.py:23-40), sharded for one process per GPU.

The reference builds `TrainingSampler(len(dataset), True, cfg.SEED)` in every process (dataloader.py:55): launched as N
data-parallel processes unchanged, every rank would read the SAME images.  Here rank r of W takes elements r, r+W, ... of
that one stream, so the ranks together consume exactly the reference's sequence, disjointly; W = 1 is the reference."""
import torch
from torch.utils.data.sampler import Sampler

__all__ = ["TrainingSampler"]


def _rank_world(rank, world_size):
    if rank is None or world_size is None:
        import torch.distributed as dist
        on = dist.is_available() and dist.is_initialized()
        rank = (dist.get_rank() if on else 0) if rank is None else rank
        world_size = (dist.get_world_size() if on else 1) if world_size is None else world_size
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return int(rank), int(world_size)


class TrainingSampler(Sampler):
    def __init__(self, size, shuffle, seed=None, rank=None, world_size=None):
        self._size = int(size)
        self._shuffle = shuffle
        self._seed = int(seed)
        self._rank, self._world = _rank_world(rank, world_size)

    def __iter__(self):
        seen = 0                                   # elements of the global stream before this epoch
        for order in self._epochs():
            # this rank's share of the epoch: stream positions p with p % world == rank (slicing keeps 0-d tensors, as
            # the reference yields them)
            yield from order[(self._rank - seen) % self._world::self._world]
            seen += self._size

    def _epochs(self):
        """one index tensor per pass over the dataset: sampler.py:32-40 (one generator seeded once, randperm per pass)"""
        g = torch.Generator()
        g.manual_seed(self._seed)
        while True:
            yield torch.randperm(self._size, generator=g) if self._shuffle else torch.arange(self._size)
