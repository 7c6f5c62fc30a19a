"""Input pipeline, CPU side: oracle/data_oracle.py against the vectors the UNMODIFIED reference
produced (tests/golden/data_pipeline.npz, made by tests/golden/make_golden_data.py), and the product's host half --
get_transforms / RandomCrop / Batch / TrainingSampler -- against the same vectors.  No GPU: the device half is
tests/test_gpu_data.py."""
import itertools
import os

import numpy as np
import pytest
import torch

from image_compression_b200 import data_utils as DU
from image_compression_b200.utils.config import get_cfg_defaults
from oracle import data_oracle as D

GOLDEN = np.load(os.path.join(os.path.dirname(__file__), "golden", "data_pipeline.npz"))
CASES = ("train_default", "train_meanstd", "val_full")


def case(name):
    n = int(GOLDEN[name + ".n"])
    return dict(imgs=[GOLDEN[f"{name}.img{i}"] for i in range(n)], out=GOLDEN[name + ".out"],
                offs=GOLDEN[name + ".offsets_yx"], size=int(GOLDEN[name + ".size"]), mean=GOLDEN[name + ".mean"].tolist(),
                std=GOLDEN[name + ".std"].tolist(), train=bool(GOLDEN[name + ".train"]))


def cfg_of(c):
    cfg = get_cfg_defaults()
    cfg.DATA.SIZE, cfg.DATA.MEAN, cfg.DATA.STD = c["size"], c["mean"], c["std"]
    return cfg


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_the_reference_vectors(name):
    c = case(name)
    np.random.seed(int(GOLDEN["seed"]))
    got, offs = D.make_batch(c["imgs"], c["size"], c["mean"], c["std"], c["train"])
    assert got.dtype == np.float32 and np.array_equal(got, c["out"]) and np.array_equal(offs, c["offs"])


@pytest.mark.parametrize("name", CASES)
def test_host_transforms_cut_the_reference_windows(name):
    """same draws from the global numpy stream, in the same order (x before y), as the reference's RandomCrop"""
    c = case(name)
    tf = DU.get_transforms(cfg_of(c), "train" if c["train"] else "val")
    np.random.seed(int(GOLDEN["seed"]))
    items = [tf(im) for im in c["imgs"]]
    assert [it.offset_yx for it in items] == [tuple(o) for o in c["offs"].tolist()]
    for it, im, (y, x), want in zip(items, c["imgs"], c["offs"], c["out"]):
        ch, cw = want.shape[1:]
        assert it.pixels.dtype == np.uint8 and it.pixels.flags.c_contiguous
        assert np.array_equal(it.pixels, im[y:y + ch, x:x + cw]) and it.shape == want.shape
        # the window the host sends, finished by the oracle, is the reference's output
        assert np.array_equal(D.crop_normalize(it.pixels, 0, 0, (cw, ch), it.mean, it.std), want)
    b = DU.Batch([(str(i), it) for i, it in enumerate(items)], None)
    assert b.image_ids == tuple(str(i) for i in range(len(items)))
    assert b.pixels.dtype == torch.uint8 and tuple(b.pixels.shape) == (len(items),) + items[0].pixels.shape
    with pytest.raises(RuntimeError, match="to\\(device\\)"):
        b.imgs                                                     # no CPU path
    with pytest.raises(RuntimeError, match="no CPU path"):
        b.to("cpu")


def test_undersized_ragged_and_foreign_inputs_fail_like_the_reference():
    with pytest.raises(AssertionError):                            # transforms.py:158-159: no padding
        DU.RandomCrop(64)(np.zeros((63, 64, 3), np.uint8))
    DU.RandomCrop(64)(np.zeros((64, 64, 3), np.uint8))             # exact fit: randint(1) twice
    tf = DU.get_transforms(get_cfg_defaults(), "val")
    a, b = tf(np.zeros((8, 9, 3), np.uint8)), tf(np.zeros((9, 8, 3), np.uint8))
    with pytest.raises(RuntimeError, match="equal size"):          # torch.stack's complaint in the reference
        DU.Batch([("a", a), ("b", b)], None)
    with pytest.raises(TypeError):
        DU.Batch([("a", torch.zeros(3, 8, 8))], None)
    with pytest.raises(TypeError):
        tf(np.zeros((8, 8, 3), np.float32))
    with pytest.raises(ValueError):
        tf(np.zeros((8, 8, 1), np.uint8))                          # 3 means for 1 channel
    assert tf(np.zeros((8, 9, 3), np.uint8)).shape == (3, 8, 9)
    with pytest.raises(RuntimeError, match="CUDA"):
        DU.DeviceImageSet(np.zeros((2, 8, 8, 3), np.uint8), (0, 0, 0), (1, 1, 1), "cpu")


@pytest.mark.parametrize("key", [k for k in GOLDEN.files if k.startswith("sampler.")])
def test_sampler_stream_is_the_references_and_shards_disjointly(key):
    _, size, shuffle, seed = key.split(".")
    want = GOLDEN[key].tolist()
    one = [int(i) for i in itertools.islice(iter(DU.TrainingSampler(int(size), bool(int(shuffle)), int(seed), 0, 1)), len(want))]
    assert one == want                                             # world of 1 == the reference
    for world in (2, 3, 8):
        per = len(want) // world
        parts = [[int(i) for i in itertools.islice(iter(DU.TrainingSampler(int(size), bool(int(shuffle)), int(seed), r, world)),
                                                   per)] for r in range(world)]
        # interleaving the ranks' streams gives the reference's stream back: disjoint, nothing skipped
        inter = [parts[k % world][k // world] for k in range(per * world)]
        assert inter == want[:per * world], (world, inter, want)
    first = next(iter(DU.TrainingSampler(5, True, 1)))
    assert isinstance(first, torch.Tensor) and first.dim() == 0    # 0-d tensors, like the reference yields
    with pytest.raises(ValueError):
        DU.TrainingSampler(5, True, 1, rank=2, world_size=2)


class _ToyDataset(torch.utils.data.Dataset):
    """the shape of the reference's datasets (imagenet_dataset.py:41-49): __getitem__ -> (image_id, transforms(img))"""

    def __init__(self, images, tf):
        self.images, self.tf = images, tf

    def __len__(self):
        return len(self.images)

    def __getitem__(self, idx):
        return f"img{int(idx)}", self.tf(self.images[int(idx)])


@pytest.mark.parametrize("workers", [0, 2])
def test_reference_dataloader_wiring_yields_uint8_batches(workers):
    """data_utils/dataloader.py:52-59 with this package's names: TrainingSampler -> BatchSampler -> DataLoader with
    collate_fn = Batch.  The items cross the worker boundary (pickle) as uint8 windows; the batch is uint8 (N,H,W,3)."""
    from torch.utils.data import BatchSampler, DataLoader
    c = case("train_default")
    cfg = cfg_of(c)
    ds = _ToyDataset(c["imgs"], DU.get_transforms(cfg, "train"))
    sampler = DU.TrainingSampler(len(ds), True, 7, rank=0, world_size=1)
    loader = DataLoader(ds, batch_sampler=BatchSampler(sampler, 3, drop_last=True), num_workers=workers,
                        collate_fn=lambda x: DU.Batch(x, cfg))
    it = iter(loader)
    want_ids = [int(i) for i in itertools.islice(iter(DU.TrainingSampler(len(ds), True, 7, 0, 1)), 6)]
    for step in range(2):
        b = next(it)
        assert b.image_ids == tuple(f"img{i}" for i in want_ids[3 * step: 3 * step + 3])
        assert b.pixels.dtype == torch.uint8 and tuple(b.pixels.shape) == (3, c["size"], c["size"], 3)
        for k, (y, x) in enumerate(b.offsets_yx):                      # every window is a crop of its own image
            src = c["imgs"][want_ids[3 * step + k]]
            assert np.array_equal(b.pixels[k].numpy(), src[y:y + c["size"], x:x + c["size"]])
    del it
