"""Device half of the input pipeline (csrc/misc_kernels.cu::hpc_u8_crop_normalize through the C ABI and through
data_utils.Batch / DeviceImageSet) against the vectors of the unmodified reference (tests/golden/data_pipeline.npz) and
against oracle/data_oracle.py at BASELINE's batch size.  fp32 arithmetic with IEEE division on both sides: bit equality."""
import ctypes

import numpy as np
import pytest
import torch

from image_compression_b200 import _lib
from image_compression_b200 import data_utils as DU
from image_compression_b200._lib import lib, ptr, stream
from oracle import data_oracle as D
from test_data_host import CASES, GOLDEN, case, cfg_of

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("name", CASES)
def test_batch_to_device_equals_the_reference(name):
    c = case(name)
    tf = DU.get_transforms(cfg_of(c), "train" if c["train"] else "val")
    np.random.seed(int(GOLDEN["seed"]))
    data = [(f"id{i}", tf(im)) for i, im in enumerate(c["imgs"])]
    for pin in (False, True):
        b = DU.Batch(data, None)
        if pin:
            b.pin_memmory()
        assert b.to(DEV) is b
        assert b.imgs.dtype == torch.float32 and b.imgs.is_cuda and b.imgs.is_contiguous()
        assert np.array_equal(b.imgs.cpu().numpy(), c["out"])      # bit for bit
    assert b.image_ids == tuple(f"id{i}" for i in range(len(data)))


def test_resident_image_set_crops_by_index_and_offset():
    rng = np.random.RandomState(3)
    bank = rng.randint(0, 256, size=(7, 90, 123, 3)).astype(np.uint8)
    mean, std = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
    ds = DU.DeviceImageSet(bank, mean, std, DEV)
    idx = [6, 0, 3, 3, 5]
    np.random.seed(21)
    b = ds.batch(idx, size=(64, 48))
    np.random.seed(21)
    want, offs = D.make_batch([bank[i] for i in idx], (64, 48), mean, std, True)
    assert b.offsets_yx == [tuple(o) for o in offs.tolist()] and b.image_ids == tuple(str(i) for i in idx)
    assert np.array_equal(b.imgs.cpu().numpy(), want)
    full = ds.batch(range(7))                                      # the "val" chain: whole images
    want, _ = D.make_batch(list(bank), None, mean, std, False)
    assert np.array_equal(full.imgs.cpu().numpy(), want)
    assert full.to(DEV) is full
    with pytest.raises(IndexError):
        ds.batch([7])
    with pytest.raises(AssertionError):
        ds.batch([0], size=124)


def test_c_abi_strides_guards_and_baseline_batch():
    # BASELINE's training batch: 64 crops of 256x256 from padded rows (row stride > w*C) with an image stride gap
    rng = np.random.RandomState(5)
    N, H, W, C, S = 64, 300, 280, 3, 256
    row, img = W * C + 8, (W * C + 8) * H + 100
    buf = rng.randint(0, 256, size=N * img).astype(np.uint8)
    offs = np.stack([rng.randint(0, H - S + 1, size=N), rng.randint(0, W - S + 1, size=N)], 1).astype(np.int32)
    offs[0] = (H - S, 0)                                           # the last legal row offset (used below)
    mean, std = (0.1, 0.2, 0.3), (0.5, 0.25, 2.0)
    src, o = torch.from_numpy(buf).to(DEV), torch.from_numpy(offs).to(DEV)
    out = torch.empty(N, C, S, S, dtype=torch.float32, device=DEV)
    fl = lambda v: (ctypes.c_float * len(v))(*v)
    call = lambda *a: lib().hpc_u8_crop_normalize(*a, stream())
    _lib.check(call(ptr(src), img, row, N, H, W, None, ptr(o), N, C, S, S, fl(mean), fl(std), ptr(out)))
    imgs = [np.lib.stride_tricks.as_strided(buf[n * img:], shape=(H, W, C), strides=(row, C, 1)) for n in range(N)]
    want = np.stack([D.crop_normalize(im, int(x), int(y), S, mean, std) for im, (y, x) in zip(imgs, offs)])
    assert np.array_equal(out.cpu().numpy(), want)
    # a window or an index that leaves the source is not read: zeros go through the normalisation, no fault
    bad = o.clone()
    bad[0] = torch.tensor([H - S + 1, 0])
    bad[1] = torch.tensor([-3, W - 10])
    _lib.check(call(ptr(src), img, row, N, H, W, None, ptr(bad), N, C, S, S, fl(mean), fl(std), ptr(out)))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    zero = (np.float32(0) / np.float32(255) - np.asarray(mean, np.float32)) / np.asarray(std, np.float32)
    assert np.array_equal(got[0, :, -1, :], np.repeat(zero[:, None], S, 1)) and np.array_equal(got[0, :, :-1], want[0, :, 1:])
    assert np.array_equal(got[1, :, :3, :], np.broadcast_to(zero[:, None, None], (C, 3, S)))
    assert np.array_equal(got[2:], want[2:])
    # NULL mean / std = 0 / 1; one channel; argument errors are reported, nothing is launched
    g = torch.from_numpy(rng.randint(0, 256, size=(2, 5, 6, 1)).astype(np.uint8)).to(DEV)
    o1 = torch.empty(2, 1, 5, 6, dtype=torch.float32, device=DEV)
    _lib.check(call(ptr(g), 30, 6, 2, 5, 6, None, None, 2, 1, 5, 6, None, None, ptr(o1)))
    # (computed on the CPU like the reference's transforms: torch's CUDA division by a scalar multiplies by 1/255 instead)
    assert torch.equal(o1.cpu(), g.cpu().permute(0, 3, 1, 2).float() / 255.0)
    for args in ((ptr(g), 30, 6, 2, 5, 6, None, None, 2, 1, 6, 6, None, None, ptr(o1)),        # output taller than source
                 (ptr(g), 30, 5, 2, 5, 6, None, None, 2, 1, 5, 6, None, None, ptr(o1)),        # row stride < row
                 (ptr(g), 30, 6, 1, 5, 6, None, None, 2, 1, 5, 6, None, None, ptr(o1)),        # 1 source, 2 outputs, no index
                 (ptr(g), 30, 6, 2, 5, 6, None, None, 2, 5, 5, 6, None, None, ptr(o1)),        # C > 4
                 (ptr(g), 30, 6, 2, 5, 6, None, None, 2, 1, 5, 6, None, fl((0.0,)), ptr(o1)),  # std 0
                 (None, 30, 6, 2, 5, 6, None, None, 2, 1, 5, 6, None, None, ptr(o1))):
        assert call(*args) == 1 and b"hpc_u8_crop_normalize" in lib().hpc_last_error()
