"""Batch: the collate object the reference's trainer and evaluator receive (data_utils/batch.py:19-29) --
`batch.image_ids`, `batch.to(device)`, `batch.imgs`, `batch.pin_memmory()` (the reference's spelling) -- holding uint8
pixels on the host and finishing ToTensor / ChannelFirst / Normalize on the GPU inside `.to(device)`.

    reference:  workers -> fp32 (C,H,W) per image -> torch.stack -> .to(device)       12 B / pixel over PCIe
    here:       workers -> uint8 (H,W,C) per image -> stack -> H2D -> one kernel       3 B / pixel over PCIe
`batch.imgs` after `.to(device)` is bit-identical to the reference's.

DeviceImageSet keeps a whole uint8 image set resident in HBM (180 GB holds ~2.4e5 decoded 500x500 RGB images; Kodak is
28 MB) and cuts batches from it by index + crop offset: 12 bytes per image cross PCIe per step."""
import numpy as np
import torch

from .. import _lib
from .transforms import RandomCrop, U8Image

__all__ = ["Batch", "DeviceImageSet"]


def _floats(v):
    import ctypes
    return (ctypes.c_float * len(v))(*v)


def _crop_normalize(src, n_src, src_h, src_w, index, offsets, n, c, oh, ow, mean, std):
    """src: device uint8 (n_src, src_h, src_w, c) contiguous -> fp32 (n, c, oh, ow) on the current stream"""
    out = torch.empty((n, c, oh, ow), dtype=torch.float32, device=src.device)
    _lib.check(_lib.lib().hpc_u8_crop_normalize(
        _lib.ptr(src), src_h * src_w * c, src_w * c, n_src, src_h, src_w, _lib.ptr(index), _lib.ptr(offsets), n, c, oh, ow,
        _floats(mean), _floats(std), _lib.ptr(out), _lib.stream()))
    return out


class Batch:
    def __init__(self, data, cfg=None):
        """data: [(image_id, U8Image)] as the datasets' __getitem__ returns with this package's get_transforms."""
        image_ids, imgs = list(zip(*data))
        self.image_ids = image_ids
        for im in imgs:
            if not isinstance(im, U8Image):
                raise TypeError("Batch takes the U8Image items of image_compression_b200.data_utils.get_transforms; a "
                                f"{type(im).__name__} came from another transform chain")
        first = imgs[0]
        for im in imgs[1:]:
            if im.pixels.shape != first.pixels.shape:      # torch.stack in the reference raises on this too
                raise RuntimeError(f"stack expects each image to be equal size, but got {first.pixels.shape} and "
                                   f"{im.pixels.shape}")
            if im.mean != first.mean or im.std != first.std:
                raise ValueError("images of one batch carry different mean / std")
        self.mean, self.std = first.mean, first.std
        self.offsets_yx = [im.offset_yx for im in imgs]   # where each window was cut (diagnostics; not used on the device)
        self.pixels = torch.from_numpy(np.stack([im.pixels for im in imgs]))    # uint8 (N,H,W,C), host
        self._imgs = None

    @property
    def imgs(self):
        if self._imgs is None:
            raise RuntimeError("Batch.imgs is produced on the GPU: call batch.to(device) first (the trainer does, "
                               "engine/trainer.py:173). There is no CPU path.")
        return self._imgs

    def to(self, device, non_blocking=True):
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError(f"Batch.to({device}): the normalisation kernel runs on a CUDA device; there is no CPU path")
        with torch.cuda.device(device):
            px = self.pixels.to(device, non_blocking=non_blocking and self.pixels.is_pinned())
            n, h, w, c = px.shape
            self._imgs = _crop_normalize(px, n, h, w, None, None, n, c, h, w, self.mean, self.std)
        return self

    def pin_memmory(self):
        self.pixels = self.pixels.pin_memory()
        return self

    pin_memory = pin_memmory       # so that DataLoader(pin_memory=True) finds it as well


class DeviceImageSet:
    """n same-sized uint8 images resident on the device; `batch(indices, train)` cuts a Batch-like object from them.
    The crop offsets are drawn on the host exactly as RandomCrop draws them (x first, then y, per image in order)."""

    def __init__(self, images_u8_nhwc, mean, std, device, image_ids=None):
        a = images_u8_nhwc if isinstance(images_u8_nhwc, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(images_u8_nhwc))
        if a.dtype != torch.uint8 or a.dim() != 4 or a.shape[3] > 4:
            raise TypeError(f"expected uint8 (N,H,W,C<=4), got {a.dtype} {tuple(a.shape)}")
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("DeviceImageSet lives on a CUDA device")
        self.pixels = a.contiguous().to(device)
        self.mean, self.std = tuple(float(m) for m in mean), tuple(float(s) for s in std)
        if len(self.mean) != a.shape[3] or len(self.std) != a.shape[3] or any(s == 0.0 for s in self.std):
            raise ValueError("mean / std must have one non-degenerate entry per channel")
        self.image_ids = list(image_ids) if image_ids is not None else [str(i) for i in range(a.shape[0])]

    def __len__(self):
        return self.pixels.shape[0]

    def batch(self, indices, size=None, rng=None):
        """size=None: whole images (the "val" chain); else a random size x size (or (cw, ch)) window per image."""
        idx = [int(i) for i in indices]
        n_src, h, w, c = self.pixels.shape
        if any(i < 0 or i >= n_src for i in idx):
            raise IndexError(f"image index outside 0..{n_src - 1}")
        if size is None:
            cw, ch, offs = w, h, [(0, 0)] * len(idx)
        else:
            crop = RandomCrop(size, rng)
            cw, ch = crop.size
            offs = [crop.offsets(w, h)[::-1] for _ in idx]         # (y, x) per image
        dev = self.pixels.device
        host = torch.tensor([[i, y, x] for i, (y, x) in zip(idx, offs)], dtype=torch.int32)
        with torch.cuda.device(dev):
            tab = host.to(dev)                                    # the only host-to-device bytes of the step
            index, offsets = tab[:, 0].contiguous(), tab[:, 1:].contiguous()
            imgs = _crop_normalize(self.pixels, n_src, h, w, index, offsets, len(idx), c, ch, cw, self.mean, self.std)
        out = _ResidentBatch()
        out.image_ids = tuple(self.image_ids[i] for i in idx)
        out.offsets_yx = offs
        out.imgs = imgs
        return out


class _ResidentBatch:
    """Batch surface for a batch that is already on its device"""

    def to(self, device):
        if torch.device(device).type != "cuda" or (torch.device(device).index not in (None, self.imgs.device.index)):
            raise RuntimeError(f"this batch lives on {self.imgs.device}")
        return self

    def pin_memmory(self):
        return self
