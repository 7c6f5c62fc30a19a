"""get_transforms(cfg, mode): the host half of the reference's transform chain (data_utils/transforms/transforms.py:23-35).

The reference's chain is  RandomCrop (train) -> ToTensor -> ChannelFirst -> Normalize  and hands the collate function a
fp32 (C,H,W) tensor per image.  Here the chain stops after the crop: the worker returns the uint8 (H,W,C) window
(`U8Image`, 4x smaller) together with cfg.DATA.MEAN / STD, and `Batch.to(device)` finishes the chain on the GPU with
results bit-identical to the reference's (tests/test_gpu_data.py against tests/golden/data_pipeline.npz)."""
import numpy as np

__all__ = ["get_transforms", "RandomCrop", "U8Image", "HostTransforms"]


class U8Image:
    """One image after the host half of the chain: `pixels` uint8 (H,W,C), C-contiguous; mean / std still to apply."""
    __slots__ = ("pixels", "mean", "std", "offset_yx")

    def __init__(self, pixels, mean, std, offset_yx=(0, 0)):
        self.pixels, self.mean, self.std, self.offset_yx = pixels, mean, std, offset_yx

    @property
    def shape(self):
        """(C, H, W): the shape the reference's tensor has at this point"""
        h, w, c = self.pixels.shape
        return (c, h, w)


def _as_u8_hwc(image):
    """PIL image / ndarray -> uint8 (H,W,C) array, what the reference's ToTensor sees through np.array (:72)."""
    a = np.asarray(image)
    if a.dtype != np.uint8:
        raise TypeError(f"the device input pipeline takes 8-bit images, got {a.dtype} (the reference's ToTensor would "
                        f"accept it; convert on the host or use the reference's transforms)")
    if a.ndim == 2:
        a = a[:, :, None]
    if a.ndim != 3 or a.shape[2] > 4:
        raise ValueError(f"expected an (H,W,C<=4) image, got shape {a.shape}")
    return a


class RandomCrop:
    """transforms.py:143-166: a (cw, ch) window at x = randint(w-cw+1), y = randint(h-ch+1) -- x drawn FIRST, both from
    the global numpy stream (pass `rng` to use another); an under-sized image is an AssertionError, there is no padding."""

    def __init__(self, size, rng=None):
        self.size = (size, size) if isinstance(size, int) else tuple(size)
        assert len(self.size) == 2
        self.rng = rng

    def offsets(self, w, h):
        cw, ch = self.size
        assert w >= cw and h >= ch, f"got under-sized image {w}x{h} for crop size {cw}x{ch}"
        rng = np.random if self.rng is None else self.rng
        x = int(rng.randint(w - cw + 1))
        y = int(rng.randint(h - ch + 1))
        return x, y

    def __call__(self, image):
        a = _as_u8_hwc(image)
        h, w = a.shape[:2]
        x, y = self.offsets(w, h)
        cw, ch = self.size
        return a[y:y + ch, x:x + cw], (y, x)


class HostTransforms:
    """What get_transforms returns: callable on one decoded image, like the reference's Compose (:38-61)."""

    def __init__(self, crop, mean, std):
        self.crop = crop
        self.mean = tuple(float(m) for m in mean)
        self.std = tuple(float(s) for s in std)
        if any(s == 0.0 for s in self.std):
            raise ValueError("cfg.DATA.STD has a zero entry")

    def __call__(self, image):
        if self.crop is None:
            a, off = _as_u8_hwc(image), (0, 0)
        else:
            a, off = self.crop(image)
        if len(self.mean) != a.shape[2] or len(self.std) != a.shape[2]:
            raise ValueError(f"cfg.DATA.MEAN / STD have {len(self.mean)} / {len(self.std)} entries for a "
                             f"{a.shape[2]}-channel image")
        return U8Image(np.ascontiguousarray(a), self.mean, self.std, off)

    def __repr__(self):
        return f"HostTransforms(crop={None if self.crop is None else self.crop.size}, mean={self.mean}, std={self.std})"


def get_transforms(cfg, mode):
    """transforms.py:23-35: the crop only in "train" mode, the rest in every mode."""
    assert mode in ("train", "val")
    crop = RandomCrop(cfg.DATA.SIZE) if mode == "train" else None
    return HostTransforms(crop, cfg.DATA.MEAN, cfg.DATA.STD)
