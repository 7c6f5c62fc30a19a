"""CPU restatement (numpy) of the reference's per-image input pipeline -- TEST INFRASTRUCTURE ONLY (tests/ import it;
the product path is csrc/misc_kernels.cu::hpc_u8_crop_normalize + image_compression_b200/data_utils).

Follows /root/reference/data_utils/transforms/transforms.py:
  get_transforms :23-35   Compose([RandomCrop(cfg.DATA.SIZE) (train only), ToTensor, ChannelFirst, Normalize(MEAN, STD)])
  RandomCrop     :143-166 x = np.random.randint(w - cw + 1); y = np.random.randint(h - ch + 1)   (x FIRST, then y);
                          image.crop((x, y, x + cw, y + ch)); an under-sized image is an AssertionError (no padding)
  ToTensor       :64-75   np.array(PIL image) -> torch.from_numpy(...).float()      (H, W, C) fp32, values 0..255
  ChannelFirst   :84-100  permute(2, 0, 1)
  Normalize      :103-127 F.normalize(image / 255., mean, std) = (image / 255 - mean) / std, all in fp32
and /root/reference/data_utils/batch.py:19-29 (Batch: torch.stack of the per-image tensors, .to(device)).

PINNED: tests/golden/make_golden_data.py runs the unmodified reference transforms on seeded images and asserts this file
reproduces them bit for bit; the vectors are committed as tests/golden/data_pipeline.npz.
"""
import numpy as np


def random_crop_offsets(w, h, size, rng=np.random):
    """(x, y) of RandomCrop.__call__ (transforms.py:160-164): two draws from the GLOBAL numpy stream, x first."""
    cw, ch = (size, size) if isinstance(size, int) else size
    if not (w >= cw and h >= ch):
        raise AssertionError(f"got under-sized image {w}x{h} for crop size {cw}x{ch}")
    x = rng.randint(w - cw + 1)
    y = rng.randint(h - ch + 1)
    return x, y


def crop_normalize(img_u8_hwc, x0, y0, size, mean=(0., 0., 0.), std=(1., 1., 1.)):
    """uint8 (H, W, C) -> fp32 (C, ch, cw): crop, /255, (v - mean) / std, every step in fp32 like torch does."""
    cw, ch = (size, size) if isinstance(size, int) else size
    a = np.asarray(img_u8_hwc)[y0:y0 + ch, x0:x0 + cw]
    if a.shape[:2] != (ch, cw):
        raise AssertionError("crop window leaves the image")
    v = a.transpose(2, 0, 1).astype(np.float32) / np.float32(255.0)
    m = np.asarray(mean, dtype=np.float32)[:, None, None]
    s = np.asarray(std, dtype=np.float32)[:, None, None]
    return (v - m) / s


def make_batch(images_u8, size, mean, std, train, rng=np.random):
    """What one collate call sees: [get_transforms(cfg, 'train' | 'val')(img) for img in images] stacked (Batch.__init__)."""
    out, offs = [], []
    for img in images_u8:
        h, w = img.shape[:2]
        x, y = random_crop_offsets(w, h, size, rng) if train else (0, 0)
        sz = size if train else (w, h)
        out.append(crop_normalize(img, x, y, sz, mean, std))
        offs.append((y, x))
    return np.stack(out), np.asarray(offs, dtype=np.int32)
