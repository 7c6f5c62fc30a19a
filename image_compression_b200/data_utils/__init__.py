"""Host side of the device input pipeline: the names of the reference's data_utils package that sit
on the path from a decoded image to `model(batch.imgs)` -- get_transforms, Batch, TrainingSampler -- with the float
conversion, the channel permute and the normalisation moved onto the GPU (csrc/misc_kernels.cu::hpc_u8_crop_normalize).
Datasets, file decoding and the DataLoader itself stay the caller's (out of scope, SURVEY 8 "out")."""
from .batch import Batch, DeviceImageSet
from .sampler import TrainingSampler
from .transforms import RandomCrop, U8Image, get_transforms

__all__ = ["Batch", "DeviceImageSet", "TrainingSampler", "RandomCrop", "U8Image", "get_transforms"]
