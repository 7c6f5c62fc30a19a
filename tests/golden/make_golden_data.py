"""Golden vectors of the reference's input pipeline and index sampler.  Run in the build container (needs
/root/reference):   python tests/golden/make_golden_data.py
Executes the UNMODIFIED /root/reference/data_utils/transforms/transforms.py (loaded from its file: the package's
__init__ pulls the dataloader machinery, pandas and yacs) on seeded uint8 images, asserts that oracle/data_oracle.py
reproduces every output bit for bit, and writes tests/golden/data_pipeline.npz."""
import importlib.util
import os
import sys

import numpy as np
import PIL.Image

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import data_oracle as D  # noqa: E402
from oracle import ref_shim  # noqa: E402

ref_shim.install()
spec = importlib.util.spec_from_file_location("ref_transforms", "/root/reference/data_utils/transforms/transforms.py")
T = importlib.util.module_from_spec(spec)
spec.loader.exec_module(T)
spec = importlib.util.spec_from_file_location("ref_sampler", "/root/reference/data_utils/sampler.py")
S = importlib.util.module_from_spec(spec)
spec.loader.exec_module(S)


class _NS:
    pass


def cfg_of(size, mean, std):
    c = _NS()
    c.DATA = _NS()
    c.DATA.SIZE, c.DATA.MEAN, c.DATA.STD = size, list(mean), list(std)
    return c


def main():
    rng = np.random.RandomState(7)
    out = {}
    cases = [("train_default", 64, (0., 0., 0.), (1., 1., 1.), True, [(80, 96), (64, 64), (70, 131), (200, 65)]),
             ("train_meanstd", 32, (0.1, 0.2, 0.3), (0.5, 0.25, 1.0), True, [(40, 33), (32, 50), (47, 47)]),
             ("val_full", 0, (0., 0., 0.), (1., 1., 1.), False, [(48, 80), (48, 80)])]
    for name, size, mean, std, train, shapes in cases:
        imgs = [rng.randint(0, 256, size=(h, w, 3)).astype(np.uint8) for h, w in shapes]
        cfg = cfg_of(size, mean, std)
        tf = T.get_transforms(cfg, "train" if train else "val")
        np.random.seed(11)
        ref = [tf(PIL.Image.fromarray(im)).numpy() for im in imgs]
        np.random.seed(11)
        got, offs = D.make_batch(imgs, size, mean, std, train)
        for i, (a, b) in enumerate(zip(ref, got)):
            assert a.dtype == np.float32 and a.shape == b.shape and np.array_equal(a, b), (name, i)
        out[name + ".n"] = np.int32(len(imgs))
        for i, im in enumerate(imgs):
            out[f"{name}.img{i}"] = im
        out[name + ".out"] = np.stack(ref)
        out[name + ".offsets_yx"] = offs
        out[name + ".size"] = np.int32(size)
        out[name + ".mean"] = np.asarray(mean, dtype=np.float32)
        out[name + ".std"] = np.asarray(std, dtype=np.float32)
        out[name + ".train"] = np.int32(train)
    out["seed"] = np.int32(11)
    # the reference's infinite index stream (data_utils/sampler.py:23-40): first 45 indices of three settings
    import itertools
    for size, shuffle, seed in ((10, True, 5), (7, False, 0), (13, True, 2020)):
        it = iter(S.TrainingSampler(size, shuffle, seed))
        out[f"sampler.{size}.{int(shuffle)}.{seed}"] = np.asarray([int(i) for i in itertools.islice(it, 45)], dtype=np.int64)
    # an under-sized image is an AssertionError in the reference (transforms.py:158-159) and in the oracle
    for fn in (lambda: T.RandomCrop(64)(PIL.Image.fromarray(np.zeros((63, 64, 3), np.uint8))),
               lambda: D.random_crop_offsets(64, 63, 64)):
        try:
            fn()
            raise SystemExit("under-sized image was accepted")
        except AssertionError:
            pass
    path = os.path.join(ROOT, "tests", "golden", "data_pipeline.npz")
    np.savez_compressed(path, **out)
    print("oracle == reference transforms on", sum(int(out[k]) for k in out if k.endswith(".n")), "images; wrote", path,
          os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
