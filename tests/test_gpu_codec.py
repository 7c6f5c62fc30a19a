"""GPU entropy coder (csrc/rans_kernels.cu through codec.py and the module surface) -- SURVEY 8f rank 3.
The reference leaves compress / decompress / "symbol" as stubs (bmshl2018.py:106-110, entropy_model.py:157-169,235-251):
PARITY UNPINNED.  Checked instead:
  * the kernels' words equal oracle/rans_oracle.py's word for word on the same symbols and tables;
  * decoded symbols == rounded symbols, bit exactly, for both latents and both conditional models;
  * Compressor2018.decompress(compress(x)) equals the evaluation forward's reconstruction bit for bit;
  * the actual bitstream length against the estimated bpp of the evaluation forward (entropy_model.py:185)."""
import numpy as np
import pytest
import torch

import image_compression_b200 as pkg
from image_compression_b200 import codec
from oracle import rans_oracle as R
from util import make_pair, perturb, record_parity

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("n_img,npos,C,A,with_level", [(1, 1, 8, 1, False), (3, 12, 192, 21, False), (2, 48, 192, 65, True),
                                                        (1, 5, 16, 400, True)])
def test_kernel_words_equal_the_oracle(n_img, npos, C, A, with_level):
    rng = np.random.default_rng(npos * 7 + A)
    T = R.SCALE_LEVELS if with_level else C
    p = rng.random((T, A)) ** 6 + 1e-9
    p[:, A // 2] += 1.0
    cum = R.quantise_pmf(p)
    vmin = -(A // 2)
    level_np = rng.integers(0, T, size=(n_img, npos, C)).astype(np.uint8) if with_level else None
    sym_np = rng.integers(0, A, size=(n_img, npos, C)) + vmin
    sym_np[..., ::3] = 0 if A > 1 else vmin
    sym = torch.from_numpy(sym_np.astype(np.float32)).to(DEV)
    level = torch.from_numpy(level_np).to(DEV) if with_level else None
    lens, words, offs = codec._encode(sym.reshape(-1, C), None if level is None else level.reshape(-1, C), cum, n_img, vmin)
    want = R.encode_latent(sym_np, level_np, cum, vmin, codec.CHANNELS_PER_STREAM)
    for n in range(n_img):
        for s in range(C // codec.CHANNELS_PER_STREAM):
            got = words[offs[n, s]:offs[n, s] + lens[n, s]].tolist()
            assert got == want[n][s], (n, s, len(got), len(want[n][s]))
    back, back_lp = codec._decode(words, lens, None if level is None else level.reshape(-1, C), cum, n_img, npos, C, vmin,
                                  torch.device(DEV), True)
    assert torch.equal(back.view(n_img, npos, C), sym) and torch.equal(back_lp.float().view(n_img, npos, C), sym)
    # a corrupted stream is reported, not decoded into garbage silently
    if words.size > 4 and A > 1:
        bad = words.copy()
        bad[1] ^= 0x5555
        with pytest.raises(RuntimeError, match="malformed"):
            codec._decode(bad, lens, None if level is None else level.reshape(-1, C), cum, n_img, npos, C, vmin,
                          torch.device(DEV), False)
    # a symbol outside the table range is refused by the encoder
    sym2 = sym.clone()
    sym2.view(-1)[0] = float(vmin + A)
    with pytest.raises(RuntimeError, match="outside"):
        codec._encode(sym2.reshape(-1, C), None if level is None else level.reshape(-1, C), cum, n_img, vmin)


def test_scale_levels_kernel_matches_the_oracle():
    g = torch.Generator().manual_seed(3)
    sigma = torch.exp(torch.rand(4, 6, 5, 192, generator=g) * 16 - 8)
    sigma.view(-1)[:4] = torch.tensor([1e-10, 1e10, 0.11, 256.0])
    got = codec.scale_levels(sigma.to(DEV)).cpu().numpy()
    want = R.scale_levels(sigma.numpy())
    # fp32 logf on the device vs numpy's may differ by an ulp exactly on a level boundary: allow one level there
    diff = np.abs(got.astype(int) - want.astype(int))
    assert diff.max() <= 1 and (diff != 0).mean() < 1e-3, (diff.max(), (diff != 0).mean())


@pytest.mark.parametrize("conditional", ["LaplacianConditionalModel", "GaussianConditionalModel"])
def test_entropy_model_module_round_trips(conditional):
    model, params, _ = make_pair("psnr_256", 192, conditional=conditional)
    model.load_state_dict(perturb(params))
    model.eval()
    g = torch.Generator(device=DEV).manual_seed(5)
    z = torch.randn(2, 192, 3, 4, device=DEV, generator=g) * 3
    y = torch.randn(2, 192, 12, 16, device=DEV, generator=g) * 4
    scale = torch.exp(torch.randn(2, 192, 12, 16, device=DEV, generator=g))
    with torch.no_grad():
        zs = model.entropy_model.compress(z)
        assert len(zs) == 2 and all(isinstance(s, bytes) for s in zs)
        assert torch.equal(model.entropy_model.decompress(zs, (3, 4)), torch.round(z))
        ys = model.conditional_model.compress(y, scale)
        assert torch.equal(model.conditional_model.decompress(ys, scale), torch.round(y))
        assert torch.equal(model.conditional_model._dequantize(torch.round(y)), torch.round(y))


@pytest.mark.parametrize("recipe,conditional,N,hw", [("psnr_256", "LaplacianConditionalModel", 2, (128, 192)),
                                                     ("psnr_256", "GaussianConditionalModel", 1, (64, 64)),
                                                     ("psnr_256", "LaplacianConditionalModel", 3, (512, 768))])
def test_model_compress_decompress_equals_the_evaluation_forward(recipe, conditional, N, hw):
    model, params, _ = make_pair(recipe, 192, conditional=conditional)
    model.load_state_dict(perturb(params))
    model.eval()
    g = torch.Generator(device=DEV).manual_seed(9)
    x = torch.rand(N, 3, *hw, device=DEV, generator=g)
    with pkg.compute_mode("bf16"), torch.no_grad():
        strings = model.compress(x)
        xt = model.decompress(strings)
        # the evaluation forward of the same batch: same y (g_a runs batched on both sides), and x~ depends on y~ only.
        # (sigma differs in principle -- compress evaluates h_s image by image so that the decoder can reproduce it --
        #  but that moves the estimate by rounding noise, not the symbols)
        xt_ref, losses = model(x)
    assert len(strings) == N and xt.shape == x.shape
    assert torch.equal(xt, xt_ref)                                  # bit-exact reconstruction from the bitstream alone
    bits = [8 * len(s) for s in strings]
    est = [float(losses["bpp"]) * N * hw[0] * hw[1]]
    # actual size vs the cross-entropy estimate (entropy_model.py:185).  Above the estimate: 16-bit table quantisation
    # (freq >= 1 floors), 64 scale levels instead of the exact sigma, 32 bits of final state per stream (48 streams per
    # image), the stream-length table (4 B per stream) and the 16-byte header.  BELOW it: the tables are renormalised over
    # the value range [vmin, vmax] the header carries, which removes the tail mass the estimate pays for -- a large
    # effect for an untrained model (factorised prior initialised with scale 10, latents in a narrow range): measured
    # 0.857 at 3 x 512x768, 0.846 at 2 x 128x192, 1.29 for one 64x64 image (fixed overheads) (profiles/r02_parity.json).  The exact statement about code length is
    # made elsewhere: the kernels' words equal the oracle's (above), whose length is checked against the ideal code length
    # of the tables on the CPU (tests/test_rans_oracle.py).
    ratio = sum(bits) / sum(est)
    record_parity(f"codec_{conditional}_{N}x{hw[0]}x{hw[1]}_actual_vs_estimated_bits",
                  {"actual_bits": bits, "estimated_bits": est, "ratio": ratio,
                   "bpp_actual": [b / (hw[0] * hw[1]) for b in bits]})
    small = hw[0] * hw[1] <= 64 * 64
    assert 0.5 <= ratio <= (1.60 if small else 1.06), (ratio, bits, est)


def test_decompress_rejects_foreign_strings():
    model, _, _ = make_pair("psnr_256", 192)
    model.eval()
    x = torch.rand(1, 3, 64, 64, device=DEV)
    with torch.no_grad():
        s = model.compress(x)[0]
        with pytest.raises(ValueError):
            model.decompress([b"XXXX" + s[4:]])
        with pytest.raises(ValueError):
            model.decompress([s[:-2]])
        with pytest.raises(ValueError):
            model.compress(torch.rand(1, 3, 72, 64, device=DEV))
