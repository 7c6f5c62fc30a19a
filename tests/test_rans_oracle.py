"""CPU tests of the entropy-coder format (oracle/rans_oracle.py) and of the host-side table code of the product
(image_compression_b200/codec.py).  Parity is unpinned by the reference (it has no coder: bmshl2018.py:106-110,
entropy_model.py:157-169,235-251), so the bar is self-consistency: decode(encode(x)) == x bit for bit, the coded length
within a stated margin of the ideal code length, and the two independent table builders agreeing exactly."""
import numpy as np
import pytest

from image_compression_b200 import codec
from oracle import rans_oracle as R


def _random_tables(rng, T, A):
    p = rng.random((T, A)) ** 4 + 1e-9
    p[:, A // 2] += 2.0               # a peaked pmf, like a latent's
    return p


def test_table_quantisation_matches_the_product_and_is_valid():
    rng = np.random.default_rng(0)
    for T, A in ((1, 1), (3, 2), (192, 37), (64, 511)):
        p = _random_tables(rng, T, A)
        p[0, :] = 0.0                 # a degenerate row: falls back to the uniform table
        cum = R.quantise_pmf(p)
        assert np.array_equal(cum, codec.quantise_pmf(p))
        assert cum.dtype == np.uint32 and cum.shape == (T, A + 1)
        assert (cum[:, 0] == 0).all() and (cum[:, -1] == R.TOTAL).all()
        assert (np.diff(cum.astype(np.int64), axis=1) >= 1).all()          # every symbol stays codable
    with pytest.raises(ValueError):
        R.quantise_pmf(np.ones((1, R.TOTAL)))


def test_scale_table_and_conditional_pmf_match_the_product():
    assert np.array_equal(R.scale_table(), codec.scale_table())
    assert (R.SCALE_MIN, R.SCALE_MAX, R.SCALE_LEVELS) == (codec.SCALE_MIN, codec.SCALE_MAX, codec.SCALE_LEVELS)
    for dist, gauss in (("laplace", False), ("gauss", True)):
        a = R.conditional_pmf(R.scale_table(), -20, 17, dist)
        b = codec.conditional_pmf(-20, 17, gauss)
        assert np.array_equal(a, b)
        assert (a >= 0).all() and (a.sum(axis=1) <= 1.0 + 1e-12).all()
        # symmetric in the value (entropy_model.py:347-348 takes abs) and widening with the scale
        assert np.allclose(a[:, 20 - 5], a[:, 20 + 5], rtol=0, atol=0)
        assert a[0, 20] > a[-1, 20]


def test_scale_levels_pick_the_smallest_covering_level():
    s = R.scale_table()
    sig = np.array([1e-10, 0.05, 0.11, 0.12, 1.0, 255.9, 256.0, 1e10], dtype=np.float32)
    lv = R.scale_levels(sig)
    assert lv[0] == 0 and lv[1] == 0 and lv[-1] == R.SCALE_LEVELS - 1 and lv[-2] >= R.SCALE_LEVELS - 2
    inside = (sig > s[0]) & (sig < s[-1] * 0.999)
    assert (s[lv[inside]] >= sig[inside] * (1 - 1e-5)).all()
    assert (s[np.maximum(lv[inside], 1) - 1] <= sig[inside] * (1 + 1e-5)).all()


@pytest.mark.parametrize("n_img,npos,C,cpg,A,with_level", [(1, 1, 8, 8, 1, False), (2, 6, 16, 8, 9, False),
                                                            (2, 16, 24, 8, 33, True), (1, 40, 8, 4, 300, True)])
def test_round_trip_and_code_length(n_img, npos, C, cpg, A, with_level):
    rng = np.random.default_rng(n_img * 1000 + npos)
    T = 5 if with_level else C
    cum = R.quantise_pmf(_random_tables(rng, T, A))
    vmin = -(A // 2)
    level = rng.integers(0, T, size=(n_img, npos, C)).astype(np.uint8) if with_level else None
    # draw the symbols from the tables they are coded with (plus a few rare ones), so the length check means something
    t = level.astype(np.int64) if with_level else np.broadcast_to(np.arange(C), (n_img, npos, C))
    u = rng.integers(0, R.TOTAL, size=(n_img, npos, C))
    idx = (cum[t] <= u[..., None]).sum(axis=-1) - 1
    idx.reshape(-1)[::7] = rng.integers(0, A, size=idx.reshape(-1)[::7].shape)
    sym = idx + vmin
    streams = R.encode_latent(sym, level, cum, vmin, cpg)
    assert len(streams) == n_img and all(len(s) == C // cpg for s in streams)
    back = R.decode_latent(streams, level, cum, vmin, cpg, npos, C)
    assert np.array_equal(back, sym)
    bits = 16 * sum(len(w) for s in streams for w in s)
    ideal = R.ideal_bits(sym, level, cum, vmin)
    nstreams = n_img * (C // cpg)
    # rANS: at most the 32-bit final state per stream above the ideal length (and never meaningfully below it)
    assert ideal - 1e-6 <= bits + 16 * nstreams and bits <= ideal + 32 * nstreams + 1e-6, (bits, ideal)


def test_malformed_stream_is_detected():
    cum = R.quantise_pmf(np.array([[0.5, 0.3, 0.2]]))
    sym = np.array([[[0], [1], [2], [0], [0], [1]]])
    words = R.encode_latent(sym, None, np.repeat(cum, 1, 0), 0, 1)[0][0]
    tab = np.zeros(6, dtype=np.int64)
    assert np.array_equal(R.decode_stream(words, tab, cum), sym.reshape(-1))
    bad = list(words)
    bad[1] ^= 0x0101
    with pytest.raises((AssertionError, IndexError)):
        R.decode_stream(bad, tab, cum)
