"""CPU restatement (numpy / plain Python) of the entropy coder behind compress / decompress -- TEST INFRASTRUCTURE ONLY:
imported by tests/ and nothing else; the product path is csrc/rans_kernels.cu + image_compression_b200/codec.py.

PARITY UNPINNED: the reference has no coder.  `compress` / `decompress` are `pass` (modelling/meta_arch/bmshl2018.py:
106-110), the entropy models raise NotImplementedError (modelling/blocks/entropy_model.py:157-169, 241-251, "symbol"
mode :235-236) and `_standardized_quantile` (:364-365, :377-378) is never called.  What IS pinned by the reference is
the probability model the tables are cut from: the factorised prior `_prob_mass` (:259-269) and the conditional
`_prob_mass` (:341-352), restated in oracle/hyperprior_oracle.py.  This file restates the FORMAT (table quantisation,
scale levels, rANS recurrences, stream layout) so that the GPU coder can be checked word for word.

rANS (Duda 2013), 32-bit state, 16-bit words, 16-bit probabilities; see csrc/rans_kernels.cu for the layout.
"""
import math

import numpy as np

PREC = 16
TOTAL = 1 << PREC
RANS_L = 1 << 16
SCALE_MIN, SCALE_MAX, SCALE_LEVELS = 0.11, 256.0, 64


def quantise_pmf(p):
    """(T, A) non-negative floats -> (T, A+1) uint32 cumulative frequencies: every symbol gets freq >= 1, the sum is 2^16;
    freq = 1 + floor(p / sum(p) * (2^16 - A)), the remainder goes to the most probable symbol (first of equals)."""
    p = np.asarray(p, dtype=np.float64)
    T, A = p.shape
    if A > TOTAL - 1:
        raise ValueError(f"alphabet of {A} symbols does not fit 16-bit probabilities")
    p = np.where(np.isfinite(p) & (p > 0), p, 0.0)
    s = p.sum(axis=1, keepdims=True)
    pn = np.where(s > 0, p / np.where(s > 0, s, 1.0), 1.0 / A)
    f = 1 + np.floor(pn * (TOTAL - A)).astype(np.int64)
    rest = TOTAL - f.sum(axis=1)
    f[np.arange(T), np.argmax(pn, axis=1)] += rest
    assert (f >= 1).all() and (f.sum(axis=1) == TOTAL).all()
    cum = np.zeros((T, A + 1), dtype=np.uint32)
    cum[:, 1:] = np.cumsum(f, axis=1)
    return cum


def scale_table(nlevels=SCALE_LEVELS, smin=SCALE_MIN, smax=SCALE_MAX):
    """The nlevels scales of the conditional coder, fp32 like the kernel: smin * exp(l * step)."""
    log_smin = np.float32(math.log(np.float32(smin)))
    step = (np.float32(math.log(np.float32(smax))) - log_smin) / np.float32(nlevels - 1)
    return np.exp(np.float64(log_smin) + np.arange(nlevels) * np.float64(step))


def scale_levels(sigma, nlevels=SCALE_LEVELS, smin=SCALE_MIN, smax=SCALE_MAX):
    """hpc_scale_levels in fp32: ceil((log sigma - log smin) / step) clamped to [0, nlevels-1]."""
    sigma = np.asarray(sigma, dtype=np.float32)
    log_smin = np.float32(math.log(np.float32(smin)))
    step = (np.float32(math.log(np.float32(smax))) - log_smin) / np.float32(nlevels - 1)
    inv = np.float32(1.0) / step
    l = np.ceil((np.log(sigma).astype(np.float32) - log_smin) * inv)
    return np.clip(l, 0, nlevels - 1).astype(np.uint8)


def conditional_pmf(scales, vmin, vmax, dist="laplace"):
    """(len(scales), vmax-vmin+1): entropy_model.py:341-352 with mean 0, bin 1, at the table scales, in float64."""
    k = np.abs(np.arange(vmin, vmax + 1, dtype=np.float64))[None, :]
    s = np.asarray(scales, dtype=np.float64)[:, None]

    def cdf(t):
        if dist == "laplace":       # torch.distributions.Laplace(0,1).cdf: 0.5 - 0.5 sign(t) expm1(-|t|)
            return 0.5 - 0.5 * np.sign(t) * np.expm1(-np.abs(t))
        return 0.5 * (1.0 + np.vectorize(math.erf)(t / math.sqrt(2.0)))
    return cdf((0.5 - k) / s) - cdf((-0.5 - k) / s)


def encode_stream(idx, tab, cum):
    """idx, tab: 1-D int arrays (symbol index, table index) in decoding order -> list of 16-bit words, decoding order."""
    x = RANS_L
    out = []
    for i in range(len(idx) - 1, -1, -1):
        start = int(cum[tab[i], idx[i]])
        freq = int(cum[tab[i], idx[i] + 1]) - start
        if x >= (freq << 16):
            out.append(x & 0xffff)
            x >>= 16
        x = ((x // freq) << 16) + (x % freq) + start
    out.append(x & 0xffff)
    out.append(x >> 16)
    return out[::-1]


def decode_stream(words, tab, cum):
    x = (int(words[0]) << 16) | int(words[1])
    r = 2
    out = np.empty(len(tab), dtype=np.int64)
    for i in range(len(tab)):
        ct = cum[tab[i]]
        slot = x & 0xffff
        j = int(np.searchsorted(ct, slot, side="right")) - 1
        start, freq = int(ct[j]), int(ct[j + 1]) - int(ct[j])
        x = freq * (x >> 16) + slot - start
        if x < RANS_L:
            x = (x << 16) | int(words[r])
            r += 1
        out[i] = j
    assert r == len(words) and x == RANS_L, "malformed stream"
    return out


def encode_latent(sym, level, cum, vmin, cpg):
    """sym: (n_img, npos, C) integer values; level: same shape uint8 or None (table = channel).
    Returns [[words of stream s] for s in range(C // cpg)] per image."""
    n_img, npos, C = sym.shape
    res = []
    for n in range(n_img):
        streams = []
        for s in range(C // cpg):
            v = sym[n, :, s * cpg:(s + 1) * cpg].reshape(-1).astype(np.int64) - vmin
            if level is None:
                t = np.tile(np.arange(s * cpg, (s + 1) * cpg), npos)
            else:
                t = level[n, :, s * cpg:(s + 1) * cpg].reshape(-1).astype(np.int64)
            streams.append(encode_stream(v, t, cum))
        res.append(streams)
    return res


def decode_latent(streams, level, cum, vmin, cpg, npos, C):
    n_img = len(streams)
    out = np.empty((n_img, npos, C), dtype=np.int64)
    for n in range(n_img):
        for s in range(C // cpg):
            if level is None:
                t = np.tile(np.arange(s * cpg, (s + 1) * cpg), npos)
            else:
                t = level[n, :, s * cpg:(s + 1) * cpg].reshape(-1).astype(np.int64)
            out[n, :, s * cpg:(s + 1) * cpg] = (decode_stream(streams[n][s], t, cum) + vmin).reshape(npos, cpg)
    return out


def ideal_bits(sym, level, cum, vmin):
    """-sum log2(freq / 2^16) of the quantised tables: what an ideal coder spends with them."""
    n_img, npos, C = sym.shape
    idx = sym.astype(np.int64) - vmin
    t = np.broadcast_to(np.arange(C), sym.shape) if level is None else level.astype(np.int64)
    f = cum[t, idx + 1].astype(np.float64) - cum[t, idx].astype(np.float64)
    return float(-(np.log2(f / TOTAL)).sum())
