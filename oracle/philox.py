"""Philox4x32-10 and the inverse-CDF multinomial used by the CUDA sampler, restated in numpy
(TEST INFRASTRUCTURE, see oracle/__init__.py).

The reference samples with torch.multinomial on torch's global generator, reseeded by
`setup_seed(0)` on every generate() call (ref stream_generator.py:38-46, 223, 296).  A device
kernel cannot reproduce that stream, so the product defines its own:
counter = (global sequence index, step, 0, 0), key = 64-bit seed; u = (x0 >> 8) * 2^-24;
token = first vocabulary index whose running sum of exp(score - max) exceeds u * total.
This file is the executable definition the CUDA kernel is checked against.
"""
import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(counter, key):
    """counter: 4 python ints (uint32), key: 2 python ints (uint32) -> 4 uint32 ints."""
    c = [np.uint64(x & 0xFFFFFFFF) for x in counter]
    k0, k1 = key[0] & 0xFFFFFFFF, key[1] & 0xFFFFFFFF
    for r in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK
        c = [hi1 ^ c[1] ^ np.uint64(k0), lo1, hi0 ^ c[3] ^ np.uint64(k1), lo0]
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return [int(x) for x in c]


def uniform(seed, seq_index, step):
    """u in [0,1) with 24 bits, fp32-exact."""
    x = philox4x32_10((int(seq_index), int(step), 0, 0), (int(seed) & 0xFFFFFFFF, (int(seed) >> 32) & 0xFFFFFFFF))
    return np.float32(x[0] >> 8) * np.float32(1.0 / 16777216.0)


def multinomial_row(scores, u):
    """scores: fp32 [V] with -inf for removed tokens.  Returns (token, margin) where margin is the
    relative distance of u*Z from the nearest CDF edge (tests skip rows with a tiny margin, where
    fp32 summation order legitimately decides)."""
    s = scores.astype(np.float32)
    m = s.max()
    e = np.exp((s - m).astype(np.float32)).astype(np.float32)
    e[~np.isfinite(s)] = 0.0
    cdf = np.cumsum(e.astype(np.float64))
    z = cdf[-1]
    target = float(u) * z
    tok = int(np.searchsorted(cdf, target, side="right"))
    nz = np.nonzero(e)[0]
    tok = min(tok, int(nz[-1]))
    lo = cdf[tok - 1] if tok > 0 else 0.0
    margin = min(abs(target - lo), abs(cdf[tok] - target)) / z
    return tok, margin


def multinomial_rows(scores, seed, seq_indices, step):
    out = np.zeros(scores.shape[0], dtype=np.int64)
    for r in range(scores.shape[0]):
        out[r], _ = multinomial_row(scores[r], uniform(seed, seq_indices[r], step))
    return out
