"""CPU twin of the in-kernel counter-based RNG (Philox4x32-10, Salmon et al., SC'11 / Random123).

TEST INFRASTRUCTURE.  The reference draws dropout masks from TensorFlow's stateful RNG
(Keras ``Dropout`` under ``training=True``, reference uq_techniques.py:22), which cannot be reproduced
outside TensorFlow; the build's kernels use Philox4x32-10 keyed by (seed) and counted by
(window, pass, layer, row, channel-chunk).  This module regenerates exactly the same keep-masks on the CPU
so that the oracle network can be driven by the masks the kernels use ("masks exported from the
reference" in BASELINE.json is realised as masks exported from this twin).

Keep-flag definition (must match csrc/common.cuh, "dropout keep-flags"); key = (seed & 0xffffffff, seed >> 32):
    The flags of (window, pass t, layer, row) are generated bit-sliced, 32 elements per 32-bit word.  Channel c belongs
    to column half h = (c >> 4) & 1, flag word i = c >> 6, bit 8*(c & 3) + 4*((c >> 5) & 1) + ((c >> 2) & 3).
    * rate 0.5 (layer 6): flag word i = output word i of ONE Philox call with counter
      (window, t, layer << 16 | row, h): one random bit per element, keep iff the bit is 1.
    * other rates: U = (hi8 << 8) | lo8 is the element's 16-bit uniform, keep <=> U >= round(rate * 65536)
      (TF: keep iff u >= rate; kept values scaled by 1/(1-rate)).  Bit k of hi8 is the element's bit of plane word P_k;
      P_0..P_7 are the 2 x 4 output words of the calls with counter (window, t, layer << 16 | row, h | i << 1 | q << 3),
      q = 0, 1.  lo8 is a tie-break byte shared by the elements of (row, h, layer): byte `layer` of output words (x, y)
      of the call with counter (window, t, 6 << 16 | row, h).
"""
from __future__ import annotations

import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0: int, k1: int, rounds: int = 10):
    """Vectorised Philox4x32-R. Counter words are broadcastable uint32 arrays; key words are ints."""
    c0 = np.asarray(c0, dtype=np.uint64)
    c1 = np.asarray(c1, dtype=np.uint64)
    c2 = np.asarray(c2, dtype=np.uint64)
    c3 = np.asarray(c3, dtype=np.uint64)
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 &= 0xFFFFFFFF
    k1 &= 0xFFFFFFFF
    for _ in range(rounds):
        p0 = M0 * c0
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
        n0 = hi1 ^ c1 ^ np.uint64(k0)
        n2 = hi0 ^ c3 ^ np.uint64(k1)
        c0, c1, c2, c3 = n0, lo1, n2, lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def drop_bitpos(c):
    c = np.asarray(c)
    return 8 * (c & 3) + 4 * ((c >> 5) & 1) + ((c >> 2) & 3)


def dropout_keep_mask(seed: int, window_idx, t: int, layer: int, cout: int, rate: float,
                      rows: int = 60) -> np.ndarray:
    """Boolean keep-mask of shape (len(window_idx), rows, cout) for pass ``t`` and ``layer``."""
    window_idx = np.asarray(window_idx, dtype=np.uint64)
    n = window_idx.shape[0]
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    thr16 = int(round(rate * 65536.0))
    nw = (cout + 63) // 64
    c0 = (window_idx & MASK32).reshape(n, 1, 1)
    c1 = np.uint64(t)
    rws = np.arange(rows, dtype=np.uint64).reshape(1, rows, 1)
    c2 = np.uint64(layer << 16) | rws
    hh = np.arange(2, dtype=np.uint64).reshape(1, 1, 2)
    words = np.zeros((n, rows, 2, nw), dtype=np.uint32)              # [window, row, h, i]
    if thr16 == 32768:
        out = philox4x32(c0, c1, c2, hh, k0, k1)
        for i in range(nw):
            words[..., i] = out[i]
    else:
        tie = philox4x32(c0, c1, np.uint64(6 << 16) | rws, hh, k0, k1)
        tb = ((tie[0] >> np.uint32(8 * layer)) if layer < 4 else tie[1]) & np.uint32(0xFF)
        g0 = np.where(tb >= (thr16 & 0xFF), np.uint32(0xFFFFFFFF), np.uint32(0))
        thi = thr16 >> 8
        for i in range(nw):
            pa = philox4x32(c0, c1, c2, hh | np.uint64(i << 1), k0, k1)
            pb = philox4x32(c0, c1, c2, hh | np.uint64((i << 1) | 8), k0, k1)
            g = g0.copy()
            for k, pl in enumerate(list(pa) + list(pb)):
                g = (pl & g) if (thi >> k) & 1 else (pl | g)
            words[..., i] = g
    c = np.arange(cout)
    sel = words[:, :, (c >> 4) & 1, c >> 6]                          # (n, rows, cout)
    return ((sel >> drop_bitpos(c).astype(np.uint32)) & np.uint32(1)).astype(bool)


def pack_mask_bits(keep: np.ndarray) -> np.ndarray:
    """(..., C) bool -> (..., 8) uint32 words, bit (c % 32) of word (c // 32); C <= 256, zero padded."""
    c = keep.shape[-1]
    pad = np.zeros(keep.shape[:-1] + (256,), dtype=np.uint64)
    pad[..., :c] = keep
    bits = pad.reshape(keep.shape[:-1] + (8, 32))
    w = (bits << np.arange(32, dtype=np.uint64)).sum(axis=-1)
    return w.astype(np.uint32)
