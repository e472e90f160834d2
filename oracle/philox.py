"""CPU twin of the in-kernel counter-based RNG (Philox4x32-10, Salmon et al., SC'11 / Random123).

TEST INFRASTRUCTURE.  The reference draws dropout masks from TensorFlow's stateful RNG
(Keras ``Dropout`` under ``training=True``, reference uq_techniques.py:22), which cannot be reproduced
outside TensorFlow; the build's kernels use Philox4x32-10 keyed by (seed) and counted by
(window, pass, layer, row, channel-chunk).  This module regenerates exactly the same keep-masks on the CPU
so that the oracle network can be driven by the masks the kernels use ("masks exported from the
reference" in BASELINE.json is realised as masks exported from this twin).

Counter mapping (must match csrc/philox.cuh):
    c0 = global window index (low 32 bits)
    c1 = pass index t
    c2 = (layer << 16) | row          layer in 0..5, row in 0..59
    c3 = channel chunk q              (8 consecutive channels share one Philox call)
    key = (seed & 0xffffffff, seed >> 32)
    element e in 0..7 of the chunk uses the 15-bit field  (out[e >> 1] >> (16 * (e & 1))) & 0x7fff
    keep  <=>  field >= round(rate * 32768)      (TF: keep iff u >= rate; kept values scaled by 1/(1-rate))
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


def dropout_keep_mask(seed: int, window_idx, t: int, layer: int, cout: int, rate: float,
                      rows: int = 60) -> np.ndarray:
    """Boolean keep-mask of shape (len(window_idx), rows, cout) for pass ``t`` and ``layer``."""
    window_idx = np.asarray(window_idx, dtype=np.uint64)
    n = window_idx.shape[0]
    nq = cout // 8
    thr = int(round(rate * 32768.0))
    c0 = (window_idx & MASK32).reshape(n, 1, 1)
    c1 = np.uint64(t)
    c2 = (np.uint64(layer << 16) | np.arange(rows, dtype=np.uint64)).reshape(1, rows, 1)
    c3 = np.arange(nq, dtype=np.uint64).reshape(1, 1, nq)
    out = philox4x32(c0, c1, c2, c3, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    words = np.stack(out, axis=-1)                                  # (n, rows, nq, 4)
    lo = words & np.uint32(0x7FFF)
    hi = (words >> np.uint32(16)) & np.uint32(0x7FFF)
    fields = np.stack([lo, hi], axis=-1).reshape(n, rows, nq * 8)  # e = 2*word + half
    return fields >= thr


def pack_mask_bits(keep: np.ndarray) -> np.ndarray:
    """(..., C) bool -> (..., 8) uint32 words, bit (c % 32) of word (c // 32); C <= 256, zero padded."""
    c = keep.shape[-1]
    pad = np.zeros(keep.shape[:-1] + (256,), dtype=np.uint64)
    pad[..., :c] = keep
    bits = pad.reshape(keep.shape[:-1] + (8, 32))
    w = (bits << np.arange(32, dtype=np.uint64)).sum(axis=-1)
    return w.astype(np.uint32)
