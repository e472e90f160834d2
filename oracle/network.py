"""CPU restatement of the 1D-CNN forward pass with Keras 2.12 layer semantics.  TEST INFRASTRUCTURE.

PARITY UNPINNED: the arithmetic lives in TensorFlow/Keras 2.12.0 (reference README.md:79), un-vendored and
not installable here; the reference ships no golden vectors.  This file restates the published layer
semantics for the layer list at reference models/cnn_baseline_train.py:55-94 and the
call semantics of uncertainty_quantification/uq_techniques.py:22 (``model(x, training=True)``: whole array is
one batch, BatchNormalization uses batch statistics, Dropout is active) and :29 (``model.predict``:
inference mode).

Implementations of the same function:
  * ``forward_f64``  -- NumPy float64, ground truth for tolerances and the label guard band.
  * ``forward_f64_torch`` -- the same arithmetic in torch float64 (multi-threaded), for the parity tests at the sizes
    the bench times (N >= 1024 windows x T = 50 passes); pinned to ``forward_f64`` by tests/test_oracle_network.py.
  * ``forward_f32_torch`` -- torch-CPU float32, the timed CPU baseline (bench.py cpu_baseline leg).

Weights are the 38 arrays of Keras ``get_weights()`` order:
per block [kernel (k,Cin,Cout), bias, gamma, beta, moving_mean, moving_variance] x6, dense kernel (96,1), bias.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np

# (Cin, Cout, k, rate): cnn_baseline_train.py:59-61, 63-65, 68-70, 73-75, 78-80, 83-86
BLOCKS = ((4, 128, 7, 0.3), (128, 192, 5, 0.3), (192, 224, 3, 0.4),
          (224, 96, 7, 0.2), (96, 256, 9, 0.3), (256, 96, 9, 0.5))
BN_EPS = 1e-3   # Keras BatchNormalization default


def _conv1d_same(x: np.ndarray, w: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Keras Conv1D(padding='same', strides=1): y[n,t,o] = b[o] + sum_j sum_c x[n,t+j-h,c] W[j,c,o],
    out-of-range x = 0, h=(k-1)//2  (cnn_baseline_train.py:59)."""
    n, L, cin = x.shape
    k, _, cout = w.shape
    h = (k - 1) // 2
    xp = np.zeros((n, L + k - 1, cin), dtype=x.dtype)
    xp[:, h:h + L, :] = x
    # one GEMM over the unfolded taps: cols[n, t, j*cin + c] = xp[n, t + j, c]
    cols = np.lib.stride_tricks.sliding_window_view(xp, k, axis=1)           # (n, L, cin, k)
    cols = np.ascontiguousarray(cols.transpose(0, 1, 3, 2)).reshape(n * L, k * cin)
    return (cols @ w.reshape(k * cin, cout)).reshape(n, L, cout) + b


def forward_f64(weights: Sequence[np.ndarray], x: np.ndarray, training: bool = False,
                masks: Optional[List[np.ndarray]] = None, bn_mode: Optional[str] = None,
                bn_stats: Optional[List] = None, return_stats: bool = False, return_acts: bool = False,
                stats_exchange=None, global_windows: Optional[int] = None):
    """One forward pass of all N windows as ONE batch.

    training=False : BN uses moving statistics, dropout is the identity (uq_techniques.py:29,
                     analyze_mcd_patient_level.py:204).
    training=True  : dropout uses ``masks`` (list of six boolean keep-masks (N,60,Cout)); BN uses
                     batch mean / biased variance over axes (N, L) (uq_techniques.py:22) unless
                     ``bn_mode='inference'`` (frozen moving stats) or ``bn_stats`` (list of six
                     (mean, var) pairs, the calibrated mode) overrides it.
    stats_exchange : host model of windows sharded over ranks (exact batch statistics): a callable
                     ``(sum, sum_sq) -> (sum, sum_sq)`` over all ranks; ``x`` is then this rank's shard and
                     ``global_windows`` the window count of the whole call.
    Returns probabilities (N, 1) float64 (and the per-layer batch (mean, var) if return_stats).
    """
    a = np.asarray(x, dtype=np.float64)
    if bn_mode is None:
        bn_mode = "batch" if training else "inference"
    stats_out = []
    acts_out = []
    for i, (cin, cout, k, rate) in enumerate(BLOCKS):
        w, b, g, be, mm, mv = [np.asarray(t, dtype=np.float64) for t in weights[6 * i:6 * i + 6]]
        a = np.maximum(_conv1d_same(a, w, b), 0.0)                       # Conv1D + ReLU
        if bn_stats is not None:
            mu, var = bn_stats[i]
        elif bn_mode == "batch" and stats_exchange is not None:
            ssum, ssq = stats_exchange(a.sum(axis=(0, 1)), (a * a).sum(axis=(0, 1)))
            cnt = float(global_windows) * a.shape[1]
            mu = ssum / cnt
            var = np.maximum(ssq / cnt - mu * mu, 0.0)                    # biased, from the exchanged moments
        elif bn_mode == "batch":
            mu = a.mean(axis=(0, 1))
            var = a.var(axis=(0, 1))                                      # biased
        else:
            mu, var = mm, mv
        stats_out.append((np.asarray(mu, dtype=np.float64), np.asarray(var, dtype=np.float64)))
        a = g * (a - mu) / np.sqrt(var + BN_EPS) + be                    # BatchNormalization
        if training and masks is not None:
            a = a * masks[i] / (1.0 - rate)                               # Dropout: keep iff u >= rate
        acts_out.append(a)
    gap = a.mean(axis=1)                                                  # GlobalAveragePooling1D
    z = gap @ np.asarray(weights[36], dtype=np.float64) + np.asarray(weights[37], dtype=np.float64)
    p = 1.0 / (1.0 + np.exp(-z))                                          # Dense(1, sigmoid)
    if return_acts:
        return p, acts_out
    if return_stats:
        return p, stats_out
    return p


def forward_f64_torch(weights: Sequence[np.ndarray], x: np.ndarray, training: bool = False,
                      masks: Optional[List[np.ndarray]] = None, bn_mode: Optional[str] = None,
                      bn_stats: Optional[List] = None) -> np.ndarray:
    """``forward_f64`` with torch float64 tensors (same semantics and argument meaning, no exchange model)."""
    import torch
    import torch.nn.functional as F

    if bn_mode is None:
        bn_mode = "batch" if training else "inference"
    with torch.no_grad():
        a = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)).permute(0, 2, 1)      # (N, C, L)
        tw = [torch.from_numpy(np.ascontiguousarray(w, dtype=np.float64)) for w in weights]
        for i, (cin, cout, k, rate) in enumerate(BLOCKS):
            w, b, g, be, mm, mv = tw[6 * i:6 * i + 6]
            a = F.relu(F.conv1d(a, w.permute(2, 1, 0).contiguous(), b, padding=(k - 1) // 2))
            if bn_stats is not None:
                mu, var = [torch.as_tensor(np.asarray(t, dtype=np.float64)) for t in bn_stats[i]]
            elif bn_mode == "batch":
                mu = a.mean(dim=(0, 2))
                var = a.var(dim=(0, 2), unbiased=False)
            else:
                mu, var = mm, mv
            a = (g / torch.sqrt(var + BN_EPS))[None, :, None] * (a - mu[None, :, None]) + be[None, :, None]
            if training and masks is not None:
                m = torch.from_numpy(np.ascontiguousarray(masks[i])).permute(0, 2, 1)
                a = a * m / (1.0 - rate)
        z = a.mean(dim=2) @ tw[36] + tw[37]
        return torch.sigmoid(z).numpy()


def forward_f32_torch(weights: Sequence[np.ndarray], x: np.ndarray, training: bool = False,
                      masks: Optional[List[np.ndarray]] = None, rng_seed: Optional[int] = None,
                      batch_size: Optional[int] = None) -> np.ndarray:
    """float32 torch-CPU restatement (the timed CPU baseline).  ``training=True`` with ``masks=None``
    draws Bernoulli keep-masks from torch's generator (statistically like Keras Dropout).
    ``batch_size`` splits inference-mode calls like Keras ``predict`` (default 32 there); training-mode
    calls always use the whole array as one batch, as the reference does."""
    import torch
    import torch.nn.functional as F

    gen = None
    if rng_seed is not None:
        gen = torch.Generator().manual_seed(rng_seed)
    xt = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).permute(0, 2, 1)  # (N, C, L)
    tw = [torch.from_numpy(np.ascontiguousarray(w, dtype=np.float32)) for w in weights]

    def run(a):
        for i, (cin, cout, k, rate) in enumerate(BLOCKS):
            w, b, g, be, mm, mv = tw[6 * i:6 * i + 6]
            a = F.relu(F.conv1d(a, w.permute(2, 1, 0).contiguous(), b, padding=(k - 1) // 2))
            if training:
                mu = a.mean(dim=(0, 2))
                var = a.var(dim=(0, 2), unbiased=False)
            else:
                mu, var = mm, mv
            a = (a - mu[None, :, None]) * (g / torch.sqrt(var + BN_EPS))[None, :, None] + be[None, :, None]
            if training:
                if masks is not None:
                    m = torch.from_numpy(np.ascontiguousarray(masks[i])).permute(0, 2, 1).to(a.dtype)
                else:
                    m = (torch.rand(a.shape, generator=gen) >= rate).to(a.dtype)
                a = a * m / (1.0 - rate)
        gap = a.mean(dim=2)
        z = gap @ tw[36] + tw[37]
        return torch.sigmoid(z)

    with torch.no_grad():
        if training or batch_size is None:
            out = run(xt)
        else:
            out = torch.cat([run(xt[s:s + batch_size]) for s in range(0, xt.shape[0], batch_size)])
    return out.numpy()
