"""Seeded synthetic inputs of the shapes the reference pipeline emits.

* windows: float64 (N, 60, 4), per-window standardised exactly as ``standardize_per_window``
  (reference data_prepocessing/prepare_numpy_datasets.py:83-95).
* weights: one float32 blob per ensemble member in Keras ``get_weights()`` order
  (seed 2025+m mirrors models/train_deep_ensemble_cnns.py:126).
* patient ids: contiguous runs, ids 200001+i (preprocess_shhs_raw.py:302, :323-325).
"""
from __future__ import annotations

import numpy as np

from . import arch


def standardize_per_window(data_3d: np.ndarray, epsilon: float = 1e-8) -> np.ndarray:
    """(x - mean_t) / (std_t + eps), population std over the time axis
    (prepare_numpy_datasets.py:83-95)."""
    mean = np.mean(data_3d, axis=1, keepdims=True)
    std = np.std(data_3d, axis=1, keepdims=True)
    return (data_3d - mean) / (std + epsilon)


def make_windows(n: int, seed: int = 2025, chunk: int = 1 << 16) -> np.ndarray:
    """AR(1)-smoothed Gaussian windows, 1 % of windows with one constant channel, standardised."""
    out = np.empty((n, arch.WINDOW_LEN, arch.WINDOW_FEATURES), dtype=np.float64)
    rng = np.random.default_rng(seed)
    rho = 0.9
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        raw = rng.standard_normal((e - s, arch.WINDOW_LEN, arch.WINDOW_FEATURES))
        for t in range(1, arch.WINDOW_LEN):
            raw[:, t, :] = rho * raw[:, t - 1, :] + np.sqrt(1 - rho * rho) * raw[:, t, :]
        const = rng.random(e - s) < 0.01
        ch = rng.integers(0, arch.WINDOW_FEATURES, size=e - s)
        idx = np.nonzero(const)[0]
        raw[idx, :, ch[idx]] = raw[idx, :1, ch[idx]]
        out[s:e] = standardize_per_window(raw)
    return out


def make_labels(n: int, seed: int = 2026, p: float = 0.2) -> np.ndarray:
    return (np.random.default_rng(seed).random(n) < p).astype(np.int64)


def make_patient_ids(n_patients: int = 2651, seed: int = 2027, lo: int = 300, hi: int = 660) -> np.ndarray:
    """Contiguous per-patient runs; windows per patient ~ U{lo..hi}; ids 200001+i."""
    rng = np.random.default_rng(seed)
    counts = rng.integers(lo, hi + 1, size=n_patients)
    return np.repeat(np.arange(n_patients, dtype=np.int64) + 200001, counts)


def make_member_weights(member: int = 0, seed_base: int = 2025, logit_std: float = 1.5,
                        n_calib: int = 256):
    """Random-init member that behaves like a trained one: Glorot-uniform kernels, small biases,
    BN gamma~U(0.8,1.2), beta~N(0,0.1), and BN moving statistics set to the batch statistics of a
    seeded calibration set (+-5 % jitter) -- as they would be after training -- so that inference-mode and
    batch-statistic BN see comparable activations.  The dense layer is scaled/shifted so that
    deterministic logits have std ``logit_std`` and mean 0 over the calibration windows; otherwise every
    probability collapses to 0.5.  Returns the 38 float32 arrays in Keras
    ``get_weights()`` order.  (torch-CPU is used only to propagate the calibration windows.)
    """
    import torch
    import torch.nn.functional as F

    rng = np.random.default_rng(seed_base + member)
    ws = []
    for b in arch.BLOCKS:
        fan_in, fan_out = b.k * b.cin, b.k * b.cout
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        ws.append(rng.uniform(-lim, lim, size=(b.k, b.cin, b.cout)).astype(np.float32))
        ws.append((0.05 * rng.standard_normal(b.cout)).astype(np.float32))
        ws.append(rng.uniform(0.8, 1.2, size=b.cout).astype(np.float32))          # gamma
        ws.append((0.1 * rng.standard_normal(b.cout)).astype(np.float32))          # beta
        ws.append(np.zeros(b.cout, dtype=np.float32))                              # moving_mean (set below)
        ws.append(np.ones(b.cout, dtype=np.float32))                               # moving_variance
    lim = np.sqrt(6.0 / (arch.DENSE_IN + 1))
    ws.append(rng.uniform(-lim, lim, size=(arch.DENSE_IN, 1)).astype(np.float32))
    ws.append(np.zeros(1, dtype=np.float32))

    xc = make_windows(n_calib, seed=424242)
    with torch.no_grad():
        a = torch.from_numpy(xc.astype(np.float32)).permute(0, 2, 1)
        for i, b in enumerate(arch.BLOCKS):
            w, bi, g, be = [torch.from_numpy(t) for t in ws[6 * i:6 * i + 4]]
            a = F.relu(F.conv1d(a, w.permute(2, 1, 0).contiguous(), bi, padding=b.halo))
            mu = a.mean(dim=(0, 2))
            var = a.var(dim=(0, 2), unbiased=False)
            ws[6 * i + 4] = (mu.numpy() * (1 + 0.05 * rng.standard_normal(b.cout))).astype(np.float32)
            ws[6 * i + 5] = (var.numpy() * (1 + 0.05 * rng.standard_normal(b.cout)) + 1e-4).astype(np.float32)
            a = (a - mu[None, :, None]) * (g / torch.sqrt(var + arch.BN_EPS))[None, :, None] + be[None, :, None]
        z = (a.mean(dim=2).numpy() @ ws[36])[:, 0]
    gain = logit_std / max(float(z.std()), 1e-6)
    ws[36] = (ws[36] * gain).astype(np.float32)
    ws[37] = np.array([-float(z.mean()) * gain], dtype=np.float32)
    return ws


def weights_to_blob(ws) -> np.ndarray:
    lay = arch.blob_layout()
    if len(ws) != len(lay):
        raise ValueError(f"expected {len(lay)} weight arrays (Keras get_weights order), got {len(ws)}")
    parts = []
    for (name, shp), w in zip(lay, ws):
        w = np.asarray(w, dtype=np.float32)
        if tuple(w.shape) != tuple(shp):
            raise ValueError(f"{name}: expected shape {shp}, got {w.shape}")
        parts.append(w.ravel())
    return np.ascontiguousarray(np.concatenate(parts))


def blob_to_weights(blob: np.ndarray):
    out, o = [], 0
    for _, shp in arch.blob_layout():
        n = int(np.prod(shp))
        out.append(np.asarray(blob[o:o + n], dtype=np.float32).reshape(shp).copy())
        o += n
    return out
