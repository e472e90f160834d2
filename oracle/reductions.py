"""NumPy restatement of the reference's per-window UQ reductions.  TEST INFRASTRUCTURE.

Pinned: every function here is checked in tests/test_oracle_reductions.py against tests/golden/*.npz, which
were produced by the reference's own code (uncertainty_quantification/uq_techniques.py imported from
/root/reference with tensorflow/matplotlib stubbed -- see oracle/reference_loader.py and
tests/golden/make_golden.py).
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np


def safe_entropy(probs: np.ndarray, axis: int = 1, epsilon: float = 1e-10) -> np.ndarray:
    """uq_techniques.py:35-38.  scipy.stats.entropy(clip(p)) = renormalise along axis, then -sum p ln p."""
    q = np.clip(probs, epsilon, 1 - epsilon)
    q = 1.0 * q / np.sum(q, axis=axis, keepdims=True)
    return np.sum(-q * np.log(q), axis=axis)


def uq_evaluation_dist(uq_predictions: np.ndarray, y_true: Optional[np.ndarray]) -> Dict[str, np.ndarray]:
    """uq_techniques.py:40-112 (same keys, same dtypes: float32 in -> float32 out)."""
    predictions = np.squeeze(uq_predictions)                         # :62
    if predictions.ndim == 1:
        predictions = predictions.reshape(1, -1)                     # :63-64
    mean_pred = np.mean(predictions, axis=0)                         # :69
    pred_variance = np.var(predictions, axis=0)                      # :72
    mean_probs = np.stack([1 - mean_pred, mean_pred], axis=-1)       # :78
    total = safe_entropy(mean_probs, axis=1)                         # :79
    ents = [safe_entropy(np.stack([1 - p, p], axis=-1), axis=1) for p in predictions]   # :83-86
    expected = np.mean(ents, axis=0)                                 # :87
    mi = np.maximum(total - expected, 0)                             # :91
    out = {"mean_pred": mean_pred, "pred_variance": pred_variance, "total_pred_entropy": total,
           "expected_aleatoric_entropy": expected, "mutual_info": mi,
           "overall_mean_variance": np.mean(pred_variance)}          # :94
    if y_true is not None:
        c0, c1 = (y_true == 0), (y_true == 1)                        # :97-98
        out["mean_variance_class_0"] = np.mean(pred_variance[c0]) if np.any(c0) else 0.0   # :100
        out["mean_variance_class_1"] = np.mean(pred_variance[c1]) if np.any(c1) else 0.0   # :101
    return out


def window_csv_metrics(probs: np.ndarray):
    """Inline per-window metrics of the drivers (analyze_mcd_patient_level.py:109-117,
    analyze_de_patient_level.py:124-132): mean, variance, log2 entropy with +1e-9, label = mean > 0.5."""
    mean_p = np.mean(probs, axis=0).flatten()
    var_p = np.var(probs, axis=0).flatten()
    eps = 1e-9
    ent = -(mean_p * np.log2(mean_p + eps) + (1 - mean_p) * np.log2(1 - mean_p + eps))
    label = (mean_p > 0.5).astype(int)
    return mean_p, var_p, ent, label


def bootstrap_aggregates_via_gather(metrics: Dict[str, np.ndarray], y_true: np.ndarray, n_bootstrap: int,
                                    random_state: Optional[int]):
    """Algebraic shortcut for uq_techniques.py:116-172: the per-window metrics of ``preds[:, idx]`` are
    ``metrics[idx]``, so each bootstrap iteration is a gather-mean.  Index stream reproduced with the
    legacy NumPy API exactly as the reference (:130-131, :142)."""
    if random_state is not None:
        np.random.seed(random_state)
    n = len(y_true)
    res = []
    for _ in range(n_bootstrap):
        idx = np.random.choice(n, n, replace=True)
        v = metrics["pred_variance"][idx]
        yb = y_true[idx]
        c0, c1 = (yb == 0), (yb == 1)
        res.append({
            "overall_mean_variance": np.mean(v),
            "mean_variance_class_0": np.mean(v[c0]) if np.any(c0) else 0.0,
            "mean_variance_class_1": np.mean(v[c1]) if np.any(c1) else 0.0,
            "mean_total_pred_entropy": np.mean(metrics["total_pred_entropy"][idx]),
            "mean_expected_aleatoric_entropy": np.mean(metrics["expected_aleatoric_entropy"][idx]),
            "mean_mutual_info": np.mean(metrics["mutual_info"][idx]),
        })
    return res
