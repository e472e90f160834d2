"""Drop-in replacement for the reference module ``uncertainty_quantification/uq_techniques.py``.

Same ten public names, signatures, return types, dict keys and stdout markers as the reference
(uq_techniques.py:12, 26, 35, 40, 116, 175, 210, 242, 258, 278), so the reference drivers
``analyze_mcd_patient_level.py`` / ``analyze_de_patient_level.py`` / ``evaluate_*_global.py`` run unchanged with
this file's directory first on ``sys.path`` (they do ``from uq_techniques import ...``).  The arithmetic runs
in libuqk.so on a B200; nothing here falls back to the CPU for the forward passes or the per-window
reductions.

Build-side knobs (the callers' positional signatures are frozen, so these are environment variables):
  UQK_BN_MODE    auto | batch | calibrated | inference      BatchNorm statistics of the MC-Dropout passes
                 auto = batch = exact Keras ``training=True`` semantics (statistics of the whole call): fp32
                 layer-synchronous kernels up to UQK_EXACT_MAX_WINDOWS windows, above that the bf16 tensor-core
                 kernel swept once per layer (same semantics, 1.9x the flops of one run).
                 calibrated = statistics of one exact pass over the call (or over a strided subset of
                 UQK_CALIB_WINDOWS windows), frozen for all passes (fastest)
  UQK_PRECISION  bf16 | fp32   arithmetic of the compat entry points.  Unset: MC-Dropout as described under UQK_BN_MODE;
                 deep_ensembles_predict, model(x, training=False) and model.predict run in fp32 like the reference
                 (one to M passes are cheap).  bf16 = tensor-core kernel + fp32 re-evaluation of the windows whose mean
                 probability lies within 5e-3 of 0.5, so that `mean > 0.5` labels are those of the fp32 path.
  UQK_SEED       fixed Philox key of the dropout masks.  Unset: the key is derived from the seed given to
                 tf.random.set_seed (else 2025) AND a per-process call counter, so that -- as with TensorFlow's
                 stateful generator (uq_techniques.py:22) -- every call draws fresh masks.
"""
from __future__ import annotations

import importlib
import os
import sys
import time
from typing import Dict, List, Optional

import numpy as np

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ROOT = os.path.dirname(_PKG_DIR)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_pkg = importlib.import_module(os.path.basename(_PKG_DIR))
_engine = importlib.import_module(_pkg.__name__ + ".engine")
_model = importlib.import_module(_pkg.__name__ + ".model")

AGG_KEYS = ("overall_mean_variance", "mean_variance_class_0", "mean_variance_class_1",
            "mean_total_pred_entropy", "mean_expected_aleatoric_entropy", "mean_mutual_info")


def _env(name: str, default: str) -> str:
    return os.environ.get(name, default)


def _seed() -> int:
    s = os.environ.get("UQK_SEED")
    return int(s) if s is not None else int(_model.UQModel._global_seed)


_calls = {"base": None, "n": 0}


def _call_seed() -> int:
    """Philox key of ONE stochastic call: UQK_SEED when set (reproducible runs), else the global seed folded with a
    call counter that restarts whenever the global seed changes (tf.random.set_seed)."""
    s = os.environ.get("UQK_SEED")
    if s is not None:
        return int(s)
    base = (int(_model.UQModel._global_seed), int(_model.UQModel._seed_epoch))
    if _calls["base"] != base:
        _calls["base"], _calls["n"] = base, 0
    base = base[0]
    k = (base * 0x9E3779B97F4A7C15 + 0xD1B54A32D192ED03 * _calls["n"]) & (2 ** 64 - 1)
    _calls["n"] += 1
    return k


def _members_key(blobs):
    return tuple(_model.fingerprint(b) for b in blobs)


# ===================== forward passes =====================
def mc_dropout_predict(model, x_test_data: np.ndarray, n_pred: int = 50) -> np.ndarray:
    """T stochastic forward passes of ``model`` over all windows -> (n_pred, N, 1) float32
    (reference uq_techniques.py:12-24: ``np.stack([model(x, training=True) for _ in range(n_pred)])``)."""
    t0 = time.time()
    eng = _engine.get_engine()
    blob = _model.weights_of(model)
    eng.ensure_members(lambda: blob[None], key=("mcd",) + _members_key([blob]))
    x = np.asarray(x_test_data)
    n = x.shape[0]
    mode = _env("UQK_BN_MODE", "auto")
    prec = _env("UQK_PRECISION", "bf16")
    if mode == "auto":
        mode = "batch"
    if mode == "batch":
        # exact batch statistics couple all windows of a layer: fp32 layer-synchronous kernels while the whole call's
        # activations fit comfortably, the swept tensor-core kernel above
        if n <= int(_env("UQK_EXACT_MAX_WINDOWS", "16384")) and "UQK_PRECISION" not in os.environ:
            prec = "fp32"
    seed = _call_seed()
    if mode == "calibrated" and not eng.calibrated:
        eng.calibrate(x, n_windows=int(_env("UQK_CALIB_WINDOWS", "0")) or None, n_passes=1, seed=seed)
    # bf16 passes: the windows whose mean probability is within the guard band of 0.5 are re-evaluated in fp32, so the
    # labels the drivers derive from the returned tensor (analyze_mcd_patient_level.py:109-117) are the fp32 path's
    out = eng.mcd(x, n_passes=int(n_pred), seed=seed, bn_mode=mode, precision=prec, return_probs=True,
                  metrics=("mean_pred",), exact_labels=True)
    mc_predictions = out["probs"].reshape(int(n_pred), n, 1)
    print(f"MC Dropout completed in {time.time()-t0:.1f}s ({n_pred} passes)")
    return mc_predictions


def deep_ensembles_predict(ensemble_models: List, x_test_data: np.ndarray) -> np.ndarray:
    """One inference-mode forward per ensemble member -> (M, N, 1) float32 (reference uq_techniques.py:26-32)."""
    t0 = time.time()
    eng = _engine.get_engine()
    models = list(ensemble_models)
    blobs = [_model.weights_of(m) for m in models]
    eng.ensure_members(lambda: np.stack(blobs), key=("de",) + _members_key(blobs))
    x = np.asarray(x_test_data)
    out = eng.de(x, n_members=len(models), precision=_env("UQK_PRECISION", "fp32"), return_probs=True,
                 metrics=("mean_pred",), exact_labels=True)
    de_predictions = out["probs"].reshape(len(models), x.shape[0], 1)
    print(f"Deep Ensemble completed in {time.time()-t0:.1f}s ({len(models)} models)")
    return de_predictions


# ===================== fused entry points (additive; no counterpart in the reference) =====================
def _finish_uq(out: Dict, y_true) -> Dict:
    res = {k: out[k] for k in ("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy",
                               "mutual_info") if k in out}
    res["predictive_entropy_bits"] = out.get("entropy_bits")      # CSV column Predictive_Entropy (log2 form)
    res["predicted_label"] = out.get("label")                     # CSV column Predicted_Label (mean > 0.5)
    if "probs" in out:
        res["probs"] = out["probs"]
    if y_true is not None and "pred_variance" in out and not hasattr(out["pred_variance"], "is_cuda"):
        y = np.asarray(y_true)
        var = out["pred_variance"]
        is0, is1 = (y == 0), (y == 1)
        res["overall_mean_variance"] = np.mean(var)
        res["mean_variance_class_0"] = np.mean(var[is0]) if is0.any() else 0.0
        res["mean_variance_class_1"] = np.mean(var[is1]) if is1.any() else 0.0
    return res


def mc_dropout_uq(model, x_test_data, n_pred: int = 50, *, y_true=None, return_probs: bool = False,
                  bn_mode: str = "calibrated", precision: str = "bf16", seed: Optional[int] = None, masks=None,
                  calib_windows: Optional[int] = None, calib_passes: int = 1, recalibrate: bool = True,
                  device_out: bool = False, window_base: int = 0, exact_labels: bool = False,
                  guard_delta: Optional[float] = None) -> Dict:
    """mc_dropout_predict + uq_evaluation_dist + the drivers' inline CSV metrics in ONE fused call: the
    (T,N) probability tensor is never written to HBM unless ``return_probs``.  Returns the per-window arrays of
    uq_techniques.py:103-108 plus ``predictive_entropy_bits`` / ``predicted_label``
    (analyze_mcd_patient_level.py:113-117).  ``bn_mode='calibrated'`` first measures the batch statistics with
    ``calib_passes`` exact pass(es) over all windows of the call (or a strided subset of ``calib_windows``), inside this
    call when ``recalibrate``, and freezes them for the T passes;
    ``bn_mode='batch'`` is the exact Keras ``training=True`` behaviour (statistics of all windows of this call):
    ``precision='bf16'`` sweeps the tensor-core kernel once per layer, ``'fp32'`` uses the layer-synchronous kernels.
    ``window_base`` = global index of the first window (multi-GPU shards: masks are keyed by the global index).
    ``exact_labels``: re-evaluate in fp32 (same masks, same statistics) every window whose bf16 mean probability lies
    within ``guard_delta`` (default 5e-3) of 0.5 and replace all of its outputs -- `predicted_label` is then the fp32
    path's label; the result carries ``guard_band = (re-evaluated windows, all windows)``.
    ``seed=None`` draws a fresh key per call (see UQK_SEED in the module docstring)."""
    eng = _engine.get_engine()
    blob = _model.weights_of(model)
    eng.ensure_members(lambda: blob[None], key=("mcd",) + _members_key([blob]))
    seed = _call_seed() if seed is None else int(seed)
    if bn_mode == "calibrated" and (recalibrate or not eng.calibrated):
        eng.calibrate(x_test_data, n_windows=calib_windows, n_passes=calib_passes, seed=seed)
    kw = {} if guard_delta is None else {"guard_delta": float(guard_delta)}
    out = eng.mcd(x_test_data, n_passes=int(n_pred), seed=seed, bn_mode=bn_mode, precision=precision, masks=masks,
                  return_probs=return_probs, device_out=device_out, window_base=int(window_base),
                  exact_labels=exact_labels, **kw)
    res = _finish_uq(out, y_true)
    if exact_labels:
        res["guard_band"] = eng.last_guard_band
    return res


def deep_ensemble_uq(ensemble_models: List, x_test_data, *, y_true=None, return_probs: bool = False,
                     precision: str = "bf16", device_out: bool = False, exact_labels: bool = False,
                     guard_delta: Optional[float] = None) -> Dict:
    """deep_ensembles_predict + uq_evaluation_dist + inline CSV metrics in one fused call (see mc_dropout_uq)."""
    eng = _engine.get_engine()
    models = list(ensemble_models)
    blobs = [_model.weights_of(m) for m in models]
    eng.ensure_members(lambda: np.stack(blobs), key=("de",) + _members_key(blobs))
    kw = {} if guard_delta is None else {"guard_delta": float(guard_delta)}
    out = eng.de(x_test_data, n_members=len(models), precision=precision, return_probs=return_probs,
                 device_out=device_out, exact_labels=exact_labels, **kw)
    res = _finish_uq(out, y_true)
    if exact_labels:
        res["guard_band"] = eng.last_guard_band
    return res


# ===================== uncertainty metrics =====================
def safe_entropy(probs: np.ndarray, axis: int = 1, epsilon: float = 1e-10) -> np.ndarray:
    """Entropy (nats) of clipped, renormalised probability rows -- scipy.stats.entropy semantics
    (reference uq_techniques.py:35-38).  Tiny helper kept on the host; the per-window entropies of the hot
    path are computed in the reduction kernels."""
    q = np.clip(probs, epsilon, 1 - epsilon)
    q = q / np.sum(q, axis=axis, keepdims=True)
    return -np.sum(q * np.log(q), axis=axis)


def _as_2d(uq_predictions: np.ndarray) -> np.ndarray:
    p = np.squeeze(uq_predictions)
    if p.ndim == 0:
        p = p.reshape(1, 1)
    if p.ndim == 1:
        p = p.reshape(1, -1)
    return p


def uq_evaluation_dist(uq_predictions: np.ndarray, y_true: np.ndarray) -> Dict[str, np.ndarray]:
    """Per-window mean / variance / H[E p] / E[H p] / MI from a (T, N[,1]) probability tensor, plus the three
    variance aggregates (reference uq_techniques.py:40-112; same dict keys :103-112).  Runs in the K4 kernel."""
    predictions = _as_2d(np.asarray(uq_predictions))
    if predictions.shape[0] == 1:
        print("Warning: Only one set of predictions provided. Variance and Mutual Info will be zero.")
    in_dtype = predictions.dtype if predictions.dtype in (np.float32, np.float64) else np.float32
    eng = _engine.get_engine()
    m = eng.reduce(predictions.astype(np.float32, copy=False),
                   metrics=("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy",
                            "mutual_info"))
    m = {k: v.astype(in_dtype, copy=False) for k, v in m.items()}
    var = m["pred_variance"]
    y = np.asarray(y_true)
    is0, is1 = (y == 0), (y == 1)
    m["overall_mean_variance"] = np.mean(var)
    m["mean_variance_class_0"] = np.mean(var[is0]) if is0.any() else 0.0
    m["mean_variance_class_1"] = np.mean(var[is1]) if is1.any() else 0.0
    return m


# ===================== confidence intervals =====================
def _aggregate_row_to_dict(row) -> Dict[str, float]:
    return {k: float(v) for k, v in zip(AGG_KEYS, row)}


def bootstrap_metrics(uq_predictions: np.ndarray, y_true: np.ndarray,
                      n_bootstrap: int = 100, random_state: Optional[int] = None) -> Optional[List[Dict]]:
    """Bootstrap over windows of the six aggregate means (reference uq_techniques.py:116-172).

    The reference resamples columns of the (T,N) tensor and recomputes every per-window metric; since those
    metrics are functions of one column only, metrics(preds[:, idx]) == metrics(preds)[idx], so each iteration
    is a gather-mean (K6 kernel).  The index stream is drawn on the host with the legacy NumPy generator in
    exactly the reference's order (:130-131, :142), so CI values are reproducible against it."""
    if random_state is not None:
        np.random.seed(random_state)
    preds = np.asarray(uq_predictions)
    n_samples = preds.shape[1]
    y = np.asarray(y_true)
    print(f"Starting bootstrap with {n_bootstrap} iterations...")
    results: List[Dict] = []
    try:
        base = uq_evaluation_dist(preds, y)
        eng = _engine.get_engine()
        chunk = max(1, min(n_bootstrap, (256 << 20) // max(8 * n_samples, 1)))
        done = 0
        while done < n_bootstrap:
            b = min(chunk, n_bootstrap - done)
            idx = np.empty((b, n_samples), dtype=np.int64)
            for i in range(b):
                if (done + i + 1) % 10 == 0:
                    print(f"  Bootstrap iteration {done+i+1}/{n_bootstrap}")
                idx[i] = np.random.choice(n_samples, n_samples, replace=True)
            rows = eng.bootstrap(base, y, idx)
            results.extend(_aggregate_row_to_dict(r) for r in rows)
            done += b
    except Exception as e:  # the reference tolerates per-iteration failures (:162-165)
        print(f"Error during bootstrap: {e}")
    if not results:
        print("Error: No bootstrap results generated.")
        return None
    print("Bootstrap finished.")
    return results


def compute_confidence_intervals(bootstrap_results: List[Dict], alpha: float = 0.05) -> Dict[str, float]:
    """Mean and percentile CI bounds per aggregate (reference uq_techniques.py:175-206; keys :200-202)."""
    out: Dict[str, float] = {}
    if not bootstrap_results:
        return out
    lo_q, hi_q = 100 * alpha / 2, 100 * (1 - alpha / 2)
    for name in bootstrap_results[0].keys():
        try:
            vals = np.array([r[name] for r in bootstrap_results], dtype=np.float64)
            out[f"{name}_mean"] = float(vals.mean())
            out[f"{name}_ci_lower"] = float(np.percentile(vals, lo_q))
            out[f"{name}_ci_upper"] = float(np.percentile(vals, hi_q))
        except Exception as e:
            print(f"Error calculating CI for metric '{name}': {e}")
    return out


# ===================== visualisation (off the hot path) =====================
def _pyplot():
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        return plt
    except Exception:
        return None


def _save_plot_data(output_dir: str, stem: str, **arrays):
    """matplotlib is absent from this image: keep the data a figure would have shown."""
    os.makedirs(output_dir, exist_ok=True)
    np.savez_compressed(os.path.join(output_dir, stem + ".npz"), **arrays)


def plot_uncertainty_metric(uncertainty_values: np.ndarray, uq_name: str, metric_name: str,
                            output_dir: str = "./uq_plots", title: Optional[str] = None, max_samples: int = 5000):
    """Line plot of a per-window metric, subsampled above max_samples (reference uq_techniques.py:210-239)."""
    os.makedirs(output_dir, exist_ok=True)
    vals = np.asarray(uncertainty_values)
    pos = np.arange(len(vals))
    sub = len(vals) > max_samples
    if sub:
        pos = np.sort(np.random.choice(len(vals), max_samples, replace=False))
        vals = vals[pos]
        print(f"Warning: Plotting line plot for {max_samples} random samples out of {len(uncertainty_values)}.")
    plt = _pyplot()
    stem = f"line_{metric_name}_{uq_name}"
    if plt is None:
        return _save_plot_data(output_dir, stem, index=pos, values=vals)
    plt.figure(figsize=(15, 4))
    plt.plot(pos, vals, alpha=0.7)
    plt.title(title or f"{metric_name} over Samples - {uq_name}")
    plt.xlabel("Sample Index (Subsampled)" if sub else "Sample Index")
    plt.ylabel(metric_name)
    plt.grid(True, alpha=0.3)
    plt.tight_layout()
    plt.savefig(os.path.join(output_dir, stem + ".png"), bbox_inches="tight")
    plt.close()


def plot_class_uncertainties(class0_unc: float, class1_unc: float, uq_name: str, output_dir: str = "./uq_plots"):
    """Bar chart of the two class-conditional mean variances (reference uq_techniques.py:242-255)."""
    os.makedirs(output_dir, exist_ok=True)
    plt = _pyplot()
    stem = f"bar_class_variance_{uq_name}"
    if plt is None:
        return _save_plot_data(output_dir, stem, values=np.array([class0_unc, class1_unc], dtype=np.float64))
    plt.figure(figsize=(6, 5))
    bars = plt.bar(["Normal (0)", "Apnea/Hypopnea (1)"], [class0_unc, class1_unc], color=["skyblue", "salmon"])
    plt.bar_label(bars, fmt="%.6f")
    plt.title(f"Mean Predictive Variance by True Class - {uq_name}")
    plt.ylabel("Mean Predictive Variance")
    plt.ylim(bottom=0)
    plt.tight_layout()
    plt.savefig(os.path.join(output_dir, stem + ".png"), bbox_inches="tight")
    plt.close()


def plot_metric_distribution(metric_values: np.ndarray, y_true: np.ndarray, uq_name: str, metric_name: str,
                             output_dir: str = "./uq_plots", bins: int = 30):
    """Per-class histograms of a per-window metric (reference uq_techniques.py:258-275)."""
    os.makedirs(output_dir, exist_ok=True)
    vals, y = np.asarray(metric_values), np.asarray(y_true)
    plt = _pyplot()
    stem = f"hist_{metric_name}_by_class_{uq_name}"
    if plt is None:
        data = {}
        for cls in (0, 1):
            if np.any(y == cls):
                h, edges = np.histogram(vals[y == cls], bins=bins, density=True)
                data[f"density_class{cls}"] = h
                data[f"edges_class{cls}"] = edges
        return _save_plot_data(output_dir, stem, **data)
    plt.figure(figsize=(10, 6))
    for cls, label in ((0, "True Normal (0)"), (1, "True Apnea/Hypopnea (1)")):
        if np.any(y == cls):
            plt.hist(vals[y == cls], bins=bins, alpha=0.6, label=label, density=True)
    plt.title(f"{metric_name} Distribution by True Class - {uq_name}")
    plt.xlabel(f"{metric_name} Value")
    plt.ylabel("Density")
    plt.legend()
    plt.grid(True, alpha=0.3)
    plt.tight_layout()
    plt.savefig(os.path.join(output_dir, stem + ".png"), bbox_inches="tight")
    plt.close()


# ===================== orchestrator =====================
def evaluate_uq_methods(predictions: np.ndarray, y_test: np.ndarray, evaluation_label: str = "UQ Evaluation",
                        n_bootstrap: int = 100, random_state: Optional[int] = None,
                        output_plot_dir: str = "./uq_plots") -> Optional[Dict]:
    """Metrics + bootstrap CIs + figures; returns the aggregated dict (reference uq_techniques.py:278-391;
    keys :200-202 and :350-355; None-in -> None-out :309-311; length mismatch -> ValueError :312-313)."""
    print(f"\n=== Evaluating Uncertainty: {evaluation_label} ===")
    if predictions is None or y_test is None:
        print("Error: Input predictions or labels are None.")
        return None
    print(f"Input prediction shape: {predictions.shape}, Samples: {len(y_test)}")
    if len(y_test) != predictions.shape[1]:
        raise ValueError(f"Mismatch between prediction samples ({predictions.shape[1]}) and label samples ({len(y_test)})")
    if predictions.ndim == 3 and predictions.shape[-1] == 1:
        predictions = predictions[..., 0]
    if predictions.ndim == 1:
        predictions = predictions.reshape(1, -1)

    print("\nCalculating base UQ metrics...")
    per_window = uq_evaluation_dist(predictions, y_test)
    if per_window is None:
        print("Error: Failed to calculate base UQ metrics.")
        return None
    point = {
        "overall_mean_variance": per_window["overall_mean_variance"],
        "mean_variance_class_0": per_window["mean_variance_class_0"],
        "mean_variance_class_1": per_window["mean_variance_class_1"],
        "mean_total_pred_entropy": np.mean(per_window["total_pred_entropy"]),
        "mean_expected_aleatoric_entropy": np.mean(per_window["expected_aleatoric_entropy"]),
        "mean_mutual_info": np.mean(per_window["mutual_info"]),
    }
    print(f"- Overall Mean Variance: {point['overall_mean_variance']:.6f}")
    print(f"- Mean Variance Class 0: {point['mean_variance_class_0']:.6f}")
    print(f"- Mean Variance Class 1: {point['mean_variance_class_1']:.6f}")
    print(f"- Mean Predictive Entropy (Total): {point['mean_total_pred_entropy']:.4f}")
    print(f"- Mean Expected Entropy (Aleatoric): {point['mean_expected_aleatoric_entropy']:.4f}")
    print(f"- Mean Mutual Info (Epistemic): {point['mean_mutual_info']:.6f}")

    print(f"\nComputing CIs (n_bootstrap={n_bootstrap})...")
    t0 = time.time()
    aggregated: Dict = {}
    boot = bootstrap_metrics(predictions, y_test, n_bootstrap, random_state)
    if boot:
        aggregated.update(compute_confidence_intervals(boot))
        print(f"CIs computed in {time.time()-t0:.2f}s")
    else:
        print("Warning: Bootstrap failed, CIs not computed.")
    aggregated.update(point)

    print("\nGenerating visualizations...")
    for key, name in (("pred_variance", "Predictive Variance"), ("total_pred_entropy", "Predictive Entropy"),
                      ("mutual_info", "Mutual Information")):
        plot_metric_distribution(per_window[key], y_test, evaluation_label, name, output_dir=output_plot_dir)
    plot_class_uncertainties(aggregated["mean_variance_class_0"], aggregated["mean_variance_class_1"],
                             evaluation_label, output_dir=output_plot_dir)
    print("\n=== Evaluation Complete ===")
    return aggregated
