#!/usr/bin/env python
"""Generates tests/golden/*.npz|csv by running the REFERENCE's own code from /root/reference.

Run in the build container only (the reference tree does not travel to the GPU box):
    python tests/golden/make_golden.py

Outputs (committed):
  uq_dummy_golden.npz         reference uq_evaluation_dist + bootstrap/CI + evaluate_uq_methods on the
                              reference's own dummy recipe (uq_techniques.py:401-417, seed 42, 5 x 1000)
  uq_f32_golden.npz           reference uq_evaluation_dist on float32 (T,N,1) inputs incl. edge cases
                              (p in {0,1,0.5}, T=1, single-class labels) + the drivers' inline CSV metrics
                              (analyze_mcd_patient_level.py:109-117 restated verbatim here)
  patient_detail_fixture.csv  synthetic per-window CSV in the format analyze_mcd_patient_level.py:134-152 writes
  patient_summary_golden.csv  output of the UNMODIFIED reference aggregate_patient_uq_metrics.py on that CSV
  forward_oracle_golden.npz   oracle/network.py float64 outputs on seeded synthetic weights/windows
                              (self-generated: parity UNPINNED for the forward, see oracle/__init__.py)
"""
import importlib
import os
import runpy
import shutil
import sys
import tempfile
import contextlib
import io

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import reference_loader, network, philox  # noqa: E402

uqk = importlib.import_module("uncertaintyquantification-sleepapnea-1dcnn_b200")


def dummy_recipe():
    # verbatim recipe of reference uq_techniques.py:401-417
    N_MODELS, N_SAMPLES, RANDOM_SEED = 5, 1000, 42
    np.random.seed(RANDOM_SEED)
    dummy_preds = []
    for _ in range(N_MODELS):
        p = np.random.rand(N_SAMPLES) * 0.6 + 0.2
        noise = np.random.randn(N_SAMPLES) * 0.1
        dummy_preds.append(np.clip(p + noise, 0.01, 0.99))
    preds = np.stack(dummy_preds)
    labels = (np.random.rand(N_SAMPLES) > 0.7).astype(int)
    return preds, labels


def f32_cases():
    rng = np.random.default_rng(1234)
    T, N = 50, 257
    p = rng.random((T, N, 1)).astype(np.float32)
    p[:, 0, 0] = 0.0
    p[:, 1, 0] = 1.0
    p[:, 2, 0] = 0.5
    p[:, 3, 0] = np.where(np.arange(T) % 3 == 0, 1.0, 0.0)
    p[:, 4, 0] = 1e-12
    p[:, 5, 0] = np.float32(1.0) - np.float32(6e-8)
    p[:, 6:40, 0] = (0.5 + 1e-4 * rng.standard_normal((T, 34))).astype(np.float32)   # tiny variances
    y = (rng.random(N) < 0.3).astype(np.int64)
    return p, y


def forward_golden():
    """Section 4 alone (`python make_golden.py forward`): it is the only part that depends on the build's own RNG
    definition (oracle/philox.py), not on the reference."""
    out = {}
    X = uqk.synthetic.make_windows(24, seed=11)
    for m_ in range(2):
        ws = uqk.synthetic.make_member_weights(m_)
        out[f"det_member{m_}"] = network.forward_f64(ws, X)
    ws = uqk.synthetic.make_member_weights(0)
    seed, T = 2025, 3
    for t in range(T):
        masks = [philox.dropout_keep_mask(seed, np.arange(24), t, i, b.cout, b.rate)
                 for i, b in enumerate(uqk.arch.BLOCKS)]
        out[f"mcd_batch_t{t}"] = network.forward_f64(ws, X, training=True, masks=masks)
        out[f"mcd_frozen_t{t}"] = network.forward_f64(ws, X, training=True, masks=masks, bn_mode="inference")
    np.savez_compressed(os.path.join(HERE, "forward_oracle_golden.npz"), **out)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "forward":
        forward_golden()
        return
    ref = reference_loader.load_reference_uq_techniques()
    sink = io.StringIO()

    # ---- 1. dummy recipe through the reference's own functions
    preds, labels = dummy_recipe()
    with contextlib.redirect_stdout(sink):
        m = ref.uq_evaluation_dist(preds, labels)
        boot = ref.bootstrap_metrics(preds, labels, n_bootstrap=50, random_state=42)
        ci = ref.compute_confidence_intervals(boot)
        tmp = tempfile.mkdtemp()
        agg = ref.evaluate_uq_methods(predictions=preds, y_test=labels, evaluation_label="Dummy Ensemble Test",
                                      n_bootstrap=50, random_state=42, output_plot_dir=tmp)
        shutil.rmtree(tmp, ignore_errors=True)
    out = {f"m_{k}": np.asarray(v) for k, v in m.items()}
    keys = sorted(boot[0].keys())
    out["boot_keys"] = np.array(keys)
    out["boot_values"] = np.array([[float(b[k]) for k in keys] for b in boot])
    out["ci_keys"] = np.array(sorted(ci.keys()))
    out["ci_values"] = np.array([ci[k] for k in sorted(ci.keys())])
    out["agg_keys"] = np.array(sorted(agg.keys()))
    out["agg_values"] = np.array([float(agg[k]) for k in sorted(agg.keys())])
    np.savez_compressed(os.path.join(HERE, "uq_dummy_golden.npz"), **out)

    # ---- 2. float32 (T,N,1) cases, as the drivers pass them
    p32, y32 = f32_cases()
    out = {"probs": p32, "y": y32}
    with contextlib.redirect_stdout(sink):
        m = ref.uq_evaluation_dist(p32, y32)
        for k, v in m.items():
            out[f"m_{k}"] = np.asarray(v)
        m1 = ref.uq_evaluation_dist(p32[:1], y32)                 # T=1 (warning path :65-66)
        for k, v in m1.items():
            out[f"t1_{k}"] = np.asarray(v)
        mz = ref.uq_evaluation_dist(p32, np.zeros_like(y32))      # class 1 absent -> 0.0 (:101)
        out["allzero_c1"] = np.asarray(mz["mean_variance_class_1"])
        out["allzero_c0"] = np.asarray(mz["mean_variance_class_0"])
        out["safe_entropy_probe"] = ref.safe_entropy(np.array([[1.0, 0.0], [0.5, 0.5], [0.3, 0.7]]))
        out["safe_entropy_probe_f32"] = ref.safe_entropy(
            np.array([[1.0, 0.0], [0.5, 0.5], [0.3, 0.7]], dtype=np.float32))
    # inline driver metrics, verbatim from analyze_mcd_patient_level.py:109-117
    mean_probs_per_window = np.mean(p32, axis=0).flatten()
    pred_variance_per_window = np.var(p32, axis=0).flatten()
    epsilon = 1e-9
    pred_entropy_per_window = - (mean_probs_per_window * np.log2(mean_probs_per_window + epsilon) +
                                 (1 - mean_probs_per_window) * np.log2(1 - mean_probs_per_window + epsilon))
    final_predicted_labels = (mean_probs_per_window > 0.5).astype(int)
    out["csv_mean"] = mean_probs_per_window
    out["csv_var"] = pred_variance_per_window
    out["csv_entropy"] = pred_entropy_per_window
    out["csv_label"] = final_predicted_labels
    np.savez_compressed(os.path.join(HERE, "uq_f32_golden.npz"), **out)

    # ---- 3. patient aggregation: run the unmodified reference script on a fixture CSV
    rng = np.random.default_rng(77)
    counts = np.concatenate([rng.integers(3, 40, size=30), [1, 1, 2, 120]])
    pid = np.repeat(np.arange(len(counts)) + 200001, counts)
    perm = rng.permutation(len(counts))          # ids not sorted in file order
    pid = np.concatenate([pid[pid == 200001 + j] for j in perm])
    n = len(pid)
    var = (rng.random(n) ** 3 * 0.05).astype(np.float32)
    meanp = rng.random(n).astype(np.float32)
    ent = (-(meanp * np.log2(meanp + 1e-9) + (1 - meanp) * np.log2(1 - meanp + 1e-9))).astype(np.float32)
    df = pd.DataFrame({"Patient_ID": pid, "Window_Index": np.arange(n), "True_Label": (rng.random(n) < 0.3).astype(int),
                       "Predicted_Label": (meanp > 0.5).astype(int), "Predicted_Probability": meanp,
                       "Predictive_Variance": var, "Predictive_Entropy": ent})
    df.to_csv(os.path.join(HERE, "patient_detail_fixture.csv"), index=False)
    work = tempfile.mkdtemp()
    cwd = os.getcwd()
    try:
        os.chdir(work)
        shutil.copy(os.path.join(HERE, "patient_detail_fixture.csv"), "detail_patient_MCD.csv")
        with contextlib.redirect_stdout(sink):
            runpy.run_path(os.path.join(reference_loader.REFERENCE_ROOT, "uncertainty_quantification",
                                        "aggregate_patient_uq_metrics.py"), run_name="__main__")
        shutil.copy(os.path.join("patient_level_uq_analysis_MCD", "patient_summary_metrics_MCD.csv"),
                    os.path.join(HERE, "patient_summary_golden.csv"))
    finally:
        os.chdir(cwd)
        shutil.rmtree(work, ignore_errors=True)

    # ---- 4. forward: oracle float64 on seeded synthetic inputs (self-generated, unpinned)
    forward_golden()

    # ---- 5. window standardisation: the reference's own standardize_per_window
    #         (data_prepocessing/prepare_numpy_datasets.py:83-95; imblearn is absent and only used elsewhere -> stubbed)
    import importlib.util, types
    saved = {k: sys.modules.get(k) for k in ("imblearn", "imblearn.over_sampling", "imblearn.under_sampling")}
    sys.modules["imblearn"] = types.ModuleType("imblearn")
    sys.modules["imblearn.over_sampling"] = types.ModuleType("imblearn.over_sampling")
    sys.modules["imblearn.over_sampling"].SMOTE = object
    sys.modules["imblearn.under_sampling"] = types.ModuleType("imblearn.under_sampling")
    sys.modules["imblearn.under_sampling"].RandomUnderSampler = object
    try:
        spec = importlib.util.spec_from_file_location(
            "_ref_prepare", os.path.join(reference_loader.REFERENCE_ROOT, "data_prepocessing", "prepare_numpy_datasets.py"))
        prep = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(prep)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    rng = np.random.default_rng(4242)
    raw = rng.standard_normal((24, 60, 4)) * np.array([3.0, 12.0, 0.5, 0.5]) + np.array([95.0, 70.0, 0.0, 0.0])
    raw[3, :, 1] = 72.0                      # constant channel -> std = 0 -> (x - mean) / 1e-8 = 0
    raw[5, :, 0] = raw[5, 0, 0]
    with contextlib.redirect_stdout(sink):
        std = prep.standardize_per_window(raw)
    np.savez_compressed(os.path.join(HERE, "standardize_golden.npz"), raw=raw, out=std)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
