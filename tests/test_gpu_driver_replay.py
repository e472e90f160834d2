"""Replay of the reference drivers' call sequence on the GPU (the reference tree does not exist on the GPU box, so the
scripts themselves cannot run there; tests/test_reference_callers_cpu.py runs them unchanged on an oracle-backed engine).
Same files, shapes and dtypes: float64 (N,60,4) windows from ./processed_datasets, models loaded through the tensorflow
shim, deterministic check, T = 50 passes, per-window columns, CSV, evaluate_uq_methods with bootstrap CIs on both data
sets (analyze_mcd_patient_level.py:196-238), the M = 5 ensemble (analyze_de_patient_level.py:203-234) and the patient
aggregation of the CSV (aggregate_patient_uq_metrics.py:34-46)."""
import contextlib
import importlib
import io
import os
import sys

import numpy as np
import pandas as pd
import pytest

from conftest import PKG_NAME, ROOT
from oracle import patient as patient_oracle

pytestmark = pytest.mark.gpu
UQ_DIR = os.path.join(ROOT, PKG_NAME, "uncertainty_quantification")
SHIM_DIR = os.path.join(ROOT, PKG_NAME, "shim")
AGG_KEYS = {"overall_mean_variance", "mean_variance_class_0", "mean_variance_class_1", "mean_total_pred_entropy",
            "mean_expected_aleatoric_entropy", "mean_mutual_info"}


def test_driver_call_sequence_on_the_gpu(tmp_path, monkeypatch):
    fixtures = importlib.import_module(PKG_NAME + ".fixtures")
    emit = importlib.import_module(PKG_NAME + ".emit")
    patient = importlib.import_module(PKG_NAME + ".patient")
    work = str(tmp_path)
    n_u, n_r = 600, 200
    fixtures.build_fixture_dir(work, n_unbalanced=n_u, n_rus=n_r, n_members=5)
    monkeypatch.chdir(work)
    monkeypatch.syspath_prepend(SHIM_DIR)
    monkeypatch.syspath_prepend(UQ_DIR)
    for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.") or k == "uq_techniques"]:
        del sys.modules[k]
    import tensorflow as tf                                   # the shim
    from tensorflow.keras.models import load_model
    uq = importlib.import_module("uq_techniques")
    tf.random.set_seed(2025)
    sets = {"Unbalanced": (np.load("processed_datasets/X_test_std_unbalanced.npy"), np.load("processed_datasets/y_test_unbalanced.npy"),
                           np.load("processed_datasets/patient_ids_test_unbalanced.npy")),
            "RUS": (np.load("processed_datasets/X_test_std_rus.npy"), np.load("processed_datasets/y_test_rus.npy"), None)}
    assert sets["Unbalanced"][0].dtype == np.float64 and sets["Unbalanced"][0].shape == (n_u, 60, 4)
    cnn = load_model("./AlCNN1D_no_pool.keras")
    ensemble = [load_model(f"./models/ensemble_cnn_no_pool/AlCNN_smote_seed{i + 5}.keras") for i in range(5)]
    sink = io.StringIO()
    for method in ("MCD", "DE"):
        for label, (X, y, pid) in sets.items():
            with contextlib.redirect_stdout(sink):
                if method == "MCD":
                    det = cnn(X, training=False).numpy().flatten()                 # analyze_mcd_patient_level.py:204
                    assert det.shape == (len(X),) and np.all((det > 0) & (det < 1))
                    preds = uq.mc_dropout_predict(cnn, X, 50)
                    assert preds.shape == (50, len(X), 1) and preds.dtype == np.float32
                else:
                    preds = uq.deep_ensembles_predict(ensemble, X)
                    assert preds.shape == (5, len(X), 1) and preds.dtype == np.float32
            mean_p = np.mean(preds, axis=0).flatten()                              # :109-117, the drivers' own NumPy lines
            var_p = np.var(preds, axis=0).flatten()
            ent = -(mean_p * np.log2(mean_p + 1e-9) + (1 - mean_p) * np.log2(1 - mean_p + 1e-9))
            lab = (mean_p > 0.5).astype(int)
            if pid is not None:
                df = pd.DataFrame({"Patient_ID": pid, "Window_Index": np.arange(len(y)), "True_Label": y, "Predicted_Label": lab,
                                   "Predicted_Probability": mean_p, "Predictive_Variance": var_p, "Predictive_Entropy": ent})
                df.to_csv(f"detail_{method}.csv", index=False)
                emit.write_detailed_csv(f"detail_{method}_native.csv", pid, y, lab, mean_p, var_p, ent)
                assert open(f"detail_{method}.csv", "rb").read() == open(f"detail_{method}_native.csv", "rb").read()
                back = pd.read_csv(f"detail_{method}.csv")
                mine = patient.aggregate_patients(back)
                ref = patient_oracle.aggregate_patients(back)
                np.testing.assert_array_equal(mine["num_windows"].values, ref["num_windows"].values)
                np.testing.assert_allclose(mine["median_entropy"].values, ref["median_entropy"].values.astype(float), rtol=1e-6, atol=1e-9)
                assert int(mine["num_windows"].sum()) == len(y)
            with contextlib.redirect_stdout(sink):
                agg = uq.evaluate_uq_methods(preds, y, f"CNN_{method}_{label}", n_bootstrap=100, random_state=2025,
                                             output_plot_dir=os.path.join(work, "uq_plots", f"{method}_{label}"))
            assert AGG_KEYS <= set(agg) and all(np.isfinite(float(agg[k])) for k in AGG_KEYS)
            for k in AGG_KEYS:
                assert f"{k}_ci_lower" in agg and agg[f"{k}_ci_lower"] <= agg[f"{k}_mean"] <= agg[f"{k}_ci_upper"] + 1e-12
    out = sink.getvalue()
    assert out.count("MC Dropout completed in") == 2 and out.count("Deep Ensemble completed in") == 2
