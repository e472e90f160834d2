"""The drop-in boundary on the GPU: uq_techniques' public functions against the reference's own outputs (golden),
the reference caller scripts run UNCHANGED on top of the build (when /root/reference is present), patient-level
aggregation (K5) and bootstrap gather-means (K6) through the C-ABI."""
import contextlib
import importlib
import io
import os
import runpy
import shutil
import sys

import numpy as np
import pandas as pd
import pytest

from conftest import PKG_NAME, ROOT
from gpu_common import engine_with_members, members
from oracle import reductions, patient as patient_oracle, network, reference_loader

pytestmark = pytest.mark.gpu
UQ_DIR = os.path.join(ROOT, PKG_NAME, "uncertainty_quantification")


@pytest.fixture(scope="module")
def uq():
    sys.path.insert(0, UQ_DIR)
    sys.modules.pop("uq_techniques", None)
    mod = importlib.import_module("uq_techniques")
    yield mod
    sys.path.remove(UQ_DIR)


def test_uq_evaluation_dist_matches_reference_golden(uq, golden_dir):
    g = np.load(os.path.join(golden_dir, "uq_f32_golden.npz"))
    m = uq.uq_evaluation_dist(g["probs"], g["y"])                      # (T,N,1) float32, as the drivers pass it
    assert set(m.keys()) == {"mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy",
                             "mutual_info", "overall_mean_variance", "mean_variance_class_0", "mean_variance_class_1"}
    for k in ("mean_pred", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info"):
        assert m[k].dtype == np.float32
        np.testing.assert_allclose(m[k], g[f"m_{k}"], rtol=0, atol=1e-6, err_msg=k)
    np.testing.assert_allclose(m["pred_variance"], g["m_pred_variance"], rtol=1e-4, atol=1e-7)
    for k in ("overall_mean_variance", "mean_variance_class_0", "mean_variance_class_1"):
        np.testing.assert_allclose(float(m[k]), float(g[f"m_{k}"]), rtol=1e-4)
    mz = uq.uq_evaluation_dist(g["probs"], np.zeros_like(g["y"]))       # class absent -> 0.0
    assert mz["mean_variance_class_1"] == 0.0
    m1 = uq.uq_evaluation_dist(g["probs"][0, :, 0], g["y"])             # 1-D input -> (1,-1) + warning path
    assert np.all(m1["pred_variance"] == 0)


def test_dummy_recipe_bootstrap_and_cis_match_reference(uq, golden_dir):
    sys.path.insert(0, golden_dir)
    from make_golden import dummy_recipe
    g = np.load(os.path.join(golden_dir, "uq_dummy_golden.npz"))
    preds, labels = dummy_recipe()                                      # float64 (5,1000), the reference's own demo
    with contextlib.redirect_stdout(io.StringIO()):
        boot = uq.bootstrap_metrics(preds, labels, n_bootstrap=50, random_state=42)
        ci = uq.compute_confidence_intervals(boot)
        agg = uq.evaluate_uq_methods(preds, labels, "Dummy Ensemble Test", n_bootstrap=50, random_state=42,
                                     output_plot_dir=os.path.join(ROOT, "gpurun_out", "plots_dummy"))
    keys = [str(k) for k in g["boot_keys"]]
    mine = np.array([[b[k] for k in keys] for b in boot])
    np.testing.assert_allclose(mine, g["boot_values"], rtol=2e-5, atol=1e-7)     # float32 kernel vs float64 reference
    for k, v in zip(g["ci_keys"], g["ci_values"]):
        np.testing.assert_allclose(ci[str(k)], v, rtol=2e-5, atol=1e-7, err_msg=str(k))
    assert sorted(agg.keys()) == [str(k) for k in g["agg_keys"]]
    for k, v in zip(g["agg_keys"], g["agg_values"]):
        np.testing.assert_allclose(float(agg[str(k)]), v, rtol=2e-5, atol=1e-7, err_msg=str(k))
    with pytest.raises(ValueError):
        uq.evaluate_uq_methods(preds, labels[:-1])
    assert uq.evaluate_uq_methods(None, labels) is None


def test_predict_functions_shapes_and_semantics(uq, uqk):
    model = importlib.import_module(PKG_NAME + ".model")
    x = uqk.synthetic.make_windows(37, seed=41)
    ms = [model.UQModel(w) for w in members()]
    with contextlib.redirect_stdout(io.StringIO()) as out:
        de = uq.deep_ensembles_predict(ms, x)
        mc = uq.mc_dropout_predict(ms[0], x, n_pred=3)                    # auto -> exact batch statistics (fp32)
    assert "Deep Ensemble completed in" in out.getvalue() and "(3 models)" in out.getvalue()
    assert "MC Dropout completed in" in out.getvalue() and "(3 passes)" in out.getvalue()
    assert de.shape == (3, 37, 1) and de.dtype == np.float32 and mc.shape == (3, 37, 1) and mc.dtype == np.float32
    ref = np.stack([network.forward_f64(w, x) for w in members()])
    assert np.abs(de - ref).max() < 2e-2
    assert np.all((mc > 0) & (mc < 1)) and np.abs(mc[0] - mc[1]).max() > 1e-4   # stochastic passes differ
    # Keras-like model surface used by the drivers (analyze_mcd_patient_level.py:204; uq_techniques.py:22,29)
    det = ms[0](x, training=False).numpy().flatten()
    assert np.abs(det - ref[0, :, 0]).max() < 2e-2
    assert ms[0].predict(x, verbose=0).shape == (37, 1)
    one = np.stack([ms[0](x, training=True) for _ in range(2)])
    assert one.shape == (2, 37, 1) and np.abs(one[0] - one[1]).max() > 1e-4
    fused = uq.mc_dropout_uq(ms[0], x, n_pred=4, y_true=np.zeros(37, int), bn_mode="inference", return_probs=True)
    m = reductions.uq_evaluation_dist(fused["probs"], np.zeros(37, int))
    np.testing.assert_allclose(fused["mean_pred"], m["mean_pred"], atol=1e-6)
    np.testing.assert_allclose(fused["mutual_info"], m["mutual_info"], atol=1e-6)
    assert fused["mean_variance_class_1"] == 0.0


def test_mc_dropout_predict_keeps_exact_semantics_above_the_fp32_limit(uq, uqk, monkeypatch):
    """`mc_dropout_predict` in its default mode always normalises with the statistics of the whole call
    (uq_techniques.py:22): fp32 kernels for small calls, the swept tensor-core kernel above the limit."""
    model = importlib.import_module(PKG_NAME + ".model")
    n = 70
    x = uqk.synthetic.make_windows(n, seed=17)
    m0 = model.UQModel(members()[0])
    monkeypatch.setenv("UQK_SEED", "123")
    with contextlib.redirect_stdout(io.StringIO()):
        small = uq.mc_dropout_predict(m0, x, n_pred=2)                     # n <= limit: fp32 layer-synchronous
        monkeypatch.setenv("UQK_EXACT_MAX_WINDOWS", "16")
        big = uq.mc_dropout_predict(m0, x, n_pred=2)                       # n > limit: bf16 sweeps, same semantics
    assert small.shape == big.shape == (2, n, 1)
    d = np.abs(small - big)
    assert 0 < d.max() < 4e-2 and d.mean() < 5e-3
    from gpu_common import twin_masks
    ref = np.stack([network.forward_f64(members()[0], x, training=True, masks=twin_masks(123, n, t)) for t in range(2)])
    assert np.abs(small - ref).max() < 1e-4 and np.abs(big - ref).max() < 4e-2


def test_patient_reduce_matches_reference_script_output(golden_dir):
    patient = importlib.import_module(PKG_NAME + ".patient")
    df = pd.read_csv(os.path.join(golden_dir, "patient_detail_fixture.csv"))
    ref = pd.read_csv(os.path.join(golden_dir, "patient_summary_golden.csv"))     # written by the reference script
    mine = patient.aggregate_patients(df)
    assert list(mine.columns) == list(ref.columns)
    np.testing.assert_array_equal(mine["Patient_ID"].values, ref["Patient_ID"].values)       # sorted ids, exact
    np.testing.assert_array_equal(mine["num_windows"].values, ref["num_windows"].values)     # counts, exact
    for c in ref.columns[1:-1]:
        np.testing.assert_allclose(mine[c].values, ref[c].values.astype(float), rtol=1e-6, atol=1e-9, err_msg=c)
    np.testing.assert_allclose(mine["patient_accuracy"].values, ref["patient_accuracy"].values, rtol=1e-12)


def test_patient_reduce_cohort_scale_properties(uqk):
    """SHHS2-scale cohort (2651 patients x ~480 windows): counts exact, means equal a NumPy segmented reduce."""
    patient = importlib.import_module(PKG_NAME + ".patient")
    pid = uqk.synthetic.make_patient_ids()
    n = len(pid)
    rng = np.random.default_rng(3)
    var = (rng.random(n) ** 2 * 0.1).astype(np.float32)
    ent = rng.random(n).astype(np.float32)
    yt = (rng.random(n) < 0.2).astype(np.int32)
    yp = (rng.random(n) < 0.3).astype(np.int32)
    out = patient.aggregate_arrays(pid, var, ent, yt, yp)
    counts = np.bincount(pid - pid.min())
    np.testing.assert_array_equal(out["num_windows"].values, counts)
    assert int(out["num_windows"].sum()) == n
    starts = np.concatenate([[0], np.cumsum(counts)[:-1]])
    np.testing.assert_allclose(out["mean_variance"].values, np.add.reduceat(var.astype(np.float64), starts) / counts, rtol=1e-9)
    df = pd.DataFrame({"Patient_ID": pid, "True_Label": yt, "Predicted_Label": yp, "Predictive_Variance": var,
                       "Predictive_Entropy": ent})
    ref = patient_oracle.aggregate_patients(df.iloc[: int(starts[40])])            # pandas on the first 40 patients
    for c in ("median_variance", "std_entropy", "median_entropy", "patient_accuracy"):
        np.testing.assert_allclose(out[c].values[:40], ref[c].values.astype(float), rtol=1e-6, atol=1e-9, err_msg=c)


@pytest.mark.skipif(not reference_loader.reference_available(), reason="reference tree not present on the GPU box")
def test_reference_drivers_run_unchanged(tmp_path):
    """Runs the reference's analyze_mcd_patient_level.py unchanged on a synthetic fixture directory."""
    fixtures = importlib.import_module(PKG_NAME + ".fixtures")
    work = str(tmp_path)
    fixtures.build_fixture_dir(work, n_unbalanced=300, n_rus=120, n_members=5)
    env_path = [UQ_DIR, os.path.join(ROOT, PKG_NAME, "shim")]
    cwd = os.getcwd()
    saved = list(sys.path)
    try:
        os.chdir(work)
        sys.path[:0] = env_path
        for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.") or k == "uq_techniques"]:
            del sys.modules[k]
        with contextlib.redirect_stdout(io.StringIO()):
            runpy.run_path(os.path.join(reference_loader.REFERENCE_ROOT, "uncertainty_quantification",
                                        "analyze_mcd_patient_level.py"), run_name="__main__")
        csv = os.path.join(work, "uq_results_patient_no_pool", "detailed_results_CNN_MCD_Unbalanced.csv")
        df = pd.read_csv(csv)
        assert len(df) == 300 and list(df.columns)[:4] == ["Patient_ID", "Window_Index", "True_Label", "Predicted_Label"]
    finally:
        os.chdir(cwd)
        sys.path[:] = saved
        for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.") or k == "uq_techniques"]:
            del sys.modules[k]


def test_standardize_matches_reference_golden(golden_dir):
    """N3: uqk_standardize vs the reference's own standardize_per_window output (golden), float64 and float32."""
    g = np.load(os.path.join(golden_dir, "standardize_golden.npz"))
    eng = engine_with_members()
    out = eng.standardize(g["raw"])
    assert out.dtype == np.float64
    # windows 3 and 5 hold a constant channel: std = 0, so (x - mean) / 1e-8 amplifies the last-bit rounding of the
    # mean by 1e8 (NumPy's own value there is summation-order noise of ~1e-6); everything else must agree tightly
    const = np.zeros(g["raw"].shape, dtype=bool)
    const[3, :, 1] = True
    const[5, :, 0] = True
    np.testing.assert_allclose(out[~const], g["out"][~const], rtol=1e-11, atol=1e-9)
    assert np.abs(out[const] - g["out"][const]).max() < 1e-4
    out32 = eng.standardize(g["raw"].astype(np.float32))
    assert out32.dtype == np.float32
    np.testing.assert_allclose(out32, g["out"], rtol=2e-3, atol=2e-3)     # float32 input quantisation of 95 +- 3 signals
    assert eng.standardize(g["raw"][:0]).shape == (0, 60, 4)


def test_stochastic_calls_draw_fresh_masks_unless_seeded(uq, uqk, monkeypatch):
    """The reference draws new dropout masks on every call (TensorFlow's stateful generator, uq_techniques.py:22):
    two calls differ; UQK_SEED (or seed=) pins the Philox key and makes a call reproducible bit for bit; re-seeding
    through tf.random.set_seed restarts the call counter."""
    model = importlib.import_module(PKG_NAME + ".model")
    x = uqk.synthetic.make_windows(40, seed=43)
    m0 = model.UQModel(members()[0])
    monkeypatch.delenv("UQK_SEED", raising=False)
    with contextlib.redirect_stdout(io.StringIO()):
        a = uq.mc_dropout_predict(m0, x, n_pred=2)
        b = uq.mc_dropout_predict(m0, x, n_pred=2)
        assert np.abs(a - b).max() > 1e-4
        fa = uq.mc_dropout_uq(m0, x, n_pred=3, bn_mode="inference")["mean_pred"]
        fb = uq.mc_dropout_uq(m0, x, n_pred=3, bn_mode="inference")["mean_pred"]
        assert np.abs(fa - fb).max() > 1e-5
        sys.path.insert(0, os.path.join(ROOT, PKG_NAME, "shim"))
        for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.")]:
            del sys.modules[k]
        try:
            import tensorflow as tf               # the shim
            tf.random.set_seed(777)
            c1 = uq.mc_dropout_predict(m0, x, n_pred=2)
            c1b = uq.mc_dropout_predict(m0, x, n_pred=2)
            tf.random.set_seed(777)               # re-seeding restarts the sequence of keys
            c2 = uq.mc_dropout_predict(m0, x, n_pred=2)
            np.testing.assert_array_equal(c1, c2)
            assert np.abs(c1 - c1b).max() > 1e-4
        finally:
            sys.path.remove(os.path.join(ROOT, PKG_NAME, "shim"))
            for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.")]:
                del sys.modules[k]
            model.UQModel._global_seed = 2025
        monkeypatch.setenv("UQK_SEED", "5")
        d1 = uq.mc_dropout_predict(m0, x, n_pred=2)
        d2 = uq.mc_dropout_predict(m0, x, n_pred=2)
    np.testing.assert_array_equal(d1, d2)
    s1 = uq.mc_dropout_uq(m0, x, n_pred=3, bn_mode="inference", seed=9)["mean_pred"]
    s2 = uq.mc_dropout_uq(m0, x, n_pred=3, bn_mode="inference", seed=9)["mean_pred"]
    np.testing.assert_array_equal(s1, s2)
