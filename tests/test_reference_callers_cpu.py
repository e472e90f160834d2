"""The reference's caller scripts, UNCHANGED, on top of the build's drop-in module and tensorflow shim.

In this container there is no GPU, so the run must proceed through the scripts' own data / model loading and then
stop inside the first GPU call with the library's loud "no CUDA device" error, which the scripts' own
``except Exception`` blocks report (reference analyze_mcd_patient_level.py:233-238).  (On the GPU box the reference
tree does not exist; the same call sequence is exercised there by tests/test_gpu_boundary.py.)"""
import contextlib
import importlib
import io
import os
import runpy
import sys

import numpy as np
import pytest
import torch

from conftest import PKG_NAME, ROOT
from oracle import reference_loader

UQ_DIR = os.path.join(ROOT, PKG_NAME, "uncertainty_quantification")
SHIM_DIR = os.path.join(ROOT, PKG_NAME, "shim")


def _purge():
    for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.") or k == "uq_techniques"]:
        del sys.modules[k]


@pytest.mark.skipif(not reference_loader.reference_available(), reason="reference tree not present")
@pytest.mark.skipif(torch.cuda.is_available(), reason="CPU-container variant of the test")
@pytest.mark.parametrize("script,marker", [
    ("analyze_mcd_patient_level.py", "Loading CNN model for MC Dropout"),
    ("analyze_de_patient_level.py", "Successfully loaded 5 ensemble models."),
])
def test_reference_driver_runs_to_first_gpu_call(tmp_path, script, marker):
    fixtures = importlib.import_module(PKG_NAME + ".fixtures")
    work = str(tmp_path)
    fixtures.build_fixture_dir(work, n_unbalanced=64, n_rus=32, n_members=5)
    cwd, saved = os.getcwd(), list(sys.path)
    buf = io.StringIO()
    try:
        os.chdir(work)
        sys.path[:0] = [UQ_DIR, SHIM_DIR]
        _purge()
        with contextlib.redirect_stdout(buf):
            try:
                runpy.run_path(os.path.join(reference_loader.REFERENCE_ROOT, "uncertainty_quantification", script),
                               run_name="__main__")
            except Exception as e:      # the DE driver does not wrap its main block
                print(f"driver raised: {e}")
    finally:
        os.chdir(cwd)
        sys.path[:] = saved
        _purge()
    out = buf.getvalue()
    assert "Test datasets loaded successfully." in out
    assert marker in out
    assert "no CUDA device" in out or "no CPU fallback" in out        # loud failure, not a silent CPU path


@pytest.mark.skipif(not reference_loader.reference_available(), reason="reference tree not present")
def test_reference_drivers_run_to_completion_on_an_oracle_backed_engine(tmp_path, monkeypatch):
    """The four driver scripts and the patient aggregation script of the reference, UNCHANGED, run to the end on top of
    the drop-in module (analyze_mcd_patient_level.py:196-238, analyze_de_patient_level.py:203-234,
    evaluate_{mcd,de}_global.py, aggregate_patient_uq_metrics.py).  The GPU engine is replaced -- in this test only -- by
    an oracle-backed fake (tests/fake_engine.py), so everything around the kernels is exercised for real: model and
    data loading through the shim, the deterministic check, T = 50 passes, DataFrame assembly, to_csv, the second (RUS)
    run, evaluate_uq_methods with bootstrap CIs, and the aggregation of the CSV the MCD driver wrote."""
    import pandas as pd
    from fake_engine import FakeEngine
    uqk = importlib.import_module(PKG_NAME)
    fixtures = importlib.import_module(PKG_NAME + ".fixtures")
    engine_mod = importlib.import_module(PKG_NAME + ".engine")
    fake = FakeEngine(uqk.arch, uqk.synthetic)
    monkeypatch.setattr(engine_mod, "get_engine", lambda device=None: fake)
    monkeypatch.setenv("UQK_SEED", "7")
    work = str(tmp_path)
    n_u, n_r = 48, 24
    fixtures.build_fixture_dir(work, n_unbalanced=n_u, n_rus=n_r, n_members=5, n_global_members=20)
    ref_uq = os.path.join(reference_loader.REFERENCE_ROOT, "uncertainty_quantification")
    cwd, saved = os.getcwd(), list(sys.path)
    outs = {}
    try:
        os.chdir(work)
        sys.path[:0] = [UQ_DIR, SHIM_DIR]
        for script in ("analyze_mcd_patient_level.py", "analyze_de_patient_level.py", "evaluate_mcd_global.py", "evaluate_de_global.py"):
            _purge()
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                runpy.run_path(os.path.join(ref_uq, script), run_name="__main__")      # must not raise
            outs[script] = buf.getvalue()
        # the CSV the MCD driver wrote is what the aggregation script reads (file-name drift: SURVEY.md Appendix B)
        mcd_csv = os.path.join("uq_results_patient_no_pool", "detailed_results_CNN_MCD_Unbalanced.csv")
        de_csv = os.path.join("uq_results_patient_DE_new", "detailed_results_CNN_DE_Unbalanced.csv")
        assert os.path.isfile(mcd_csv) and os.path.isfile(de_csv)
        df = pd.read_csv(mcd_csv)
        assert list(df.columns) == ["Patient_ID", "Window_Index", "True_Label", "Predicted_Label", "Predicted_Probability",
                                    "Predictive_Variance", "Predictive_Entropy"] and len(df) == n_u
        assert len(pd.read_csv(de_csv)) == n_u
        import shutil
        shutil.copy(mcd_csv, "detail_patient_MCD.csv")
        _purge()
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            runpy.run_path(os.path.join(ref_uq, "aggregate_patient_uq_metrics.py"), run_name="__main__")
        outs["aggregate"] = buf.getvalue()
        summ = pd.read_csv(os.path.join("patient_level_uq_analysis_MCD", "patient_summary_metrics_MCD.csv"))
        assert int(summ["num_windows"].sum()) == n_u and len(summ) == df["Patient_ID"].nunique()
    finally:
        os.chdir(cwd)
        sys.path[:] = saved
        _purge()
    o = outs["analyze_mcd_patient_level.py"]
    assert "MC Dropout completed in" in o and "(50 passes)" in o and "ERROR" not in o
    o = outs["analyze_de_patient_level.py"]
    assert "Deep Ensemble completed in" in o and "(5 models)" in o
    assert "(20 models)" in outs["evaluate_de_global.py"] and "(50 passes)" in outs["evaluate_mcd_global.py"]
    kinds = [c[0] for c in fake.calls]
    # what the drivers drove through the engine: deterministic check, T passes on both data sets, ensembles, reductions, bootstrap
    assert kinds.count("mcd") >= 4 and kinds.count("de") >= 4 and "forward_det" in kinds and "bootstrap" in kinds and "reduce" in kinds
    assert ("mcd", n_u, 50, "batch", "fp32") in fake.calls and ("mcd", n_r, 50, "batch", "fp32") in fake.calls


# ---- Keras-model adapter ("read model.get_weights() in layer order, check layer class names/rate/epsilon")
def _fake_keras(uqk, **override):
    """Duck-typed stand-in for the Sequential built at models/cnn_baseline_train.py:55-94 (TensorFlow is absent)."""
    def mk(cls, **kw):
        return type(cls, (), {"get_config": lambda self, kw=kw: dict(kw), "name": kw.get("name", cls)})()
    layers = [mk("InputLayer")]
    for i, b in enumerate(uqk.arch.BLOCKS):
        layers.append(mk("Conv1D", filters=b.cout, kernel_size=(b.k,), padding="same", activation="relu", strides=(1,),
                         dilation_rate=(1,), use_bias=True, name=f"conv1d_{i+1}"))
        layers.append(mk("BatchNormalization", epsilon=override.get("epsilon", 1e-3), axis=[2], center=True, scale=True))
        layers.append(mk("Dropout", rate=override.get("rate", b.rate) if i == 2 else b.rate, noise_shape=None))
    layers.append(mk("GlobalAveragePooling1D"))
    layers.append(mk("Dense", units=1, activation=override.get("dense_act", "sigmoid"), use_bias=True))
    if override.get("drop_layer"):
        del layers[5]
    ws = uqk.synthetic.make_member_weights(0)
    return type("Sequential", (), {"layers": layers, "get_weights": lambda self: ws})()


def test_keras_model_adapter_validates_architecture(uqk):
    import pytest
    model = importlib.import_module(PKG_NAME + ".model")
    good = _fake_keras(uqk)
    blob = model.weights_of(good)
    assert blob.shape == (853441,) and blob.dtype == np.float32
    np.testing.assert_array_equal(blob, uqk.synthetic.weights_to_blob(uqk.synthetic.make_member_weights(0)))
    for bad, msg in ((dict(rate=0.25), "Dropout rate"), (dict(epsilon=1e-5), "epsilon"), (dict(dense_act="linear"), "Dense"),
                     (dict(drop_layer=True), "layer sequence")):
        with pytest.raises(ValueError, match=msg):
            model.weights_of(_fake_keras(uqk, **bad))
