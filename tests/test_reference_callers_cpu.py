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
