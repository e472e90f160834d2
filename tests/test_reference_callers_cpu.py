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
