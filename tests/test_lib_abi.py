"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol include/uqk.h declares, and
fails loudly (no CPU fallback) when no GPU is present.  No compute calls here."""
import ctypes as C
import importlib
import os
import re

import numpy as np
import pytest
import torch

from conftest import PKG_NAME, ROOT


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "uqk.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(uqk_[a-z_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = importlib.import_module(PKG_NAME + "._lib")
    cdll = lib.load()
    syms = _declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(cdll, s), f"{s} declared in include/uqk.h but not exported by libuqk.so"
    assert set(lib.EXPORTS) == set(syms)
    assert cdll.uqk_version() == 101


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU error path")
def test_no_cpu_fallback_without_gpu():
    lib = importlib.import_module(PKG_NAME + "._lib")
    cdll = lib.load()
    h = C.c_void_p()
    rc = cdll.uqk_create(0, C.byref(h))
    assert rc < 0 and b"no CPU fallback" in cdll.uqk_last_error()
    engine = importlib.import_module(PKG_NAME + ".engine")
    with pytest.raises(lib.UqkError):
        engine.Engine(0)
    # the drop-in module raises too instead of silently computing on the CPU
    import sys
    sys.path.insert(0, os.path.join(ROOT, PKG_NAME, "uncertainty_quantification"))
    uq = importlib.import_module("uq_techniques")
    with pytest.raises(lib.UqkError):
        uq.uq_evaluation_dist(np.random.rand(3, 5).astype(np.float32), np.zeros(5, dtype=int))


def test_argument_errors_without_gpu_calls():
    lib = importlib.import_module(PKG_NAME + "._lib")
    cdll = lib.load()
    assert cdll.uqk_reduce(None, 0, 10, None, None) == -1          # n_passes < 1
    assert cdll.uqk_set_model(None, 1, None, 0) == -1
    assert cdll.uqk_patient_reduce(None, None, None, None, None, -1, None, None) == -1
    with pytest.raises(ValueError):
        lib.check(cdll.uqk_reduce(None, 0, 10, None, None))


def test_model_container_and_shim(tmp_path, uqk):
    import sys
    model = importlib.import_module(PKG_NAME + ".model")
    m = model.synthetic_model(1)
    assert len(m.get_weights()) == 38 and m.count_params() == 853441
    path = str(tmp_path / "AlCNN1D_no_pool.keras")
    m.save(path)
    sys.path.insert(0, os.path.join(ROOT, PKG_NAME, "shim"))
    try:
        for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.")]:
            del sys.modules[k]
        import tensorflow as tf
        from tensorflow.keras.models import load_model
        tf.random.set_seed(123)
        m2 = load_model(path)
        assert model.UQModel._global_seed == 123
        assert isinstance(m2, tf.keras.Model)
        np.testing.assert_array_equal(m2.blob(), m.blob())
    finally:
        sys.path.remove(os.path.join(ROOT, PKG_NAME, "shim"))
        for k in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.")]:
            del sys.modules[k]
        model.UQModel._global_seed = 2025
    # the engine's upload cache is keyed by weight content, not object identity
    fp = model.fingerprint(m.blob())
    ws = m.get_weights()
    ws[1] = ws[1] + 1.0
    m.set_weights(ws)
    assert model.fingerprint(m.blob()) != fp and model.fingerprint(model.synthetic_model(1).blob()) == fp
    with pytest.raises(ValueError):
        model.UQModel(m.get_weights()[:-1])
    bad = tmp_path / "bad.keras"
    bad.write_bytes(b"not a zip")
    with pytest.raises(ValueError):
        model.load_model(str(bad))


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package (nor the ctypes layer) may import it; bench.py may do
    so only inside its CPU-baseline legs."""
    import re
    pat = re.compile(r"^\s*(from\s+oracle\b|import\s+oracle\b|from\s+tests\b|import\s+fake_engine\b)", re.M)
    pkg_dir = os.path.join(ROOT, PKG_NAME)
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert not pat.search(text), f"{os.path.join(dirpath, f)} imports test infrastructure"
    bench = open(os.path.join(ROOT, "bench.py")).read()
    for m in pat.finditer(bench):
        before = bench[:m.start()]
        fn = re.findall(r"^def (\w+)\(", before, re.M)[-1]
        assert fn in ("cpu_mcd_sample", "cpu_config0_sample"), f"bench.py imports the oracle inside {fn}()"
    import subprocess, sys
    code = (f"import sys, importlib; sys.path.insert(0, {ROOT!r}); importlib.import_module({PKG_NAME!r} + '.engine'); "
            f"importlib.import_module({PKG_NAME!r} + '.model'); importlib.import_module({PKG_NAME!r} + '.sharding'); "
            "assert not [m for m in sys.modules if m == 'oracle' or m.startswith('oracle.')], 'oracle got imported'")
    subprocess.run([sys.executable, "-c", code], check=True)
