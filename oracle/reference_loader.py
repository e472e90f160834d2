"""Import the reference's own ``uq_techniques.py`` from /root/reference.  TEST INFRASTRUCTURE.

Only usable in the build container (``/root/reference`` does not exist on the GPU box): used by
tests/golden/make_golden.py to generate the committed golden vectors and by the ``not gpu`` tests to
cross-check the oracle restatement when the reference tree is present.

TensorFlow and matplotlib are absent from this image, and the reference module imports them at the top
(uq_techniques.py:3,6,7) although none of the reduction functions uses them; they are stubbed in
``sys.modules`` for the duration of the import.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("UQK_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "uncertainty_quantification", "uq_techniques.py"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    return m


def load_reference_uq_techniques():
    """Returns the reference module object (its own code, unmodified)."""
    if not reference_available():
        raise FileNotFoundError(f"reference tree not found under {REFERENCE_ROOT}")
    noop = lambda *a, **k: None

    class _Anything:
        def __getattr__(self, name):
            return noop

    saved = {k: sys.modules.get(k) for k in
             ("tensorflow", "tensorflow.keras", "tensorflow.keras.models", "matplotlib", "matplotlib.pyplot")}
    keras_models = _stub("tensorflow.keras.models", Model=object, load_model=noop)
    keras = _stub("tensorflow.keras", models=keras_models, Model=object)
    tf = _stub("tensorflow", keras=keras)
    plt = _stub("matplotlib.pyplot")
    plt.__getattr__ = lambda name: noop  # type: ignore[attr-defined]
    mpl = _stub("matplotlib", pyplot=plt)
    sys.modules.update({"tensorflow": tf, "tensorflow.keras": keras, "tensorflow.keras.models": keras_models,
                        "matplotlib": mpl, "matplotlib.pyplot": plt})
    try:
        path = os.path.join(REFERENCE_ROOT, "uncertainty_quantification", "uq_techniques.py")
        spec = importlib.util.spec_from_file_location("_reference_uq_techniques", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod
