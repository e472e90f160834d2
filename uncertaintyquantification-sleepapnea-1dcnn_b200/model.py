"""Model objects the reference callers hand to the UQ functions.

The reference passes whatever ``tensorflow.keras.models.load_model(path)`` returned
(analyze_mcd_patient_level.py:199; analyze_de_patient_level.py:48) and uses it in three ways:
``model(x, training=True)`` (uq_techniques.py:22), ``model(x, training=False).numpy()``
(analyze_mcd_patient_level.py:204) and ``model.predict(x, verbose=0)`` (uq_techniques.py:29).
``UQModel`` supports exactly those on the GPU; ``weights_of`` also accepts a genuine Keras model
(duck-typed through ``get_weights()``) when TensorFlow exists.

File formats: ``load_model`` reads genuine Keras v3 archives (zip of config.json + model.weights.h5, parsed by
``keras_archive.py`` without h5py/TensorFlow) and this build's own container: an ``.npz`` payload of the 38
``get_weights()`` arrays stored under the ``*.keras`` file name the callers expect, which is
what ``save_model`` writes.
"""
from __future__ import annotations

import io
import zipfile
from typing import List, Sequence

import numpy as np

from . import arch, synthetic

_MAGIC_KEY = "uqk_format"


class EagerArray(np.ndarray):
    """ndarray that also answers ``.numpy()`` like a TF EagerTensor (analyze_mcd_patient_level.py:204)."""

    def numpy(self):
        return np.asarray(self)


class UQModel:
    """The 6-block 1D-CNN (models/cnn_baseline_train.py:55-94) as a bag of weights + GPU forward."""

    _global_seed = 2025   # set by tf.random.set_seed in the shim (analyze_mcd_patient_level.py:31)
    _seed_epoch = 0       # bumped by every tf.random.set_seed: the per-call key counter restarts, like TF's generator

    def __init__(self, weights: Sequence[np.ndarray], name: str = "Alarcon_1D_CNN_Model"):
        self._blob = synthetic.weights_to_blob(weights)   # validates the 38 shapes
        self.name = name
        self.input_shape = (None, arch.WINDOW_LEN, arch.WINDOW_FEATURES)
        self._stochastic_calls = 0

    # Keras-like surface -------------------------------------------------------------------------------------------
    def get_weights(self) -> List[np.ndarray]:
        return synthetic.blob_to_weights(self._blob)

    def set_weights(self, weights):
        self._blob = synthetic.weights_to_blob(weights)

    def count_params(self) -> int:
        return int(self._blob.size)

    def save(self, path: str):
        save_model(self, path)

    def blob(self) -> np.ndarray:
        return self._blob

    def __call__(self, x, training: bool = False):
        from . import engine as _engine
        eng = _engine.get_engine()
        eng.ensure_members(lambda: self._blob[None], key=("single", fingerprint(self._blob)))
        x = np.asarray(x)
        if training:
            # ONE stochastic pass with exact Keras semantics (batch-statistic BN + dropout); every call draws
            # new masks: the call counter is folded into the Philox key.
            seed = (UQModel._global_seed * 0x9E3779B97F4A7C15 + self._stochastic_calls) & (2 ** 64 - 1)
            self._stochastic_calls += 1
            # fp32 layer-synchronous kernels while the call's activations fit comfortably in HBM, the swept tensor-core
            # kernel above that (same semantics, bf16 operands) -- the threshold mc_dropout_predict uses
            import os
            small = len(x) <= int(os.environ.get("UQK_EXACT_MAX_WINDOWS", "16384"))
            out = eng.mcd(x, n_passes=1, seed=seed, bn_mode="batch", precision="fp32" if small else "bf16",
                          return_probs=True, metrics=())
            p = out["probs"].reshape(-1, 1)
        else:
            p = eng.forward_det(x, member=0, precision=_default_precision()).reshape(-1, 1)
        return p.astype(np.float32).view(EagerArray)

    def predict(self, x, verbose=0, batch_size=32, **_):
        # Keras batches by 32 (uq_techniques.py:29); windows are independent in inference mode, so batching
        # does not change the result and the GPU path takes the whole array.
        return np.asarray(self(x, training=False))


def _default_precision() -> str:
    """Single deterministic passes (``model(x, training=False)``, ``model.predict``) run in fp32 like the reference
    unless UQK_PRECISION says otherwise: one pass is cheap, and labels / accuracy checks derived from it
    (analyze_mcd_patient_level.py:204-207, evaluate_classification.py:49) then need no guard band."""
    import os
    return os.environ.get("UQK_PRECISION", "fp32")


def fingerprint(blob: np.ndarray) -> int:
    """Content hash of a weight blob: the engine re-uploads a model when this changes (object identity is not a safe
    cache key -- ids are reused after garbage collection and weights can be changed with ``set_weights``)."""
    import zlib
    return zlib.crc32(memoryview(np.ascontiguousarray(blob)).cast("B"))


def _cfg(layer, name, default=None):
    """Attribute of a Keras layer, falling back to ``get_config()`` (both exist on real layers)."""
    if hasattr(layer, name):
        return getattr(layer, name)
    if hasattr(layer, "get_config"):
        return layer.get_config().get(name, default)
    return default


def _first(v):
    return v[0] if isinstance(v, (tuple, list)) else v


def _act_name(a) -> str:
    return a if isinstance(a, str) else getattr(a, "__name__", str(a))


def validate_keras_model(model) -> None:
    """Checks that a genuine Keras model has exactly the architecture the kernels implement
    (models/cnn_baseline_train.py:55-94): layer classes in order, filters / kernel_size / padding / activation of
    every Conv1D, BatchNormalization epsilon, Dropout rates, the pooling layer and the sigmoid Dense(1).  The weight
    SHAPES are validated separately by ``weights_to_blob``; this catches what shapes cannot (a different dropout rate,
    epsilon, activation or padding would silently change the numbers).  Raises ValueError with the first mismatch."""
    layers = [l for l in getattr(model, "layers", []) if type(l).__name__ != "InputLayer"]
    want = []
    for b in arch.BLOCKS:
        want += [("Conv1D", b), ("BatchNormalization", b), ("Dropout", b)]
    want += [("GlobalAveragePooling1D", None), ("Dense", None)]
    names = [type(l).__name__ for l in layers]
    if names != [w[0] for w in want]:
        raise ValueError(f"layer sequence {names} is not the Alarcon 1D-CNN ({[w[0] for w in want]})")
    for i, (layer, (kind, b)) in enumerate(zip(layers, want)):
        where = f"layer {i} ({getattr(layer, 'name', kind)})"
        if kind == "Conv1D":
            got = (int(_cfg(layer, "filters")), int(_first(_cfg(layer, "kernel_size"))), str(_cfg(layer, "padding")).lower(),
                   _act_name(_cfg(layer, "activation")), int(_first(_cfg(layer, "strides", 1))),
                   int(_first(_cfg(layer, "dilation_rate", 1))), bool(_cfg(layer, "use_bias", True)))
            exp = (b.cout, b.k, "same", "relu", 1, 1, True)
            if got != exp:
                raise ValueError(f"{where}: Conv1D (filters, kernel, padding, activation, stride, dilation, bias) = {got}, need {exp}")
        elif kind == "BatchNormalization":
            eps = float(_cfg(layer, "epsilon"))
            if abs(eps - arch.BN_EPS) > 1e-12:
                raise ValueError(f"{where}: BatchNormalization epsilon {eps}, need {arch.BN_EPS}")
            axis = _cfg(layer, "axis", -1)
            axis = _first(list(axis) if isinstance(axis, (list, tuple)) else [axis])
            if axis not in (-1, 2):
                raise ValueError(f"{where}: BatchNormalization axis {axis}, need the channel axis")
            if not (_cfg(layer, "center", True) and _cfg(layer, "scale", True)):
                raise ValueError(f"{where}: BatchNormalization needs center=True and scale=True")
        elif kind == "Dropout":
            rate = float(_cfg(layer, "rate"))
            if abs(rate - b.rate) > 1e-9:
                raise ValueError(f"{where}: Dropout rate {rate}, need {b.rate}")
            if _cfg(layer, "noise_shape", None) is not None:
                raise ValueError(f"{where}: Dropout noise_shape is not supported (masks are per element)")
        elif kind == "Dense":
            got = (int(_cfg(layer, "units")), _act_name(_cfg(layer, "activation")), bool(_cfg(layer, "use_bias", True)))
            if got != (1, "sigmoid", True):
                raise ValueError(f"{where}: Dense (units, activation, bias) = {got}, need (1, 'sigmoid', True)")


def weights_of(model) -> np.ndarray:
    """Blob (853441,) float32 of a ``UQModel`` or of any object with Keras' ``get_weights()`` (a genuine Keras model,
    whose layer list is validated first when it exposes one)."""
    if isinstance(model, UQModel):
        return model.blob()
    if hasattr(model, "get_weights"):
        if hasattr(model, "layers"):
            validate_keras_model(model)
        return synthetic.weights_to_blob(model.get_weights())
    raise TypeError(f"cannot extract weights from {type(model).__name__}: need UQModel or a Keras model")


def save_model(model, path: str):
    ws = synthetic.blob_to_weights(weights_of(model))
    buf = io.BytesIO()
    np.savez(buf, **{_MAGIC_KEY: np.array([1]), **{f"w{i:02d}": w for i, w in enumerate(ws)}})
    with open(path, "wb") as f:
        f.write(buf.getvalue())


def load_model(path: str, compile: bool = True, **_) -> UQModel:   # noqa: A002 - Keras signature
    with open(path, "rb") as f:
        data = f.read()
    try:
        with zipfile.ZipFile(io.BytesIO(data)) as z:
            names = set(z.namelist())
    except zipfile.BadZipFile:
        raise ValueError(f"{path}: not a model container this build can read")
    if "model.weights.h5" in names or "config.json" in names:
        # a genuine Keras v3 archive (model.save("*.keras"), cnn_baseline_train.py:230): config.json is validated against
        # the architecture the kernels implement, the variables come from model.weights.h5 (keras_archive.py)
        from . import keras_archive
        layers, ws = keras_archive.weights_from_keras_archive(data)
        validate_keras_model(type("KerasArchive", (), {"layers": layers})())
        return UQModel(ws)
    z = np.load(io.BytesIO(data), allow_pickle=False)
    ws = [z[f"w{i:02d}"] for i in range(len(arch.blob_layout()))]
    return UQModel(ws)


def synthetic_model(member: int = 0) -> UQModel:
    return UQModel(synthetic.make_member_weights(member))
