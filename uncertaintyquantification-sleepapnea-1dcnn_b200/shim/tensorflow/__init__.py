"""Minimal stand-in for the ``tensorflow`` package, for images that have no TensorFlow.

The unchanged reference drivers import TensorFlow only for three things
(analyze_mcd_patient_level.py:5-6,31,199; analyze_de_patient_level.py:5-6,31,48):
``tf.random.set_seed``, ``tf.keras.Model`` (type annotations) and ``tensorflow.keras.models.load_model``.
Put this directory on ``sys.path`` ONLY when the real package is absent; ``load_model`` then returns the GPU-backed
``UQModel`` of this build.
"""
import importlib
import os
import sys

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
_ROOT = os.path.dirname(_PKG_DIR)
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
_model = importlib.import_module(os.path.basename(_PKG_DIR) + ".model")

__version__ = "0.0-uqk-shim"

from . import keras  # noqa: E402,F401


class _Random:
    @staticmethod
    def set_seed(seed):
        """Seeds the Philox key of the dropout masks (the reference seeds TF's global generator)."""
        _model.UQModel._global_seed = int(seed)
        _model.UQModel._seed_epoch += 1


random = _Random()
