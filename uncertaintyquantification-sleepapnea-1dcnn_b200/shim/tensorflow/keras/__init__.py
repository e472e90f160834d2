import importlib
import os

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
_model = importlib.import_module(os.path.basename(_PKG_DIR) + ".model")
Model = _model.UQModel

from . import models  # noqa: E402,F401
