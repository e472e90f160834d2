import importlib
import os

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))))
_model = importlib.import_module(os.path.basename(_PKG_DIR) + ".model")
Model = _model.UQModel
load_model = _model.load_model
save_model = _model.save_model
