"""B200-native UQ inference hot path for the sleep-apnea 1D-CNN (MC-Dropout + Deep-Ensemble forward
passes and the per-window UQ reductions), behind the call surface of the reference's
``uncertainty_quantification/uq_techniques.py``.

The directory name contains hyphens, so import it with::

    import importlib
    uqk = importlib.import_module("uncertaintyquantification-sleepapnea-1dcnn_b200")

Sub-modules (``arch``, ``synthetic``, ``model`` are pure host code; ``_lib`` / ``engine`` need the built
CUDA library ``libuqk.so`` and fail loudly without it -- there is no CPU fallback).
"""
from . import arch, synthetic  # noqa: F401

__all__ = ["arch", "synthetic"]


def __getattr__(name):
    # Lazy so that host-only users (oracle, fixtures) do not need the CUDA library.
    if name in ("model", "engine", "_lib", "sharding", "patient", "tf_shim", "emit"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
