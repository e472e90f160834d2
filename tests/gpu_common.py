"""Shared helpers of the -m gpu parity tests."""
import importlib
import functools

import numpy as np

from conftest import PKG_NAME


@functools.lru_cache(maxsize=None)
def members(n=3):
    pkg = importlib.import_module(PKG_NAME)
    return [pkg.synthetic.make_member_weights(m) for m in range(n)]


@functools.lru_cache(maxsize=None)
def engine_with_members(n=3):
    pkg = importlib.import_module(PKG_NAME)
    engine = importlib.import_module(PKG_NAME + ".engine")
    eng = engine.Engine(0)
    eng.set_members(np.stack([pkg.synthetic.weights_to_blob(w) for w in members(n)]))
    return eng


def twin_masks(seed, n, t, window_base=0):
    from oracle import philox
    pkg = importlib.import_module(PKG_NAME)
    return [philox.dropout_keep_mask(seed, np.arange(n) + window_base, t, i, b.cout, b.rate)
            for i, b in enumerate(pkg.arch.BLOCKS)]


@functools.lru_cache(maxsize=None)
def trained_weights():
    """The pseudo-trained member (tests/golden/trained_member0.keras, made by make_trained_fixture.py)."""
    import os
    from conftest import GOLDEN
    model = importlib.import_module(PKG_NAME + ".model")
    return model.load_model(os.path.join(GOLDEN, "trained_member0.keras")).get_weights()


@functools.lru_cache(maxsize=None)
def engine_trained():
    """Engine whose member 0 is the pseudo-trained model and member 1 the Glorot stress model."""
    pkg = importlib.import_module(PKG_NAME)
    engine = importlib.import_module(PKG_NAME + ".engine")
    eng = engine.Engine(0)
    eng.set_members(np.stack([pkg.synthetic.weights_to_blob(trained_weights()),
                              pkg.synthetic.weights_to_blob(members()[0])]))
    return eng


def twin_masks_all(seed, n, n_passes, window_base=0):
    """twin_masks for every pass, generated on a thread pool (NumPy releases the GIL inside its loops)."""
    from concurrent.futures import ThreadPoolExecutor
    import os
    with ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 4)) as ex:
        return list(ex.map(lambda t: twin_masks(seed, n, t, window_base), range(n_passes)))


def error_summary(err):
    e = np.abs(np.asarray(err, dtype=np.float64)).ravel()
    return {"p50": float(np.quantile(e, 0.5)), "p90": float(np.quantile(e, 0.9)), "p99": float(np.quantile(e, 0.99)),
            "max": float(e.max()), "mean": float(e.mean()), "count": int(e.size)}
