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
