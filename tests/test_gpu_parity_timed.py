"""Parity on the configuration bench.py times (MC-Dropout T = 50, bf16 tensor-core kernel, BatchNorm statistics
"calibrated" and exact "batch") at N = 1024 windows against the float64 oracle driven by the SAME Philox masks and --
in calibrated mode -- the same frozen statistics, on a pseudo-trained model whose probabilities spread over (0, 1)
(tests/golden/trained_member0.keras) and on the Glorot model with a scaled head as a stress case.

Bounds (SURVEY.md section 8c states max |dp| <= 2e-2 and mean |dp| <= 3e-3 for the bf16 mode, "to be tightened by
measurement"): the measured distributions are written to gpurun_out/r02_parity_distribution.json (a copy is committed
under profiles/); the asserted bounds sit ~2x above the measured values.  For the trained model that is 3-5x TIGHTER
than the survey's numbers (max 6e-3, mean 6e-4 per pass under dropout).  The Glorot stress model amplifies rounding
through its scaled dense head (logit gain ~100): its dropout bounds (max 3.5e-2, mean 5e-3) stay above the survey's.
"""
import json
import os

import numpy as np
import pytest

from conftest import ROOT
from gpu_common import engine_trained, trained_weights, members, twin_masks_all, error_summary
from oracle import network, reductions

pytestmark = pytest.mark.gpu
N, T, SEED = 1024, 50, 20252
OUT = os.path.join(ROOT, "gpurun_out", "r02_parity_distribution.json")

# (max |dp| per pass, mean |dp| per pass, max |d mean_pred|): trained model / Glorot stress model
# measured (profiles/r02_parity_distribution.json): trained det 1.9e-3 / 2.4e-4, dropout T=50 2.8e-3 / 2.9e-4 / 1.5e-3;
# Glorot det 1.2e-2 / 2.6e-3, dropout T=50 2.3e-2 / 3.1e-3 / 2.3e-3
BOUNDS = {"trained": {"det": (4e-3, 5e-4, 4e-3), "calibrated": (6e-3, 6e-4, 3e-3), "batch": (6e-3, 6e-4, 3e-3)},
          "glorot": {"det": (2e-2, 4e-3, 2e-2), "calibrated": (3.5e-2, 5e-3, 5e-3), "batch": (3.5e-2, 5e-3, 5e-3)}}


def _record(key, value):
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    data = json.load(open(OUT)) if os.path.isfile(OUT) else {}
    data[key] = value
    json.dump(data, open(OUT, "w"), indent=1, sort_keys=True)


@pytest.fixture(scope="module")
def windows(uqk):
    return uqk.synthetic.make_windows(N, seed=909)


@pytest.fixture(scope="module")
def masks():
    return twin_masks_all(SEED, N, T)


def _weights(which):
    return trained_weights() if which == "trained" else members()[0]


def _n(which):
    return N if which == "trained" else 256       # the Glorot stress case runs on the first 256 windows


def _use(eng, which):
    """member 0 of engine_trained() is the trained model, member 1 the Glorot one; MC-Dropout runs member 0."""
    import importlib
    from conftest import PKG_NAME
    pkg = importlib.import_module(PKG_NAME)
    order = [trained_weights(), members()[0]] if which == "trained" else [members()[0], trained_weights()]
    eng.set_members(np.stack([pkg.synthetic.weights_to_blob(w) for w in order]))


def _check(tag, which, mode, probs, ref, out):
    e_pass = error_summary(probs - ref)
    ref_mean = ref.mean(axis=0)
    e_mean = error_summary(out["mean_pred"] - ref_mean)
    rm = reductions.uq_evaluation_dist(ref.astype(np.float32), None)
    e_var = error_summary(out["pred_variance"] - rm["pred_variance"])
    e_mi = error_summary(out["mutual_info"] - rm["mutual_info"])
    flips = int(np.sum(out["label"] != (ref_mean > 0.5)))
    band = float(np.abs(out["mean_pred"] - ref_mean).max())
    in_band = int(np.sum(np.abs(ref_mean - 0.5) <= band))
    _record(tag, {"model": which, "bn_mode": mode, "windows": int(ref.shape[1]), "passes": T, "abs_dp_per_pass": e_pass,
                  "abs_d_mean_pred": e_mean, "abs_d_variance": e_var, "abs_d_mutual_info": e_mi,
                  "label_flips_vs_f64": flips, "windows_within_max_error_of_0.5": in_band})
    b = BOUNDS[which][mode]
    assert e_pass["max"] < b[0] and e_pass["mean"] < b[1] and e_mean["max"] < b[2], (tag, e_pass, e_mean)
    outside = np.abs(ref_mean - 0.5) > band
    np.testing.assert_array_equal(out["label"][outside], (ref_mean > 0.5).astype(np.int32)[outside])
    assert flips <= in_band


@pytest.mark.parametrize("which", ["trained", "glorot"])
def test_deterministic_distribution(uqk, windows, which):
    eng = engine_trained()
    _use(eng, which)
    p = eng.forward_det(windows, member=0)
    ref = network.forward_f64_torch(_weights(which), windows)[:, 0]
    e = error_summary(p - ref)
    _record(f"det_{which}", {"model": which, "windows": N, "abs_dp": e,
                              "label_flips_vs_f64": int(np.sum((p > 0.5) != (ref > 0.5)))})
    b = BOUNDS[which]["det"]
    assert e["max"] < b[0] and e["mean"] < b[1], e


@pytest.mark.parametrize("which", ["trained", "glorot"])
def test_calibrated_T50_vs_oracle_with_the_same_statistics(uqk, windows, masks, which):
    eng = engine_trained()
    _use(eng, which)
    n = _n(which)
    windows, masks = windows[:n], [[m[:n] for m in mt] for mt in masks]
    eng.calibrate(windows, n_windows=n, n_passes=2, seed=SEED)       # what bench.py does (there: one pass over all windows)
    mean, var = eng.get_bn_stats()
    stats = [(mean[l, :b.cout].astype(np.float64), var[l, :b.cout].astype(np.float64)) for l, b in enumerate(uqk.arch.BLOCKS)]
    out = eng.mcd(windows, n_passes=T, seed=SEED, bn_mode="calibrated", precision="bf16", return_probs=True)
    ws = _weights(which)
    ref = np.stack([network.forward_f64_torch(ws, windows, training=True, masks=masks[t], bn_stats=stats)[:, 0] for t in range(T)])
    _check(f"calibrated_{which}", which, "calibrated", out["probs"], ref, out)


@pytest.mark.parametrize("which", ["trained", "glorot"])
def test_exact_batch_T50_vs_oracle(uqk, windows, masks, which):
    eng = engine_trained()
    _use(eng, which)
    n = _n(which)
    windows, masks = windows[:n], [[m[:n] for m in mt] for mt in masks]
    out = eng.mcd(windows, n_passes=T, seed=SEED, bn_mode="batch", precision="bf16", return_probs=True)
    ws = _weights(which)
    ref = np.stack([network.forward_f64_torch(ws, windows, training=True, masks=masks[t])[:, 0] for t in range(T)])
    _check(f"batch_{which}", which, "batch", out["probs"], ref, out)
