"""Bit-exact labels (BASELINE.json north_star: "argmax labels ... must be bit-exact"; the reference's label is
`mean_p > 0.5`, analyze_mcd_patient_level.py:117, evaluate_classification.py:49).  The bf16 tensor-core path flags every
window whose mean probability lies within the guard band of 0.5 and re-evaluates it on the fp32 path with the same
masks and the same BatchNorm statistics (uqk_refine_band).  Checked against the float64 oracle: labels are equal for
every window outside the fp32 noise band (|mean - 0.5| <= 1e-5), re-evaluated windows carry fp32-accurate outputs and
all other windows are untouched."""
import numpy as np
import pytest

from gpu_common import engine_trained, trained_weights, members, twin_masks_all
from oracle import network, reductions

pytestmark = pytest.mark.gpu
FP32_NOISE = 1e-5


def _engine(uqk):
    eng = engine_trained()
    eng.set_members(np.stack([uqk.synthetic.weights_to_blob(w) for w in (trained_weights(), members()[0])]))
    return eng


def _check(plain, exact, ref_probs, band, delta, tol_mean):
    ref_mean = ref_probs.mean(axis=0)
    refined = np.abs(plain["mean_pred"] - 0.5) < delta
    assert band == (int(refined.sum()), len(ref_mean)) and refined.sum() > 0
    # windows outside the band are untouched, bit for bit
    for k in ("mean_pred", "pred_variance", "mutual_info", "entropy_bits", "label"):
        np.testing.assert_array_equal(exact[k][~refined], plain[k][~refined], err_msg=k)
    np.testing.assert_array_equal(exact["probs"][:, ~refined], plain["probs"][:, ~refined])
    # re-evaluated windows: fp32 accuracy, all outputs consistent with their probabilities
    assert np.abs(exact["mean_pred"][refined] - ref_mean[refined]).max() < tol_mean
    m = reductions.uq_evaluation_dist(exact["probs"], None)
    for k in ("mean_pred", "pred_variance", "mutual_info"):
        np.testing.assert_allclose(exact[k], m[k], atol=1e-6, err_msg=k)
    # labels: exact outside the fp32 noise band
    clear = np.abs(ref_mean - 0.5) > max(FP32_NOISE, tol_mean)
    np.testing.assert_array_equal(exact["label"][clear], (ref_mean > 0.5).astype(np.int32)[clear])
    return int(np.sum(plain["label"] != (ref_mean > 0.5))), int(np.sum(exact["label"] != (ref_mean > 0.5)))


def test_calibrated_mcd_labels_are_exact(uqk):
    eng = _engine(uqk)
    n, T, seed, delta = 2048, 8, 77, 5e-3
    x = uqk.synthetic.make_windows(n, seed=515)
    eng.calibrate(x, n_windows=n, n_passes=2, seed=seed)
    mean, var = eng.get_bn_stats()
    stats = [(mean[l, :b.cout].astype(np.float64), var[l, :b.cout].astype(np.float64)) for l, b in enumerate(uqk.arch.BLOCKS)]
    plain = eng.mcd(x, n_passes=T, seed=seed, bn_mode="calibrated", return_probs=True)
    exact = eng.mcd(x, n_passes=T, seed=seed, bn_mode="calibrated", return_probs=True, exact_labels=True, guard_delta=delta)
    masks = twin_masks_all(seed, n, T)
    ref = np.stack([network.forward_f64_torch(trained_weights(), x, training=True, masks=masks[t], bn_stats=stats)[:, 0] for t in range(T)])
    before, after = _check(plain, exact, ref, eng.last_guard_band, delta, FP32_NOISE)
    assert after <= before
    # a shard with a window base re-evaluates with the masks of the global positions
    part = eng.mcd(x[1000:1300], n_passes=T, seed=seed, bn_mode="calibrated", return_probs=True, exact_labels=True,
                   guard_delta=delta, window_base=1000)
    np.testing.assert_array_equal(part["probs"], exact["probs"][:, 1000:1300])


def test_ensemble_and_inference_mode_labels_are_exact(uqk):
    eng = _engine(uqk)
    n, delta = 3000, 5e-3
    x = uqk.synthetic.make_windows(n, seed=516)
    plain = eng.de(x, n_members=2, return_probs=True)
    exact = eng.de(x, n_members=2, return_probs=True, exact_labels=True, guard_delta=delta)
    ref = np.stack([network.forward_f64_torch(w, x)[:, 0] for w in (trained_weights(), members()[0])])
    _check(plain, exact, ref, eng.last_guard_band, delta, FP32_NOISE)
    fp32 = eng.de(x, n_members=2, precision="fp32", return_probs=True)
    band = np.abs(plain["mean_pred"] - 0.5) < delta
    np.testing.assert_array_equal(exact["probs"][:, band], fp32["probs"][:, band])      # the re-evaluation IS the fp32 path


def test_exact_batch_mode_labels_follow_the_fp32_path(uqk):
    """UQK_BN_BATCH: the re-evaluation uses the per-pass batch statistics the swept bf16 run measured (statistics of
    bf16-rounded activations), so against the float64 oracle with ITS batch statistics the re-evaluated means agree to
    the accuracy of those statistics (5e-4) rather than to fp32 rounding."""
    eng = _engine(uqk)
    n, T, seed, delta = 1024, 6, 78, 5e-3
    x = uqk.synthetic.make_windows(n, seed=517)
    plain = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
    exact = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True, exact_labels=True, guard_delta=delta)
    masks = twin_masks_all(seed, n, T)
    ref = np.stack([network.forward_f64_torch(trained_weights(), x, training=True, masks=masks[t])[:, 0] for t in range(T)])
    _check(plain, exact, ref, eng.last_guard_band, delta, 5e-4)


def test_refine_band_argument_checks(uqk):
    eng = _engine(uqk)
    x = uqk.synthetic.make_windows(64, seed=518)
    eng.mcd(x, n_passes=2, seed=1, bn_mode="batch", precision="bf16")
    # a batch-mode re-evaluation must follow ITS OWN swept call: another seed is refused
    import importlib
    from conftest import PKG_NAME
    lib = importlib.import_module(PKG_NAME + "._lib")
    out = eng.mcd(x, n_passes=2, seed=2, bn_mode="inference", device_out=True)
    mo = lib.MetricsOut()
    mo.mean_pred = out["mean_pred"].data_ptr()
    xt, dt = eng.to_device(x)
    with pytest.raises(lib.UqkError):
        eng._refine(0, xt, dt, 64, 2, 2, lib.BN_BATCH, 0, None, 0.01, None, mo)
    all_in = eng.mcd(x, n_passes=2, seed=3, bn_mode="inference", exact_labels=True, guard_delta=1.0, return_probs=True)
    assert eng.last_guard_band == (64, 64)
    fp32 = eng.mcd(x, n_passes=2, seed=3, bn_mode="inference", precision="fp32", return_probs=True)
    np.testing.assert_array_equal(all_in["probs"], fp32["probs"])
    none = eng.mcd(x, n_passes=2, seed=3, bn_mode="inference", exact_labels=True, guard_delta=0.0)
    assert eng.last_guard_band == (0, 64) and none["label"].shape == (64,)
