"""The oracle's reduction restatement vs golden vectors produced by the reference's own code
(tests/golden/make_golden.py), and -- when /root/reference is present -- vs the reference module live."""
import os

import numpy as np
import pandas as pd
import pytest

from oracle import reductions, patient, reference_loader


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=False)


def test_dummy_recipe_matches_reference(golden_dir):
    g = _load(golden_dir, "uq_dummy_golden.npz")
    import sys
    sys.path.insert(0, golden_dir)
    from make_golden import dummy_recipe
    preds, labels = dummy_recipe()
    m = reductions.uq_evaluation_dist(preds, labels)
    for k, v in m.items():
        np.testing.assert_allclose(np.asarray(v), g[f"m_{k}"], rtol=1e-12, atol=1e-14, err_msg=k)
    boot = reductions.bootstrap_aggregates_via_gather(m, labels, 50, 42)
    keys = [str(k) for k in g["boot_keys"]]
    mine = np.array([[float(b[k]) for k in keys] for b in boot])
    np.testing.assert_allclose(mine, g["boot_values"], rtol=1e-10, atol=1e-13)


def test_f32_cases_match_reference(golden_dir):
    g = _load(golden_dir, "uq_f32_golden.npz")
    p, y = g["probs"], g["y"]
    m = reductions.uq_evaluation_dist(p, y)
    for k, v in m.items():
        ref = g[f"m_{k}"]
        assert np.asarray(v).dtype == ref.dtype, k
        np.testing.assert_allclose(np.asarray(v), ref, rtol=0, atol=2e-7, err_msg=k)
    m1 = reductions.uq_evaluation_dist(p[:1], y)
    np.testing.assert_allclose(m1["pred_variance"], g["t1_pred_variance"], atol=0)
    np.testing.assert_allclose(m1["mutual_info"], g["t1_mutual_info"], atol=1e-7)
    mz = reductions.uq_evaluation_dist(p, np.zeros_like(y))
    assert float(mz["mean_variance_class_1"]) == float(g["allzero_c1"]) == 0.0
    mean_p, var_p, ent, label = reductions.window_csv_metrics(p)
    np.testing.assert_array_equal(label, g["csv_label"])
    np.testing.assert_allclose(mean_p, g["csv_mean"], atol=0)
    np.testing.assert_allclose(var_p, g["csv_var"], atol=0)
    np.testing.assert_allclose(ent, g["csv_entropy"], atol=1e-7)
    np.testing.assert_allclose(reductions.safe_entropy(np.array([[1.0, 0.0], [0.5, 0.5], [0.3, 0.7]])),
                               g["safe_entropy_probe"], rtol=1e-12)


def test_patient_aggregation_matches_reference_script(golden_dir):
    df = pd.read_csv(os.path.join(golden_dir, "patient_detail_fixture.csv"))
    ref = pd.read_csv(os.path.join(golden_dir, "patient_summary_golden.csv"))
    mine = patient.aggregate_patients(df)
    assert list(mine.columns) == list(ref.columns)
    np.testing.assert_array_equal(mine["Patient_ID"].values, ref["Patient_ID"].values)
    np.testing.assert_array_equal(mine["num_windows"].values, ref["num_windows"].values)
    for c in ref.columns[1:-1]:
        np.testing.assert_allclose(mine[c].values.astype(float), ref[c].values.astype(float), rtol=1e-12, atol=1e-15)


@pytest.mark.skipif(not reference_loader.reference_available(), reason="reference tree not present (GPU box)")
def test_live_reference_agrees_with_restatement():
    ref = reference_loader.load_reference_uq_techniques()
    rng = np.random.default_rng(3)
    p = rng.random((7, 33, 1)).astype(np.float32)
    y = (rng.random(33) < 0.5).astype(int)
    a = ref.uq_evaluation_dist(p, y)
    b = reductions.uq_evaluation_dist(p, y)
    for k in a:
        np.testing.assert_allclose(np.asarray(a[k]), np.asarray(b[k]), atol=2e-7, err_msg=k)
