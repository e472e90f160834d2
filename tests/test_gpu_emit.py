"""N2 on the GPU: ``window_level_summary`` against the reference script's own pandas code
(uncertainty_quantification/analyze_window_level_uncertainty.py:30-64), restated here on a DataFrame because the script
has no functions to import (it runs at module level on a hard-coded CSV path)."""
import importlib

import numpy as np
import pandas as pd
import pytest

from conftest import PKG_NAME

pytestmark = pytest.mark.gpu
NUM_BINS = 10


def _reference_summary(df):
    """Lines 30-64 of the reference script, prints removed."""
    df = df.copy()
    df['Correct'] = (df['True_Label'] == df['Predicted_Label'])
    cols = ['Predictive_Entropy', 'Predictive_Variance']
    out = {"windows": len(df), "accuracy": df['Correct'].mean(),
           "correct": df[df['Correct'] == True][cols].describe(),
           "incorrect": df[df['Correct'] == False][cols].describe()}
    metric = 'Predictive_Entropy'
    bins = np.linspace(df[metric].min(), df[metric].max() + 1e-9, NUM_BINS + 1)
    labels = [f'{bins[i]:.3f}-{bins[i+1]:.3f}' for i in range(NUM_BINS)]
    df[f'{metric}_Bin'] = pd.cut(df[metric], bins=bins, labels=labels, right=False)
    binned = df.groupby(f'{metric}_Bin', observed=False).agg(window_count=('Correct', 'size'), accuracy=('Correct', 'mean'))
    binned['error_rate'] = 1 - binned['accuracy']
    out.update(bins=bins, labels=labels, binned=binned)
    return out


def _frame(n, seed, all_correct=False):
    rng = np.random.default_rng(seed)
    probs = rng.random((9, n)).astype(np.float32) ** 2
    mean = probs.mean(axis=0)
    var = probs.var(axis=0)
    ent = -(mean * np.log2(mean + 1e-9) + (1 - mean) * np.log2(1 - mean + 1e-9))
    lab = (mean > 0.5).astype(int)
    y = lab.copy() if all_correct else np.where(rng.random(n) < 0.8, lab, 1 - lab)
    return pd.DataFrame({'Patient_ID': 200001 + np.arange(n) // 500, 'Window_Index': np.arange(n), 'True_Label': y,
                         'Predicted_Label': lab, 'Predicted_Probability': mean, 'Predictive_Variance': var,
                         'Predictive_Entropy': ent})


def _check(got, ref, rtol):
    assert got["windows"] == ref["windows"]
    np.testing.assert_allclose(got["accuracy"], ref["accuracy"], rtol=1e-12)
    for cname in ("correct", "incorrect"):
        for col in ("Predictive_Entropy", "Predictive_Variance"):
            d = ref[cname][col]
            for row in ("count", "mean", "std", "min", "25%", "50%", "75%", "max"):
                want = float(d[row])
                np.testing.assert_allclose(got[cname][col][row], want, rtol=rtol, atol=1e-300, equal_nan=True,
                                           err_msg=f"{cname}/{col}/{row}")
    np.testing.assert_allclose(got["bin_edges"], ref["bins"], rtol=rtol)
    assert got["bin_labels"] == ref["labels"]
    np.testing.assert_array_equal(got["window_count"], ref["binned"]["window_count"].values)
    np.testing.assert_allclose(got["bin_accuracy"], ref["binned"]["accuracy"].values, rtol=1e-12, equal_nan=True)
    np.testing.assert_allclose(got["error_rate"], ref["binned"]["error_rate"].values, rtol=1e-12, atol=1e-15, equal_nan=True)


@pytest.mark.parametrize("n", [2, 7, 1000, 300_001])     # (the reference itself fails on one window: duplicate bin labels)
def test_window_summary_matches_the_reference_pandas_code(n):
    emit = importlib.import_module(PKG_NAME + ".emit")
    df = _frame(n, seed=n)
    got = emit.window_level_summary(df["Predictive_Entropy"].values, df["Predictive_Variance"].values,
                                    df["True_Label"].values, df["Predicted_Label"].values, n_bins=NUM_BINS)
    df64 = df.astype({"Predictive_Entropy": np.float64, "Predictive_Variance": np.float64})
    _check(got, _reference_summary(df64), rtol=1e-11)


def test_window_summary_through_the_csv_like_the_reference(tmp_path):
    """The reference reads the detail CSV back (decimal text -> float64): same numbers within the float32 text
    round-trip (6e-8 relative); bin populations are identical for this fixture."""
    emit = importlib.import_module(PKG_NAME + ".emit")
    df = _frame(50_000, seed=5)
    path = str(tmp_path / "detail.csv")
    emit.write_detailed_csv(path, df["Patient_ID"].values, df["True_Label"].values, df["Predicted_Label"].values,
                            df["Predicted_Probability"].values, df["Predictive_Variance"].values,
                            df["Predictive_Entropy"].values)
    ref = _reference_summary(pd.read_csv(path))
    got = emit.window_level_summary(df["Predictive_Entropy"].values, df["Predictive_Variance"].values,
                                    df["True_Label"].values, df["Predicted_Label"].values, n_bins=NUM_BINS)
    _check(got, ref, rtol=2e-7)


def test_window_summary_degenerate_classes():
    emit = importlib.import_module(PKG_NAME + ".emit")
    df = _frame(500, seed=2, all_correct=True)            # no incorrect window: describe() of an empty frame is NaN
    got = emit.window_level_summary(df["Predictive_Entropy"].values, df["Predictive_Variance"].values,
                                    df["True_Label"].values, df["Predicted_Label"].values)
    df64 = df.astype({"Predictive_Entropy": np.float64, "Predictive_Variance": np.float64})
    _check(got, _reference_summary(df64), rtol=1e-11)
    assert got["accuracy"] == 1.0 and got["incorrect"]["Predictive_Entropy"]["count"] == 0
    with pytest.raises(ValueError):
        emit.window_level_summary(df["Predictive_Entropy"].values[:-1], df["Predictive_Variance"].values,
                                  df["True_Label"].values, df["Predicted_Label"].values)
    with pytest.raises(ValueError):
        emit.window_level_summary(df["Predictive_Entropy"].values, df["Predictive_Variance"].values,
                                  df["True_Label"].values, df["Predicted_Label"].values, n_bins=65)
