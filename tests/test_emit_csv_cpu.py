"""N2: the native detail-CSV writer against pandas' own ``DataFrame.to_csv(index=False)`` on the DataFrame the reference
builds (analyze_mcd_patient_level.py:134-152) -- byte for byte.  Host code only (no GPU)."""
import importlib
import os

import numpy as np
import pandas as pd

from conftest import PKG_NAME


def _reference_csv(path, pid, y, lab, prob, var, ent):
    df = pd.DataFrame({                                   # same constructor as the reference, same dtypes
        'Patient_ID': pid,
        'Window_Index': np.arange(len(y)),
        'True_Label': y,
        'Predicted_Label': lab,
        'Predicted_Probability': prob,
        'Predictive_Variance': var,
        'Predictive_Entropy': ent,
    })
    df.to_csv(path, index=False)


def _columns(n, seed):
    rng = np.random.default_rng(seed)
    probs = rng.random((7, n)).astype(np.float32) ** 3                      # (T, N) like mc_probabilities
    mean = np.mean(probs, axis=0).flatten()
    var = np.var(probs, axis=0).flatten()
    eps = 1e-9
    ent = -(mean * np.log2(mean + eps) + (1 - mean) * np.log2(1 - mean + eps))   # reference :113-115 (stays float32)
    lab = (mean > 0.5).astype(int)
    y = rng.integers(0, 2, n)
    pid = 200001 + np.sort(rng.integers(0, 40, n)).astype(np.int64)
    assert mean.dtype == var.dtype == ent.dtype == np.float32
    return pid, y, lab, mean, var, ent


def test_csv_bytes_equal_pandas(tmp_path):
    emit = importlib.import_module(PKG_NAME + ".emit")
    pid, y, lab, mean, var, ent = _columns(200_000, 3)
    # formatting corner cases: zeros, ones, subnormals, the 1e-4 / 1e16 notation switches, huge and tiny exponents
    special = np.array([0.0, -0.0, 1.0, 0.5, 1e-4, 9.9999e-5, 1e-5, 1.17549435e-38, 1e-45, 3.4028235e38, 1e16, 9.999999e15,
                        123456.79, 0.1, 1 / 3, 2 / 3, 1e-7, 16777216.0, 0.001, 65504.0, np.inf, -np.inf, np.nan],
                       dtype=np.float32)
    mean[:len(special)] = special
    var[100:100 + len(special)] = special
    ent[200:200 + len(special)] = -special
    ref, got = str(tmp_path / "ref.csv"), str(tmp_path / "got.csv")
    _reference_csv(ref, pid, y, lab, mean, var, ent)
    emit.write_detailed_csv(got, pid, y, lab, mean, var, ent)
    a, b = open(ref, "rb").read(), open(got, "rb").read()
    if a != b:
        la, lb = a.split(b"\n"), b.split(b"\n")
        bad = [(i, x, z) for i, (x, z) in enumerate(zip(la, lb)) if x != z][:5]
        raise AssertionError(f"{len(la)} vs {len(lb)} lines; first differences: {bad}")
    # sharded emission: header once, then appended blocks with global window indices
    got2 = str(tmp_path / "got2.csv")
    cut = 70_001
    emit.write_detailed_csv(got2, pid[:cut], y[:cut], lab[:cut], mean[:cut], var[:cut], ent[:cut], n_threads=3)
    emit.write_detailed_csv(got2, pid[cut:], y[cut:], lab[cut:], mean[cut:], var[cut:], ent[cut:], window_index_base=cut,
                            header=False, n_threads=1)
    assert open(got2, "rb").read() == a


def test_csv_round_trip_and_errors(tmp_path):
    emit = importlib.import_module(PKG_NAME + ".emit")
    pid, y, lab, mean, var, ent = _columns(1000, 9)
    path = str(tmp_path / "d.csv")
    emit.write_detailed_csv(path, pid, y, lab, mean, var, ent)
    df = pd.read_csv(path)                                 # what aggregate_patient_uq_metrics.py:22 does next
    assert list(df.columns) == ["Patient_ID", "Window_Index", "True_Label", "Predicted_Label", "Predicted_Probability",
                                "Predictive_Variance", "Predictive_Entropy"]
    np.testing.assert_array_equal(df["Predictive_Variance"].values.astype(np.float32), var)   # shortest repr round-trips
    np.testing.assert_array_equal(df["Patient_ID"].values, pid)
    emit.write_detailed_csv(path, pid[:0], y[:0], lab[:0], mean[:0], var[:0], ent[:0])        # empty: header only
    assert open(path).read().count("\n") == 1
    import pytest
    with pytest.raises(ValueError):
        emit.write_detailed_csv(path, pid, y[:-1], lab, mean, var, ent)
    with pytest.raises(Exception):
        emit.write_detailed_csv(os.path.join(str(tmp_path), "no_such_dir", "x.csv"), pid, y, lab, mean, var, ent)


def test_csv_float_text_on_random_bit_patterns(tmp_path):
    """Every float32 exponent range: 300k random bit patterns (NaNs included -> empty fields) format like pandas."""
    emit = importlib.import_module(PKG_NAME + ".emit")
    rng = np.random.default_rng(11)
    n = 300_000
    cols = [rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32).view(np.float32) for _ in range(3)]
    pid = np.full(n, 200001, dtype=np.int64)
    y = np.zeros(n, dtype=np.int64)
    lab = np.ones(n, dtype=int)
    ref, got = str(tmp_path / "ref.csv"), str(tmp_path / "got.csv")
    _reference_csv(ref, pid, y, lab, *cols)
    emit.write_detailed_csv(got, pid, y, lab, *cols)
    a, b = open(ref, "rb").read(), open(got, "rb").read()
    if a != b:
        la, lb = a.split(b"\n"), b.split(b"\n")
        bad = [(i, x, z) for i, (x, z) in enumerate(zip(la, lb)) if x != z][:5]
        raise AssertionError(f"{len(la)} vs {len(lb)} lines; first differences: {bad}")
