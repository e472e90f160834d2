"""pandas restatement of the patient-level aggregation.  TEST INFRASTRUCTURE.

Follows reference uncertainty_quantification/aggregate_patient_uq_metrics.py:34-46.  Pinned against
tests/golden/patient_summary_golden.csv, which the reference script itself produced (run unchanged in a
scratch directory by tests/golden/make_golden.py).
"""
from __future__ import annotations

import pandas as pd


def aggregate_patients(df: pd.DataFrame) -> pd.DataFrame:
    df = df.copy()
    df["Correct"] = (df["True_Label"] == df["Predicted_Label"])                     # :34
    s = df.groupby("Patient_ID").agg(                                                # :35-44
        mean_variance=("Predictive_Variance", "mean"),
        median_variance=("Predictive_Variance", "median"),
        std_variance=("Predictive_Variance", "std"),
        mean_entropy=("Predictive_Entropy", "mean"),
        median_entropy=("Predictive_Entropy", "median"),
        std_entropy=("Predictive_Entropy", "std"),
        patient_accuracy=("Correct", "mean"),
        num_windows=("Patient_ID", "size"),
    ).reset_index()
    one = s["num_windows"] <= 1                                                      # :45-46
    s.loc[one, "std_variance"] = 0
    s.loc[one, "std_entropy"] = 0
    return s
