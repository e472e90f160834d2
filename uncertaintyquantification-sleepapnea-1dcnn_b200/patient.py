"""Patient-level aggregation on the GPU (K5): the twin of the reference's pandas groupby
(uncertainty_quantification/aggregate_patient_uq_metrics.py:34-46).

Host work: one stable argsort of the patient ids (pandas sorts the group keys too) to obtain contiguous
segments; the per-patient mean / median / std(ddof=1) of Predictive_Variance and Predictive_Entropy, the accuracy
and the window count are computed by ``uqk_patient_reduce``.
"""
from __future__ import annotations

import numpy as np
import pandas as pd

from . import engine as _engine

COLUMNS = ["Patient_ID", "mean_variance", "median_variance", "std_variance", "mean_entropy", "median_entropy",
           "std_entropy", "patient_accuracy", "num_windows"]


def segments(patient_ids: np.ndarray):
    """Sorted unique ids, the permutation that groups rows by id (stable), and segment offsets (P+1,)."""
    ids = np.asarray(patient_ids)
    order = np.argsort(ids, kind="stable")
    uniq, counts = np.unique(ids, return_counts=True)
    offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    return uniq, order, offs


def aggregate_arrays(patient_ids, variance, entropy, true_label, pred_label, device=None) -> pd.DataFrame:
    uniq, order, offs = segments(patient_ids)
    contiguous = bool(np.all(order == np.arange(len(order))))
    take = (lambda a: np.asarray(a)) if contiguous else (lambda a: np.asarray(a)[order])
    eng = _engine.get_engine(device)
    rows = eng.patient_reduce(take(variance), take(entropy), take(true_label), take(pred_label), offs)
    df = pd.DataFrame(rows[:, :7], columns=COLUMNS[1:8])
    df.insert(0, "Patient_ID", uniq)
    df["num_windows"] = rows[:, 7].astype(np.int64)
    return df


def aggregate_patients(df: pd.DataFrame, device=None) -> pd.DataFrame:
    """DataFrame in, DataFrame out, same columns as the reference's ``patient_summary``
    (aggregate_patient_uq_metrics.py:35-46)."""
    return aggregate_arrays(df["Patient_ID"].values, df["Predictive_Variance"].values, df["Predictive_Entropy"].values,
                            df["True_Label"].values, df["Predicted_Label"].values, device)
