"""N2: the two steps right after the hot path.

``write_detailed_csv``  -- the per-window detail CSV of analyze_mcd_patient_level.py:134-152 /
                           analyze_de_patient_level.py, byte-identical to ``DataFrame.to_csv(path, index=False)``,
                           formatted by host threads in libuqk.so (no GPU involved).
``window_level_summary`` -- GPU twin of analyze_window_level_uncertainty.py:30-64 (accuracy, describe() of entropy and
                           variance for correct / incorrect windows, accuracy per entropy bin).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import numpy as np

from . import _lib

DESCRIBE_ROWS = ("count", "mean", "std", "min", "25%", "50%", "75%", "max")


def _host(a, dtype):
    if hasattr(a, "is_cuda"):
        a = a.detach().cpu().numpy()
    return np.ascontiguousarray(np.asarray(a).reshape(-1), dtype=dtype)


def write_detailed_csv(path: str, patient_ids, y_true, predicted_label, predicted_probability, predictive_variance,
                       predictive_entropy, window_index_base: int = 0, header: bool = True, n_threads: int = 0) -> None:
    """Columns and order of the reference's ``detailed_results_df`` (analyze_mcd_patient_level.py:134-145).
    ``header=False`` appends (multi-GPU shards written one after the other, ``window_index_base`` = first global row)."""
    pid = _host(patient_ids, np.int64)
    y = _host(y_true, np.int64)
    lab = _host(predicted_label, np.int32)
    cols = [_host(c, np.float32) for c in (predicted_probability, predictive_variance, predictive_entropy)]
    n = len(pid)
    if not (len(y) == len(lab) == len(cols[0]) == len(cols[1]) == len(cols[2]) == n):
        raise ValueError("Mismatch in array lengths!")           # the reference asserts the same (:127-130)
    lib = _lib.load()
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    _lib.check(lib.uqk_write_detail_csv(str(path).encode(), n, p(pid), int(window_index_base), p(y), p(lab), p(cols[0]),
                                        p(cols[1]), p(cols[2]), 1 if header else 0, int(n_threads)), "uqk_write_detail_csv")


def window_level_summary(predictive_entropy, predictive_variance, y_true, predicted_label, n_bins: int = 10,
                         device: Optional[int] = None) -> Dict:
    """Returns {'windows', 'accuracy', 'correct': {'Predictive_Entropy': {count, mean, ...}, 'Predictive_Variance': ...},
    'incorrect': {...}, 'bin_edges', 'bin_labels', 'window_count', 'bin_accuracy', 'error_rate'} -- the numbers the
    reference script prints (analyze_window_level_uncertainty.py:30-64)."""
    import torch
    from . import engine as _engine
    eng = _engine.get_engine(device)
    dev = torch.device("cuda", eng.device)

    def to_dev(a, dtype):
        if isinstance(a, torch.Tensor):
            return a.to(device=dev, dtype=dtype).contiguous().reshape(-1)
        return torch.from_numpy(np.array(np.asarray(a).reshape(-1), copy=True)).to(device=dev, dtype=dtype)

    with torch.cuda.device(eng.device):
        ent, var = to_dev(predictive_entropy, torch.float32), to_dev(predictive_variance, torch.float32)
        y, lab = to_dev(y_true, torch.int32), to_dev(predicted_label, torch.int32)
        n = ent.numel()
        if not (var.numel() == y.numel() == lab.numel() == n):
            raise ValueError("Mismatch in array lengths!")
        out = torch.empty(35 + 3 * n_bins, dtype=torch.float64, device=dev)
        _lib.check(eng.lib.uqk_window_summary(ent.data_ptr(), var.data_ptr(), y.data_ptr(), lab.data_ptr(), n, int(n_bins),
                                              out.data_ptr(), eng._stream()), "uqk_window_summary")
        o = out.cpu().numpy()
    res = {"windows": int(o[0]), "accuracy": float(o[1])}
    for c, cname in enumerate(("correct", "incorrect")):
        res[cname] = {}
        for m, mname in enumerate(("Predictive_Entropy", "Predictive_Variance")):
            res[cname][mname] = dict(zip(DESCRIBE_ROWS, o[2 + 16 * c + 8 * m: 10 + 16 * c + 8 * m].tolist()))
    edges = o[34:35 + n_bins]
    res["bin_edges"] = edges
    res["bin_labels"] = [f"{edges[i]:.3f}-{edges[i + 1]:.3f}" for i in range(n_bins)]   # reference line 53
    res["window_count"] = o[35 + n_bins:35 + 2 * n_bins].astype(np.int64)
    res["bin_accuracy"] = o[35 + 2 * n_bins:35 + 3 * n_bins]
    res["error_rate"] = 1.0 - res["bin_accuracy"]
    return res
