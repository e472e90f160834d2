"""Builds a synthetic working directory with the file names the reference's CONSUMER scripts read
(SURVEY.md Appendix B): ./processed_datasets/*.npy, ./AlCNN1D_no_pool.keras,
./models/ensemble_cnn_no_pool/AlCNN_smote_seed{5..}.keras."""
from __future__ import annotations

import os

import numpy as np

from . import model, synthetic


def build_fixture_dir(root: str, n_unbalanced: int = 2000, n_rus: int = 500, n_members: int = 5, seed: int = 2025,
                      n_global_members: int = 0):
    ds = os.path.join(root, "processed_datasets")
    os.makedirs(ds, exist_ok=True)
    xu = synthetic.make_windows(n_unbalanced, seed=seed)
    yu = synthetic.make_labels(n_unbalanced, seed=seed + 1)
    per = max(1, n_unbalanced // 12)
    pid = (200001 + np.arange(n_unbalanced) // per).astype(np.int64)
    np.save(os.path.join(ds, "X_test_std_unbalanced.npy"), xu)           # analyze_mcd_patient_level.py:43
    np.save(os.path.join(ds, "y_test_unbalanced.npy"), yu)
    np.save(os.path.join(ds, "patient_ids_test_unbalanced.npy"), pid)
    np.save(os.path.join(ds, "X_test_std_rus.npy"), synthetic.make_windows(n_rus, seed=seed + 2))
    np.save(os.path.join(ds, "y_test_rus.npy"), (np.arange(n_rus) % 2).astype(np.int64))
    model.synthetic_model(0).save(os.path.join(root, "AlCNN1D_no_pool.keras"))   # analyze_mcd_patient_level.py:19
    ens = os.path.join(root, "models", "ensemble_cnn_no_pool")
    os.makedirs(ens, exist_ok=True)
    members = [model.synthetic_model(i) for i in range(n_members)]
    for i, m in enumerate(members):                                           # analyze_de_patient_level.py:45 (i+5)
        m.save(os.path.join(ens, f"AlCNN_smote_seed{i + 5}.keras"))
    if n_global_members:                                                      # evaluate_de_global.py:11,18,29 (seed0..19)
        glob = os.path.join(root, "models", "ensemble_cnn")
        os.makedirs(glob, exist_ok=True)
        for i in range(n_global_members):
            members[i % len(members)].save(os.path.join(glob, f"AlCNN_smote_seed{i}.keras"))
    return root
