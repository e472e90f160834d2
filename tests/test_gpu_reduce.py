"""K4 reductions kernel (through the C-ABI) vs the reference's own uq_evaluation_dist outputs (golden) and
the oracle restatement.  Tolerances: abs 1e-6 on mean / entropies / MI,
abs 1e-7 + rel 1e-4 on variance, labels exact."""
import os

import numpy as np
import pytest

from gpu_common import engine_with_members
from oracle import reductions

pytestmark = pytest.mark.gpu
KEYS = ("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info")


def _check(m, ref):
    for k in KEYS:
        if k == "pred_variance":
            np.testing.assert_allclose(m[k], ref[k], rtol=1e-4, atol=1e-7, err_msg=k)
        else:
            np.testing.assert_allclose(m[k], ref[k], rtol=0, atol=1e-6, err_msg=k)


def test_reduce_matches_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "uq_f32_golden.npz"))
    eng = engine_with_members()
    m = eng.reduce(g["probs"][..., 0])
    _check(m, {k: g[f"m_{k}"] for k in KEYS})
    np.testing.assert_array_equal(m["label"], g["csv_label"])
    np.testing.assert_allclose(m["entropy_bits"], g["csv_entropy"], atol=1e-6)
    np.testing.assert_allclose(m["mean_pred"], g["csv_mean"], atol=1e-7)
    m1 = eng.reduce(g["probs"][:1, :, 0])          # T = 1
    assert np.all(m1["pred_variance"] == 0) and np.all(m1["mutual_info"] <= 1e-6)


@pytest.mark.parametrize("t,n", [(1, 2), (3, 2), (5, 1000), (50, 4096), (50, 4099), (100, 777), (25, 128), (700, 300)])
def test_reduce_matches_oracle_shapes(t, n):
    rng = np.random.default_rng(t * 1000 + n)
    p = rng.random((t, n)).astype(np.float32)
    p[:, : min(n, 3)] = np.array([0.0, 1.0, 0.5])[: min(n, 3)]
    eng = engine_with_members()
    m = eng.reduce(p)
    ref = reductions.uq_evaluation_dist(p, None)
    _check(m, ref)
    mean_p, var_p, ent, label = reductions.window_csv_metrics(p)
    outside = np.abs(mean_p - 0.5) > 1e-6
    np.testing.assert_array_equal(m["label"][outside], label[outside])
    np.testing.assert_allclose(m["entropy_bits"], ent, atol=2e-6)


def test_reduce_large_property():
    """Full benchmark size: property checks that do not need the CPU oracle on 1M x 50."""
    import torch
    t, n = 50, 1_000_000
    gen = torch.Generator(device="cuda").manual_seed(1)
    p = torch.rand((t, n), generator=gen, device="cuda")
    eng = engine_with_members()
    m = eng.reduce(p, device_out=True)
    mean = p.double().mean(dim=0)
    var = p.double().var(dim=0, unbiased=False)
    assert float((m["mean_pred"].double() - mean).abs().max()) < 1e-6
    assert float((m["pred_variance"].double() - var).abs().max()) < 1e-6
    assert float(m["mutual_info"].min()) >= 0.0
    assert float((m["total_pred_entropy"] - m["expected_aleatoric_entropy"] - m["mutual_info"]).abs().max()) < 1e-6
    assert float(m["total_pred_entropy"].max()) <= 0.6931472 + 1e-6
    lab = (m["mean_pred"] > 0.5).int()
    assert bool((lab == m["label"]).all())
