"""Exact Keras ``training=True`` semantics (batch-statistic BatchNorm over the whole call, uq_techniques.py:22) on the
tensor-core kernel: one statistics sweep per layer + the head kernel (csrc/api.cu run_fused_exact).

Oracle: float64 restatement with the same dropout masks.  Tolerances are the bf16 ones of tests/test_gpu_fused.py
(max |dp| <= 3.5e-2 under dropout; mean |dp| <= 5e-3 for n >= 64); the per-layer statistics themselves are checked much
tighter against the fp32 layer-synchronous path."""
import numpy as np
import pytest
import torch

from gpu_common import engine_with_members, members, twin_masks
from oracle import network, philox

pytestmark = pytest.mark.gpu


def _x(uqk, n, seed=11):
    return uqk.synthetic.make_windows(n, seed=seed)


@pytest.mark.parametrize("n", [1, 2, 5, 37, 130])
def test_exact_batch_philox_vs_oracle(uqk, n):
    eng = engine_with_members()
    x = _x(uqk, n)
    seed, T = 99, 3
    out = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
    ref = np.stack([network.forward_f64(members()[0], x, training=True, masks=twin_masks(seed, n, t))[:, 0]
                    for t in range(T)])
    err = np.abs(out["probs"] - ref)
    assert err.max() < 3.5e-2, err.max()
    if n >= 64:
        assert err.mean() < 5e-3, err.mean()
    m2 = eng.reduce(out["probs"])                       # fused reductions of the final run = stand-alone kernel
    for k in ("mean_pred", "pred_variance", "mutual_info", "label"):
        np.testing.assert_array_equal(out[k], m2[k])
    again = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
    np.testing.assert_array_equal(again["probs"], out["probs"])      # integer statistics: bit-reproducible


def test_exact_batch_mask_injection_vs_oracle(uqk):
    eng = engine_with_members()
    n, T = 21, 2
    x = _x(uqk, n, seed=4)
    rng = np.random.default_rng(1)
    masks = [[rng.random((n, 60, b.cout)) >= b.rate for b in uqk.arch.BLOCKS] for _ in range(T)]
    bits = np.zeros((T, n, 6, 60, 8), dtype=np.uint32)
    for t in range(T):
        for l in range(6):
            bits[t, :, l] = philox.pack_mask_bits(masks[t][l])
    out = eng.mcd(x, n_passes=T, seed=0, bn_mode="batch", precision="bf16", masks=bits, return_probs=True)
    ref = np.stack([network.forward_f64(members()[0], x, training=True, masks=masks[t])[:, 0] for t in range(T)])
    assert np.abs(out["probs"] - ref).max() < 3.5e-2


def test_exact_batch_matches_the_fp32_exact_path(uqk):
    """Same Philox masks, same semantics, two implementations (bf16 tensor-core sweeps vs fp32 layer-synchronous)."""
    eng = engine_with_members()
    n, T, seed = 200, 4, 31
    x = _x(uqk, n, seed=6)
    a = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)["probs"]
    b = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="fp32", return_probs=True)["probs"]
    d = np.abs(a - b)
    assert d.max() < 4e-2 and d.mean() < 5e-3, (d.max(), d.mean())


def test_exact_batch_differs_from_inference_statistics(uqk):
    """The sweeps really install batch statistics: with the synthetic moving statistics (mean 0.2, var ~1) the
    inference-mode result is measurably different, and frozen statistics taken from the exact fp32 calibration of the
    SAME windows and passes reproduce the exact mode closely."""
    eng = engine_with_members()
    n, T, seed = 256, 2, 8
    x = _x(uqk, n, seed=9)
    exact = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)["probs"]
    inf = eng.mcd(x, n_passes=T, seed=seed, bn_mode="inference", precision="bf16", return_probs=True)["probs"]
    assert np.abs(exact - inf).mean() > 1e-3
    eng.calibrate(x, n_windows=n, n_passes=T, seed=seed)          # pass-averaged statistics of the same masks
    cal = eng.mcd(x, n_passes=T, seed=seed, bn_mode="calibrated", precision="bf16", return_probs=True)["probs"]
    assert np.abs(exact - cal).mean() < np.abs(exact - inf).mean()


def test_exact_batch_sharded_playback_is_bit_identical(uqk):
    """Two shards fed with the all-reduced statistics of the full call reproduce their slices bit for bit, and their
    own fixed-point sums add up to the full call's exactly (one exchange per sweep, T x 512 int64 values)."""
    eng = engine_with_members()
    n, cut, T, seed = 41, 18, 3, 5
    x = _x(uqk, n, seed=2)

    def run(xs, fn, base=0):
        eng.set_stats_exchange(fn, global_windows=n)
        try:
            return eng.mcd(xs, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True,
                           metrics=(), window_base=base)["probs"]
        finally:
            eng.set_stats_exchange(None)

    full_stats = []
    want = run(x, lambda t: full_stats.append(t.clone()))
    assert len(full_stats) == 6 and all(t.numel() == T * 512 and t.dtype == torch.int64 for t in full_stats)
    plain = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True, metrics=())["probs"]
    np.testing.assert_array_equal(plain, want)

    def shard(xs, base):
        local, it = [], iter(full_stats)

        def fn(t):
            local.append(t.clone())
            t.copy_(next(it))
        return run(xs, fn, base), local

    pa, la = shard(x[:cut], 0)
    pb, lb = shard(x[cut:], cut)
    pe, le = shard(x[:0], 0)
    np.testing.assert_array_equal(np.concatenate([pa, pb], axis=1), want)
    assert pe.shape == (T, 0) and len(le) == 6 and all(int(t.abs().sum()) == 0 for t in le)
    for a, b, f in zip(la, lb, full_stats):
        assert torch.equal(a + b, f)


def test_exact_batch_large_t_goes_through_hbm(uqk):
    eng = engine_with_members()
    n, T = 6, 300                                   # > 256 passes: probabilities via HBM + the stand-alone reduction
    x = _x(uqk, n, seed=12)
    out = eng.mcd(x, n_passes=T, seed=3, bn_mode="batch", precision="bf16", return_probs=True)
    assert out["probs"].shape == (T, n) and np.isfinite(out["probs"]).all()
    np.testing.assert_array_equal(out["mean_pred"], eng.reduce(out["probs"])["mean_pred"])


def test_tensor_core_calibration_matches_the_fp32_calibration(uqk):
    """uqk_calibrate above `calib_fp32_max` windows runs the statistics sweeps: same definition (pass-averaged batch
    mean / biased variance of the exact training=True forward, same Philox masks) as the fp32 path, bf16 operands."""
    eng = engine_with_members()
    n, passes, seed = 1500, 2, 12
    x = _x(uqk, n, seed=15)
    eng.calibrate(x, n_windows=n, n_passes=passes, seed=seed)
    m32, v32 = eng.get_bn_stats()
    eng.option("calib_fp32_max", 0)
    try:
        eng.calibrate(x, n_windows=n, n_passes=passes, seed=seed)
        mtc, vtc = eng.get_bn_stats()
        eng.calibrate(x, n_windows=n, n_passes=passes, seed=seed)
        mtc2, vtc2 = eng.get_bn_stats()
    finally:
        eng.option("calib_fp32_max", 8192)
    for l, b in enumerate(uqk.arch.BLOCKS):
        a, c = m32[l, :b.cout], mtc[l, :b.cout]
        assert np.abs(a - c).max() < 2e-2 * max(1.0, np.abs(a).max()), (l, np.abs(a - c).max())
        a, c = v32[l, :b.cout], vtc[l, :b.cout]
        assert np.abs(a - c).max() < 4e-2 * np.abs(a).max(), (l, np.abs(a - c).max())
    np.testing.assert_array_equal(mtc, mtc2)          # integer statistics: bit-reproducible
    np.testing.assert_array_equal(vtc, vtc2)


def test_exact_batch_is_independent_of_the_pass_chunking(uqk):
    """The exact mode processes the passes in chunks sized by a memory budget; Philox pass counters, injected masks and
    output rows are offset per chunk, so the result must not depend on the chunk size."""
    eng = engine_with_members()
    n, T, seed = 23, 5, 17
    x = _x(uqk, n, seed=19)
    rng = np.random.default_rng(4)
    bits = rng.integers(0, 2 ** 32, (T, n, 6, 60, 8), dtype=np.uint64).astype(np.uint32)
    whole = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
    whole_m = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", masks=bits, return_probs=True)
    try:
        for chunk in (1, 2):
            eng.option("exact_chunk_max", chunk)
            part = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
            part_m = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", masks=bits, return_probs=True)
            for k in ("probs", "mean_pred", "pred_variance", "label"):
                np.testing.assert_array_equal(part[k], whole[k])
                np.testing.assert_array_equal(part_m[k], whole_m[k])
    finally:
        eng.option("exact_chunk_max", 0)


def test_pipelined_sweep_variant_is_bit_identical(uqk):
    """Option exact_pipelined (default 1): the resumed sweeps as a three-stage pipeline over two buffers (loader warps /
    MMAs / statistics warps) against the lock-step form (0: both tiles of a CTA walk the layer together).  Same arithmetic
    and the same integer statistics, so every output must be identical."""
    eng = engine_with_members()
    n, T, seed = 203, 3, 23
    x = _x(uqk, n, seed=29)
    got = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
    tiny = eng.mcd(x[:3], n_passes=2, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
    eng.option("exact_pipelined", 0)
    try:
        want = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
        ref = eng.mcd(x[:3], n_passes=2, seed=seed, bn_mode="batch", precision="bf16", return_probs=True)
    finally:
        eng.option("exact_pipelined", 1)
    for k in ("probs", "mean_pred", "pred_variance", "label"):
        np.testing.assert_array_equal(got[k], want[k])
    np.testing.assert_array_equal(tiny["probs"], ref["probs"])
