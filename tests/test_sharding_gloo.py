"""Host logic of the N>1 path on CPU: patient-aligned shards and the final gather over a world_size-2 gloo group
(the per-window "metrics" are a deterministic function of the global window index, standing in for the kernel)."""
import importlib
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import PKG_NAME, ROOT


def test_patient_aligned_shards_properties(uqk):
    sh = importlib.import_module(PKG_NAME + ".sharding")
    pid = uqk.synthetic.make_patient_ids(n_patients=101, seed=3)
    n = len(pid)
    for world in (1, 2, 4, 8):
        shards = sh.patient_aligned_shards(pid, world)
        assert shards[0][0] == 0 and shards[-1][1] == n and len(shards) == world
        for (s0, e0), (s1, e1) in zip(shards[:-1], shards[1:]):
            assert e0 == s1 and s0 <= e0
            assert e0 == n or e0 == 0 or pid[e0 - 1] != pid[e0]            # cut at a patient boundary
        sizes = np.array([e - s for s, e in shards])
        assert sizes.max() - sizes.min() <= 2 * 660                        # at most one patient per cut
    assert sh.window_shards(10, 3) == [(0, 3), (3, 6), (6, 10)]
    assert sh.patient_aligned_shards(np.array([1, 1, 2, 1, 1, 2]), 2) == [(0, 3), (3, 6)]   # non-contiguous ids
    assert sh.patient_aligned_shards([], 2) == [(0, 0), (0, 0)]


def _worker(rank, world, port, n, tmp):
    import sys
    sys.path.insert(0, ROOT)
    sh = importlib.import_module(PKG_NAME + ".sharding")
    syn = importlib.import_module(PKG_NAME).synthetic
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pid = syn.make_patient_ids(n_patients=7, seed=5, lo=5, hi=9)[:n]
    shards = sh.patient_aligned_shards(pid, world)
    s, e = shards[rank]
    idx = torch.arange(s, e, dtype=torch.float32)
    local = {"mean_pred": torch.sin(idx), "label": (torch.arange(s, e) % 3 == 0).to(torch.int32)}
    full = sh.gather_metrics(local, shards)
    torch.save({k: v for k, v in full.items()}, os.path.join(tmp, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_gather_metrics_world2_gloo(tmp_path):
    n, world, port = 41, 2, 29533
    mp.spawn(_worker, args=(world, port, n, str(tmp_path)), nprocs=world, join=True)
    want_mean = torch.sin(torch.arange(n, dtype=torch.float32))
    want_label = (torch.arange(n) % 3 == 0).to(torch.int32)
    for r in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"rank{r}.pt"))
        assert torch.equal(got["mean_pred"], want_mean)
        assert torch.equal(got["label"], want_label)


def _bn_worker(rank, world, port, n, tmp):
    """Exact batch-statistic MC-Dropout pass with the windows sharded over `world` ranks: every layer's moments go
    through sharding.StatsExchange (int64 fixed point, gloo all-reduce); the float64 oracle does the arithmetic."""
    import sys
    sys.path.insert(0, ROOT)
    sh = importlib.import_module(PKG_NAME + ".sharding")
    pkg = importlib.import_module(PKG_NAME)
    from oracle import network, philox
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x = pkg.synthetic.make_windows(n, seed=11)
    w = pkg.synthetic.make_member_weights(0)
    s, e = [(0, 0), (0, n)][rank] if n < 4 else sh.window_shards(n, world)[rank]      # n < 4: rank 0 holds nothing
    masks = [philox.dropout_keep_mask(77, np.arange(s, e), 0, i, b.cout, b.rate) for i, b in enumerate(pkg.arch.BLOCKS)]
    ex = sh.StatsExchange()
    p = network.forward_f64(w, x[s:e], training=True, masks=masks, stats_exchange=ex.exchange_moments, global_windows=n)
    assert ex.calls == len(pkg.arch.BLOCKS)
    np.save(os.path.join(tmp, f"p{rank}.npy"), p[:, 0])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n,port", [(12, 29541), (3, 29543)])
def test_exact_batch_stats_exchange_world2_gloo(uqk, tmp_path, n, port):
    from oracle import network, philox
    mp.spawn(_bn_worker, args=(2, port, n, str(tmp_path)), nprocs=2, join=True)
    x = uqk.synthetic.make_windows(n, seed=11)
    w = uqk.synthetic.make_member_weights(0)
    masks = [philox.dropout_keep_mask(77, np.arange(n), 0, i, b.cout, b.rate) for i, b in enumerate(uqk.arch.BLOCKS)]
    want = network.forward_f64(w, x, training=True, masks=masks)[:, 0]          # one rank holding all windows
    got = np.concatenate([np.load(os.path.join(str(tmp_path), f"p{r}.npy")) for r in range(2)])
    assert got.shape == want.shape
    # the exchanged moments are quantised to 2^-24 per rank and the variance comes from E[x^2]-mu^2
    assert np.abs(got - want).max() < 1e-6


def _sharded_worker(rank, world, port, n, tmp):
    """sharding.sharded_mc_dropout_uq as the product calls it, over gloo, with the GPU engine replaced by the oracle-backed
    fake (tests/fake_engine.py): shards, one key for the whole call, calibration of every rank's shard through the
    statistics exchange, window_base, final gather."""
    import sys
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, os.path.join(ROOT, PKG_NAME, "uncertainty_quantification"))
    pkg = importlib.import_module(PKG_NAME)
    sh = importlib.import_module(PKG_NAME + ".sharding")
    engine_mod = importlib.import_module(PKG_NAME + ".engine")
    model_mod = importlib.import_module(PKG_NAME + ".model")
    from fake_engine import FakeEngine
    fake = FakeEngine(pkg.arch, pkg.synthetic)
    engine_mod.get_engine = lambda device=None: fake
    uq = importlib.import_module("uq_techniques")
    if world > 1:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
    x = pkg.synthetic.make_windows(n, seed=21)
    pid = pkg.synthetic.make_patient_ids(n_patients=6, seed=5, lo=3, hi=6)[:n]
    model = model_mod.UQModel(pkg.synthetic.make_member_weights(0))
    full = sh.sharded_mc_dropout_uq(uq, model, x, n_pred=2, patient_ids=pid, seed=31, bn_mode="calibrated")
    assert ("calibrate", len(x) if world == 1 else sh.patient_aligned_shards(pid, world)[rank][1] - sh.patient_aligned_shards(pid, world)[rank][0]) in fake.calls
    assert fake._xchg is None                                     # the exchange is removed again after the calibration
    np.savez(os.path.join(tmp, f"sharded_w{world}_r{rank}.npz"), **{k: v for k, v in full.items()})
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def test_sharded_calibrated_call_world2_gloo(tmp_path):
    n = 21
    mp.spawn(_sharded_worker, args=(2, 29551, n, str(tmp_path)), nprocs=2, join=True)
    mp.spawn(_sharded_worker, args=(1, 0, n, str(tmp_path)), nprocs=1, join=True)     # one rank holding all windows (own process:
    # the worker replaces engine.get_engine)
    one = np.load(os.path.join(str(tmp_path), "sharded_w1_r0.npz"))
    for r in range(2):
        two = np.load(os.path.join(str(tmp_path), f"sharded_w2_r{r}.npz"))
        assert set(two.files) == set(one.files) and len(two["mean_pred"]) == n
        for k in one.files:
            # same masks (global window index), statistics of ALL windows through the fixed-point exchange
            np.testing.assert_allclose(two[k], one[k], rtol=0, atol=2e-6, err_msg=k)
