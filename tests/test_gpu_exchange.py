"""Exact batch-statistic BatchNorm with the windows of one call sharded over GPUs (include/uqk.h
uqk_set_stats_exchange; SURVEY.md 8e "the exception"; reference semantics uq_techniques.py:22).

On one GPU the second rank is emulated: the statistics of the full call are recorded through the exchange hook and
played back to each shard, which must then reproduce its slice of the full call bit for bit, and the shards' own
fixed-point sums must add up to the recorded ones exactly.  With two GPUs the real NCCL all-reduce runs."""
import importlib
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import PKG_NAME, ROOT
import gpu_common

pytestmark = pytest.mark.gpu
T, SEED = 2, 4242


def _run(eng, x, reduce_fn, global_n, window_base=0):
    eng.set_stats_exchange(reduce_fn, global_windows=global_n)
    try:
        out = eng.mcd(x, n_passes=T, seed=SEED, bn_mode="batch", precision="fp32", return_probs=True, metrics=(),
                      window_base=window_base)
    finally:
        eng.set_stats_exchange(None)
    return out["probs"]


def test_identity_exchange_is_a_no_op(uqk):
    eng = gpu_common.engine_with_members()
    x = uqk.synthetic.make_windows(37, seed=3)
    want = eng.mcd(x, n_passes=T, seed=SEED, bn_mode="batch", precision="fp32", return_probs=True, metrics=())["probs"]
    calls = []
    got = _run(eng, x, lambda t: calls.append((t.dtype, t.numel(), t.is_cuda)), 37)
    assert np.array_equal(got, want)
    assert calls == [(torch.int64, 2 * b.cout, True) for _ in range(T) for b in uqk.arch.BLOCKS]


def test_two_shards_played_back_match_the_full_call(uqk):
    eng = gpu_common.engine_with_members()
    n, cut = 50, 19
    x = uqk.synthetic.make_windows(n, seed=8)
    full_stats = []
    want = _run(eng, x, lambda t: full_stats.append(t.clone()), n)

    def shard(xs, base):
        local, it = [], iter(full_stats)

        def fn(t):
            local.append(t.clone())
            t.copy_(next(it))              # what an all-reduce over both shards would leave here
        return _run(eng, xs, fn, n, window_base=base), local

    pa, la = shard(x[:cut], 0)
    pb, lb = shard(x[cut:], cut)
    pe, le = shard(x[:0], 0)               # a rank with an empty shard still makes every exchange call
    assert np.array_equal(np.concatenate([pa, pb], axis=1), want)
    assert pe.shape == (T, 0) and len(le) == len(full_stats) and all(int(t.abs().sum()) == 0 for t in le)
    for a, b, f in zip(la, lb, full_stats):
        assert torch.equal(a + b, f)       # integer fixed point: the sum over shards is exact


def test_exchange_failure_is_reported(uqk):
    eng = gpu_common.engine_with_members()
    x = uqk.synthetic.make_windows(5, seed=1)

    def boom(t):
        raise KeyError("peer gone")
    with pytest.raises(RuntimeError, match="exchange failed"):
        _run(eng, x, boom, 5)
    with pytest.raises(ValueError, match="global_windows"):
        eng.set_stats_exchange(lambda t: None, global_windows=0)
    # the handle still works afterwards and carries no stale hook
    p = eng.mcd(x, n_passes=1, seed=1, bn_mode="batch", precision="fp32", return_probs=True, metrics=())["probs"]
    assert p.shape == (1, 5) and np.isfinite(p).all()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_nccl_two_gpus_bit_identical_to_one():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29561", os.path.join(ROOT, "tools", "multigpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    import json
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    assert res["all_identical"] and res["batch_all_identical"], res
