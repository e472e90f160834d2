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
