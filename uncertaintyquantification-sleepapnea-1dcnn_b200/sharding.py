"""Multi-GPU partitioning of the hot path.

Windows are independent units (ensemble / deterministic / calibrated MC-Dropout), so the path shards with no
data-path collective: every rank owns one contiguous range of windows, cut at patient boundaries so that the
per-patient aggregation needs no cross-rank median, and the only exchange is the final gather of the per-window
metrics (24-28 B per window) and of the per-patient rows.  Dropout masks are keyed by the GLOBAL window index
(``window_base``), so results do not depend on the number of ranks.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist


def window_shards(n: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous, near-equal ranges covering [0, n)."""
    cuts = [(n * r) // world for r in range(world + 1)]
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def patient_aligned_shards(patient_ids: Sequence, world: int) -> List[Tuple[int, int]]:
    """Like ``window_shards`` but every cut is moved to the nearest patient boundary (patients are contiguous
    runs of rows: reference preprocess_shhs_raw.py:323-325, prepare_numpy_datasets.py:142-152).  Imbalance is at
    most one patient per cut.  Falls back to plain cuts when ids are not contiguous runs."""
    ids = np.asarray(patient_ids)
    n = len(ids)
    if n == 0:
        return [(0, 0)] * world
    bounds = np.concatenate([[0], np.nonzero(ids[1:] != ids[:-1])[0] + 1, [n]])
    if len(np.unique(ids)) != len(bounds) - 1:      # an id appears in two separate runs
        return window_shards(n, world)
    cuts = [0]
    for r in range(1, world):
        target = (n * r) // world
        j = int(np.argmin(np.abs(bounds - target)))
        cuts.append(max(int(bounds[j]), cuts[-1]))
    cuts.append(n)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def _world(group=None) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def gather_metrics(local: Dict[str, torch.Tensor], shards: List[Tuple[int, int]], group=None) -> Dict[str, torch.Tensor]:
    """All-gathers per-window arrays of unequal shard length (padded to the longest shard) and returns the
    concatenation in global window order on every rank.  Works with NCCL (CUDA tensors) and gloo (CPU tensors)."""
    rank, world = _world(group)
    if world == 1:
        return dict(local)
    sizes = [e - s for s, e in shards]
    mx = max(sizes)
    out = {}
    for k, v in local.items():
        v = v.reshape(-1)
        pad = torch.zeros(mx, dtype=v.dtype, device=v.device)
        pad[: v.numel()] = v
        buf = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(buf, pad, group=group)
        out[k] = torch.cat([b[:sz] for b, sz in zip(buf, sizes)])
    return out


STAT_SCALE = float(1 << 24)   # fixed point of the batch-statistic accumulators (csrc/layersync.cu kStatScale)


class StatsExchange:
    """The one real exchange step of the path ("the exception"): exact batch-statistic BatchNorm
    (Keras ``training=True``, uq_techniques.py:22) normalises with the statistics of ALL windows of the call, so with
    the windows sharded every layer of every pass SUM-all-reduces its per-channel (sum, sum of squares).  The sums
    travel as int64 fixed point (2^-24): integer addition is associative, so the result does not depend on the number
    of ranks or on the reduction order.  Called by the library through ``Engine.set_stats_exchange`` with a CUDA
    tensor (NCCL); the gloo test calls it with CPU tensors."""

    def __init__(self, group=None):
        self.group = group
        self.calls = 0

    def __call__(self, t: torch.Tensor) -> torch.Tensor:
        if t.dtype != torch.int64:
            raise TypeError("batch statistics are exchanged as int64 fixed point")
        self.calls += 1
        if _world(self.group)[1] > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    # host model of what the library does around the exchange (used by the CPU tests with the float64 oracle)
    def exchange_moments(self, ssum: np.ndarray, ssq: np.ndarray):
        """(sum, sum of squares) per channel of this rank's windows -> the same over all ranks, via fixed point."""
        q = torch.from_numpy(np.concatenate([np.rint(np.asarray(ssum, np.float64) * STAT_SCALE),
                                             np.rint(np.asarray(ssq, np.float64) * STAT_SCALE)]).astype(np.int64))
        q = self(q).numpy().astype(np.float64) / STAT_SCALE
        c = len(q) // 2
        return q[:c], q[c:]


def sharded_mc_dropout_uq(uq_module, model, x, n_pred: int = 50, patient_ids=None, seed: Optional[int] = None,
                          bn_mode: str = "calibrated", calib_windows: Optional[int] = None, group=None, **kw) -> Dict[str, np.ndarray]:
    """Each rank runs the MC-Dropout path on its shard of ``x`` (host array, identical on all ranks) and the
    per-window metrics are gathered on every rank.  The dropout masks are keyed by the global window index.
    ``bn_mode='calibrated'``: one exact pass over the whole call measures the statistics -- every rank sweeps its own
    shard and the per-layer sums are all-reduced (``StatsExchange``, 6 messages of 4 KB); with ``calib_windows`` every
    rank instead uses the SAME strided subset of the global windows (no collective at all).  ``bn_mode='batch'``: exact Keras semantics -- the per-layer
    statistics are all-reduced over the ranks (``StatsExchange``).  Either way the gathered result is bit-identical
    to a single-GPU run over all of ``x``."""
    rank, world = _world(group)
    n = len(x)
    shards = patient_aligned_shards(patient_ids, world) if patient_ids is not None else window_shards(n, world)
    s, e = shards[rank]
    if seed is None:
        seed = uq_module._call_seed()      # one key for the whole sharded call (the ranks' call counters advance in step)
    if bn_mode == "batch":
        uq_module.mc_dropout_uq(model, x[:0], n_pred=1, seed=seed, bn_mode="inference")   # uploads the model
        eng = uq_module._engine.get_engine()
        eng.set_stats_exchange(StatsExchange(group), global_windows=n)
        try:
            out = uq_module.mc_dropout_uq(model, x[s:e], n_pred=n_pred, seed=seed, bn_mode="batch", device_out=True,
                                          window_base=s, **kw)
        finally:
            eng.set_stats_exchange(None)
        local = {k: v for k, v in out.items() if isinstance(v, torch.Tensor)}
        return {k: v.cpu().numpy() for k, v in gather_metrics(local, shards, group).items()}
    if bn_mode == "calibrated":
        uq_module.mc_dropout_uq(model, x[:0], n_pred=1, seed=seed, bn_mode="inference")  # uploads the model
        eng = uq_module._engine.get_engine()
        if calib_windows is not None and calib_windows < n:
            eng.calibrate(x, n_windows=calib_windows, n_passes=1, seed=int(seed))
        else:
            eng.set_stats_exchange(StatsExchange(group), global_windows=n)
            try:
                eng.calibrate(x[s:e], n_passes=1, seed=int(seed), window_base=s)
            finally:
                eng.set_stats_exchange(None)
    out = uq_module.mc_dropout_uq(model, x[s:e], n_pred=n_pred, seed=seed, bn_mode=bn_mode, recalibrate=False,
                                  device_out=True, window_base=s, **kw)
    local = {k: v for k, v in out.items() if isinstance(v, torch.Tensor)}
    full = gather_metrics(local, shards, group)
    return {k: v.cpu().numpy() for k, v in full.items()}
