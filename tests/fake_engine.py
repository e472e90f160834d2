"""TEST INFRASTRUCTURE: an oracle-backed stand-in for ``engine.Engine`` so that the reference's driver scripts can run to
completion in the CPU container (no GPU there).  It answers the calls the host modules make (``uq_techniques``,
``model``, ``patient``) with the CPU oracle; it is installed by monkey-patching ``engine.get_engine`` INSIDE A TEST ONLY --
the product never imports this file or ``oracle/``."""
import numpy as np

from oracle import network, reductions


class FakeEngine:
    device = 0

    def __init__(self, arch, synthetic):
        self.arch, self.syn = arch, synthetic
        self.members, self._key = [], None
        self.calibrated, self.last_guard_band, self.calls = False, (0, 0), []
        self._stats = None
        self._xchg, self._global_n = None, 0

    def set_stats_exchange(self, reduce_fn, global_windows=0):
        """Host model of Engine.set_stats_exchange: ``reduce_fn`` is a sharding.StatsExchange (its exchange_moments
        quantises the moments to the library's 2^-24 fixed point and all-reduces them)."""
        self._xchg, self._global_n = reduce_fn, int(global_windows)

    # model ------------------------------------------------------------------------------------------------------------
    def ensure_members(self, blobs_fn, key):
        if key != self._key:
            self.set_members(blobs_fn(), key)

    def set_members(self, blobs, key=None):
        blobs = np.asarray(blobs, dtype=np.float32)
        blobs = blobs[None] if blobs.ndim == 1 else blobs
        self.members = [self.syn.blob_to_weights(b) for b in blobs]
        self.n_members, self._key, self.calibrated = len(self.members), key, False

    def option(self, *_):
        pass

    def launch_count(self):
        return len(self.calls)

    # forward passes ---------------------------------------------------------------------------------------------------
    def _check_x(self, x):
        x = np.asarray(x)
        assert x.ndim == 3 and x.shape[1:] == (60, 4), x.shape
        return x

    def forward_det(self, x, member=0, precision="fp32", device_out=False):
        self.calls.append(("forward_det", len(x), precision))
        return network.forward_f32_torch(self.members[member], self._check_x(x))[:, 0]

    def calibrate(self, x, n_windows=None, n_passes=1, seed=2025, sampling="strided", window_base=0):
        self.calls.append(("calibrate", len(x)))
        masks = self._masks(len(x), 0, seed, window_base)
        kw = {} if self._xchg is None else {"stats_exchange": self._xchg.exchange_moments, "global_windows": self._global_n}
        _, self._stats = network.forward_f64(self.members[0], self._check_x(x), training=True, masks=masks, return_stats=True, **kw)
        self.calibrated = True

    def _masks(self, n, t, seed, window_base):
        from oracle import philox
        return [philox.dropout_keep_mask(seed, np.arange(n) + window_base, t, i, b.cout, b.rate) for i, b in enumerate(self.arch.BLOCKS)]

    def _metrics(self, probs, want):
        m = reductions.uq_evaluation_dist(probs.astype(np.float32), None)
        mean = m["mean_pred"]
        m["entropy_bits"] = (-(mean * np.log2(mean + np.float32(1e-9)) + (1 - mean) * np.log2((1 - mean) + np.float32(1e-9)))).astype(np.float32)
        m["label"] = (mean > 0.5).astype(np.int32)
        return {k: v for k, v in m.items() if k in want}

    def mcd(self, x, n_passes=50, seed=2025, bn_mode="calibrated", precision="bf16", masks=None, return_probs=False,
            metrics=("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info", "entropy_bits", "label"),
            window_base=0, device_out=False, exact_labels=False, guard_delta=5e-3):
        x = self._check_x(x)
        self.calls.append(("mcd", len(x), n_passes, bn_mode, precision))
        kw = {"bn_mode": "inference"} if bn_mode in ("inference", "frozen") else ({"bn_stats": self._stats} if bn_mode == "calibrated" else {})
        probs = np.stack([network.forward_f32_torch(self.members[0], x, training=True, masks=self._masks(len(x), t, seed, window_base))[:, 0]
                          if not kw else network.forward_f64(self.members[0], x, training=True, masks=self._masks(len(x), t, seed, window_base), **kw)[:, 0]
                          for t in range(n_passes)]).astype(np.float32)
        out = self._metrics(probs, metrics)
        if return_probs:
            out["probs"] = probs
        return self._dev(out, device_out)

    @staticmethod
    def _dev(out, device_out):
        if not device_out:
            return out
        import torch
        return {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in out.items()}

    def de(self, x, n_members=None, precision="fp32", return_probs=False,
           metrics=("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info", "entropy_bits", "label"),
           device_out=False, exact_labels=False, guard_delta=5e-3):
        x = self._check_x(x)
        m = self.n_members if n_members is None else n_members
        self.calls.append(("de", len(x), m, precision))
        probs = np.stack([network.forward_f32_torch(w, x)[:, 0] for w in self.members[:m]]).astype(np.float32)
        out = self._metrics(probs, metrics)
        if return_probs:
            out["probs"] = probs
        return out

    # reductions -------------------------------------------------------------------------------------------------------
    def reduce(self, probs, metrics=("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info", "entropy_bits", "label"),
               device_out=False):
        probs = np.asarray(probs, dtype=np.float32)
        self.calls.append(("reduce",) + probs.shape)
        return self._metrics(probs, metrics)

    def bootstrap(self, metrics, y_true, idx):
        self.calls.append(("bootstrap",) + tuple(np.shape(idx)))
        y = np.asarray(y_true)
        out = np.zeros((len(idx), 6))
        for b, ix in enumerate(np.asarray(idx)):
            v, yy = metrics["pred_variance"][ix], y[ix]
            out[b] = [v.mean(), v[yy == 0].mean() if (yy == 0).any() else 0.0, v[yy == 1].mean() if (yy == 1).any() else 0.0,
                      metrics["total_pred_entropy"][ix].mean(), metrics["expected_aleatoric_entropy"][ix].mean(),
                      metrics["mutual_info"][ix].mean()]
        return out

    def patient_reduce(self, variance, entropy, y_true, y_pred, seg_offsets, device_out=False):
        self.calls.append(("patient_reduce", len(seg_offsets) - 1))
        v, e = np.asarray(variance, np.float64), np.asarray(entropy, np.float64)
        ok = (np.asarray(y_true) == np.asarray(y_pred)).astype(np.float64)
        rows = np.zeros((len(seg_offsets) - 1, 8))
        for p in range(len(seg_offsets) - 1):
            a, b = int(seg_offsets[p]), int(seg_offsets[p + 1])
            sd = lambda z: float(np.std(z, ddof=1)) if b - a > 1 else 0.0
            rows[p] = [v[a:b].mean(), np.median(v[a:b]), sd(v[a:b]), e[a:b].mean(), np.median(e[a:b]), sd(e[a:b]), ok[a:b].mean(), b - a]
        return rows
