"""Host-side driver of libuqk.so: device buffers (torch is used only as the allocator / stream provider),
model upload, and the calls behind ``uq_techniques.mc_dropout_predict`` / ``deep_ensembles_predict`` /
``uq_evaluation_dist`` (reference uncertainty_quantification/uq_techniques.py:12-112).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import _lib, arch

METRIC_KEYS = ("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info",
               "entropy_bits", "label")
_PREC = {"bf16": _lib.PREC_BF16_TC, "bf16_tc": _lib.PREC_BF16_TC, "fp32": _lib.PREC_FP32}
# Guard band of the labels: half-width around 0.5 inside which a bf16 mean probability is re-evaluated in fp32.  The
# measured |d mean_pred| of the bf16 path against the float64 oracle is <= 2.4e-3 over both test models
# (profiles/r02_parity_distribution.json), so 5e-3 leaves a factor 2.
GUARD_DELTA = 5e-3
# Largest calibration set swept in one go (scratch: 53,760 B per window, held by the handle); UQK_CALIB_MAX_WINDOWS overrides.
CALIB_MAX_WINDOWS = int(os.environ.get("UQK_CALIB_MAX_WINDOWS", "2000000"))
_BN = {"inference": _lib.BN_INFERENCE, "frozen": _lib.BN_INFERENCE, "batch": _lib.BN_BATCH,
       "calibrated": _lib.BN_CALIBRATED}


def _require_cuda():
    if not torch.cuda.is_available():
        raise _lib.UqkError("no CUDA device: this package runs on B200 (sm_100a) only and has no CPU fallback")


class _DevicePtr:
    """Minimal ``__cuda_array_interface__`` carrier so torch can alias memory owned by libuqk (no copy)."""

    def __init__(self, ptr: int, count: int):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": "<i8", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def _device_view_i64(ptr: int, count: int, device: int) -> "torch.Tensor":
    return torch.as_tensor(_DevicePtr(ptr, count), device=torch.device("cuda", device))


class Engine:
    """One engine per GPU (wraps one ``uqk_handle``).  Not thread-safe."""

    def __init__(self, device: Optional[int] = None):
        _require_cuda()
        self.lib = _lib.load()
        self.device = torch.cuda.current_device() if device is None else int(device)
        h = C.c_void_p()
        _lib.check(self.lib.uqk_create(self.device, C.byref(h)), "uqk_create")
        self._h = h
        self.n_members = 0
        self._model_key = None
        self.calibrated = False
        self._xchg_cb = None
        self._xchg_error = None
        self.last_guard_band = (0, 0)

    def close(self):
        if getattr(self, "_h", None):
            self.lib.uqk_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing ---------------------------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _dev(self):
        return torch.device("cuda", self.device)

    def to_device(self, x):
        """(N,60,4) float32/float64 host array, pinned torch tensor or CUDA tensor -> (CUDA tensor, dtype code)."""
        if isinstance(x, torch.Tensor):
            t = x
        else:
            a = np.asarray(x)
            if a.dtype not in (np.float32, np.float64):
                a = a.astype(np.float32)
            t = torch.from_numpy(np.ascontiguousarray(a))
        if t.dim() != 3 or t.shape[1] != arch.WINDOW_LEN or t.shape[2] != arch.WINDOW_FEATURES:
            raise ValueError(f"windows must have shape (N, {arch.WINDOW_LEN}, {arch.WINDOW_FEATURES}), got {tuple(t.shape)}")
        if t.dtype not in (torch.float32, torch.float64):
            t = t.float()
        if not t.is_cuda:
            t = t.to(self._dev(), non_blocking=True)
        t = t.contiguous()
        return t, (_lib.F64 if t.dtype == torch.float64 else _lib.F32)

    def option(self, key: str, value: int):
        _lib.check(self.lib.uqk_set_option(self._h, key.encode(), int(value)), "uqk_set_option")

    def launch_count(self) -> int:
        return int(self.lib.uqk_launch_count(self._h))

    # ---- model ------------------------------------------------------------------------------------------------------
    def set_members(self, blobs: np.ndarray, key=None):
        """blobs: (M, 853441) float32, each row one network in Keras ``get_weights()`` order."""
        blobs = np.ascontiguousarray(blobs, dtype=np.float32)
        if blobs.ndim == 1:
            blobs = blobs[None]
        with torch.cuda.device(self.device):
            _lib.check(self.lib.uqk_set_model(self._h, blobs.shape[0], blobs.ctypes.data_as(C.c_void_p),
                                              blobs.shape[1]), "uqk_set_model")
        self.n_members = blobs.shape[0]
        self._model_key = key
        self.calibrated = False

    def ensure_members(self, blobs_fn, key):
        """Uploads only when ``key`` (identity of the model list) changed."""
        if self._model_key != key or self.n_members == 0:
            self.set_members(blobs_fn(), key)

    # ---- outputs ----------------------------------------------------------------------------------------------------
    def _alloc_metrics(self, n: int, want: Sequence[str]):
        bufs, mo = {}, _lib.MetricsOut()
        for k in METRIC_KEYS:
            if k in want:
                t = torch.empty(n, dtype=torch.int32 if k == "label" else torch.float32, device=self._dev())
                bufs[k] = t
                setattr(mo, k, t.data_ptr())
        return bufs, mo

    def poll_error(self):
        """Raises if a device-side self-check of earlier work fired (include/uqk.h ``uqk_poll_error``): call after
        synchronising when results were left on the device."""
        _lib.check(self.lib.uqk_poll_error(self._h), "uqk_poll_error")

    def _finish(self, bufs, probs, device_out):
        if device_out:
            out = dict(bufs)
            if probs is not None:
                out["probs"] = probs
            return out
        out = {k: v.cpu().numpy() for k, v in bufs.items()}   # D2H (synchronises)
        if probs is not None:
            out["probs"] = probs.cpu().numpy()
        self.poll_error()                                      # the work is complete: a watchdog / self-check is visible now
        return out

    # ---- compute ----------------------------------------------------------------------------------------------------
    def forward_det(self, x, member: int = 0, precision: str = "bf16", device_out: bool = False):
        """Deterministic forward (BN inference mode, no dropout) -> (N,) float32."""
        with torch.cuda.device(self.device):
            xt, dt = self.to_device(x)
            n = xt.shape[0]
            probs = torch.empty(n, dtype=torch.float32, device=self._dev())
            _lib.check(self.lib.uqk_forward_det(self._h, int(member), xt.data_ptr(), dt, n, _PREC[precision],
                                                probs.data_ptr(), self._stream()), "uqk_forward_det")
            return probs if device_out else probs.cpu().numpy()

    def calibrate(self, x, n_windows: Optional[int] = None, n_passes: int = 1, seed: int = 2025, sampling: str = "strided",
                  window_base: int = 0):
        """Freezes the batch statistics Keras ``training=True`` would use (uq_techniques.py:22), measured on ``x``:
        per-pass batch mean / biased variance of the exact forward, averaged over ``n_passes`` passes (fp32
        layer-synchronous kernels up to 8192 windows, the statistics sweeps of the tensor-core kernel above).
        Default: ALL windows of the call, one pass -- the statistics the reference normalises pass 0 with, at the cost of
        one exact pass (~3 % of a T=50 run).  ``n_windows`` < len(x) calibrates on a subset instead (measured: mean
        |d mean_pred| vs the exact mode 1.2e-3 for 8192..32768 of 65536 windows, 6e-5 for all of them --
        profiles/r02_calibration_compare.json): ``sampling='strided'`` takes every (N // n_windows)-th window of the WHOLE
        call -- rows of a real data set are patient-contiguous (preprocess_shhs_raw.py:323-325), so a prefix would be
        the first few dozen patients, a biased sample; ``'prefix'`` takes ``x[:n_windows]`` (tests).
        With a statistics exchange installed, ``x`` is this rank's shard (``window_base`` = its first global index)
        and the statistics are those of the windows of all ranks."""
        with torch.cuda.device(self.device):
            n = len(x)
            xs = x
            if n_windows is None and n > CALIB_MAX_WINDOWS:
                # the sweeps hand one pass of activations through HBM (54 KB per window): beyond ~2 M windows that does
                # not fit next to the data, so the statistics come from a strided subset of that size (error ~1e-4)
                n_windows = CALIB_MAX_WINDOWS
            if n_windows is not None and 0 < int(n_windows) < n:
                if sampling == "strided":
                    xs = x[::max(1, n // int(n_windows))][:int(n_windows)]
                elif sampling == "prefix":
                    xs = x[:int(n_windows)]
                else:
                    raise ValueError(f"unknown calibration sampling '{sampling}'")
            if not isinstance(xs, torch.Tensor):
                xs = np.asarray(xs)
            xt, dt = self.to_device(xs)
            _lib.check(self.lib.uqk_calibrate(self._h, xt.data_ptr(), dt, xt.shape[0], int(n_passes),
                                              C.c_uint64(seed & (2 ** 64 - 1)), int(window_base), self._stream()), "uqk_calibrate")
            self.calibrated = True

    def set_stats_exchange(self, reduce_fn, global_windows: int = 0):
        """Installs (``reduce_fn`` callable) or removes (``None``) the cross-rank exchange of the batch statistics
        (include/uqk.h ``uqk_set_stats_exchange``).  ``reduce_fn(t)`` receives a 1-D int64 CUDA tensor aliasing the
        library's (sum, sum of squares) accumulators of one layer and must SUM-all-reduce it in place over the ranks
        on the current stream (``torch.distributed.all_reduce`` does).  ``global_windows`` = windows of the whole call."""
        if reduce_fn is None:
            _lib.check(self.lib.uqk_set_stats_exchange(self._h, C.cast(None, _lib.STATS_EXCHANGE_FN), None, 0), "uqk_set_stats_exchange")
            self._xchg_cb = None
            return
        dev = self.device
        self._xchg_error = None

        def _cb(_user, ptr, count, stream):
            try:
                cur = torch.cuda.current_stream(dev).cuda_stream
                if (stream or 0) != cur:
                    raise RuntimeError("the exchange must run on the stream the pass was enqueued on")
                reduce_fn(_device_view_i64(ptr, count, dev))
                return 0
            except Exception as e:       # never let an exception cross the C boundary
                self._xchg_error = e
                return 1

        cb = _lib.STATS_EXCHANGE_FN(_cb)
        _lib.check(self.lib.uqk_set_stats_exchange(self._h, cb, None, int(global_windows)), "uqk_set_stats_exchange")
        self._xchg_cb = cb               # keep the ctypes thunk alive while the library holds its address

    def get_bn_stats(self):
        mean = np.empty((arch.N_BLOCKS, arch.MAX_CH), dtype=np.float32)
        var = np.empty_like(mean)
        _lib.check(self.lib.uqk_get_bn_stats(self._h, mean.ctypes.data_as(C.c_void_p), var.ctypes.data_as(C.c_void_p)),
                   "uqk_get_bn_stats")
        return mean, var

    def set_bn_stats(self, mean: np.ndarray, var: np.ndarray):
        mean = np.ascontiguousarray(mean, dtype=np.float32).reshape(arch.N_BLOCKS, arch.MAX_CH)
        var = np.ascontiguousarray(var, dtype=np.float32).reshape(arch.N_BLOCKS, arch.MAX_CH)
        _lib.check(self.lib.uqk_set_bn_stats(self._h, mean.ctypes.data_as(C.c_void_p), var.ctypes.data_as(C.c_void_p)),
                   "uqk_set_bn_stats")
        self.calibrated = True

    def mcd(self, x, n_passes: int = 50, seed: int = 2025, bn_mode: str = "calibrated", precision: str = "bf16",
            masks=None, return_probs: bool = False, metrics: Sequence[str] = METRIC_KEYS, window_base: int = 0,
            device_out: bool = False, exact_labels: bool = False, guard_delta: float = GUARD_DELTA) -> Dict:
        """T stochastic passes of member 0 + per-window reductions (uq_techniques.py:12-24 fused with :40-112).
        ``masks``: optional (T,N,6,60,8) uint32 keep-bits (mask injection).
        ``exact_labels``: windows whose bf16 mean probability lies within ``guard_delta`` of 0.5 are re-evaluated on
        the fp32 path with the same masks and statistics (``uqk_refine_band``); ``self.last_guard_band`` then holds
        (re-evaluated windows, all windows)."""
        with torch.cuda.device(self.device):
            xt, dt = self.to_device(x)
            n = xt.shape[0]
            bufs, mo = self._alloc_metrics(n, metrics)
            probs = torch.empty((n_passes, n), dtype=torch.float32, device=self._dev()) if return_probs else None
            mt = None
            if masks is not None:
                mt = masks if isinstance(masks, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(masks, dtype=np.uint32).view(np.int32))
                if tuple(mt.shape) != (n_passes, n, arch.N_BLOCKS, arch.WINDOW_LEN, 8):
                    raise ValueError(f"masks must have shape (T,N,6,60,8), got {tuple(mt.shape)}")
                mt = mt.to(self._dev()).contiguous()
            rc = self.lib.uqk_mcd(self._h, xt.data_ptr(), dt, n, int(n_passes), C.c_uint64(seed & (2 ** 64 - 1)),
                                  _BN[bn_mode], _PREC[precision], int(window_base),
                                  mt.data_ptr() if mt is not None else None,
                                  probs.data_ptr() if probs is not None else None,
                                  C.byref(mo) if bufs else None, self._stream())
            if rc != 0 and getattr(self, "_xchg_error", None) is not None:
                err, self._xchg_error = self._xchg_error, None
                raise RuntimeError("batch-statistics exchange failed") from err
            _lib.check(rc, "uqk_mcd")
            if exact_labels and precision in ("bf16", "bf16_tc") and "mean_pred" in bufs and n > 0:
                self._refine(0, xt, dt, n, int(n_passes), seed, _BN[bn_mode], int(window_base), mt, guard_delta, probs, mo)
            return self._finish(bufs, probs, device_out)

    def de(self, x, n_members: Optional[int] = None, precision: str = "bf16", return_probs: bool = False,
           metrics: Sequence[str] = METRIC_KEYS, device_out: bool = False, exact_labels: bool = False,
           guard_delta: float = GUARD_DELTA) -> Dict:
        """M ensemble members in inference mode + per-window reductions (uq_techniques.py:26-32 fused with :40-112).
        ``exact_labels``: see ``mcd``."""
        with torch.cuda.device(self.device):
            m = self.n_members if n_members is None else int(n_members)
            xt, dt = self.to_device(x)
            n = xt.shape[0]
            bufs, mo = self._alloc_metrics(n, metrics)
            probs = torch.empty((m, n), dtype=torch.float32, device=self._dev()) if return_probs else None
            _lib.check(self.lib.uqk_de(self._h, xt.data_ptr(), dt, n, m, _PREC[precision],
                                       probs.data_ptr() if probs is not None else None,
                                       C.byref(mo) if bufs else None, self._stream()), "uqk_de")
            if exact_labels and precision in ("bf16", "bf16_tc") and "mean_pred" in bufs and n > 0:
                self._refine(1, xt, dt, n, m, 0, _lib.BN_INFERENCE, 0, None, guard_delta, probs, mo)
            return self._finish(bufs, probs, device_out)

    def _refine(self, mode_de, xt, dt, n, n_groups, seed, bn_code, window_base, mt, delta, probs, mo):
        cnt = C.c_int64(0)
        _lib.check(self.lib.uqk_refine_band(self._h, int(mode_de), xt.data_ptr(), dt, n, int(n_groups),
                                            C.c_uint64(seed & (2 ** 64 - 1)), int(bn_code), int(window_base),
                                            mt.data_ptr() if mt is not None else None, C.c_float(delta),
                                            probs.data_ptr() if probs is not None else None, C.byref(mo),
                                            C.byref(cnt), self._stream()), "uqk_refine_band")
        self.last_guard_band = (int(cnt.value), int(n))

    def reduce(self, probs, metrics: Sequence[str] = METRIC_KEYS, device_out: bool = False) -> Dict:
        """(T,N) float32 probabilities -> per-window metrics (uq_techniques.py:40-112)."""
        with torch.cuda.device(self.device):
            pt = probs if isinstance(probs, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(probs, dtype=np.float32))
            if pt.dim() != 2:
                raise ValueError(f"probabilities must be (T, N), got {tuple(pt.shape)}")
            pt = pt.to(self._dev(), dtype=torch.float32).contiguous()
            t, n = pt.shape
            bufs, mo = self._alloc_metrics(n, metrics)
            _lib.check(self.lib.uqk_reduce(pt.data_ptr(), int(t), int(n), C.byref(mo), self._stream()), "uqk_reduce")
            return self._finish(bufs, None, device_out)

    def bootstrap(self, metrics: Dict[str, np.ndarray], y_true: np.ndarray, idx: np.ndarray) -> np.ndarray:
        """Gather-mean aggregates for bootstrap index vectors ``idx`` (B,N) int64 -> (B,6) float64
        (uq_techniques.py:137-158)."""
        with torch.cuda.device(self.device):
            dev = self._dev()
            f = lambda k: torch.as_tensor(np.ascontiguousarray(metrics[k], dtype=np.float32)).to(dev)
            v, tt, ee, mi = f("pred_variance"), f("total_pred_entropy"), f("expected_aleatoric_entropy"), f("mutual_info")
            y = torch.as_tensor(np.ascontiguousarray(y_true).astype(np.int32)).to(dev)
            it = torch.as_tensor(np.ascontiguousarray(idx, dtype=np.int64)).to(dev)
            b, n = it.shape
            out = torch.empty((b, 6), dtype=torch.float64, device=dev)
            _lib.check(self.lib.uqk_bootstrap(v.data_ptr(), tt.data_ptr(), ee.data_ptr(), mi.data_ptr(), y.data_ptr(),
                                              it.data_ptr(), int(b), int(n), out.data_ptr(), self._stream()), "uqk_bootstrap")
            return out.cpu().numpy()

    def patient_reduce(self, variance, entropy, y_true, y_pred, seg_offsets, device_out: bool = False):
        """Per-patient rows (P,8) float64 from patient-sorted columns (aggregate_patient_uq_metrics.py:34-46).
        Columns may be host arrays or CUDA tensors of this device (float32 / int32 / int64 offsets)."""
        with torch.cuda.device(self.device):
            dev = self._dev()

            def col(a, np_dtype, t_dtype):
                if isinstance(a, torch.Tensor):
                    return a.to(dev, dtype=t_dtype).contiguous()
                return torch.as_tensor(np.ascontiguousarray(np.asarray(a).astype(np_dtype))).to(dev)
            v, e = col(variance, np.float32, torch.float32), col(entropy, np.float32, torch.float32)
            yt, yp = col(y_true, np.int32, torch.int32), col(y_pred, np.int32, torch.int32)
            seg = col(seg_offsets, np.int64, torch.int64)
            p = seg.shape[0] - 1
            out = torch.empty((p, 8), dtype=torch.float64, device=dev)
            _lib.check(self.lib.uqk_patient_reduce(v.data_ptr(), e.data_ptr(), yt.data_ptr(), yp.data_ptr(),
                                                   seg.data_ptr(), int(p), out.data_ptr(), self._stream()),
                       "uqk_patient_reduce")
            return out if device_out else out.cpu().numpy()

    def standardize(self, x_raw, device_out: bool = False):
        """Per-window z-score of raw (N,60,4) windows (prepare_numpy_datasets.py:83-95), same dtype out."""
        with torch.cuda.device(self.device):
            xt, dt = self.to_device(x_raw)
            out = torch.empty_like(xt)
            _lib.check(self.lib.uqk_standardize(xt.data_ptr(), dt, xt.shape[0], out.data_ptr(), self._stream()),
                       "uqk_standardize")
            return out if device_out else out.cpu().numpy()

    def debug_dump(self) -> np.ndarray:
        out = np.zeros((arch.N_BLOCKS, 128, 256), dtype=np.float32)
        _lib.check(self.lib.uqk_debug_read_dump(self._h, out.ctypes.data_as(C.c_void_p)), "uqk_debug_read_dump")
        return out


_engines: Dict[int, Engine] = {}


def get_engine(device: Optional[int] = None) -> Engine:
    """Process-wide engine cache, one per device."""
    _require_cuda()
    d = torch.cuda.current_device() if device is None else int(device)
    if d not in _engines:
        _engines[d] = Engine(d)
    return _engines[d]
