"""ctypes binding of libuqk.so (C-ABI declared in include/uqk.h).

There is no CPU fallback: if the shared library has not been built, or no sm_100 device is present, every
compute call raises.  Build with ``python __graft_entry__.py`` (or ``make -C <pkg>/csrc``).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libuqk.so")

OK = 0
F32, F64 = 0, 1
BN_INFERENCE, BN_BATCH, BN_CALIBRATED = 0, 1, 2
PREC_BF16_TC, PREC_FP32 = 0, 1

EXPORTS = [
    "uqk_version", "uqk_last_error", "uqk_create", "uqk_destroy", "uqk_set_model", "uqk_forward_det",
    "uqk_calibrate", "uqk_get_bn_stats", "uqk_set_bn_stats", "uqk_mcd", "uqk_de", "uqk_reduce",
    "uqk_patient_reduce", "uqk_standardize", "uqk_bootstrap", "uqk_bootstrap_noalloc", "uqk_launch_count", "uqk_set_option",
    "uqk_set_stats_exchange", "uqk_write_detail_csv", "uqk_window_summary", "uqk_refine_band",
    "uqk_poll_error", "uqk_debug_trip_watchdog",
]

# int fn(void* user, long long* stats_dev, int count, void* stream)  -- include/uqk.h uqk_stats_exchange_fn
STATS_EXCHANGE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p)


class MetricsOut(C.Structure):
    """Mirror of ``uqk_metrics_out``: device pointers, NULL = not produced."""
    _fields_ = [("mean_pred", C.c_void_p), ("pred_variance", C.c_void_p), ("total_pred_entropy", C.c_void_p),
                ("expected_aleatoric_entropy", C.c_void_p), ("mutual_info", C.c_void_p),
                ("entropy_bits", C.c_void_p), ("label", C.c_void_p)]


class UqkError(RuntimeError):
    pass


_lib = None


def load() -> C.CDLL:
    """Loads libuqk.so once; raises loudly when it is missing (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise UqkError(
            f"{LIB_PATH} not found: the CUDA library has not been built. Run `python __graft_entry__.py` "
            "(nvcc, sm_100a). This package has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
    lib.uqk_version.restype = i32
    lib.uqk_last_error.restype = C.c_char_p
    lib.uqk_create.argtypes = [i32, C.POINTER(vp)]
    lib.uqk_destroy.argtypes = [vp]
    lib.uqk_set_model.argtypes = [vp, i32, vp, C.c_size_t]
    lib.uqk_forward_det.argtypes = [vp, i32, vp, i32, i64, i32, vp, vp]
    lib.uqk_calibrate.argtypes = [vp, vp, i32, i64, i32, u64, i64, vp]
    lib.uqk_get_bn_stats.argtypes = [vp, vp, vp]
    lib.uqk_set_bn_stats.argtypes = [vp, vp, vp]
    lib.uqk_mcd.argtypes = [vp, vp, i32, i64, i32, u64, i32, i32, i64, vp, vp, C.POINTER(MetricsOut), vp]
    lib.uqk_de.argtypes = [vp, vp, i32, i64, i32, i32, vp, C.POINTER(MetricsOut), vp]
    lib.uqk_reduce.argtypes = [vp, i32, i64, C.POINTER(MetricsOut), vp]
    lib.uqk_refine_band.argtypes = [vp, i32, vp, i32, i64, i32, u64, i32, i64, vp, C.c_float, vp, C.POINTER(MetricsOut), C.POINTER(i64), vp]
    lib.uqk_poll_error.argtypes = [vp]
    lib.uqk_debug_trip_watchdog.argtypes = [vp]
    lib.uqk_patient_reduce.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp]
    lib.uqk_standardize.argtypes = [vp, i32, i64, vp, vp]
    lib.uqk_bootstrap.argtypes = [vp, vp, vp, vp, vp, vp, i32, i64, vp, vp]
    lib.uqk_bootstrap_noalloc.argtypes = [vp, vp, vp, vp, vp, vp, i32, i64, vp, vp]
    lib.uqk_launch_count.argtypes = [vp]
    lib.uqk_launch_count.restype = i64
    lib.uqk_set_option.argtypes = [vp, C.c_char_p, i64]
    lib.uqk_set_stats_exchange.argtypes = [vp, STATS_EXCHANGE_FN, vp, i64]
    lib.uqk_write_detail_csv.argtypes = [C.c_char_p, i64, vp, i64, vp, vp, vp, vp, vp, i32, i32]
    lib.uqk_window_summary.argtypes = [vp, vp, vp, vp, i64, i32, vp, vp]
    if hasattr(lib, "uqk_debug_read_dump"):
        lib.uqk_debug_read_dump.argtypes = [vp, vp]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if fn.restype is C.c_int and name not in ("uqk_version",):
            fn.restype = C.c_int
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    """Maps the int error convention of the C-ABI onto Python exceptions (ValueError for bad
    arguments / shapes, UqkError otherwise) -- the callers' own ``except Exception`` blocks
    (reference analyze_mcd_patient_level.py:236) keep working."""
    if rc == OK:
        return
    msg = load().uqk_last_error().decode("utf-8", "replace")
    text = f"{what}: {msg}" if what else msg
    if rc in (-1, -2):
        raise ValueError(text)
    raise UqkError(f"{text} (code {rc})")
