#!/usr/bin/env python
"""Device-side timing of the kernels either side of the hot path: window
standardisation, bootstrap gather-means, patient aggregation, window-level summary.  CUDA events, device-resident
inputs, algorithmic bytes against the measured HBM copy bandwidth.  Prints one JSON line."""
import ctypes as C, importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
import numpy as np, torch
pkg = importlib.import_module(PKG)
engine = importlib.import_module(PKG + ".engine")
_lib = importlib.import_module(PKG + "._lib")


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    eng = engine.Engine(0)
    lib, st = eng.lib, eng._stream()
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.isfile(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6548.5
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(3)
    res = {"peak_gbs": peak}

    def entry(name, ms, nbytes, n):
        res[name] = {"ms": ms, "windows": n, "windows_per_s": n / (ms * 1e-3), "algorithmic_bytes": nbytes,
                     "achieved_gbs": nbytes / (ms * 1e-3) / 1e9, "frac": nbytes / (ms * 1e-3) / 1e9 / peak}

    # N3: standardize, raw float64 (N,60,4) -> float64
    n = 1_000_000
    x = torch.randn((n, 60, 4), device=dev, generator=g, dtype=torch.float64)
    out = torch.empty_like(x)
    ms = timed(lambda: _lib.check(lib.uqk_standardize(x.data_ptr(), _lib.F64, n, out.data_ptr(), st)))
    entry("standardize_f64", ms, 2 * x.numel() * 8, n)
    del x, out

    # per-window columns of a cohort (config 4 size)
    n = 1_430_000
    var = torch.rand(n, device=dev, generator=g) * 0.2
    ent = torch.rand(n, device=dev, generator=g)
    tot, exp_, mi = ent * 0.69, ent * 0.5, ent * 0.19
    y = (torch.rand(n, device=dev, generator=g) < 0.2).to(torch.int32)
    lab = (torch.rand(n, device=dev, generator=g) < 0.3).to(torch.int32)

    # N1: bootstrap gather-means, B = 100 resamples
    b = 100
    idx = torch.randint(0, n, (b, n), device=dev, generator=g, dtype=torch.int64)
    o6 = torch.empty((b, 6), dtype=torch.float64, device=dev)
    ms = timed(lambda: _lib.check(lib.uqk_bootstrap(var.data_ptr(), tot.data_ptr(), exp_.data_ptr(), mi.data_ptr(), y.data_ptr(),
                                                    idx.data_ptr(), b, n, o6.data_ptr(), st)), reps=3)
    entry("bootstrap_B100", ms, b * n * (8 + 4 * 4 + 4), n)
    del idx

    # K5: patient aggregation, 2651 contiguous patients
    cuts = np.sort(np.random.default_rng(1).choice(np.arange(1, n), 2650, replace=False))
    seg = torch.from_numpy(np.concatenate([[0], cuts, [n]]).astype(np.int64)).to(dev)
    p = seg.numel() - 1
    o8 = torch.empty((p, 8), dtype=torch.float64, device=dev)
    ms = timed(lambda: _lib.check(lib.uqk_patient_reduce(var.data_ptr(), ent.data_ptr(), y.data_ptr(), lab.data_ptr(), seg.data_ptr(),
                                                         p, o8.data_ptr(), st)))
    entry("patient_reduce_2651", ms, n * 16, n)

    # N2: window-level summary (describe + 10 entropy bins)
    osum = torch.empty(35 + 30, dtype=torch.float64, device=dev)
    ms = timed(lambda: _lib.check(lib.uqk_window_summary(ent.data_ptr(), var.data_ptr(), y.data_ptr(), lab.data_ptr(), n, 10,
                                                         osum.data_ptr(), st)))
    entry("window_summary_10bins", ms, n * 16, n)
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
