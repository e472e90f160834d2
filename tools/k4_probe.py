#!/usr/bin/env python
"""Bandwidth of the stand-alone reduction kernel K4 (uqk_reduce: (T,N) float32 probabilities -> 7 per-window arrays).
Algorithmic bytes per window: 4*T read + 28 written; peak = measured HBM copy bandwidth.
usage: tools/k4_probe.py [N] [T]   -- prints one JSON line"""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
import numpy as np, torch
engine = importlib.import_module(PKG + ".engine")


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    eng = engine.Engine(0)
    g = torch.Generator(device="cuda").manual_seed(1)
    probs = torch.rand((T, n), device="cuda", generator=g, dtype=torch.float32)      # 800 MB at the defaults: > L2
    eng.reduce(probs, device_out=True); torch.cuda.synchronize()
    reps = 5
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        eng.reduce(probs, device_out=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    by = n * (4 * T + 28)
    peak = 6548.5
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    res = {"kernel": "reduce_kernel_smem (K4)", "windows": n, "passes": T, "ms": ms, "windows_per_s": n / (ms * 1e-3),
           "algorithmic_bytes": by, "achieved_gbs": by / (ms * 1e-3) / 1e9, "peak_gbs": peak, "frac": by / (ms * 1e-3) / 1e9 / peak}
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
