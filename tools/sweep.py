#!/usr/bin/env python
"""Scaling sweep of BASELINE configs[4]: T in {10,25,50,100}, M in {5,10,20} on one GPU (device-resident, fused path).
Writes gpurun_out/sweep.json; copy into profiles/ to keep."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
import numpy as np, torch
pkg = importlib.import_module(PKG)
engine = importlib.import_module(PKG + ".engine")
syn = pkg.synthetic
FLOP = 98_844_864

def timed(fn):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)

def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 400_000
    eng = engine.Engine(0)
    blobs = np.stack([syn.weights_to_blob(syn.make_member_weights(m)) for m in range(20)])
    xu = torch.from_numpy(syn.make_windows(1 << 16, seed=6))
    x = xu.repeat((n >> 16) + 1, 1, 1)[:n].contiguous().cuda()
    res = []
    eng.set_members(blobs[:1])
    eng.calibrate(x, n_passes=1)
    for T in (10, 25, 50, 100):
        ms = timed(lambda: eng.mcd(x, n_passes=T, bn_mode="calibrated", device_out=True))
        res.append({"mode": "mcd", "T": T, "windows": n, "ms": ms, "windows_per_s": n / ms * 1e3, "tflops_valid": n * T * FLOP / ms / 1e9})
        print(json.dumps(res[-1]), flush=True)
    eng.set_members(blobs)
    for M in (5, 10, 20):
        ms = timed(lambda: eng.de(x, n_members=M, device_out=True))
        res.append({"mode": "de", "M": M, "windows": n, "ms": ms, "windows_per_s": n / ms * 1e3, "tflops_valid": n * M * FLOP / ms / 1e9})
        print(json.dumps(res[-1]), flush=True)
    ms = timed(lambda: eng.forward_det(x, member=0, device_out=True))
    res.append({"mode": "deterministic", "windows": n, "ms": ms, "windows_per_s": n / ms * 1e3, "tflops_valid": n * FLOP / ms / 1e9})
    print(json.dumps(res[-1]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "sweep.json"), "w"), indent=1)

if __name__ == "__main__":
    main()
