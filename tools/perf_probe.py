#!/usr/bin/env python
"""Throughput probe of the fused kernel for a list of (variant, mode) pairs; prints TFLOP/s (valid-tap count).
Bring-up tool, not part of the product or of bench.py."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
import numpy as np, torch
pkg = importlib.import_module(PKG)
if os.environ.get("UQK_PROBE_LIB"):      # A/B runs of two builds on the same box (bring-up only)
    importlib.import_module(PKG + "._lib").LIB_PATH = os.path.join(ROOT, PKG, os.environ["UQK_PROBE_LIB"])
engine = importlib.import_module(PKG + ".engine")
syn = pkg.synthetic

def main():
    variants = [int(v) for v in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["0", "16"])]
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 40000
    T = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    eng = engine.Engine(0)
    for kv in os.environ.get("UQK_PROBE_OPTS", "").split(","):      # e.g. UQK_PROBE_OPTS=exact_pipelined=1
        if "=" in kv:
            eng.option(kv.split("=")[0], int(kv.split("=")[1]))
    eng.set_members(np.stack([syn.weights_to_blob(syn.make_member_weights(m)) for m in range(2)]))
    x = torch.from_numpy(syn.make_windows(4096, seed=6)).cuda()
    x = x.repeat((n + 4095) // 4096, 1, 1)[:n].contiguous()
    res = []
    for v in variants:
        eng.option("variant", v)
        for mode in (sys.argv[4].split(",") if len(sys.argv) > 4 else ("mcd", "det")):
            def run():
                if mode == "mcd":
                    eng.mcd(x, n_passes=T, bn_mode="inference", device_out=True, metrics=("mean_pred",))
                elif mode == "exact":     # exact batch-statistic BatchNorm: 6 statistics sweeps + the final run
                    eng.mcd(x, n_passes=T, bn_mode="batch", device_out=True, metrics=("mean_pred",))
                else:
                    eng.de(x, n_members=1, device_out=True, metrics=("mean_pred",))
            run(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); run(); e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            passes = T if mode in ("mcd", "exact") else 1
            tf = n * passes * 98844864 / (ms * 1e-3) / 1e12
            res.append({"variant": v, "mode": mode, "ms": ms, "tflops": tf, "windows_per_s": n / (ms * 1e-3)})
            print(json.dumps(res[-1]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "perf_probe.json"), "w"), indent=1)

if __name__ == "__main__":
    main()
