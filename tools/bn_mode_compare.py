#!/usr/bin/env python
"""How far the calibrated-statistics mode (bench headline) is from the exact batch-statistic mode (the reference's
`model(x, training=True)` semantics) on the same windows, passes and dropout masks.  Prints one JSON line."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
import numpy as np, torch
pkg = importlib.import_module(PKG)
engine = importlib.import_module(PKG + ".engine")
syn = pkg.synthetic


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    seed = 2025
    eng = engine.Engine(0)
    eng.set_members(syn.weights_to_blob(syn.make_member_weights(0))[None])
    x = torch.from_numpy(syn.make_windows(n, seed=7)).cuda()
    exact = eng.mcd(x, n_passes=T, seed=seed, bn_mode="batch", precision="bf16")
    res = {"windows": n, "passes": T}
    for calib, passes in ((4096, 2), (32768, 2), (65536, 2), (65536, 4), (131072, 2)):
        if calib > n:
            continue
        eng.calibrate(x, n_windows=calib, n_passes=passes, seed=seed)      # warm (allocations)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng.calibrate(x, n_windows=calib, n_passes=passes, seed=seed); e1.record(); torch.cuda.synchronize()
        cal = eng.mcd(x, n_passes=T, seed=seed, bn_mode="calibrated", precision="bf16")
        r = {"calibrate_ms": e0.elapsed_time(e1), "path": "fp32 layer-synchronous" if calib <= 8192 else "tensor-core sweeps"}
        for k in ("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info", "entropy_bits"):
            d = np.abs(cal[k].astype(np.float64) - exact[k])
            r[k] = {"mean_abs_diff": float(d.mean()), "max_abs_diff": float(d.max()), "exact_mean": float(np.mean(exact[k]))}
        r["label_flips"] = int((cal["label"] != exact["label"]).sum())
        r["label_flip_fraction"] = r["label_flips"] / n
        res[f"calibrated_{calib}x{passes}"] = r
    inf = eng.mcd(x, n_passes=T, seed=seed, bn_mode="inference", precision="bf16")
    res["inference_stats"] = {"mean_pred_mean_abs_diff": float(np.abs(inf["mean_pred"] - exact["mean_pred"]).mean()),
                              "label_flip_fraction": float((inf["label"] != exact["label"]).mean())}
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
