#!/usr/bin/env python
"""Schedule of one unit of the default fused kernel (fused_pair2_kernel), from clock64 stamps taken inside the kernel.

Bring-up tool.  Needs the instrumented build (`make -C <pkg>/csrc ../libuqk_timeline.so`, -DUQK_TIMELINE), which the
package never loads by itself.  CTA 0 (leader of pair 0) records, for its third unit and the first 16 passes, per
tile-pair and layer:
  ev0  issuer: "activations ready" observed, first MMA of the layer about to be issued
  ev1  issuer: last MMA + commit of the layer issued
  ev6  epilogue warp 0: entered the layer (previous layer's epilogue done)
  ev2  epilogue warp 0: keep-flags of the layer generated, starts polling the accumulator barrier
  ev3  epilogue warp 0: accumulator complete (after the tile's named barrier)
  ev4  epilogue warp 0 / ev5 warp 7: layer's epilogue done ("activations ready" signalled)

  python tools/timeline_probe.py [windows] [passes] [mcd|det] [variant] [--all]
(variant bits 12-15 select the unit of CTA 0 that is recorded; --all prints every pass instead of passes 4-5)
"""
import ctypes, importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
import numpy as np, torch
pkg = importlib.import_module(PKG)
_lib = importlib.import_module(PKG + "._lib")
_lib.LIB_PATH = os.path.join(ROOT, PKG, os.environ.get("UQK_PROBE_LIB", "libuqk_timeline.so"))
engine = importlib.import_module(PKG + ".engine")
syn = pkg.synthetic


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 148 * 4 * 8
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    mode = sys.argv[3] if len(sys.argv) > 3 else "mcd"
    eng = engine.Engine(0)
    if len(sys.argv) > 4:
        eng.option("variant", int(sys.argv[4]))
    lib = _lib.load()
    lib.uqk_debug_set_timeline.argtypes = [ctypes.c_void_p]
    eng.set_members(np.stack([syn.weights_to_blob(syn.make_member_weights(m)) for m in range(16)]))
    x = torch.from_numpy(syn.make_windows(n, seed=6)).cuda()
    buf = torch.zeros(2 * 2 * 16 * 6 * 8, dtype=torch.int64, device="cuda")

    def run():
        if mode == "mcd":
            eng.mcd(x, n_passes=T, bn_mode="inference", device_out=True, metrics=("mean_pred",))
        else:
            eng.de(x, n_members=min(T, 16), device_out=True, metrics=("mean_pred",))
    run(); torch.cuda.synchronize()
    lib.uqk_debug_set_timeline(ctypes.c_void_p(buf.data_ptr()))
    run(); torch.cuda.synchronize()
    lib.uqk_debug_set_timeline(ctypes.c_void_p(0))
    both = buf.cpu().numpy().reshape(2, 2, 16, 6, 8)
    t, seg = both[0], both[1]
    g_n = min(T, 16)
    t0 = t[:, :g_n][t[:, :g_n] > 0].min()
    rel = np.where(t > 0, t - t0, -1)
    out = {"mode": mode, "windows": n, "passes": T, "events": ["issue_start", "issue_done", "flags_done", "acc_ready", "epi_done_w0", "epi_done_w7", "layer_enter", "input_signalled"]}
    rows = []
    for g in range(g_n):
        for L in range(6):
            for tl in range(2):
                e = rel[tl, g, L]
                rows.append({"pass": g, "layer": L + 1, "tile_pair": tl, **{out["events"][k]: int(e[k]) for k in range(8)}})
    out["stamps"] = rows
    # per-layer means over passes 2.. (steady state) and both tile-pairs
    summ = []
    sl = slice(2, g_n)
    for L in range(6):
        e = t[:, sl, L].astype(np.float64)
        mma = e[..., 3] - e[..., 0]                   # activations ready -> accumulator complete (incl. queueing behind the other tile-pair)
        issue = e[..., 1] - e[..., 0]
        epi = np.maximum(e[..., 4], e[..., 5]) - e[..., 3]
        flags = e[..., 2] - e[..., 6]
        poll = e[..., 3] - e[..., 2]
        summ.append({"layer": L + 1, "act_to_acc": mma.mean(), "issue": issue.mean(), "epilogue": epi.mean(), "flags": flags.mean(), "epi_warp_idle": poll.mean()})
    out["per_layer_mean_clk"] = summ
    # issuer segments (clk per layer, mean over passes 2..): waiting for the stage's weights / issuing MMAs + commit / loop overhead
    for L in range(6):
        e = seg[:, sl, L].astype(np.float64).mean(axis=(0, 1))
        print(json.dumps({"layer": L + 1, "issuer_wait_full": round(e[0]), "issuer_issue_commit": round(e[1]), "issuer_other": round(e[2])}))
    per_pass = (t[:, g_n - 1, 5, 4] - t[:, 2, 5, 4]).astype(np.float64) / max(g_n - 3, 1)
    out["clk_per_pass"] = per_pass.tolist()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"timeline_{mode}_{sys.argv[4] if len(sys.argv) > 4 else 0}.json"), "w"), indent=1)
    for s in summ:
        print(json.dumps({k: (round(v, 0) if isinstance(v, float) else v) for k, v in s.items()}))
    print("clk per pass (tile-pair 0, 1):", per_pass)
    # a compact chart of passes 4..5
    for g in (range(0, g_n) if "--all" in sys.argv else range(4, min(6, g_n))):
        for L in range(6):
            for tl in range(2):
                e = rel[tl, g, L]
                print(f"g{g} L{L+1} tp{tl}: issue {e[0]:>8d}..{e[1]:>8d}  acc {e[3]:>8d}  epi_done {max(e[4], e[5]):>8d}  (enter {e[6]:>8d} flags {e[2]:>8d})")


if __name__ == "__main__":
    main()
