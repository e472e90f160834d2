#!/usr/bin/env python
"""GPU bring-up diagnostics: runs each check in its own subprocess (a trap or hang in one kernel variant must
not take the others down) and writes gpurun_out/diag_*.json.  Not part of the product or the test-suite."""
import importlib, json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
OUT = os.path.join(ROOT, "gpurun_out")


def child(which):
    import numpy as np
    pkg = importlib.import_module(PKG)
    engine = importlib.import_module(PKG + ".engine")
    from oracle import network, philox
    syn, arch = pkg.synthetic, pkg.arch
    res = {"which": which}
    eng = engine.Engine(0)
    n = 10
    x = syn.make_windows(n, seed=5)
    ws = [syn.make_member_weights(m) for m in range(2)]
    eng.set_members(np.stack([syn.weights_to_blob(w) for w in ws]))
    ref, acts = network.forward_f64(ws[0], x, return_acts=True)
    ref = ref[:, 0]
    if which == "ls":
        p = eng.forward_det(x, 0, precision="fp32")
        res["ls_det_maxerr"] = float(np.abs(p - ref).max())
        seed = 99
        out = eng.mcd(x, n_passes=2, seed=seed, bn_mode="batch", precision="fp32", return_probs=True)
        r = []
        for t in range(2):
            masks = [philox.dropout_keep_mask(seed, np.arange(n), t, i, b.cout, b.rate) for i, b in enumerate(arch.BLOCKS)]
            r.append(network.forward_f64(ws[0], x, training=True, masks=masks)[:, 0])
        res["ls_mcd_batch_maxerr"] = float(np.abs(out["probs"] - np.stack(r)).max())
        res["ls_probs"] = out["probs"].tolist()
    else:
        variant = int(which.split("_")[1])
        eng.option("variant", variant)
        eng.option("dbg_dump", 1)
        t0 = time.time()
        p = eng.forward_det(x, 0, precision="bf16")
        res["time_s"] = time.time() - t0
        res["fused_det"] = p.tolist()
        res["ref_det"] = ref.tolist()
        res["fused_det_maxerr"] = float(np.abs(p - ref).max())
        d = eng.debug_dump()
        lay = []
        for l, b in enumerate(arch.BLOCKS):
            got = np.concatenate([d[l, 0:60, :b.cout], d[l, 64:124, :b.cout]])
            want = np.concatenate([acts[l][0], acts[l][1]])
            err = np.abs(got - want)
            lay.append({"layer": l, "maxerr": float(err.max()), "meanerr": float(err.mean()), "ref_absmean": float(np.abs(want).mean()),
                        "got_absmean": float(np.abs(got).mean()), "pad_rows_absmax": float(np.abs(d[l, 60:64]).max()),
                        "nan": int(np.isnan(got).sum())})
        res["layers"] = lay
        # sample values of layer 0 for manual inspection
        res["l0_got_row0"] = d[0, 0, :8].tolist()
        res["l0_ref_row0"] = acts[0][0][0, :8].tolist()
        eng.option("dbg_dump", 0)
        # ensemble + mcd quick
        de = eng.de(x, n_members=2, return_probs=True)
        ref2 = np.stack([network.forward_f64(w, x)[:, 0] for w in ws])
        res["de_maxerr"] = float(np.abs(de["probs"] - ref2).max())
        seed = 2025
        mcd = eng.mcd(x, n_passes=3, seed=seed, bn_mode="inference", return_probs=True)
        r = []
        for t in range(3):
            masks = [philox.dropout_keep_mask(seed, np.arange(n), t, i, b.cout, b.rate) for i, b in enumerate(arch.BLOCKS)]
            r.append(network.forward_f64(ws[0], x, training=True, masks=masks, bn_mode="inference")[:, 0])
        res["mcd_maxerr"] = float(np.abs(mcd["probs"] - np.stack(r)).max())
        res["mcd_probs"] = mcd["probs"].tolist()
        res["mcd_ref"] = np.stack(r).tolist()
        # quick throughput probe
        import torch
        nb = 20000
        xb = torch.from_numpy(syn.make_windows(nb, seed=6)).cuda()
        for T in (10,):
            eng.mcd(xb, n_passes=T, bn_mode="inference", device_out=True)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); eng.mcd(xb, n_passes=T, bn_mode="inference", device_out=True); e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            res[f"mcd_T{T}_win_per_s"] = nb / (ms * 1e-3)
            res[f"mcd_T{T}_tflops"] = nb * T * 98844864 / (ms * 1e-3) / 1e12
    print("DIAG_RESULT " + json.dumps(res))


def main():
    os.makedirs(OUT, exist_ok=True)
    which_list = sys.argv[1:] or ["ls", "fused_0", "fused_1", "fused_16"]
    for w in which_list:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", w], capture_output=True, text=True, timeout=240)
            out, err, rc = r.stdout, r.stderr, r.returncode
        except subprocess.TimeoutExpired as e:
            out, err, rc = (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or ""), "TIMEOUT", -999
        rec = {"which": w, "rc": rc, "secs": time.time() - t0, "stderr_tail": err[-3000:] if err else ""}
        for line in out.splitlines():
            if line.startswith("DIAG_RESULT "):
                rec["result"] = json.loads(line[len("DIAG_RESULT "):])
        if "result" not in rec:
            rec["stdout_tail"] = out[-2000:]
        with open(os.path.join(OUT, f"diag_{w}.json"), "w") as f:
            json.dump(rec, f, indent=1)
        summ = {k: v for k, v in rec.get("result", {}).items() if not isinstance(v, list) or k == "layers"}
        print(w, "rc", rc, json.dumps(summ)[:3000])
        if rc != 0:
            print(rec["stderr_tail"][-1500:])


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        child(sys.argv[2])
    else:
        main()
