#!/usr/bin/env python
"""BASELINE configs[3]: SHHS2-scale synthetic cohort (2651 patients x full-night windows), MC-Dropout T=50 and Deep
Ensemble M=5 with patient-level aggregation; one process per GPU under torchrun (patient-aligned shards, NCCL gather
of the per-window metrics), or a single GPU.  Prints one JSON line with timings and cohort-level summaries."""
import importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
import numpy as np, torch
import torch.distributed as dist

def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module(PKG)
    sh = importlib.import_module(PKG + ".sharding")
    model_mod = importlib.import_module(PKG + ".model")
    patient = importlib.import_module(PKG + ".patient")
    sys.path.insert(0, os.path.join(ROOT, PKG, "uncertainty_quantification"))
    uq = importlib.import_module("uq_techniques")
    syn = pkg.synthetic
    n_pat = int(sys.argv[1]) if len(sys.argv) > 1 else 2651
    pid = syn.make_patient_ids(n_patients=n_pat)
    n = len(pid)
    y = syn.make_labels(n)
    shards = sh.patient_aligned_shards(pid, world)
    s, e = shards[rank]
    xu = syn.make_windows(min(e - s, 1 << 16), seed=100 + rank)
    x = torch.from_numpy(xu).repeat((e - s) // len(xu) + 1, 1, 1)[: e - s].contiguous().cuda()
    model = model_mod.synthetic_model(0)
    members = [model_mod.synthetic_model(m) for m in range(5)]
    res = {"patients": n_pat, "windows": n, "world": world, "shard_sizes": [b - a for a, b in shards]}
    for name, fn in (("mcd_T50", lambda: uq.mc_dropout_uq(model, x, n_pred=50, device_out=True, window_base=s)),
                     ("de_M5", lambda: uq.deep_ensemble_uq(members, x, device_out=True))):
        fn(); torch.cuda.synchronize()
        if world > 1: dist.barrier()
        t0 = time.perf_counter()
        out = fn()
        local_m = {k: v for k, v in out.items() if isinstance(v, torch.Tensor)}
        full = sh.gather_metrics(local_m, shards)
        torch.cuda.synchronize()
        t_fwd = time.perf_counter() - t0
        if rank == 0:
            t1 = time.perf_counter()
            summ = patient.aggregate_arrays(pid, full["pred_variance"].cpu().numpy(), full["predictive_entropy_bits"].cpu().numpy(),
                                            y, full["predicted_label"].cpu().numpy(), device=local)
            res[name] = {"forward_plus_gather_s": t_fwd, "windows_per_s": n / t_fwd, "patient_reduce_s": time.perf_counter() - t1,
                         "n_patient_rows": int(len(summ)), "sum_num_windows": int(summ["num_windows"].sum()),
                         "mean_patient_entropy": float(summ["mean_entropy"].mean()), "mean_patient_accuracy": float(summ["patient_accuracy"].mean())}
            assert int(summ["num_windows"].sum()) == n and len(summ) == n_pat
    if rank == 0:
        print(json.dumps(res), flush=True)
    if world > 1:
        dist.destroy_process_group()

if __name__ == "__main__":
    main()
