#!/usr/bin/env python
"""N>1 invariance on real GPUs (torchrun, one rank per GPU): the sharded MC-Dropout result gathered over NCCL must be
bit-identical to the same call on one GPU (masks keyed by global window index, statistics calibrated on the same
global windows).  Also checks exact batch-statistic BatchNorm with the per-layer statistics all-reduced over NCCL
(sharding.StatsExchange) against one GPU holding all windows.  Prints one JSON line on rank 0."""
import importlib, json, os, sys
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
    sys.path.insert(0, os.path.join(ROOT, PKG, "uncertainty_quantification"))
    uq = importlib.import_module("uq_techniques")
    syn = pkg.synthetic
    pid = syn.make_patient_ids(n_patients=40, seed=9, lo=300, hi=660)
    n = len(pid)
    x = syn.make_windows(n, seed=321)
    model = model_mod.synthetic_model(0)
    full = sh.sharded_mc_dropout_uq(uq, model, x, n_pred=12, patient_ids=pid, seed=77)
    res = {"world": world, "windows": n}
    if rank == 0:
        eng = uq._engine.get_engine()
        eng.calibrate(x, n_passes=1, seed=77)
        single = uq.mc_dropout_uq(model, x, n_pred=12, seed=77, bn_mode="calibrated", recalibrate=False)
        res["identical"] = {k: bool(np.array_equal(full[k], single[k])) for k in single if single[k] is not None and k in full}
        res["all_identical"] = all(res["identical"].values())
    # exact batch statistics: sharded + all-reduced vs one GPU (fp32 layer-synchronous path, smaller problem)
    nb = min(n, 3000)
    xb, pb = x[:nb], pid[:nb]
    fullb = sh.sharded_mc_dropout_uq(uq, model, xb, n_pred=3, patient_ids=pb, seed=78, bn_mode="batch")
    if rank == 0:
        singleb = uq.mc_dropout_uq(model, xb, n_pred=3, seed=78, bn_mode="batch")
        res["batch_windows"] = nb
        res["batch_identical"] = {k: bool(np.array_equal(fullb[k], singleb[k])) for k in singleb
                                  if singleb[k] is not None and k in fullb}
        res["batch_all_identical"] = all(res["batch_identical"].values())
        print(json.dumps(res), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()

if __name__ == "__main__":
    main()
