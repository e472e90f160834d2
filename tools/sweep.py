#!/usr/bin/env python
"""Scaling sweep of BASELINE configs[4]: T in {10,25,50,100}, M in {5,10,20} at 1/2/4/8 GPUs (weak scaling: `n` windows
per GPU, device-resident, fused path, calibration inside the MC-Dropout step, NCCL all-gather of the per-window metrics
when N > 1), next to the CPU port on the host cores (a bounded sample, rank 0, `--cpu`).

  python tools/sweep.py [windows_per_gpu] [--cpu]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/sweep.py [windows_per_gpu]

Writes gpurun_out/sweep_<N>gpu.json; copy into profiles/ to keep."""
import importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"
import numpy as np, torch
import torch.distributed as dist
pkg = importlib.import_module(PKG)
engine = importlib.import_module(PKG + ".engine")
syn = pkg.synthetic
FLOP = 98_844_864
KEYS = ("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info", "entropy_bits")


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    n = int(args[0]) if args else 400_000
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = engine.Engine(local)
    blobs = np.stack([syn.weights_to_blob(syn.make_member_weights(m)) for m in range(20)])
    xu = torch.from_numpy(syn.make_windows(1 << 16, seed=6 + rank))
    x = xu.repeat((n >> 16) + 1, 1, 1)[:n].contiguous().cuda()
    bufs = {}

    def gather(out):
        if world > 1:
            packed = torch.stack([out[k] for k in KEYS])
            if "g" not in bufs:
                bufs["g"] = torch.empty((world,) + tuple(packed.shape), dtype=packed.dtype, device=packed.device)
            dist.all_gather_into_tensor(bufs["g"], packed)

    def timed(fn):
        fn(); torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms[0])

    res = []
    eng.set_members(blobs[:1])

    def mcd(T):
        eng.calibrate(x, n_passes=1)
        gather(eng.mcd(x, n_passes=T, bn_mode="calibrated", device_out=True))
    for T in (10, 25, 50, 100):
        ms = timed(lambda: mcd(T))
        res.append({"mode": "mcd", "T": T, "n_gpus": world, "windows_per_gpu": n, "ms": ms, "windows_per_s": world * n / ms * 1e3,
                    "tflops_valid_per_gpu": n * T * FLOP / ms / 1e9})
    eng.set_members(blobs)
    for M in (5, 10, 20):
        ms = timed(lambda: gather(eng.de(x, n_members=M, device_out=True)))
        res.append({"mode": "de", "M": M, "n_gpus": world, "windows_per_gpu": n, "ms": ms, "windows_per_s": world * n / ms * 1e3,
                    "tflops_valid_per_gpu": n * M * FLOP / ms / 1e9})
    if rank == 0 and "--cpu" in sys.argv:
        # the CPU port on the host cores, bounded samples (cost is linear in windows x passes)
        from oracle import network
        os.environ.pop("OMP_NUM_THREADS", None)
        torch.set_num_threads(len(os.sched_getaffinity(0)))
        ws = [syn.make_member_weights(m) for m in range(2)]
        xs = syn.make_windows(512, seed=1)
        t0 = time.perf_counter()
        for t in range(5):
            network.forward_f32_torch(ws[0], xs, training=True, rng_seed=t)
        per_pass = (time.perf_counter() - t0) / 5 / 512                       # seconds per window-pass, batch-stat BN
        t0 = time.perf_counter()
        network.forward_f32_torch(ws[1], xs, batch_size=32)
        per_member = (time.perf_counter() - t0) / 512                         # seconds per window, predict at batch 32
        for T in (10, 25, 50, 100):
            res.append({"mode": "cpu_mcd", "T": T, "cores": torch.get_num_threads(), "windows_per_s": 1.0 / (per_pass * T), "sample": "512 windows x 5 passes"})
        for M in (5, 10, 20):
            res.append({"mode": "cpu_de", "M": M, "cores": torch.get_num_threads(), "windows_per_s": 1.0 / (per_member * M), "sample": "512 windows x 1 member"})
    if rank == 0:
        for r in res:
            print(json.dumps(r), flush=True)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"sweep_{world}gpu.json"), "w"), indent=1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
