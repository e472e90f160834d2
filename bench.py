#!/usr/bin/env python
"""bench.py -- UQ windows/s of the fused MC-Dropout (T=50) hot path on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--windows N] [--passes T]
  torchrun --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU, weak scaling)

A "step" = one pass of the hot path over one batch of synthetic windows: calibration of the BatchNorm
statistics (32768 windows x 2 passes, statistics sweeps of the tensor-core kernel) + the fused tcgen05 kernel
(T stochastic passes + per-window reductions) [+ NCCL all-gather of the per-window metrics when N > 1].
Prints ONE JSON line (rank 0).  Timing: CUDA events on the launching stream, barrier + synchronize on both
sides, max over ranks; inputs (1.9 GB float64 per GPU) are larger than L2 (126 MB).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "uncertaintyquantification-sleepapnea-1dcnn_b200"

FLOP_VALID = 98_844_864      # per window-pass, valid taps only -- roofline numerator
FLOP_DENSE = 101_806_272


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return {"burst": d["bf16_tflops"], "sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "hbm_gbs": d["hbm_gbs"], "source": "measured (MEASURED_PEAKS.json)"}
    return {"burst": 1590.0, "sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """Samples nvidia-smi SM clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [s for s, p in zip(sm, pw) if p > 300] or sm
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def cpu_mcd_sample(n_windows, n_passes, threads=None):
    """The reference's MC-Dropout semantics on the host cores (oracle torch-f32 port): whole sample = one batch,
    batch-statistic BN, Bernoulli masks (reference uq_techniques.py:22) + the reference reductions (:40-112)."""
    import torch
    from oracle import network, reductions
    pkg = importlib.import_module(PKG)
    if threads:
        torch.set_num_threads(threads)
    ws = pkg.synthetic.make_member_weights(0)
    x = pkg.synthetic.make_windows(n_windows, seed=2025)
    t0 = time.perf_counter()
    probs = np.stack([network.forward_f32_torch(ws, x, training=True, rng_seed=t) for t in range(n_passes)])
    reductions.uq_evaluation_dist(probs, None)
    dt = time.perf_counter() - t0
    return n_windows / dt, dt, torch.get_num_threads()


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the path is TensorFlow/Keras (absent, not
    installable); the timed stand-in is the oracle port on all host threads, on a bounded sample per step."""
    if rank != 0:
        return
    n_s = args.ref_windows
    for _ in range(args.warmup):
        cpu_mcd_sample(max(8, n_s // 8), 2)
    vals, times = [], []
    for _ in range(args.steps):
        v, dt, cores = cpu_mcd_sample(n_s, args.passes)
        vals.append(v); times.append(dt)
    value = float(np.mean(vals))
    line = {"impl": "reference", "metric": "uq_windows_per_sec_mcd_T%d" % args.passes, "value": value, "unit": "windows/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * float(np.mean(times)),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"MC Dropout T={args.passes} with entropy/MI/variance, {args.windows} synthetic windows per GPU (configs[1])",
                       "note": "reference Keras/TF unavailable on this box; oracle torch-CPU port of uq_techniques.py:22 semantics"},
            "cpu_baseline": {"value": value, "unit": "windows/s", "cores": cores, "kind": "port",
                             "sample": f"{n_s} windows x {args.passes} passes per step (whole sample = one batch, batch-stat BN)"},
            "e2e": {"value": value, "unit": "windows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--windows", type=int, default=1_000_000, help="windows per GPU (configs[1]: 1M)")
    ap.add_argument("--passes", type=int, default=50)
    ap.add_argument("--ref-windows", type=int, default=512, help="windows per step of the CPU reference arm")
    ap.add_argument("--cpu-sample", type=int, default=768, help="windows of the cpu_baseline sample (0 = skip)")
    ap.add_argument("--variant", type=int, default=None)
    ap.add_argument("--skip-e2e", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = importlib.import_module(PKG)
    engine = importlib.import_module(PKG + ".engine")
    model_mod = importlib.import_module(PKG + ".model")
    sys.path.insert(0, os.path.join(ROOT, PKG, "uncertainty_quantification"))
    uq = importlib.import_module("uq_techniques")
    syn = pkg.synthetic
    n, T = args.windows, args.passes
    seed = 2025

    # ---- synthetic inputs (host, pinned) and model
    n_unique = min(n, 1 << 18)
    xu = syn.make_windows(n_unique, seed=2025 + rank)
    x_host = torch.from_numpy(xu).repeat((n + n_unique - 1) // n_unique, 1, 1)[:n].contiguous().pin_memory()
    model = model_mod.synthetic_model(0)
    eng = engine.get_engine(local)
    if args.variant is not None:
        eng.option("variant", args.variant)
    x_dev = x_host.to(torch.device("cuda", local), non_blocking=True)
    torch.cuda.synchronize()
    gather_bufs = None

    def step_device():
        out = uq.mc_dropout_uq(model, x_dev, n_pred=T, seed=seed, bn_mode="calibrated", device_out=True)
        if world > 1:   # the path's only exchange: final gather of the per-window metrics (28 B / window)
            nonlocal gather_bufs
            keys = ("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info",
                    "predictive_entropy_bits")
            packed = torch.stack([out[k] for k in keys])
            if gather_bufs is None:
                gather_bufs = torch.empty((world,) + tuple(packed.shape), dtype=packed.dtype, device=packed.device)
            dist.all_gather_into_tensor(gather_bufs, packed)
        return out

    def step_e2e():
        out = uq.mc_dropout_uq(model, x_host, n_pred=T, seed=seed, bn_mode="calibrated", device_out=False)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms, wall * 1e3], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1]) / 1e3
        return ms, wall

    for _ in range(args.warmup):
        step_device()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = eng.launch_count()
    ms_total, _ = timed(step_device, args.steps)
    launches = eng.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    value = world * n / (ms_step * 1e-3)

    # ---- dominant kernel alone: fused tcgen05 kernel (statistics already calibrated), CUDA events on its stream
    def fused_only():
        eng.mcd(x_dev, n_passes=T, seed=seed, bn_mode="calibrated", device_out=True)
    ms_fused, _ = timed(fused_only, max(1, min(args.steps, 2)))
    ms_fused /= max(1, min(args.steps, 2))
    peaks = measured_peaks()
    ach = n * T * FLOP_VALID / (ms_fused * 1e-3) / 1e12
    traffic = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.isfile(tp):
        try:
            tj = json.load(open(tp))
            traffic = tj.get("dram_bytes_per_window", 0) * n
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": "fused_pair2_kernel (tcgen05 cta_group::2 conv chain, two tile-pairs per CTA pair + Philox dropout + UQ epilogue)",
                "achieved": ach, "peak": peaks["sustained"], "unit": "TFLOP/s", "frac": ach / peaks["sustained"],
                "peak_source": peaks["source"] + ", sustained bf16 (kernel timed inside a multi-second step)",
                "frac_of_burst": ach / peaks["burst"], "achieved_dense_count": n * T * FLOP_DENSE / (ms_fused * 1e-3) / 1e12,
                "kernel_ms": ms_fused, "flop_per_window_pass": FLOP_VALID, "traffic": traffic}

    # ---- e2e: public API with HOST (pinned) buffers, H2D + D2H inside the timed region
    e2e = None
    if not args.skip_e2e:
        step_e2e()
        k2 = max(1, min(args.steps, 2))
        ms_e2e, wall_e2e = timed(step_e2e, k2)
        t_e2e = max(ms_e2e * 1e-3, wall_e2e) / k2     # D2H reads are synchronous: wall clock covers them
        e2e = {"value": world * n / t_e2e, "unit": "windows/s", "h2d_bytes_per_step": int(world * x_host.numel() * x_host.element_size()),
               "d2h_bytes_per_step": int(world * n * 28), "api": "uq_techniques.mc_dropout_uq(model, x_host_pinned_float64, n_pred=50)"}

    # ---- secondary: Deep Ensemble M=10 (configs[2]) device-resident, one step
    de_info = None
    if world == 1:
        members = [model_mod.synthetic_model(m) for m in range(10)]
        uq.deep_ensemble_uq(members, x_dev[: min(n, 200_000)], device_out=True)
        nd = min(n, 1_000_000)
        ms_de, _ = timed(lambda: uq.deep_ensemble_uq(members, x_dev[:nd], device_out=True), 1)
        de_info = {"metric": "uq_windows_per_sec_de_M10", "value": nd / (ms_de * 1e-3), "windows": nd,
                   "tflops_valid": nd * 10 * FLOP_VALID / (ms_de * 1e-3) / 1e12}

    # ---- secondary: the same MC-Dropout workload with EXACT batch-statistic BatchNorm (Keras training=True semantics:
    # six statistics sweeps of the tensor-core kernel + a head kernel), device-resident, on a quarter of the windows
    exact_info = None
    if world == 1:
        ne = min(n, 250_000)
        run_exact = lambda: uq.mc_dropout_uq(model, x_dev[:ne], n_pred=T, seed=seed, bn_mode="batch", device_out=True)
        run_exact()
        ms_ex, _ = timed(run_exact, 1)
        exact_info = {"metric": "uq_windows_per_sec_mcd_T%d_exact_batch_bn" % T, "value": ne / (ms_ex * 1e-3), "windows": ne,
                      "kernel_runs_per_step": 6, "flops_vs_one_run": 1.89}

    cpu = None
    if rank == 0 and world == 1 and args.cpu_sample > 0:
        v, dt, cores = cpu_mcd_sample(args.cpu_sample, T)
        cpu = {"value": v, "unit": "windows/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_sample} windows x {T} passes, whole sample = one batch (batch-stat BN), {dt:.1f} s; "
                         "oracle torch-f32 port (reference Keras/TF unavailable on this box)"}

    if rank == 0:
        line = {"metric": "uq_windows_per_sec_mcd_T%d" % T, "value": value, "unit": "windows/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"MC Dropout T={T} passes with entropy/MI/variance on {n} synthetic windows per GPU (BASELINE configs[1])",
                           "windows_per_gpu": n, "passes": T, "bn_mode": "calibrated (32768 windows x 2 passes, inside the step)",
                           "rng": "Philox4x32-10 in-kernel", "input": "float64 (N,60,4), 1.92 GB per GPU > L2, no L2 flush needed",
                           "unique_windows": n_unique, "collective": "nccl all_gather of per-window metrics" if world > 1 else "none"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
                "also": de_info, "also_exact_bn": exact_info}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
