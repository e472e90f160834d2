#!/usr/bin/env python
"""bench.py -- UQ windows/s of the fused MC-Dropout (T=50) hot path on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--windows N] [--passes T]
  torchrun --nproc-per-node N ... bench.py --gpus N ...        (one rank per GPU, weak scaling)

A "step" = one pass of the hot path over one batch of synthetic windows: calibration of the BatchNorm
statistics (one exact pass over all windows of the call: one tensor-core sweep per layer) + the fused tcgen05 kernel
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


def _host_threads():
    """All host cores for the CPU legs: torchrun exports OMP_NUM_THREADS=1 to its workers, which would silently
    turn the reference arm into a one-thread measurement."""
    import torch
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    os.environ.pop("OMP_NUM_THREADS", None)
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def cpu_mcd_sample(n_windows, n_passes):
    """The reference's MC-Dropout semantics on the host cores (oracle torch-f32 port): whole sample = one batch,
    batch-statistic BN, Bernoulli masks (reference uq_techniques.py:22) + the reference reductions (:40-112).
    Cost is linear in windows x passes (every layer is a dense contraction), so windows/s of a sample extrapolates
    to the full workload; at 512..1024 windows the per-call overheads are already below 2 %."""
    from oracle import network, reductions
    pkg = importlib.import_module(PKG)
    cores = _host_threads()
    ws = pkg.synthetic.make_member_weights(0)
    x = pkg.synthetic.make_windows(n_windows, seed=2025)
    t0 = time.perf_counter()
    probs = np.stack([network.forward_f32_torch(ws, x, training=True, rng_seed=t) for t in range(n_passes)])
    reductions.uq_evaluation_dist(probs, None)
    dt = time.perf_counter() - t0
    return n_windows / dt, dt, cores


def cpu_config0_sample(n_windows=2048):
    """BASELINE configs[0]: deterministic inference (`model.predict`, evaluation/evaluate_classification.py:30-31) on
    the host cores, Keras' default batch 32 and batch 1024; oracle torch-f32 port."""
    from oracle import network
    pkg = importlib.import_module(PKG)
    cores = _host_threads()
    ws = pkg.synthetic.make_member_weights(0)
    x = pkg.synthetic.make_windows(n_windows, seed=2025)
    out = {"cores": cores, "sample_windows": n_windows, "of_windows": 10_000}
    for bs in (1024, 32):
        network.forward_f32_torch(ws, x[:bs], batch_size=bs)
        t0 = time.perf_counter()
        network.forward_f32_torch(ws, x, batch_size=bs)
        out[f"batch_{bs}_windows_per_s"] = n_windows / (time.perf_counter() - t0)
    return out


def run_reference(args, rank):
    """--impl reference: the reference's own CPU implementation of the path is TensorFlow/Keras (absent, not
    installable); the timed stand-in is the oracle port on all host threads, on a bounded sample per step."""
    if rank != 0:
        return
    n_s = args.cpu_sample
    for _ in range(min(args.warmup, 1)):
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
                       "note": "reference Keras/TF unavailable on this box; oracle torch-CPU port of uq_techniques.py:22 semantics; "
                               "windows/s of the sample, cost linear in windows x passes"},
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
    ap.add_argument("--cpu-sample", type=int, default=768,
                    help="windows of the CPU sample: cpu_baseline leg and every step of the reference arm (0 = skip the leg)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default): --windows per GPU; strong: the configs[3] cohort split over the GPUs at patient boundaries")
    ap.add_argument("--patients", type=int, default=2651, help="patients of the strong-scaling cohort (SHHS2: 2651)")
    ap.add_argument("--bn-mode", default="calibrated", choices=["calibrated", "batch"],
                    help="--scaling strong only: batch = the reference's exact batch-statistic BatchNorm over the sharded cohort "
                         "(per-layer statistics all-reduced over the ranks in every sweep)")
    ap.add_argument("--skip-extras", action="store_true", help="headline + roofline + e2e only")
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
    # MC-Dropout model: the pseudo-trained member (tests/golden/trained_member0.keras, SURVEY.md 8d "pseudo-trained
    # members"): its probabilities spread over (0, 1), so the guard-band population is that of a trained network.  The
    # weight values do not change the arithmetic work.  The Deep Ensemble below uses random-init members (configs[2]).
    trained_path = os.path.join(ROOT, "tests", "golden", "trained_member0.keras")
    model = model_mod.load_model(trained_path) if os.path.isfile(trained_path) else model_mod.synthetic_model(0)
    eng = engine.get_engine(local)
    if args.variant is not None:
        eng.option("variant", args.variant)
    dev = torch.device("cuda", local)
    sh = importlib.import_module(PKG + ".sharding")
    KEYS = ("mean_pred", "pred_variance", "total_pred_entropy", "expected_aleatoric_entropy", "mutual_info",
            "predictive_entropy_bits")

    def windows_on_host(count, seed_):
        nu = min(count, 1 << 18)
        xu = syn.make_windows(nu, seed=seed_)
        return torch.from_numpy(xu).repeat((count + nu - 1) // nu, 1, 1)[:count].contiguous().pin_memory(), nu

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

    # ---- strong scaling (BASELINE configs[3]): the SHHS2-scale cohort, one patient-aligned shard per GPU; the step
    # holds the calibration pass with its statistics all-reduce, the fused kernel, the per-patient reduction of the
    # rank's own patients and BOTH gathers (per-window metrics, per-patient rows)
    pid = syn.make_patient_ids(n_patients=args.patients)
    n_cohort = len(pid)
    shards = sh.patient_aligned_shards(pid, world)
    cs, ce = shards[rank]
    strong = {}

    def setup_strong():
        xs_host, _ = windows_on_host(ce - cs, 4000 + rank)
        strong["x"] = xs_host.to(dev, non_blocking=True)
        y = syn.make_labels(n_cohort)[cs:ce]
        strong["y"] = torch.from_numpy(y.astype(np.int32)).to(dev)
        ids = pid[cs:ce]
        cuts = np.concatenate([[0], np.nonzero(ids[1:] != ids[:-1])[0] + 1, [len(ids)]]).astype(np.int64) if len(ids) else np.zeros(1, np.int64)
        strong["seg"] = torch.from_numpy(cuts).to(dev)
        strong["patients"] = len(cuts) - 1
        strong["max_shard"] = max(e - s_ for s_, e in shards)
        pcounts = [len(np.unique(pid[s_:e])) for s_, e in shards]
        strong["max_patients"] = max(pcounts)
        strong["patient_counts"] = pcounts
        uq.mc_dropout_uq(model, strong["x"][:0], n_pred=1, seed=seed, bn_mode="inference", device_out=True)   # uploads the model
        torch.cuda.synchronize()

    def step_strong(collect=None):
        xs = strong["x"]
        if world > 1:
            eng.set_stats_exchange(sh.StatsExchange(), global_windows=n_cohort)
        try:
            if args.bn_mode == "batch":     # exact Keras semantics: every sweep's statistics are all-reduced over the ranks
                out = uq.mc_dropout_uq(model, xs, n_pred=T, seed=seed, bn_mode="batch", device_out=True, window_base=cs)
            else:
                eng.calibrate(xs, n_passes=1, seed=seed, window_base=cs)
        finally:
            if world > 1:
                eng.set_stats_exchange(None)
        if args.bn_mode != "batch":
            out = uq.mc_dropout_uq(model, xs, n_pred=T, seed=seed, bn_mode="calibrated", recalibrate=False, device_out=True,
                                   window_base=cs)
        rows = eng.patient_reduce(out["pred_variance"], out["predictive_entropy_bits"], strong["y"], out["predicted_label"],
                                  strong["seg"], device_out=True)
        if world > 1:
            packed = torch.zeros((len(KEYS), strong["max_shard"]), dtype=torch.float32, device=dev)
            packed[:, : ce - cs] = torch.stack([out[k] for k in KEYS])
            if "gw" not in strong:
                strong["gw"] = torch.empty((world,) + tuple(packed.shape), dtype=packed.dtype, device=dev)
                strong["gp"] = torch.empty((world, strong["max_patients"], 8), dtype=torch.float64, device=dev)
            dist.all_gather_into_tensor(strong["gw"], packed)
            prow = torch.zeros((strong["max_patients"], 8), dtype=torch.float64, device=dev)
            prow[: rows.shape[0]] = rows
            dist.all_gather_into_tensor(strong["gp"], prow)
        if collect is not None:
            collect["rows"] = rows

    def measure_strong(k):
        if "x" not in strong:
            setup_strong()
        got = {}
        step_strong(got)
        ms, _ = timed(step_strong, k)
        ms /= k
        counts = got["rows"][:, 7].sum()
        if world > 1:
            dist.all_reduce(counts)
        assert int(counts.item()) == n_cohort, "patient-level window counts must add up to the cohort"
        return {"metric": "uq_windows_per_sec_mcd_T%d" % T, "scaling": "strong", "value": n_cohort / (ms * 1e-3), "ms_per_step": ms,
                "windows": n_cohort, "patients": int(args.patients), "shard_windows": [e - s_ for s_, e in shards],
                "shard_patients": strong["patient_counts"], "sum_patient_window_counts": int(counts.item()),
                "step": "calibration pass (statistics all-reduced over the ranks) + fused kernel on the rank's patient-aligned "
                        "shard + per-patient reduction + all_gather of per-window metrics and per-patient rows"}

    if args.scaling == "strong":
        setup_strong()
        for _ in range(max(1, args.warmup - 1)):       # measure_strong runs one more untimed step
            step_strong()
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        launches0 = eng.launch_count()
        info = measure_strong(args.steps)
        launches = (eng.launch_count() - launches0) // (args.steps + 1) * args.steps
        clocks = sampler.stop() if rank == 0 else None
        if rank == 0:
            line = {"metric": info["metric"], "value": info["value"], "unit": "windows/s", "n_gpus": world, "steps": args.steps,
                    "warmup": args.warmup, "ms_per_step": info["ms_per_step"], "higher_is_better": True, "scaling": "strong",
                    "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                    "config": {"workload": f"SHHS2-scale synthetic cohort ({args.patients} patients, {n_cohort} windows) MC Dropout T={T} with "
                                           "patient-level aggregation, split over the GPUs at patient boundaries (BASELINE configs[3])",
                               "bn_mode": "calibrated (one exact pass over the whole cohort, inside the step)" if args.bn_mode != "batch" else
                                          "batch: exact batch-statistic BatchNorm of the whole cohort (one tensor-core sweep per layer, statistics all-reduced over the ranks)",
                               "input": "float64 (N,60,4) resident per shard, larger than L2", "detail": info},
                    "clocks": clocks, "e2e": None, "gpu_launches": int(launches), "roofline": None, "cpu_baseline": None}
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    n_unique = min(n, 1 << 18)
    x_host, _ = windows_on_host(n, 2025 + rank)
    x_dev = x_host.to(dev, non_blocking=True)
    torch.cuda.synchronize()
    gather_bufs = {}

    def gather(out):   # the path's only exchange: final gather of the per-window metrics (24 B / window)
        if world > 1:
            packed = torch.stack([out[k] for k in KEYS])
            if packed.shape not in gather_bufs:
                gather_bufs[packed.shape] = torch.empty((world,) + tuple(packed.shape), dtype=packed.dtype, device=packed.device)
            dist.all_gather_into_tensor(gather_bufs[packed.shape], packed)

    def step_device():
        out = uq.mc_dropout_uq(model, x_dev, n_pred=T, seed=seed, bn_mode="calibrated", device_out=True)
        gather(out)
        return out

    def step_e2e():
        return uq.mc_dropout_uq(model, x_host, n_pred=T, seed=seed, bn_mode="calibrated", device_out=False)

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
    kf = max(1, min(args.steps, 2))
    ms_fused, _ = timed(fused_only, kf)
    ms_fused /= kf
    peaks = measured_peaks()
    ach = n * T * FLOP_VALID / (ms_fused * 1e-3) / 1e12
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.isfile(tp):
        try:
            tj = json.load(open(tp))
            traffic = tj.get("dram_bytes_per_window", 0) * n
            traffic_src = tj.get("source")
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": "fused_pair2_kernel (tcgen05 cta_group::2 conv chain, two tile-pairs per CTA pair + Philox dropout + UQ epilogue)",
                "achieved": ach, "peak": peaks["sustained"], "unit": "TFLOP/s", "frac": ach / peaks["sustained"],
                "peak_source": peaks["source"] + ", sustained bf16 (kernel timed inside a multi-second step)",
                "frac_of_burst": ach / peaks["burst"], "achieved_dense_count": n * T * FLOP_DENSE / (ms_fused * 1e-3) / 1e12,
                "kernel_ms": ms_fused, "flop_per_window_pass": FLOP_VALID, "traffic": traffic, "traffic_source": traffic_src}

    # ---- e2e: public API with HOST (pinned) buffers, H2D + D2H inside the timed region
    e2e = None
    if not args.skip_e2e:
        step_e2e()
        k2 = max(1, min(args.steps, 2))
        ms_e2e, wall_e2e = timed(step_e2e, k2)
        t_e2e = max(ms_e2e * 1e-3, wall_e2e) / k2     # D2H reads are synchronous: wall clock covers them
        e2e = {"value": world * n / t_e2e, "unit": "windows/s", "h2d_bytes_per_step": int(world * x_host.numel() * x_host.element_size()),
               "d2h_bytes_per_step": int(world * n * 28), "api": "uq_techniques.mc_dropout_uq(model, x_host_pinned_float64, n_pred=50)"}

    extras = not args.skip_extras
    # ---- secondary: Deep Ensemble M=10 (configs[2]) device-resident, every rank its own windows + the same gather
    de_info = None
    if extras:
        members = [model_mod.synthetic_model(m) for m in range(10)]
        nd = min(n, 1_000_000)

        def step_de():
            gather(uq.deep_ensemble_uq(members, x_dev[:nd], device_out=True))
        step_de()
        ms_de, _ = timed(step_de, 1)
        de_info = {"metric": "uq_windows_per_sec_de_M10", "value": world * nd / (ms_de * 1e-3), "windows_per_gpu": nd, "n_gpus": world,
                   "tflops_valid_per_gpu": nd * 10 * FLOP_VALID / (ms_de * 1e-3) / 1e12,
                   "frac_of_sustained": nd * 10 * FLOP_VALID / (ms_de * 1e-3) / 1e12 / peaks["sustained"]}
        model_mod.synthetic_model(0)   # (the engine re-uploads member 0 below through its content key)

    # ---- secondary: the same MC-Dropout workload with EXACT batch-statistic BatchNorm (Keras training=True semantics:
    # one statistics sweep per layer, per-layer bf16 hand-off through HBM, head kernel), device-resident, all windows
    exact_info = None
    if extras and world == 1:
        run_exact = lambda: uq.mc_dropout_uq(model, x_dev, n_pred=T, seed=seed, bn_mode="batch", device_out=True)
        run_exact()
        ms_ex, _ = timed(run_exact, 1)
        ach_ex = n * T * FLOP_VALID / (ms_ex * 1e-3) / 1e12
        hbm_alg = n * T * 2 * (23040 + 26880 + 11520 + 30720)      # per-layer hand-off: written once, read once (bf16, 60 rows)
        exact_info = {"metric": "uq_windows_per_sec_mcd_T%d_exact_batch_bn" % T, "value": n / (ms_ex * 1e-3), "windows": n, "ms": ms_ex,
                      "kernel_runs_per_pass_chunk": 6, "flops_vs_one_run": 1.004,
                      "roofline": {"bound": "tensor", "achieved": ach_ex, "peak": peaks["sustained"], "unit": "TFLOP/s",
                                   "frac": ach_ex / peaks["sustained"], "hbm_algorithmic_bytes": hbm_alg,
                                   "hbm_gbs_algorithmic": hbm_alg / (ms_ex * 1e-3) / 1e9, "hbm_peak_gbs": peaks["hbm_gbs"]}}

    # ---- secondary: labels made exact by the fp32 guard band (mean_p > 0.5, analyze_mcd_patient_level.py:117)
    guard_info = None
    if extras and world == 1:
        run_guard = lambda: uq.mc_dropout_uq(model, x_dev, n_pred=T, seed=seed, bn_mode="calibrated", device_out=True, exact_labels=True)
        run_guard()
        ms_g, _ = timed(run_guard, 1)
        band = eng.last_guard_band
        guard_info = {"value": n / (ms_g * 1e-3), "ms": ms_g, "guard_delta": engine.GUARD_DELTA, "windows_reevaluated_fp32": band[0],
                      "guard_band_fraction": band[0] / max(1, band[1])}

    # ---- secondary: strong scaling on the configs[3] cohort (see step_strong)
    strong_info = measure_strong(1) if extras else None
    strong.clear()

    # ---- secondary: BASELINE configs[0] on the GPU (deterministic 10k windows) next to its CPU baseline
    cfg0 = None
    if extras and rank == 0:
        x10 = x_dev[:10_000]
        cfg0 = {}
        for prec in ("fp32", "bf16"):
            eng.forward_det(x10, member=0, precision=prec, device_out=True)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); eng.forward_det(x10, member=0, precision=prec, device_out=True); e1.record()
            torch.cuda.synchronize()
            cfg0[f"gpu_{prec}_windows_per_s"] = 10_000 / (e0.elapsed_time(e1) * 1e-3)

    cpu = None
    if rank == 0 and args.cpu_sample > 0:
        v, dt, cores = cpu_mcd_sample(args.cpu_sample, T)
        cpu = {"value": v, "unit": "windows/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_sample} windows x {T} passes, whole sample = one batch (batch-stat BN), {dt:.1f} s; windows/s of the "
                         "sample (cost is linear in windows x passes); oracle torch-f32 port (reference Keras/TF unavailable on this box)"}
        if extras and cfg0 is not None:
            cfg0["cpu"] = cpu_config0_sample()

    if rank == 0:
        line = {"metric": "uq_windows_per_sec_mcd_T%d" % T, "value": value, "unit": "windows/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"MC Dropout T={T} passes with entropy/MI/variance on {n} synthetic windows per GPU (BASELINE configs[1])",
                           "windows_per_gpu": n, "passes": T, "bn_mode": "calibrated (one exact pass over all windows of the call, inside the step)",
                           "rng": "Philox4x32-10 in-kernel, bit-sliced keep-flags", "input": "float64 (N,60,4), 1.92 GB per GPU > L2, no L2 flush needed",
                           "unique_windows": n_unique, "model": "pseudo-trained member (tests/golden/trained_member0.keras)" if os.path.isfile(trained_path) else "random-init member 0",
                           "collective": "nccl all_gather of per-window metrics" if world > 1 else "none",
                           "guard_band_fraction": guard_info["guard_band_fraction"] if guard_info else None},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
                "also": de_info, "also_exact_bn": exact_info, "also_exact_labels": guard_info, "also_strong": strong_info,
                "also_config0": cfg0}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
