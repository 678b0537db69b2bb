#!/usr/bin/env python
"""bench.py -- trajectory-steps/s of the fused rollout path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c3s|c3l|c4|c5|c5g|c5m|c5gm|c4m] [--precision fp32|tf32x3|tf32|bf16] [--impl reference]

A "step" is one whole-horizon rollout of the workload's batch of control sequences (one kernel
launch).  Default workload = BASELINE.json configs[1] (C2): Basic_64_4_32, 4096 control samples x
horizon 50, FP32.  Under torchrun every rank rolls out its own batch of the same size (weak
scaling, no collective inside the horizon) and the step ends with ONE NCCL all-gather of the
per-trajectory costs.  Rank 0 prints one JSON line.

  value     trajectory-steps/s, inputs resident in HBM, per-step CUDA-event durations (L2 flushed
            between steps, outside the event pairs), max over ranks
  e2e       same metric through the public API with HOST (pinned) buffers: H2D of (init, U),
            the rollout, D2H of the predicted states [N,H,6], every step
  roofline  algorithmic FLOPs of the dominant (only) kernel / its mean launch duration against the
            FP32-FMA peak measured live by the FFMA micro-kernel (the path is compute-bound:
            32 B vs 1e3..1e7 FLOP per trajectory-step); the HBM figure is carried
            beside it against MEASURED_PEAKS.json
  cpu_baseline   oracle port (reference nn.Modules in the python loop, oracle/rollout_ref.py) on
            the host cores, rank 0, N=1 only
--impl reference times that same oracle port as the reference arm (the reference is pure Python
and /root/reference does not exist on the GPU box).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))

import numpy as np   # noqa: E402
import torch         # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
METRIC, UNIT = "trajectory-steps/sec", "traj-steps/s"

# FLOP per trajectory-step = 2*MAC of the weight matrices; bytes = 8 (u_t) + 24 (s_t+1)
WORKLOADS = {
    "c2": dict(desc="C2 MLP Basic_64_4_32 rollout, 4096 control samples x horizon 50, FP32", ckpt="Basic_64_4_32",
               N=4096, H=50, noise=False, states=True, flop=34560, cpu_n=4096),
    "c3": dict(desc="C3 GRU_64_4_4_16_False (window 4, stateless) + residual-noise net, 65536 x horizon 100, FP32",
               ckpt="GRU_64_4_4_16_False", N=65536, H=100, noise=True, states=True, flop=701184 + 4352, cpu_n=2048),
    "c3s": dict(desc="C3 stateful GRU_64_4_4_64_True + residual-noise net, 65536 x horizon 100, FP32",
                ckpt="GRU_64_4_4_64_True", N=65536, H=100, noise=True, states=True, flop=701184 + 4352, cpu_n=2048),
    "c3l": dict(desc="LSTM_64_4_4_16_False (window 4, stateless), 65536 x horizon 100, FP32",
                ckpt="LSTM_64_4_4_16_False", N=65536, H=100, noise=False, states=True, flop=934656, cpu_n=2048),
    "c4": dict(desc="C4 LSTM hidden 256 x 4 layers (seeded random init), window 4, 65536 x horizon 100, FP32",
               ckpt="LSTM_256_4_4_0_False", N=65536, H=100, noise=False, states=True, flop=14748672, cpu_n=256),
    "c5": dict(desc="C5 MLP Basic_64_4_32, 1048576 x horizon 50, cost-only", ckpt="Basic_64_4_32",
               N=1048576, H=50, noise=False, states=False, flop=34560, cpu_n=65536),
    "c5g": dict(desc="C5 GRU_64_4_4_16_False, 1048576 x horizon 100, cost-only", ckpt="GRU_64_4_4_16_False",
                N=1048576, H=100, noise=False, states=False, flop=701184, cpu_n=2048),
    # MPPI planner iteration: controls sampled in-kernel around a nominal sequence, cost-only rollout,
    # softmin update; host traffic per iteration = one state + the nominal sequence in, the new sequence out
    "c5m": dict(desc="MPPI iteration, MLP Basic_64_4_32, 1048576 samples x horizon 50 (in-kernel control sampling + softmin update)",
                ckpt="Basic_64_4_32", N=1048576, H=50, noise=False, states=False, flop=34560, cpu_n=65536, mppi=True),
    "c5gm": dict(desc="MPPI iteration, GRU_64_4_4_16_False, 1048576 samples x horizon 100 (in-kernel control sampling + softmin update)",
                 ckpt="GRU_64_4_4_16_False", N=1048576, H=100, noise=False, states=False, flop=701184, cpu_n=2048, mppi=True),
    "c4m": dict(desc="MPPI iteration, LSTM hidden 256 x 4 layers (seeded random init), window 4, 65536 samples x horizon 50 "
                     "(in-kernel control sampling + softmin update)",
                ckpt="LSTM_256_4_4_0_False", N=65536, H=50, noise=False, states=False, flop=14748672, cpu_n=256, mppi=True),
}


def load_ckpt(name):
    path = os.path.join(GOLD, "checkpoints", f"{name}.npz")
    if os.path.exists(path):
        return {k: torch.from_numpy(v) for k, v in np.load(path).items()}
    kind, hidden, layers = name.split("_")[0], int(name.split("_")[1]), int(name.split("_")[2])
    torch.manual_seed(0)                                      # seeded random init
    if kind == "LSTM":
        ref = torch.nn.ModuleDict(dict(lstm=torch.nn.LSTM(8, hidden, layers, batch_first=True), full_c=torch.nn.Linear(hidden, 6)))
    elif kind == "GRU":
        ref = torch.nn.ModuleDict(dict(gru=torch.nn.GRU(8, hidden, layers, batch_first=True), full_c=torch.nn.Linear(hidden, 6)))
    else:
        raise ValueError(name)
    return {k: v.detach().clone() for k, v in ref.state_dict().items()}


def load_stats(window):
    z = np.load(os.path.join(GOLD, "stats.npz"))
    pre = "mlp_" if window == 0 else "rnn4_"
    return {k[len(pre):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(pre)}


def name_meta(name):
    p = name.split("_")
    return dict(kind=p[0], hidden=int(p[1]), layers=int(p[2]), window=0 if p[0] == "Basic" else int(p[3]),
                state_f=p[-1] == "True")


def make_problem(w, seed):
    """CPU tensors of the workload (seeded, BARC rosbag shape; rows tiled beyond 65536 trajectories)."""
    from mpcdl import synth
    meta = name_meta(w["ckpt"])
    rows = synth.load_rows()
    base = min(w["N"], 65536)
    init, U = synth.make_inputs(rows, base, w["H"], window=meta["window"], seed=seed)
    if w["N"] > base:
        r = w["N"] // base
        init = init.repeat(r, *([1] * (init.dim() - 1)))
        U = U.repeat(r, 1, 1)
    cost = synth.make_cost(rows, w["H"])
    return meta, init, U, cost


def noise_state():
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(GOLD, "checkpoints", "noise_seed0_h3_s32_l2.npz")).items()}


# --------------------------------------------------------------------------------------- CPU side
def cpu_rollout_fn(w, n_cpu, threads):
    """The oracle port on `n_cpu` trajectories of the workload -> callable doing one full-horizon rollout."""
    from oracle import rollout_ref as R
    torch.set_num_threads(threads)
    meta, init, U, cost = make_problem(dict(w, N=n_cpu), seed=0)
    ref, _ = R.build_from_name(w["ckpt"], load_ckpt(w["ckpt"]))
    st = load_stats(meta["window"])
    noise, eps = None, None
    if w["noise"]:
        from mpcdl import synth
        noise = R.RefNoise(6, 2, 3, 2, 32, "tanh", "identity")
        noise.model.load_state_dict(noise_state())
        noise.eval()
        eps = synth.make_eps(n_cpu, w["H"], seed=1)
    if meta["window"] == 0:
        return lambda: R.rollout_mlp(ref, st, init, U, noise=noise, eps=eps, cost=cost)
    return lambda: R.rollout_rnn(ref, st, init, U, stateful=meta["state_f"], noise=noise, eps=eps, cost=cost)


def time_cpu(w, steps, warmup, budget_s=25.0):
    threads = os.cpu_count() or 1
    n_cpu = min(w["cpu_n"], w["N"])
    fn = cpu_rollout_fn(w, n_cpu, threads)
    for _ in range(max(warmup, 1)):
        fn()
    ts, t_start = [], time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
        if time.perf_counter() - t_start > budget_s and len(ts) >= 3:
            break
    med = statistics.median(ts)
    sample = (f"{n_cpu} of {w['N']} trajectories x full horizon {w['H']}, oracle port (reference nn.Modules in the python "
              f"closed loop), torch {torch.__version__} CPU FP32, {len(ts)} rollouts, median")
    return n_cpu * w["H"] / med, med, threads, sample, len(ts)


def run_reference(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, med, threads, sample, done = time_cpu(w, args.steps, args.warmup, budget_s=120.0)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
            "warmup": args.warmup, "ms_per_step": med * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": w["desc"], "N_per_step": min(w["cpu_n"], w["N"]), "H": w["H"]},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """SM clock + throttle reasons of one GPU, sampled through NVML while the timed region runs."""
    BAD = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}
    NOTE = {"sw_power_cap": 0x4, "hw_power_brake": 0x80, "sync_boost": 0x10}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.stop_flag, self.ok = [], set(), threading.Event(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            phys = index
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            if vis:
                ent = vis.split(",")[index].strip()
                phys = int(ent) if ent.isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        while not self.stop_flag.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in {**self.BAD, **self.NOTE}.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.01)

    def summary(self):
        self.stop_flag.set()
        if self.is_alive():
            self.join(timeout=1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"], "samples": 0}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# --------------------------------------------------------------------------------------- GPU side
def run_native(args, w):
    import torch.distributed as dist

    import mpcdl
    from mpcdl import ops, parallel, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the fused rollout path has no CPU fallback "
                           "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N, H = w["N"], w["H"]

    meta, init_h, U_h, cost = make_problem(dict(w, N=1), seed=0) if w.get("mppi") else make_problem(w, seed=rank)
    model, _ = mpcdl.from_checkpoint_name(w["ckpt"], load_ckpt(w["ckpt"]), dev)
    stats = load_stats(meta["window"])
    noise = None
    if w["noise"]:
        noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
        noise.model.load_state_dict(noise_state())
        noise.to(dev)
    init_p, U_p = init_h.pin_memory(), U_h.pin_memory()
    init_d, U_d = init_p.to(dev), U_p.to(dev)
    cost_d = tuple(c.to(dev) for c in cost)
    prec = args.precision
    kw = dict(noise=noise, seed=1, traj_offset=rank * N, cost=cost_d, return_states=w["states"], precision=prec)
    costs_all = torch.empty(world * N, device=dev) if world > 1 else None
    mppi = bool(w.get("mppi"))
    if mppi:
        from mpcdl.mppi import MPPI
        SIG, LO, HI = (0.06, 0.03), (-0.875, -0.285), (0.237, 0.279)      # 0.25 sigma_u and the recorded control range
        planner = MPPI(model, stats, world * N, SIG, LO, HI, cost_d, lam=5.0, precision=prec, seed=1)
        init_h, U_h = init_h[0].contiguous(), U_h[0].contiguous()         # ONE start state and the nominal sequence
        init_p, U_p = init_h.pin_memory(), U_h.pin_memory()
        init_d, U_d = init_p.to(dev), U_p.to(dev)
        kw.update(sample=dict(n=N, sigma=SIG, lo=LO, hi=HI, seed=1), shared_init=True)

    def step(init, U):
        if mppi:                                       # rollout + (all-gather of costs) + softmin update
            u_new, info = planner.step(init, U)
            return None, u_new
        out = model.rollout(init, U, stats, **kw)
        if world > 1:                                  # the path's only collective: all-gather of costs
            dist.all_gather_into_tensor(costs_all, out[1])
        return out

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    for _ in range(max(args.warmup, 3)):
        step(init_d, U_d)
    torch.cuda.synchronize(dev)
    fp32_peak = ops.ffma_peak_tflops(dev) if rank == 0 else None

    # ---- timed region 1: inputs resident in HBM, per-step device time
    sampler = ClockSampler(local)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t_wall0 = time.perf_counter()
    for e0, e1 in ev:
        flush.zero_()
        e0.record()
        step(init_d, U_d)
        e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - t_wall0
    dev_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev)

    # ---- kernel-only duration (no collective) for the roofline, same flush discipline
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for e0, e1 in kev:
        flush.zero_()
        e0.record()
        model.rollout(init_d, U_d, stats, **kw)
        e1.record()
    torch.cuda.synchronize(dev)
    k_ms = sum(e0.elapsed_time(e1) for e0, e1 in kev) / args.steps

    # ---- timed region 2: end to end with host buffers (H2D inputs, D2H result every step)
    res_shape = (N, H, 6) if w["states"] else ((H, 2) if mppi else (N,))
    res_p = torch.empty(res_shape, dtype=torch.float32).pin_memory()
    e2e_steps = max(3, min(args.steps, 200))

    def e2e_step():
        a = init_p.to(dev, non_blocking=True)
        b = U_p.to(dev, non_blocking=True)
        out = step(a, b)
        res_p.copy_(out[0] if w["states"] else out[1], non_blocking=True)

    def timed(fn, after=None):
        for _ in range(3):
            fn()
        if after:
            after()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(e2e_steps):
            fn()
        if after:
            after()                     # the current stream waits for the last copy-out before the closing event
        t1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        return t0.elapsed_time(t1) / e2e_steps

    # (a) one stream: H2D, kernel, D2H of a batch strictly one after the other
    e2e_serial_ms = timed(e2e_step)
    # (b) the public host-buffer driver (mpcdl.parallel.HostPipeline): the same bytes per batch, with the copy-in of
    #     batch i+1 and the copy-out of batch i-1 overlapping the kernel of batch i (two device input buffers)
    pipe = parallel.HostPipeline(lambda a, b: (lambda o: o[0] if w["states"] else o[1])(step(a, b)), depth=2, device=dev)
    e2e_ms = timed(lambda: pipe.submit((init_p, U_p), res_p), after=pipe.drain)
    clocks = sampler.summary()

    t = torch.tensor([dev_ms, e2e_ms, k_ms, e2e_serial_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, k_ms, e2e_serial_ms = (float(v) for v in t.tolist())
    ms_per_step = dev_ms / args.steps

    if rank == 0:
        units = world * N * H
        value = units / (ms_per_step * 1e-3)
        kind = {"Basic": 0, "GRU": 1, "LSTM": 2}[meta["kind"]]
        plan = ops.plan(kind, meta["hidden"], meta["layers"], meta["window"], N, H, stateful=meta["state_f"],
                        noise_cfg=noise.cfg() if noise is not None else None, precision=prec)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        alg_flop = N * H * w["flop"]
        alg_bytes = N * H * ((24 if w["states"] else 0) + (0 if mppi else 8)) + init_h.numel() * 4 + N * 4
        ach_tf = alg_flop / (k_ms * 1e-3) / 1e12
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload if prec == "fp32" else f"{args.workload}_{prec}")
        except Exception:
            pass
        if prec == "fp32":
            bound, peak = "fp32", fp32_peak
            peak_source = ("FP32 FFMA peak measured live by mpcdl_ffma_peak (MEASURED_PEAKS.json has no FP32 figure; "
                           "nominal 148 SM x 128 lanes x 2 x 1.965 GHz = 74.4)")
        elif prec == "bf16":
            bound, peak = "tensor", peaks.get("bf16_tflops", 1590.0)
            peak_source = "dense BF16 peak: MEASURED_PEAKS.json bf16_tflops (burst)" if peaks else "fallback 1590"
        else:
            # TF32 tensor peak = half the measured dense BF16 rate (nominal 1.1 vs 2.25 PFLOP/s); tf32x3 issues 3 MMA
            # passes per algorithmic FLOP, so its ceiling against this peak is 1/3
            bound, peak = "tensor", peaks.get("bf16_tflops", 1590.0) / 2
            peak_source = ("TF32 dense peak taken as MEASURED_PEAKS.json bf16_tflops / 2" if peaks else "fallback 1590 / 2") + \
                          ("; tf32x3 executes 3 tensor passes per algorithmic FLOP" if prec == "tf32x3" else "")
        roofline = {"bound": bound, "achieved": ach_tf, "peak": peak, "unit": "TFLOP/s", "frac": ach_tf / peak,
                    "traffic": traffic, "peak_source": peak_source, "fp32_ffma_peak": fp32_peak,
                    "kernel_ms": k_ms, "flop_per_traj_step": w["flop"],
                    "hbm": {"achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                            "frac": alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak,
                            "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": {"fp32": "f32", "tf32x3": "tf32x3 (3-pass TF32 operands, FP32 accumulate/state)", "tf32": "tf32 (FP32 accumulate/state)",
                          "bf16": "bf16 (BF16 gate-GEMM operands, FP32 accumulate / cell state / head / feedback)"}[prec],
                "data": "synthetic",
                "config": {"workload": w["desc"], "N_per_gpu": N, "H": H, "l2": "flushed between steps (256 MiB memset outside the event pairs)",
                           "parallelism": f"trajectory shards x{world}, one all-gather of costs per step" if world > 1 else "single GPU",
                           "plan": plan},
                "roofline": roofline,
                "e2e": {"value": units / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": (init_h.numel() + U_h.numel()) * 4,
                        "d2h_bytes_per_step": res_p.numel() * 4, "ms_per_step": e2e_ms, "steps": e2e_steps,
                        "how": "mpcdl.parallel.HostPipeline(depth=2): pinned host buffers, every batch copied in and its result copied out, "
                               "copies of neighbouring batches overlap the kernel",
                        "serial_value": units / (e2e_serial_ms * 1e-3), "serial_ms_per_step": e2e_serial_ms},
                "gpu_launches": args.steps * (6 if mppi else 1), "gpu_launches_total": max(args.warmup, 3) + 2 * args.steps + 2 * (e2e_steps + 3) + 6, "clocks": clocks, "wall_s_region1": wall}
        if world == 1 and not args.skip_cpu:
            val, med, threads, sample, _ = time_cpu(w, 9, 1)
            line["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                                    "ms_per_rollout": med * 1e3}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--skip-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32x3", "tf32", "bf16"],
                    help="gate-GEMM arithmetic: fp32 = FFMA kernels (parity path), tf32x3 / tf32 = tcgen05 kernels (hidden 64), "
                         "bf16 = tcgen05 kernel of the hidden-256 LSTM (workload c4)")
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, w)
    else:
        run_native(args, w)


if __name__ == "__main__":
    main()
