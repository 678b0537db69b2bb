#!/usr/bin/env python
"""bench.py -- trajectory-steps/s of the fused rollout path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c2|c3|c3s|c3l|c4|c5g|c5m|c5gm|c4m]
                    [--precision fp32|tf32x3|tf32|bf16] [--impl reference] [--no-table]

A "step" is one whole-horizon rollout of the workload's batch of control sequences (one kernel launch).

Default workload = BASELINE.json configs[4] (C5): Basic_64_4_32, 1 048 576 trajectories x horizon 50, cost-only,
FP32, SHARDED BY TRAJECTORY over the ranks (strong scaling: the job is the same 1 M trajectories on 1, 2, 4 or 8
GPUs; mpcdl.parallel.shard_range), no collective inside the horizon; the step ends with ONE NCCL all-gather of the
per-trajectory costs (4 MB in total).  It fits one GPU, so N = 1 runs the whole job.  The 1 M-trajectory workloads
(c5, c5g and the MPPI iterations) are strong-scaled; c2 / c3 / c4 are per-GPU batches (weak scaling).  Rank 0
prints one JSON line.

  value     trajectory-steps/s of the whole job, inputs resident in HBM, per-step CUDA-event durations (L2 flushed
            between steps, outside the event pairs), max over ranks
  e2e       same metric through the public API with HOST (pinned) buffers: every step copies its (init, U) shard in
            and the result (costs, or states [N,H,6]) out; mpcdl.parallel.HostPipeline overlaps the copies of
            neighbouring batches with the kernel.  Timed over >= 0.5 s per repeat, median of 5 repeats
  roofline  FLOPs the dominant (only) kernel EXECUTES / its mean launch duration against the FP32-FMA peak measured
            live by the FFMA micro-kernel, or the tensor peak of MEASURED_PEAKS.json (the path is compute-bound:
            8..32 B vs 1e3..1e7 FLOP per trajectory-step).  `frac_survey_flop` uses SURVEY 8d's
            figure, which also counts the recurrent GEMM of window frame 0 that a stateless model never runs (h0 = 0)
  workloads (N = 1, default workload only) the same device-timed measurement for every other BASELINE configuration:
            c2, c3 (fp32 / tf32x3 / tf32), c3s, c4 (bf16, fp32), c5 on the tcgen05 kernel (tf32x3 / tf32), c5g -- about a minute
  training  (same run) one device-timed epoch of the fused training kernels for 148 models side by side: MLP / GRU / LSTM 64 x 4,
            batches of 16 -> Adam model-steps/s (a side table, not the headline metric)
  cpu_baseline   oracle port (reference nn.Modules in the python loop, oracle/rollout_ref.py) on the host cores,
            rank 0, N = 1 only, bounded sample
--impl reference times that same oracle port as the reference arm (the reference is pure Python and /root/reference
does not exist on the GPU box); it does not import the mpcdl package (no CUDA library is mapped).
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import math
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "mpc-deep-learning-sp24_b200")
sys.path.insert(0, ROOT)

import numpy as np   # noqa: E402
import torch         # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
METRIC, UNIT = "trajectory-steps/sec", "traj-steps/s"
DEFAULT = "c5"


def _rnn_flop(h, L, T, G, stateless):
    """(SURVEY 8d figure, executed): 2*MAC of the weight matrices per trajectory-step; a stateless model skips the
    recurrent GEMM of window frame 0 in every layer (h0 = 0): L * 2 * h * G*h fewer FLOP."""
    full = 2 * (T * (G * h * (8 + h) + (L - 1) * G * h * 2 * h) + 6 * h)
    return full, full - (L * 2 * h * G * h if stateless else 0)


_GRU, _GRU_X = _rnn_flop(64, 4, 4, 3, True)          # 701184, 602880
_LSTM, _LSTM_X = _rnn_flop(64, 4, 4, 4, True)        # 934656, 803584
_L256, _L256_X = _rnn_flop(256, 4, 4, 4, True)       # 14748672, 12651520
_MLP = 2 * (8 * 64 + 4 * 64 * 64 + 6 * 64)           # 34560
_NOISE = 2 * (24 * 32 + 32 * 32 + 32 * 12)           # 4352
assert (_GRU, _LSTM, _L256, _MLP) == (701184, 934656, 14748672, 34560)

# flop = SURVEY.md 8d figure, flop_exec = what the kernels execute; bytes = 8 (u_t) + 24 (s_t+1) per trajectory-step
WORKLOADS = {
    "c2": dict(desc="C2 MLP Basic_64_4_32 rollout, 4096 control samples x horizon 50, FP32", ckpt="Basic_64_4_32",
               N=4096, H=50, noise=False, states=True, flop=_MLP, flop_exec=_MLP, cpu_n=4096, scaling="weak"),
    "c3": dict(desc="C3 GRU_64_4_4_16_False (window 4, stateless) + residual-noise net, 65536 x horizon 100",
               ckpt="GRU_64_4_4_16_False", N=65536, H=100, noise=True, states=True, flop=_GRU + _NOISE,
               flop_exec=_GRU_X + _NOISE, cpu_n=2048, scaling="weak"),
    "c3s": dict(desc="C3 stateful GRU_64_4_4_64_True + residual-noise net, 65536 x horizon 100",
                ckpt="GRU_64_4_4_64_True", N=65536, H=100, noise=True, states=True, flop=_GRU + _NOISE,
                flop_exec=_GRU + _NOISE, cpu_n=2048, scaling="weak"),
    "c3l": dict(desc="LSTM_64_4_4_16_False (window 4, stateless), 65536 x horizon 100",
                ckpt="LSTM_64_4_4_16_False", N=65536, H=100, noise=False, states=True, flop=_LSTM, flop_exec=_LSTM_X,
                cpu_n=2048, scaling="weak"),
    "c4": dict(desc="C4 LSTM hidden 256 x 4 layers (seeded random init), window 4, 65536 x horizon 100",
               ckpt="LSTM_256_4_4_0_False", N=65536, H=100, noise=False, states=True, flop=_L256, flop_exec=_L256_X,
               cpu_n=256, scaling="weak"),
    "c5": dict(desc="C5 MLP Basic_64_4_32, 1048576 trajectories x horizon 50, cost-only, sharded by trajectory + all-gather of costs",
               ckpt="Basic_64_4_32", N=1048576, H=50, noise=False, states=False, flop=_MLP, flop_exec=_MLP, cpu_n=65536,
               scaling="strong"),
    "c5g": dict(desc="C5 GRU_64_4_4_16_False, 1048576 trajectories x horizon 100, cost-only, sharded by trajectory + all-gather of costs",
                ckpt="GRU_64_4_4_16_False", N=1048576, H=100, noise=False, states=False, flop=_GRU, flop_exec=_GRU_X,
                cpu_n=2048, scaling="strong"),
    # MPPI planner iteration: controls sampled in-kernel around a nominal sequence, cost-only rollout,
    # softmin update; host traffic per iteration = one state + the nominal sequence in, the new sequence out
    "c5m": dict(desc="MPPI iteration, MLP Basic_64_4_32, 1048576 samples x horizon 50 (in-kernel control sampling + softmin update)",
                ckpt="Basic_64_4_32", N=1048576, H=50, noise=False, states=False, flop=_MLP, flop_exec=_MLP, cpu_n=65536,
                mppi=True, scaling="strong"),
    "c5gm": dict(desc="MPPI iteration, GRU_64_4_4_16_False, 1048576 samples x horizon 100 (in-kernel control sampling + softmin update)",
                 ckpt="GRU_64_4_4_16_False", N=1048576, H=100, noise=False, states=False, flop=_GRU, flop_exec=_GRU_X,
                 cpu_n=2048, mppi=True, scaling="strong"),
    "c4m": dict(desc="MPPI iteration, LSTM hidden 256 x 4 layers (seeded random init), window 4, 65536 samples x horizon 50 "
                     "(in-kernel control sampling + softmin update)",
                ckpt="LSTM_256_4_4_0_False", N=65536, H=50, noise=False, states=False, flop=_L256, flop_exec=_L256_X,
                cpu_n=256, mppi=True, scaling="strong"),
}
# the other BASELINE configurations, measured beside the default one at N = 1: (workload, precision, max timed steps)
TABLE = [("c2", "fp32", 200), ("c3", "fp32", 5), ("c3", "tf32x3", 10), ("c3", "tf32", 10), ("c3s", "fp32", 5),
         ("c4", "bf16", 8), ("c4", "fp32", 2), ("c5", "tf32x3", 10), ("c5", "tf32", 10), ("c5g", "fp32", 2), ("c5g", "tf32x3", 3)]
DTYPE = {"fp32": "f32", "tf32x3": "tf32x3 (3-pass TF32 operands, FP32 accumulate/state)", "tf32": "tf32 (FP32 accumulate/state)",
         "bf16": "bf16 (BF16 gate-GEMM operands, FP32 accumulate / cell state / head / feedback)"}


def synth_module():
    """mpcdl/synth.py loaded by path: the input generators are shared by both arms, but the reference arm must not
    import the mpcdl package (whose __init__ maps libmpcdl.so)."""
    spec = importlib.util.spec_from_file_location("mpcdl_synth_standalone", os.path.join(PKG, "mpcdl", "synth.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


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


BASE_N = 65536      # seeded rows are generated for at most this many trajectories and tiled beyond it


def make_problem(w, seed, n=None):
    """CPU tensors of the workload's first min(n, 65536) trajectories (seeded, BARC rosbag shape) + the cost."""
    synth = synth_module()
    meta = name_meta(w["ckpt"])
    rows = synth.load_rows()
    n = w["N"] if n is None else n
    init, U = synth.make_inputs(rows, min(n, BASE_N), w["H"], window=meta["window"], seed=seed)
    cost = synth.make_cost(rows, w["H"])
    return meta, init, U, cost


def noise_state():
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(GOLD, "checkpoints", "noise_seed0_h3_s32_l2.npz")).items()}


# --------------------------------------------------------------------------------------- CPU side
def cpu_rollout_fn(w, n_cpu, threads):
    """The oracle port on `n_cpu` trajectories of the workload -> callable doing one full-horizon rollout."""
    from oracle import rollout_ref as R
    torch.set_num_threads(threads)
    meta, init, U, cost = make_problem(w, seed=0, n=n_cpu)
    ref, _ = R.build_from_name(w["ckpt"], load_ckpt(w["ckpt"]))
    st = load_stats(meta["window"])
    noise, eps = None, None
    if w["noise"]:
        noise = R.RefNoise(6, 2, 3, 2, 32, "tanh", "identity")
        noise.model.load_state_dict(noise_state())
        noise.eval()
        eps = synth_module().make_eps(n_cpu, w["H"], seed=1)
    if meta["window"] == 0:
        return lambda: R.rollout_mlp(ref, st, init, U, noise=noise, eps=eps, cost=cost)
    return lambda: R.rollout_rnn(ref, st, init, U, stateful=meta["state_f"], noise=noise, eps=eps, cost=cost)


def time_cpu(w, steps, warmup, budget_s=25.0):
    threads = os.cpu_count() or 1
    n_cpu = min(w["cpu_n"], w["N"], BASE_N)
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
            "warmup": args.warmup, "ms_per_step": med * 1e3, "higher_is_better": True, "scaling": w["scaling"],
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": w["desc"], "N": w["N"], "H": w["H"], "sample_per_step": min(w["cpu_n"], w["N"], BASE_N)},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "native_modules_loaded": sorted(m for m in sys.modules if m == "mpcdl" or m.startswith("mpcdl."))}
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
class Problem:
    """One workload resident on this rank's GPU: model, this rank's shard of the inputs, the step closure."""
    SIG, LO, HI = (0.06, 0.03), (-0.875, -0.285), (0.237, 0.279)      # 0.25 sigma_u and the recorded control range

    def __init__(self, key, prec, dev, rank, world):
        import torch.distributed as dist

        import mpcdl
        from mpcdl import parallel
        self.w = w = WORKLOADS[key]
        self.key, self.prec, self.dev, self.rank, self.world = key, prec, dev, rank, world
        self.mppi = bool(w.get("mppi"))
        self.strong = w["scaling"] == "strong"
        self.N_total = w["N"] if self.strong else w["N"] * world
        self.lo, self.hi = parallel.shard_range(self.N_total, rank, world) if self.strong else (rank * w["N"], (rank + 1) * w["N"])
        self.N, self.H = self.hi - self.lo, w["H"]
        self.meta, init_b, U_b, cost = make_problem(w, seed=0, n=1 if self.mppi else None)
        self.model, _ = mpcdl.from_checkpoint_name(w["ckpt"], load_ckpt(w["ckpt"]), dev)
        self.stats = load_stats(self.meta["window"])
        self.noise = None
        if w["noise"]:
            self.noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
            self.noise.model.load_state_dict(noise_state())
            self.noise.to(dev)
        self.cost_d = tuple(c.to(dev) for c in cost)
        self.kw = dict(noise=self.noise, seed=1, traj_offset=self.lo, cost=self.cost_d, return_states=w["states"], precision=prec)
        if self.mppi:
            from mpcdl.mppi import MPPI
            self.planner = MPPI(self.model, self.stats, self.N_total, self.SIG, self.LO, self.HI, self.cost_d, lam=5.0,
                                precision=prec, seed=1)
            self.init_d, self.U_d = init_b[0].contiguous().to(dev), U_b[0].contiguous().to(dev)   # ONE start state + the nominal sequence
            self.kw.update(sample=dict(n=self.N, sigma=self.SIG, lo=self.LO, hi=self.HI, seed=1), shared_init=True)
        else:
            # this rank's trajectories [lo, hi) of the job: the seeded base block tiled along the trajectory axis
            idx = (torch.arange(self.lo, self.hi, device=dev) % init_b.shape[0])
            self.init_d = init_b.to(dev)[idx].contiguous()
            self.U_d = U_b.to(dev)[idx].contiguous()
            del idx
        if self.meta["state_f"]:
            # the reference's h_state, kept across steps (every step continues the recurrent state of the previous one;
            # the work per step is the same and no per-step memset runs inside the timed region)
            planes = self.meta["layers"] * (2 if self.meta["kind"] == "LSTM" else 1)
            self.kw["carry"] = torch.zeros(planes, self.N, self.meta["hidden"], device=dev)
        self.res_d = torch.empty((self.N, self.H, 6) if w["states"] else (self.N,), device=dev)
        self.costs_all = torch.empty(self.N_total, device=dev) if (world > 1 and not self.mppi) else None
        self.dist = dist
        self.launches_per_step = 6 if self.mppi else 1

    def rollout(self, init, U, out=None):
        """the kernel alone (no collective): roofline timing"""
        o = (out, None) if self.w["states"] else (None, out)
        return self.model.rollout(init, U, self.stats, out=o if out is not None else None, **self.kw)

    def step(self, init, U, out=None):
        if self.mppi:                                  # rollout + (all-gather of costs) + softmin update
            u_new, _ = self.planner.step(init, U)
            if out is not None and out.shape == u_new.shape:
                out.copy_(u_new)
                return out
            return u_new
        res = self.rollout(init, U, out)
        if self.world > 1:                             # the path's only collective: all-gather of costs
            self.dist.all_gather_into_tensor(self.costs_all, res[1])
        return res[0] if self.w["states"] else res[1]

    def host_buffers(self):
        init_p, U_p = self.init_d.cpu().pin_memory(), self.U_d.cpu().pin_memory()
        res_shape = (self.H, 2) if self.mppi else tuple(self.res_d.shape)
        return init_p, U_p, torch.empty(res_shape, dtype=torch.float32).pin_memory()

    def alg_bytes(self):
        w = self.w
        return self.N * self.H * ((24 if w["states"] else 0) + (0 if self.mppi else 8)) + self.init_d.numel() * 4 + self.N * 4


def time_device(fn, steps, flush, dev, world=1, dist=None):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for e0, e1 in ev:
        flush.zero_()
        e0.record()
        fn()
        e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - t0
    return sum(e0.elapsed_time(e1) for e0, e1 in ev), wall


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def load_traffic(key, prec):
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(key if prec == "fp32" else f"{key}_{prec}")
    except Exception:
        return None


def roofline_of(pb, k_ms, fp32_peak, peaks):
    w, prec = pb.w, pb.prec
    units = pb.N * pb.H
    ach = units * w["flop_exec"] / (k_ms * 1e-3) / 1e12
    if prec == "fp32":
        bound, peak = "fp32", fp32_peak
        src = ("FP32 FMA peak measured live: the better of the scalar-FFMA and the packed-FFMA2 register micro-kernels "
               "(mpcdl_ffma_peak / mpcdl_ffma2_peak; MEASURED_PEAKS.json has no FP32 figure; nominal 148 SM x 128 lanes x 2 x "
               "1.965 GHz = 74.4)")
    elif prec == "bf16":
        bound, peak = "tensor", peaks.get("bf16_tflops", 1590.0)
        src = "dense BF16 peak: MEASURED_PEAKS.json bf16_tflops (burst)" if peaks else "fallback 1590 (B200_PROFILING.md)"
    else:
        # TF32 tensor peak = half the measured dense BF16 rate (nominal 1.1 vs 2.25 PFLOP/s); tf32x3 issues 3 MMA
        # passes per algorithmic FLOP, so its ceiling against this peak is 1/3
        bound, peak = "tensor", peaks.get("bf16_tflops", 1590.0) / 2
        src = ("TF32 dense peak taken as MEASURED_PEAKS.json bf16_tflops / 2" if peaks else "fallback 1590 / 2") + \
              ("; tf32x3 executes 3 tensor passes per algorithmic FLOP" if prec == "tf32x3" else "")
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    hbm = pb.alg_bytes() / (k_ms * 1e-3) / 1e9
    return {"bound": bound, "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
            "traffic": load_traffic(pb.key, prec), "peak_source": src, "kernel_ms": k_ms,
            "flop_per_traj_step": w["flop_exec"], "flop_per_traj_step_survey": w["flop"],
            "frac_survey_flop": ach * w["flop"] / w["flop_exec"] / peak,
            "hbm": {"achieved": hbm, "peak": hbm_peak, "unit": "GB/s", "frac": hbm / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}


def training_entries(dev, n_models=148):
    """Device-timed epoch of the fused training kernels (csrc/train_kernels.cu, train_rnn_kernels.cu) for
    n_models models side by side (one CTA each) on the validation bag: the reference's 80 % split in batches of 16 -> Adam
    model-steps / s.  A side table of the default run, not part of the headline metric."""
    import mpcdl
    from mpcdl import _lib, synth
    bags = synth.load_bags()
    out = {}
    for name, kind, hidden, layers, window in (("mlp_64_4", 0, 64, 4, None), ("gru_64_4_4", _lib.GRU, 64, 4, 4), ("lstm_64_4_4", _lib.LSTM, 64, 4, 4)):
        ds = mpcdl.RobotData(bags, window)
        f, l = ds.featrs.float().to(dev).contiguous(), ds.labels.float().to(dev).contiguous()
        n_train, batch = int(f.shape[0] * 0.8), 16
        steps = (n_train + batch - 1) // batch
        nm = n_models if kind == 0 else 2 * n_models          # recurrent kernel: two models per SM beyond one per SM
        torch.manual_seed(0)
        m = (mpcdl.Basic_Model(8, hidden, 6, layers, False) if kind == 0 else
             (mpcdl.GRU_Model if kind == _lib.GRU else mpcdl.LSTM_Model)(8, hidden, 6, layers, False))
        flat = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).to(dev)
        stride = flat.numel() if kind == 0 else (flat.numel() + 3) // 4 * 4
        theta = torch.zeros(nm, stride, device=dev)
        theta[:, :flat.numel()] = flat
        am, av = torch.zeros_like(theta), torch.zeros_like(theta)
        g = torch.Generator().manual_seed(0)
        orders = torch.stack([torch.randperm(f.shape[0], generator=g)[:n_train] for _ in range(nm)]).int().to(dev)
        if kind == 0:
            run = lambda s0: torch.ops.mpcdl.train_mlp_epoch(theta, am, av, f, l, orders, hidden, layers, batch, s0, 1e-3, 0.9, 0.999, 1e-8)  # noqa: E731
        else:
            work = torch.empty(nm * _lib.lib.mpcdl_train_rnn_work_floats(kind, hidden, layers, window), device=dev)
            run = lambda s0: torch.ops.mpcdl.train_rnn_epoch(theta, am, av, f, l, orders, work, kind, hidden, layers, batch, s0, 1e-3, 0.9, 0.999, 1e-8)  # noqa: E731
        run(0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        losses = run(steps)
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        out[name] = {"value": nm * steps / (ms * 1e-3), "unit": "Adam model-steps/s", "n_models": nm, "steps_per_epoch": steps,
                     "batch": batch, "epoch_ms": ms, "us_per_step": ms / steps * 1e3, "gpu_launches": 2,
                     "mean_loss_second_epoch": float(losses.mean().item())}
    return out


def table_entry(key, prec, max_steps, dev, flush, fp32_peak, peaks, local):
    """device-timed measurement of one more BASELINE configuration (N = 1): >= 3 warm-ups, L2 flushed between steps"""
    pb = Problem(key, prec, dev, 0, 1)
    fn = lambda: pb.rollout(pb.init_d, pb.U_d, pb.res_d)          # noqa: E731
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    fn()
    e0.record()
    fn()
    e1.record()
    fn()
    torch.cuda.synchronize(dev)
    est = max(e0.elapsed_time(e1), 1e-3)
    steps = int(max(2, min(max_steps, math.ceil(1000.0 / est))))
    sampler = ClockSampler(local)
    sampler.start()
    ms, _ = time_device(fn, steps, flush, dev)
    clocks = sampler.summary()
    k_ms = ms / steps
    r = roofline_of(pb, k_ms, fp32_peak, peaks)
    out = {"workload": pb.w["desc"], "precision": prec, "value": pb.N * pb.H / (k_ms * 1e-3), "unit": UNIT, "kernel_ms": k_ms,
           "steps": steps, "warmup": 3, "roofline": {k: r[k] for k in ("bound", "achieved", "peak", "unit", "frac", "frac_survey_flop", "traffic", "peak_source")},
           "clocks": clocks, "gpu_launches": steps}
    del pb
    torch.cuda.empty_cache()
    return out


def run_native(args, key):
    import torch.distributed as dist

    sys.path.insert(0, PKG)
    from mpcdl import ops, parallel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the fused rollout path has no CPU fallback "
                           "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))   # a protocol bug fails in minutes, not hours
    prec = args.precision
    pb = Problem(key, prec, dev, rank, world)
    w, N, H = pb.w, pb.N, pb.H
    warm = max(args.warmup, 3)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    for _ in range(warm):
        pb.step(pb.init_d, pb.U_d, pb.res_d)
    torch.cuda.synchronize(dev)
    fp32_peaks = ops.ffma_peak_tflops(dev, both=True)
    fp32_peak = max(fp32_peaks)

    # ---- timed region 1: inputs resident in HBM, per-step device time (step = kernel [+ all-gather of costs])
    sampler = ClockSampler(local)
    sampler.start()
    dev_ms, wall = time_device(lambda: pb.step(pb.init_d, pb.U_d, pb.res_d), args.steps, flush, dev, world, dist)
    # ---- kernel-only duration (no collective) for the roofline, same flush discipline
    k_ms = time_device(lambda: pb.rollout(pb.init_d, pb.U_d, pb.res_d), args.steps, flush, dev, world, dist)[0] / args.steps

    # ---- timed region 2: end to end with host buffers (H2D inputs, D2H result every step)
    init_p, U_p, res_p = pb.host_buffers()
    est_t = torch.tensor([dev_ms / args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(est_t, op=dist.ReduceOp.MAX)      # every rank must run the SAME number of steps (each step ends in a collective)
    est_ms = max(float(est_t.item()), 0.05)
    e2e_steps = int(max(args.steps, min(math.ceil(500.0 / est_ms), 5000)))
    if e2e_steps * est_ms > 4000.0:                       # keep the five repeats of a slow workload within ~20 s
        e2e_steps = max(3, int(4000.0 / est_ms))
    REPEATS = 5

    def timed(fn, after=None, repeats=REPEATS):
        for _ in range(3):
            fn()
        if after:
            after()
        out = []
        for _ in range(repeats):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for _ in range(e2e_steps):
                fn()
            if after:
                after()                 # the current stream waits for the last copy-out before the closing event
            t1.record()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)
            out.append(t0.elapsed_time(t1) / e2e_steps)
        return out

    def e2e_serial():
        a = init_p.to(dev, non_blocking=True)
        b = U_p.to(dev, non_blocking=True)
        res_p.copy_(pb.step(a, b, pb.res_d), non_blocking=True)

    # (a) one stream: H2D, kernel, D2H of a batch strictly one after the other
    serial = timed(e2e_serial, repeats=3)
    # (b) the public host-buffer driver (mpcdl.parallel.HostPipeline): the same bytes per batch, with the copy-in of
    #     batch i+1 and the copy-out of batch i-1 overlapping the kernel of batch i (slot-owned device buffers)
    pipe = parallel.HostPipeline(lambda a, b, out: pb.step(a, b, out), depth=2, device=dev)
    piped = timed(lambda: pipe.submit((init_p, U_p), res_p), after=pipe.drain)
    clocks = sampler.summary()

    t = torch.tensor([dev_ms, k_ms] + piped + serial, device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)          # every figure: max over ranks
    v = [float(x) for x in t.tolist()]
    dev_ms, k_ms, piped, serial = v[0], v[1], v[2:2 + REPEATS], v[2 + REPEATS:]
    e2e_ms, e2e_serial_ms = statistics.median(piped), statistics.median(serial)
    ms_per_step = dev_ms / args.steps

    if rank == 0:
        units = pb.N_total * H
        meta = pb.meta
        kind = {"Basic": 0, "GRU": 1, "LSTM": 2}[meta["kind"]]
        plan = ops.plan(kind, meta["hidden"], meta["layers"], meta["window"], N, H, stateful=meta["state_f"],
                        noise_cfg=pb.noise.cfg() if pb.noise is not None else None, precision=prec)
        peaks = load_peaks()
        roofline = roofline_of(pb, k_ms, fp32_peak, peaks)
        roofline["fp32_ffma_peak"] = fp32_peak
        roofline["fp32_peaks_measured"] = {"ffma": fp32_peaks[0], "ffma2": fp32_peaks[1]}
        lps = pb.launches_per_step
        line = {"metric": METRIC, "value": units / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": w["scaling"],
                "vs_baseline": None, "dtype": DTYPE[prec], "data": "synthetic",
                "config": {"workload": w["desc"], "N_total": pb.N_total, "N_per_gpu": N, "H": H,
                           "l2": "flushed between steps (256 MiB memset outside the event pairs)",
                           "parallelism": (f"trajectory shards x{world} ({w['scaling']} scaling), one all-gather of costs per step"
                                           if world > 1 else "single GPU"),
                           "plan": plan},
                "roofline": roofline,
                "e2e": {"value": units / (e2e_ms * 1e-3), "unit": UNIT,
                        "h2d_bytes_per_step": (init_p.numel() + U_p.numel()) * 4 * world,
                        "d2h_bytes_per_step": res_p.numel() * 4 * world, "ms_per_step": e2e_ms, "steps": e2e_steps,
                        "repeats_ms": piped,
                        "how": f"mpcdl.parallel.HostPipeline(depth=2): pinned host buffers, every batch copied in and its result copied out, "
                               f"copies of neighbouring batches overlap the kernel; median of {REPEATS} repeats of {e2e_steps} steps, max over ranks",
                        "serial_value": units / (e2e_serial_ms * 1e-3), "serial_ms_per_step": e2e_serial_ms},
                "gpu_launches": args.steps * lps,
                "gpu_launches_total": (warm + 2 * args.steps + (REPEATS + 3) * e2e_steps + 6 * 2) * lps + 6,
                "clocks": clocks, "wall_s_region1": wall}
    del pipe, init_p, U_p, res_p
    if rank == 0 and world == 1 and key == DEFAULT and prec == "fp32" and not args.no_table:
        del pb
        torch.cuda.empty_cache()
        table = {}
        for k2, p2, mx in TABLE:
            try:
                table[f"{k2}_{p2}"] = table_entry(k2, p2, mx, dev, flush, fp32_peak, peaks, local)
            except Exception as e:          # one configuration failing must not take the headline line with it
                table[f"{k2}_{p2}"] = {"error": f"{type(e).__name__}: {e}"}
        line["workloads"] = table
        line["gpu_launches_total"] += sum(v.get("gpu_launches", 0) + 3 for v in table.values())
        try:
            line["training"] = training_entries(dev)
            line["gpu_launches_total"] += sum(v["gpu_launches"] for v in line["training"].values())
        except Exception as e:              # a side table: never at the price of the headline line
            line["training"] = {"error": f"{type(e).__name__}: {e}"}
    if rank == 0:
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default=DEFAULT, choices=sorted(WORKLOADS))
    ap.add_argument("--skip-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-table", action="store_true", help="skip the per-workload table of the default N = 1 run")
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32x3", "tf32", "bf16"],
                    help="gate-GEMM arithmetic: fp32 = FFMA kernels (parity path), tf32x3 / tf32 = tcgen05 kernels (hidden 64), "
                         "bf16 = tcgen05 kernel of the hidden-256 LSTM (workload c4)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args, WORKLOADS[args.workload])
    else:
        run_native(args, args.workload)


if __name__ == "__main__":
    main()
