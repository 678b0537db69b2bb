"""Development aid: accuracy table + timing of the BF16 tcgen05 LSTM-256 rollout (BASELINE.json config 4) against the
FP32 FFMA kernel and the FP32 / FP64 oracle.   python tools/lt_check.py [acc] [time]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import numpy as np, torch
import mpcdl
from mpcdl import synth
from oracle import rollout_ref as R
import bench

dev = torch.device("cuda:0")
rst = bench.load_stats(4)
rows = synth.load_rows()
what = sys.argv[1:] or ["acc", "time"]
NAME = "LSTM_256_4_4_0_False"
sd = bench.load_ckpt(NAME)

def q(e):
    e = e.flatten().double()
    return " ".join(f"{torch.quantile(e, x).item():.2e}" for x in (0.5, 0.9, 0.99)) + f" max {e.max().item():.2e}"

if "acc" in what:
    m, meta = mpcdl.from_checkpoint_name(NAME, sd, dev)
    for N, H in ((300, 1), (300, 20), (148 * 128 + 77, 3)):
        init, U = synth.make_inputs(rows, N, H, window=4, seed=3)
        yk32 = m.rollout(init.to(dev), U.to(dev), rst, precision="fp32")[0].cpu()
        ykbf = m.rollout(init.to(dev), U.to(dev), rst, precision="bf16")[0].cpu()
        torch.cuda.synchronize()
        print(f"== {NAME} N={N} H={H}: bf16 kernel vs fp32 kernel: step1 {R.rel_err(ykbf, yk32)[:, 0].max().item():.2e} | {q(R.rel_err(ykbf, yk32))}"
              f" | nan {int(torch.isnan(ykbf).sum())}", flush=True)
        if N <= 300:
            ref, _ = R.build_from_name(NAME, sd)
            y32 = R.rollout_rnn(ref, rst, init, U, stateful=False)[0]
            y64 = R.rollout_rnn(R.as_double(ref), {k: v.double() for k, v in rst.items()}, init.double(), U.double(), stateful=False)[0]
            print(f"   rel err vs FP64 oracle, p50 p90 p99 max (step 1 max | full horizon)")
            print(f"   torch fp32   : {R.rel_err(y32, y64)[:, 0].max().item():.2e} | {q(R.rel_err(y32, y64))}")
            print(f"   kernel fp32  : {R.rel_err(yk32, y64)[:, 0].max().item():.2e} | {q(R.rel_err(yk32, y64))}")
            print(f"   kernel bf16  : {R.rel_err(ykbf, y64)[:, 0].max().item():.2e} | {q(R.rel_err(ykbf, y64))}", flush=True)
    # costs + determinism
    N, H = 1000, 5
    init, U = synth.make_inputs(rows, N, H, window=4, seed=5)
    cost = tuple(c.to(dev) for c in synth.make_cost(rows, H))
    a = m.rollout(init.to(dev), U.to(dev), rst, precision="bf16", cost=cost)
    b = m.rollout(init.to(dev), U.to(dev), rst, precision="bf16", cost=cost)
    c = m.rollout(init.to(dev), U.to(dev), rst, precision="fp32", cost=cost)
    print("bitwise repeatable:", bool((a[0] == b[0]).all() and (a[1] == b[1]).all()),
          "| cost rel err vs fp32 kernel:", ((a[1] - c[1]).abs() / c[1].abs()).max().item(), flush=True)

if "time" in what:
    def timeit(fn, reps=3):
        fn(); torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record(); fn(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
        return min(ts)
    m, meta = mpcdl.from_checkpoint_name(NAME, sd, dev)
    for N, H, precs in ((148 * 128, 100, ("bf16",)), (65536, 100, ("bf16", "fp32"))):
        init, U = synth.make_inputs(rows, N, H, window=4, seed=0)
        init, U = init.to(dev), U.to(dev)
        for prec in precs:
            ms = timeit(lambda: m.rollout(init, U, rst, precision=prec), reps=3 if prec == "bf16" else 1)
            print(f"{NAME} N={N} H={H} {prec:5s}: {ms:.2f} ms  {N*H/ms/1e3:.3e} traj-steps/s  {N*H*14748672/ms/1e9:.1f} TFLOP/s", flush=True)
