"""Development aid: accuracy + timing of the tensor-core MLP rollout.   python tools/mc_check.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import numpy as np, torch
import mpcdl
from mpcdl import synth, ops
from oracle import rollout_ref as R
dev = torch.device("cuda:0")
GOLD = os.path.join(ROOT, "tests", "golden")
ck = lambda n: {k: torch.from_numpy(v) for k, v in np.load(f"{GOLD}/checkpoints/{n}.npz").items()}
z = np.load(f"{GOLD}/stats.npz")
st = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("mlp_")}
rows = synth.load_rows()
def q(e):
    e = e.flatten().double()
    return " ".join(f"{torch.quantile(e, x).item():.2e}" for x in (0.5, 0.9, 0.99)) + f" max {e.max().item():.2e}"
for name in ("Basic_64_4_32", "Basic_64_1_32", "Basic_64_2_64"):
    m, meta = mpcdl.from_checkpoint_name(name, ck(name), dev)
    N, H = 1000, 50
    init, U = synth.make_inputs(rows, N, H, window=0, seed=3)
    ref, _ = R.build_from_name(name, ck(name))
    y32 = R.rollout_mlp(ref, st, init, U)[0]
    y64 = R.rollout_mlp(R.as_double(ref), {k: v.double() for k, v in st.items()}, init.double(), U.double())[0]
    print(f"== {name} N={N} H={H}: rel err vs FP64 (step-1 max | p50 p90 p99 max)")
    print(f"   torch fp32   : {R.rel_err(y32, y64)[:, 0].max().item():.2e} | {q(R.rel_err(y32, y64))}")
    for prec in ("fp32", "tf32x3", "tf32"):
        for tune in ((), (0, 1, 0, 0), (0, 2, 0, 0)):
            if prec == "fp32" and tune: continue
            try:
                yk = m.rollout(init.to(dev), U.to(dev), st, precision=prec, tune=tune)[0].cpu()
                e = R.rel_err(yk, y64)
                print(f"   kernel {prec:7s} {str(tune):14s}: {e[:, 0].max().item():.2e} | {q(e)} | vs torch fp32 step1 {R.rel_err(yk, y32)[:, 0].max().item():.2e}", flush=True)
            except Exception as ex:
                print(f"   kernel {prec} {tune}: FAILED {ex}", flush=True)
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts)
m, meta = mpcdl.from_checkpoint_name("Basic_64_4_32", ck("Basic_64_4_32"), dev)
cost = synth.make_cost(rows, 50)
for N in (4096, 65536, 1048576):
    init, U = synth.make_inputs(rows, min(N, 65536), 50, window=0, seed=0)
    if N > 65536:
        init = init.repeat(N // 65536, 1); U = U.repeat(N // 65536, 1, 1)
    init, U = init.to(dev), U.to(dev)
    for prec in ("fp32", "tf32x3", "tf32"):
        for tune in ((), (0, 1, 0, 0), (0, 2, 0, 0)):
            if prec == "fp32" and tune: continue
            for rs in (True, False):
                try:
                    ms = timeit(lambda: m.rollout(init, U, st, precision=prec, tune=tune, cost=cost, return_states=rs))
                    pl = ops.plan(0, 64, 4, 0, N, 50, precision=prec, tune=tune)
                    print(f"N={N} {prec:7s} tune={str(tune):14s} states={rs} grid={pl['grid']} tiles/cta={pl['ksplit']}: {ms:.3f} ms {N*50/ms/1e3:.3e} traj-steps/s {N*50*34560/ms/1e9:.1f} TFLOP/s", flush=True)
                except Exception as ex:
                    print(N, prec, tune, "FAILED", ex, flush=True)
