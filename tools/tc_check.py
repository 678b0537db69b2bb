"""Development aid: accuracy table + timing of the tensor-core (tcgen05) GRU/LSTM rollout against the FP32
kernel and the FP32 / FP64 oracle.   python tools/tc_check.py [acc] [time]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import numpy as np, torch
import mpcdl
from mpcdl import synth
from oracle import rollout_ref as R

dev = torch.device("cuda:0")
GOLD = os.path.join(ROOT, "tests", "golden")
def ckpt(n): return {k: torch.from_numpy(v) for k, v in np.load(f"{GOLD}/checkpoints/{n}.npz").items()}
z = np.load(f"{GOLD}/stats.npz")
rst = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("rnn4_")}
rows = synth.load_rows()
what = sys.argv[1:] or ["acc", "time"]

def q(e):
    e = e.flatten().double()
    return " ".join(f"{torch.quantile(e, x).item():.2e}" for x in (0.5, 0.9, 0.99)) + f" max {e.max().item():.2e}"

if "acc" in what:
    for name in ("GRU_64_4_4_16_False", "GRU_64_4_4_64_True", "LSTM_64_4_4_16_False", "LSTM_64_4_4_64_True"):
        m, meta = mpcdl.from_checkpoint_name(name, ckpt(name), dev)
        N, H = 300, 50
        init, U = synth.make_inputs(rows, N, H, window=4, seed=3)
        ref, _ = R.build_from_name(name, ckpt(name))
        y32 = R.rollout_rnn(ref, rst, init, U, stateful=meta["state_f"])[0]
        y64 = R.rollout_rnn(R.as_double(ref), {k: v.double() for k, v in rst.items()}, init.double(), U.double(), stateful=meta["state_f"])[0]
        print(f"== {name} N={N} H={H}: rel err vs FP64 oracle, p50 p90 p99 max (step 1 max | full horizon)")
        print(f"   torch fp32   : {R.rel_err(y32, y64)[:, 0].max().item():.2e} | {q(R.rel_err(y32, y64))}")
        for prec in ("fp32", "tf32x3", "tf32"):
            try:
                yk = m.rollout(init.to(dev), U.to(dev), rst, precision=prec)[0].cpu()
                torch.cuda.synchronize()
                e = R.rel_err(yk, y64)
                print(f"   kernel {prec:7s}: {e[:, 0].max().item():.2e} | {q(e)}   | vs torch fp32 step1 {R.rel_err(yk, y32)[:, 0].max().item():.2e}", flush=True)
            except Exception as ex:
                print(f"   kernel {prec}: FAILED {ex}", flush=True)

if "time" in what:
    def timeit(fn, reps=3):
        fn(); torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record(); fn(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
        return min(ts)
    FLOP = {"GRU": 701184, "LSTM": 934656}
    for name, N, H in (("GRU_64_4_4_16_False", 65536, 100), ("LSTM_64_4_4_16_False", 65536, 100), ("GRU_64_4_4_16_False", 148 * 128, 100)):
        m, meta = mpcdl.from_checkpoint_name(name, ckpt(name), dev)
        init, U = synth.make_inputs(rows, N, H, window=4, seed=0)
        init, U = init.to(dev), U.to(dev)
        for prec in ("fp32", "tf32x3", "tf32"):
            for tune in ((), (0, 0, 0, -1)):
                if prec == "fp32" and tune: continue
                try:
                    kind = {"gru": 1, "lstm": 2}[meta["kind"]]
                    pl = mpcdl.ops.plan(kind, 64, 4, 4, N, H, precision=prec, tune=tune)
                    ms = timeit(lambda: m.rollout(init, U, rst, precision=prec, tune=tune))
                    print(f"{name} N={N} H={H} {prec:7s} resident={pl['resident']} nslots={pl['nslots']} smem={pl['smem_bytes']}: {ms:.2f} ms  "
                          f"{N*H/ms/1e3:.3e} traj-steps/s  {N*H*FLOP[name.split('_')[0]]/ms/1e9:.1f} TFLOP/s", flush=True)
                except Exception as ex:
                    print(name, prec, tune, "FAILED", ex, flush=True)
