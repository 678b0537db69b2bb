"""Quick GPU sanity + timing sweep (development aid, not a test): python tools/gpu_quick.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import numpy as np, torch
import mpcdl
from mpcdl import synth, ops

dev = torch.device("cuda:0")
GOLD = os.path.join(ROOT, "tests", "golden")
def ckpt(n): return {k: torch.from_numpy(v) for k, v in np.load(f"{GOLD}/checkpoints/{n}.npz").items()}
z = np.load(f"{GOLD}/stats.npz")
mst = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("mlp_")}
rst = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("rnn4_")}
rows = synth.load_rows()
print("ffma peak TFLOP/s:", ops.ffma_peak_tflops())

def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts)

FLOP = {"Basic_64_4_32": 34560, "GRU_64_4_4_16_False": 701184, "LSTM_64_4_4_16_False": 934656, "Basic_16_1_32": 960}
cases = [("Basic_64_4_32", 4096, 50, 0), ("Basic_64_4_32", 65536, 50, 0), ("Basic_64_4_32", 1048576, 50, 0),
         ("Basic_16_1_32", 1048576, 50, 0),
         ("GRU_64_4_4_16_False", 4096, 20, 4), ("GRU_64_4_4_16_False", 65536, 100, 4), ("LSTM_64_4_4_16_False", 65536, 100, 4)]
sel = sys.argv[1:] 
for name, N, H, w in cases:
    if sel and name.split("_")[0] not in sel: continue
    m, meta = mpcdl.from_checkpoint_name(name, ckpt(name), dev)
    st = mst if w == 0 else rst
    init, U = synth.make_inputs(rows, min(N, 65536), H, window=w, seed=0)
    if N > 65536:
        r = N // 65536; init = init.repeat(r, *([1] * (init.dim() - 1))); U = U.repeat(r, 1, 1)
    init, U = init.to(dev), U.to(dev)
    for tune in ([()] + [tuple(int(v) for v in t.split(",")) for t in os.environ.get("TUNES", "").split(";") if t]):
        try:
            kind = {"mlp": 0, "gru": 1, "lstm": 2}[meta["kind"]]
            pl = ops.plan(kind, meta["hidden"], meta["layers"], w, N, H, tune=tune)
            ms = timeit(lambda: m.rollout(init, U, st, tune=tune))
            print(f"{name} N={N} H={H} tune={tune} rt={pl['rt']} TN={pl['tile_traj']} thr={pl['threads']} grid={pl['grid']} smem={pl['smem_bytes']}: "
                  f"{ms:.3f} ms  {N*H/ms/1e6:.3f} G traj-steps/s  {N*H*FLOP[name]/ms/1e9:.2f} TFLOP/s", flush=True)
        except Exception as e:
            print(name, N, H, tune, "FAILED", e, flush=True)
