"""Profiling target: a few launches of one rollout kernel.  python tools/tc_prof.py NAME PREC N H [noise]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import numpy as np, torch
import mpcdl
from mpcdl import synth
name, prec, N, H = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
noise_on = len(sys.argv) > 5
dev = torch.device("cuda:0")
GOLD = os.path.join(ROOT, "tests", "golden")
ck = lambda n: {k: torch.from_numpy(v) for k, v in np.load(f"{GOLD}/checkpoints/{n}.npz").items()}
z = np.load(f"{GOLD}/stats.npz")
import bench
m, meta = mpcdl.from_checkpoint_name(name, bench.load_ckpt(name), dev)    # hidden-256 LSTM: seeded random init
pre = "mlp_" if meta["window"] == 0 else "rnn4_"
st = {k[len(pre):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(pre)}
init, U = synth.make_inputs(synth.load_rows(), N, H, window=meta["window"], seed=0)
init, U = init.to(dev), U.to(dev)
noise = None
if noise_on:
    noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    noise.model.load_state_dict(ck("noise_seed0_h3_s32_l2")); noise.to(dev)
for _ in range(3):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record(); m.rollout(init, U, st, precision=prec, noise=noise, seed=1); e1.record(); e1.synchronize()
    print(name, prec, N, H, "ms", e0.elapsed_time(e1), flush=True)
