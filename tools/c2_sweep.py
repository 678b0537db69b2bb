"""C2 tile sweep (development aid): python tools/c2_sweep.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import numpy as np, torch
import mpcdl
from mpcdl import synth, ops
dev = torch.device("cuda:0")
GOLD = os.path.join(ROOT, "tests", "golden")
ck = lambda n: {k: torch.from_numpy(v) for k, v in np.load(f"{GOLD}/checkpoints/{n}.npz").items()}
z = np.load(f"{GOLD}/stats.npz")
st = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("mlp_")}
m, meta = mpcdl.from_checkpoint_name("Basic_64_4_32", ck("Basic_64_4_32"), dev)
init, U = synth.make_inputs(synth.load_rows(), 4096, 50, window=0, seed=0)
init, U = init.to(dev), U.to(dev)
ref = m.rollout(init, U, st, tune=(4, 8, 0, 2))[0]
for tune in ((), (4, 5, 0, 2), (4, 6, 0, 2), (4, 7, 0, 2), (4, 8, 0, 2)):
    out = m.rollout(init, U, st, tune=tune)[0]
    torch.cuda.synchronize()
    ts = []
    for _ in range(20):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); m.rollout(init, U, st, tune=tune); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    pl = ops.plan(0, 64, 4, 0, 4096, 50, tune=tune)
    print(tune, "TN", pl["tile_traj"], "grid", pl["grid"], "thr", pl["threads"], "min ms %.4f med %.4f" % (min(ts), sorted(ts)[10]), "bit-equal to TN=32:", torch.equal(out, ref))
