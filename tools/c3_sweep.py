"""C3 (GRU + noise, FP32 FFMA kernel) launch-parameter sweep (development aid): tune = (rt, n_tg, nslots, kc_rows).
One wave-balanced slice of the workload: N trajectories x H = 20 steps."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import numpy as np, torch
import mpcdl
from mpcdl import synth, ops, _lib
dev = torch.device("cuda:0")
GOLD = os.path.join(ROOT, "tests", "golden")
ck = lambda n: {k: torch.from_numpy(v) for k, v in np.load(f"{GOLD}/checkpoints/{n}.npz").items()}
z = np.load(f"{GOLD}/stats.npz")
st = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("rnn4_")}
name = sys.argv[1] if len(sys.argv) > 1 else "GRU_64_4_4_16_False"
with_noise = (sys.argv[2] != "0") if len(sys.argv) > 2 else True
m, meta = mpcdl.from_checkpoint_name(name, ck(name), dev)
noise = None
if with_noise:
    noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    noise.model.load_state_dict(ck("noise_seed0_h3_s32_l2")); noise.to(dev)
N, H = 65536, 20
init, U = synth.make_inputs(synth.load_rows(), N, H, window=4, seed=0)
init, U = init.to(dev), U.to(dev)
kind = _lib.GRU if name.startswith("GRU") else _lib.LSTM
ref = None
cfgs = [()]
for rt in (8, 4):
    for ntg in ((0, 14, 12, 10, 8) if rt == 8 else (0, 30, 28, 24, 16)):
        for ns in (0, 2, 4, 5):
            for kc in (0, 8, 16, 32):
                cfgs.append((rt, ntg, ns, kc))
res = []
for tune in cfgs:
    try:
        pl = ops.plan(kind, 64, 4, 4, N, H, noise_cfg=noise.cfg() if noise else None, tune=tune)
        out = m.rollout(init, U, st, noise=noise, seed=1, tune=tune)[0]
        torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
            e0.record(); m.rollout(init, U, st, noise=noise, seed=1, tune=tune); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
        if ref is None:
            ref = out
        res.append((min(ts), tune, pl, torch.equal(out, ref)))
    except Exception as e:
        pass
res.sort(key=lambda r: r[0])
for t, tune, pl, eq in res[:25]:
    print(f"{t:8.3f} ms  tune={str(tune):18s} TN {pl['tile_traj']:3d} thr {pl['threads']:3d} grid {pl['grid']} ntiles {pl['ntiles']} slots {pl['nslots']} x {pl['slot_floats']} floats smem {pl['smem_bytes']} resident {pl['resident']} equal {eq}")
print("default:", [r for r in res if r[1] == ()][0][0], "ms; configurations tried:", len(res))
