"""Pin the oracle against the UNMODIFIED reference and (re)generate tests/golden/.

Runs only in the build container (needs /root/reference; the GPU box does not have it):

    python oracle/pin_reference.py

What it does
  1. imports the reference modules with the scratch-layout recipe of SURVEY.md 8c
     (``../mpc_data_clean.pkl`` and ``fwd_models/`` are resolved relative to the cwd at import,
     model_scripts/basic_model.py:9, eval_models.py:18,28); matplotlib / seaborn / mpclab_common
     are absent and are stubbed in ``sys.modules`` (they are never called on this path);
  2. exports the data assets the tests need on the GPU box as plain arrays
     (tests/golden/data/*.npy, tests/golden/checkpoints/*.npz) -- data, not source;
  3. checks bit-identity oracle == reference for: dataset build + z-score (a1, a3), forward of all
     16 checkpoints on every validation row/window (a2, a4, a5, stateless and stateful), the noise
     model forward under a shared seed (a6), the history buffering of DynamicsNN.get_prediction
     (a7), calc_error (a9);
  4. drives the REFERENCE modules in a python loop (closed loop of DESIGN.md) and stores the
     trajectories as golden vectors; checks oracle.rollout_* reproduces them bit-for-bit.
  5.-7. DynamicsDataset resampling, the second validation pickle + the validation sweep, and the training step
     (reference Basic_Model + nn.L1Loss + optim.Adam over explicit batches == oracle/train_ref.py, golden train.npz;
     7b: the same for the reference's GRU_Model / LSTM_Model, state_f False and True, golden train_rnn.npz).
Exit code 0 and "PINNED" on the last line when everything holds.
"""
from __future__ import annotations

import contextlib
import io
import os
import pickle
import sys
import tempfile
import types

import numpy as np
import torch

sys.dont_write_bytecode = True
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))

from oracle import rollout_ref as R  # noqa: E402

torch.set_num_threads(1)
checks = []


def check(name, ok):
    checks.append((name, bool(ok)))
    print(("ok   " if ok else "FAIL ") + name, flush=True)


def bit_equal(a, b):
    return a.shape == b.shape and torch.equal(a, b)


# ---------------------------------------------------------------------------- 1. import
def import_reference():
    scratch = tempfile.mkdtemp(prefix="mpcdl_ref_")
    os.makedirs(os.path.join(scratch, "model_data"))
    os.symlink(f"{REF}/data/validation/mpc_data_clean.pkl", f"{scratch}/mpc_data_clean.pkl")
    for d in ("fwd_models", "rnn_models"):
        os.symlink(f"{REF}/model_data/{d}", f"{scratch}/model_data/{d}")
    os.chdir(os.path.join(scratch, "model_data"))
    for name in ("matplotlib", "matplotlib.pyplot", "seaborn", "mpclab_common",
                 "mpclab_common.track", "mpclab_common.pytypes", "mpclab_common.models",
                 "mpclab_common.models.model_types", "mpclab_common.models.dynamics_models"):
        sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["mpclab_common.track"].get_track = lambda *a, **k: None

    class _VS:  # VehicleState(t=0) is evaluated as a default argument, noise_model.py:69
        def __init__(self, *a, **k):
            pass
    sys.modules["mpclab_common.pytypes"].VehicleState = _VS
    sys.modules["mpclab_common.models.model_types"].DynamicBicycleConfig = object
    sys.modules["mpclab_common.models.dynamics_models"].CasadiDynamicCLBicycle = object
    sys.path.insert(0, f"{REF}/model_scripts")
    sys.path.insert(0, REF)
    import basic_model, deep_models, eval_models          # noqa: E401
    import utils.data_utils as du
    du.q_labels = ["v_long", "v_tran", "w_psi", "x", "y", "psi"]     # utils/data_utils.py:100
    du.u_labels = ["u_a", "u_steer"]                                 # utils/data_utils.py:101
    import noise_model
    return basic_model, deep_models, eval_models, noise_model


fwd, rnn, ev, nm = import_reference()


# ---------------------------------------------------------------------------- 2. fixtures
def export_fixtures():
    os.makedirs(f"{GOLD}/data", exist_ok=True)
    os.makedirs(f"{GOLD}/checkpoints", exist_ok=True)
    for name in ("mpc_data_clean", "pid_data_clean"):
        bags = pickle.load(open(f"{REF}/data/validation/{name}.pkl", "rb"))
        assert len(bags) == 1
        np.save(f"{GOLD}/data/{name}.npy", np.asarray(bags[0], dtype=np.float64))
    for d in ("fwd_models", "rnn_models"):
        for f in sorted(os.listdir(f"{REF}/model_data/{d}")):
            sd = torch.load(f"{REF}/model_data/{d}/{f}", map_location="cpu")
            np.savez(f"{GOLD}/checkpoints/{f[:-3]}.npz", **{k: v.numpy() for k, v in sd.items()})


export_fixtures()
from mpcdl import synth  # noqa: E402  (needs the data fixture written above)

bags = synth.load_bags("mpc_data_clean")
check("fixture rows == pickle rows (float64 exact)", bags == fwd.DATA)


def ckpt(name):
    return {k: torch.from_numpy(v) for k, v in np.load(f"{GOLD}/checkpoints/{name}.npz").items()}


# ---------------------------------------------------------------------------- 3a. datasets
mf, ml = R.mlp_dataset(bags)
mst = R.zscore_stats(mf, ml)
mfn, mln = R.apply_zscore(mf, ml, mst)
ref_ds = fwd.dataset
check("a1 MLP dataset+zscore bit-identical", bit_equal(mfn, ref_ds.featrs.cpu()) and bit_equal(mln, ref_ds.labels.cpu()))
rf, rl = R.rnn_dataset(bags, 4)
rst = R.zscore_stats(rf, rl)
rfn, rln = R.apply_zscore(rf, rl, rst)
ref_rds = rnn.dataset
check("a3 RNN dataset+zscore bit-identical (T=4)", bit_equal(rfn, ref_rds.featrs.cpu()) and bit_equal(rln, ref_rds.labels.cpu()))
check("a3 stats are per window position [4,8]", tuple(rst["f_mean"].shape) == (4, 8))
np.savez(f"{GOLD}/stats.npz", **{f"mlp_{k}": v.numpy() for k, v in mst.items()},
         **{f"rnn4_{k}": v.numpy() for k, v in rst.items()})

# ---------------------------------------------------------------------------- 3b. forwards
onestep = {}
names_fwd = sorted(f[:-3] for f in os.listdir(f"{REF}/model_data/fwd_models"))
names_rnn = sorted(f[:-3] for f in os.listdir(f"{REF}/model_data/rnn_models"))
for name in names_fwd:
    p = name.split("_")
    ref_m = fwd.Basic_Model(8, int(p[1]), 6, int(p[2]), False).cpu()
    ref_m.load_state_dict(torch.load(f"{REF}/model_data/fwd_models/{name}.pt", map_location="cpu"))
    ref_m.eval()
    mine, _ = R.build_from_name(name, ckpt(name))
    with torch.no_grad():
        yr, ym = ref_m(mfn), mine(mfn)
    check(f"a2 {name} forward bit-identical on all {len(mfn)} rows", bit_equal(yr, ym))
    onestep[name] = yr.numpy()
    # a9: reference calc_error over the whole dataset (SAMPLE=False), batch 1
    ev.SAMPLE = False
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ev.calc_error(ref_ds, ref_m, name)
    printed = [float(l.split(":")[1]) for l in buf.getvalue().splitlines() if ":" in l]
    mae, rmse = R.calc_error(ym, mln)
    check(f"a9 {name} calc_error {printed} vs oracle ({mae:.4f},{rmse:.4f})",
          abs(printed[0] - mae) < 1e-4 and abs(printed[1] - rmse) < 1e-4)
    onestep[name + "_err"] = np.asarray([mae, rmse])

rnn.device = torch.device("cpu")   # init_hidden allocates on the module-global (deep_models.py:58)
for name in names_rnn:
    p = name.split("_")
    cls = rnn.GRU_Model if p[0] == "GRU" else rnn.LSTM_Model
    state_f = p[-1] == "True"
    ref_m = cls(8, int(p[1]), 6, int(p[2]), state_f).cpu()
    ref_m.load_state_dict(torch.load(f"{REF}/model_data/rnn_models/{name}.pt", map_location="cpu"))
    ref_m.eval()
    mine, meta = R.build_from_name(name, ckpt(name))
    with torch.no_grad():
        # stateless call: the reference zero-inits when not state_f or on the first call
        ref_m.h_state = None
        yr = ref_m(rfn)
        ym, st1 = mine(rfn, mine.zero_state(len(rfn)))
        ok = bit_equal(yr, ym)
        if state_f:   # second call carries h_state (deep_models.py:61-65)
            yr2 = ref_m(rfn)
            ym2, _ = mine(rfn, st1)
            ok = ok and bit_equal(yr2, ym2)
            onestep[name + "_second"] = yr2.numpy()
    check(f"a4/a5 {name} forward bit-identical on all {len(rfn)} windows (state_f={state_f})", ok)
    onestep[name] = yr.numpy()
    mae, rmse = R.calc_error(ym, rln)
    onestep[name + "_err"] = np.asarray([mae, rmse])
np.savez_compressed(f"{GOLD}/onestep.npz", **onestep)

# ---------------------------------------------------------------------------- 3c. noise model
import utils.pytorch_utils as ptu  # noqa: E402
ptu.device = torch.device("cpu")
torch.manual_seed(0)
ref_noise = nm.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh",
                                     output_activation="identity").eval()
ref_noise.logger.setLevel("WARNING")   # forward() logs whole tensors at DEBUG (noise_model.py:38)
my_noise = R.RefNoise(6, 2, 3, 2, 32, "tanh", "identity")
my_noise.model.load_state_dict(ref_noise.model.state_dict())
my_noise.eval()
np.savez(f"{GOLD}/checkpoints/noise_seed0_h3_s32_l2.npz",
         **{k: v.numpy() for k, v in ref_noise.model.state_dict().items()})
qh = torch.as_tensor(np.asarray(bags[0]), dtype=torch.float32)[:512, R.Q_PERM].repeat(1, 3)
uh = torch.as_tensor(np.asarray(bags[0]), dtype=torch.float32)[:512, 6:8].repeat(1, 3)
with torch.no_grad():
    torch.manual_seed(123)
    nr = ref_noise(qh, uh)
    torch.manual_seed(123)
    eps = torch.randn(512, 6)
    nmine = my_noise(qh, uh, eps)
check("a6 noise forward bit-identical under a shared seed (eps = randn)", bit_equal(nr, nmine))

dyn = nm.DynamicsNN(nominal_model=None, noise_model=ref_noise, dynamics_uses_frenet=False, history=3)
hist = R._NoiseHistory(3)
ok = True
rows32 = torch.as_tensor(np.asarray(bags[0]), dtype=torch.float32)
with torch.no_grad():
    for t in range(6):
        q, u = rows32[t, R.Q_PERM], rows32[t, 6:8]
        torch.manual_seed(t)
        pr = dyn.get_prediction(q.numpy(), u.numpy())
        torch.manual_seed(t)
        e = torch.randn(1, 6)
        qq, uu = hist.push(q[None], u[None])
        ok = ok and np.array_equal(pr, my_noise(qq, uu, e).numpy())
check("a7 DynamicsNN.get_prediction history/left-pad semantics bit-identical (6 calls)", ok)

# ---------------------------------------------------------------------------- 3d. run_dynamics_model (a8)
sys.modules["mpclab_common.models.dynamics_models"].DynamicBicycleConfig = object
import scripts.test_model as tm  # noqa: E402   (the reference's own loop, scripts/test_model.py:67-87)
import model_scripts.noise_model as nm2  # noqa: E402  (test_model imports the package path: a second module object)


class _IdentityNominal:
    """stand-in for the missing CasADi bicycle (mpclab_common): next q = q.  Only q_nominal uses it."""
    def qu2state(self, state, q, u): state.q = q
    def step(self, state): pass
    def state2q(self, state): return state.q


aux = {}
Tn = 24
q_data = rows32[:Tn, R.Q_PERM]
u_data = rows32[:Tn, 6:8]
dataset_h1 = [(q_data[k][None].numpy(), u_data[k][None].numpy(), None) for k in range(Tn)]
for open_loop in (True, False):
    dyn2 = nm2.DynamicsNN(_IdentityNominal(), ref_noise, False, history=3)
    with torch.no_grad():
        torch.manual_seed(11)
        (qd, qn, qz), _ = tm.run_dynamics_model(dyn2, dataset_h1, open_loop)
        torch.manual_seed(11)
        eps_dyn = torch.stack([torch.randn(1, 6) for _ in range(Tn - 1)])[:, 0]
    mine_tr = R.run_dynamics_model(my_noise, q_data, u_data, eps_dyn, open_loop)
    ref_tr = torch.from_numpy(np.stack([np.asarray(x, dtype=np.float32) for x in qz]))
    check(f"a8 run_dynamics_model(open_loop={open_loop}) 'with noise' trajectory: oracle == reference (bit)", bit_equal(ref_tr, mine_tr))
    aux["rdm_open" if open_loop else "rdm_closed"] = ref_tr.numpy()
aux["rdm_q"], aux["rdm_u"], aux["rdm_eps"] = q_data.numpy(), u_data.numpy(), eps_dyn.numpy()
# the batched closed loop restates the same recursion (q0 = q_data[0], U[t] = u_data[t+1])
tr_b = R.rollout_dynnn(my_noise, q_data[:1], u_data[None, 1:], eps_dyn[None])
check("a7/a8 rollout_dynnn (batched) == run_dynamics_model(open_loop=True)", bit_equal(tr_b[0], torch.from_numpy(aux["rdm_open"][1:])))
aux["noise_q"], aux["noise_u"], aux["noise_eps"], aux["noise_out"] = qh.numpy(), uh.numpy(), eps.numpy(), nr.numpy()
mu_, sc_ = my_noise.mu_scale(qh, uh)
aux["noise_mu"], aux["noise_scale"] = mu_.detach().numpy(), sc_.detach().numpy()
Nd, Hd = 300, 40
qd0 = rows32[torch.arange(Nd) * 7 % 4000][:, R.Q_PERM].contiguous()
_, Ud = synth.make_inputs(synth.load_rows(), Nd, Hd, window=0, seed=4)
epsd = synth.make_eps(Nd, Hd, seed=5)
aux["dyn_q0"], aux["dyn_U"], aux["dyn_eps"] = qd0.numpy(), Ud.numpy(), epsd.numpy()
aux["dyn_out"] = R.rollout_dynnn(my_noise, qd0, Ud, epsd).numpy()

# a9 SAMPLE=True branch under a seeded python RNG (eval_models.py:70-75)
import random  # noqa: E402
ev.SAMPLE, ev.EPOCHS = True, 2000
for name, ds_ref, feats, labs in (("Basic_64_4_32", ref_ds, mfn, mln), ("GRU_64_4_4_16_True", ref_rds, rfn, rln)):
    p = name.split("_")
    if p[0] == "Basic":
        ref_m = fwd.Basic_Model(8, int(p[1]), 6, int(p[2]), False).cpu().eval()
        ref_m.load_state_dict(torch.load(f"{REF}/model_data/fwd_models/{name}.pt", map_location="cpu"))
        state_f = False
    else:
        ref_m = rnn.GRU_Model(8, int(p[1]), 6, int(p[2]), True).cpu().eval()
        ref_m.load_state_dict(torch.load(f"{REF}/model_data/rnn_models/{name}.pt", map_location="cpu"))
        state_f = True
    random.seed(5)
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ev.calc_error(ds_ref, ref_m, name, state_f)
    printed = [float(l.split(":")[1]) for l in buf.getvalue().splitlines() if ":" in l]
    random.seed(5)
    mae_s, rmse_s = R.calc_error_sampled(torch.from_numpy(onestep[name]), labs, 2000, random)
    check(f"a9 {name} calc_error SAMPLE=True seed 5 {printed} vs oracle ({mae_s:.4f},{rmse_s:.4f})",
          abs(printed[0] - mae_s) < 1e-4 and abs(printed[1] - rmse_s) < 1e-4)
    aux[name + "_err_sampled"] = np.asarray([mae_s, rmse_s])
np.savez_compressed(f"{GOLD}/aux.npz", **aux)

# ---------------------------------------------------------------------------- 4. golden rollouts
rows = synth.load_rows()
gold = {}


@torch.no_grad()
def ref_loop_mlp(ref_m, st, s0, U, noise=None, seed=None):
    """closed loop through the reference modules themselves (independent of oracle.rollout_*)."""
    s, out = s0, []
    dyn = nm.DynamicsNN(None, noise, False, history=3) if noise is not None else None
    if seed is not None:
        torch.manual_seed(seed)
    for t in range(U.shape[1]):
        x = (torch.cat([s, U[:, t]], -1) - st["f_mean"]) / st["f_std"]
        nxt = ref_m(x) * st["l_std"] + st["l_mean"]
        if noise is not None:   # batch 1 only: the reference's wrapper "does not support parallel"
            res = dyn.get_prediction(s[0, R.Q_PERM].numpy(), U[0, t].numpy())
            nxt = nxt + torch.from_numpy(res)[:, R.Q_INV]
        out.append(nxt)
        s = nxt
    return torch.stack(out, 1)


@torch.no_grad()
def ref_loop_rnn(ref_m, st, win0, U):
    win, out = win0, []
    ref_m.h_state = None
    for t in range(U.shape[1]):
        y = ref_m((win - st["f_mean"]) / st["f_std"])
        nxt = y * st["l_std"] + st["l_mean"]
        out.append(nxt)
        win = torch.cat([win[:, 1:], torch.cat([nxt, U[:, t]], -1)[:, None]], 1)
    return torch.stack(out, 1)


N, H = 64, 24
s0, U = synth.make_inputs(rows, N, H, window=0, seed=0)
win0, Uw = synth.make_inputs(rows, N, H, window=4, seed=0)
gold["s0"], gold["U"], gold["win0"], gold["Uw"] = s0.numpy(), U.numpy(), win0.numpy(), Uw.numpy()
for name in names_fwd:
    p = name.split("_")
    ref_m = fwd.Basic_Model(8, int(p[1]), 6, int(p[2]), False).cpu().eval()
    ref_m.load_state_dict(torch.load(f"{REF}/model_data/fwd_models/{name}.pt", map_location="cpu"))
    tr = ref_loop_mlp(ref_m, mst, s0, U)
    mine, _ = R.build_from_name(name, ckpt(name))
    to, _ = R.rollout_mlp(mine, mst, s0, U)
    check(f"rollout {name} N={N} H={H}: oracle == reference loop (bit)", bit_equal(tr, to))
    gold[name] = tr.numpy()
for name in names_rnn:
    p = name.split("_")
    cls = rnn.GRU_Model if p[0] == "GRU" else rnn.LSTM_Model
    state_f = p[-1] == "True"
    ref_m = cls(8, int(p[1]), 6, int(p[2]), state_f).cpu().eval()
    ref_m.load_state_dict(torch.load(f"{REF}/model_data/rnn_models/{name}.pt", map_location="cpu"))
    tr = ref_loop_rnn(ref_m, rst, win0, Uw)
    mine, _ = R.build_from_name(name, ckpt(name))
    to, _, _ = R.rollout_rnn(mine, rst, win0, Uw, stateful=state_f)
    check(f"rollout {name} N={N} H={H} stateful={state_f}: oracle == reference loop (bit)", bit_equal(tr, to))
    gold[name] = tr.numpy()

# noise-injected MLP rollout, batch 1 (the reference wrapper is batch-1 only), shared seed
ref_m = fwd.Basic_Model(8, 64, 6, 4, False).cpu().eval()
ref_m.load_state_dict(torch.load(f"{REF}/model_data/fwd_models/Basic_64_4_32.pt", map_location="cpu"))
tr = ref_loop_mlp(ref_m, mst, s0[:1], U[:1], noise=ref_noise, seed=7)
torch.manual_seed(7)
eps1 = torch.stack([torch.randn(1, 6) for _ in range(H)], 1)
mine, _ = R.build_from_name("Basic_64_4_32", ckpt("Basic_64_4_32"))
to, _ = R.rollout_mlp(mine, mst, s0[:1], U[:1], noise=my_noise, eps=eps1)
check("rollout Basic_64_4_32 + noise (batch 1, seed 7): oracle == reference loop (bit)", bit_equal(tr, to))
gold["Basic_64_4_32_noise_b1"], gold["eps_b1"] = tr.numpy(), eps1.numpy()

# batched noise + cost goldens come from the (now pinned) oracle
eps = synth.make_eps(N, H, seed=1)
cost = synth.make_cost(rows, H)
gold["eps"] = eps.numpy()
gold["cost_ref"], gold["cost_qw"], gold["cost_rw"] = (c.numpy() for c in cost)
to, co = R.rollout_mlp(mine, mst, s0, U, noise=my_noise, eps=eps, cost=cost)
gold["Basic_64_4_32_noise"], gold["Basic_64_4_32_noise_cost"] = to.numpy(), co.numpy()
g, _ = R.build_from_name("GRU_64_4_4_16_False", ckpt("GRU_64_4_4_16_False"))
to, co, _ = R.rollout_rnn(g, rst, win0, Uw, noise=my_noise, eps=eps, cost=cost)
gold["GRU_64_4_4_16_False_noise"], gold["GRU_64_4_4_16_False_noise_cost"] = to.numpy(), co.numpy()
np.savez_compressed(f"{GOLD}/rollouts.npz", **gold)

# ---------------------------------------------------------------------------- 5. DynamicsDataset (f3)
from collections import OrderedDict  # noqa: E402

from oracle import dataset_ref as D  # noqa: E402

rows_np = np.asarray(bags[0], dtype=np.float64)
rng = np.random.default_rng(0)
Md = 900
ts_ns = (1_700_000_000_000_000_000 + np.cumsum(rng.integers(15_000_000, 26_000_000, Md))).astype(np.int64)   # 15..26 ms apart
raw = OrderedDict()
for k in range(Md):
    r = rows_np[k]
    ua, us_ = float(r[6]), float(r[7])
    if k % 97 == 5:
        ua, us_ = 1.0, 0.0          # "manual control" marker, dropped by from_dict (data_utils.py:53-54)
    if k % 101 == 7:
        ua, us_ = 0.0, 0.0          # idle, dropped
    raw[int(ts_ns[k])] = {"v": {"v_long": float(r[4]), "v_tran": float(r[5])}, "w": {"w_psi": float(r[3])},
                          "x": {"x": float(r[0]), "y": float(r[1])}, "e": {"psi": float(r[2])},
                          "u": {"u_a": ua, "u_steer": us_}}
import utils.data_utils as du  # noqa: E402
dyn_gold = {}
td, qd_, ud_ = D.arrays_from_dict(raw)
dyn_gold["t"], dyn_gold["q"], dyn_gold["u"] = td, qd_, ud_
for dt_ in (0.1, 0.04, 0.02):
    ref_dd = du.DynamicsDataset.from_dict(raw, dt_, history=3)
    q_ds, u_ds = D.resample_ref(td, qd_, ud_, dt_)
    ok = np.array_equal(ref_dd.q, q_ds) and np.array_equal(ref_dd.u, u_ds)
    it_r, it_m = ref_dd[4], D.get_item(q_ds, u_ds, 4, 3)
    ok = ok and all(np.array_equal(a, b) for a, b in zip(it_r, it_m)) and len(ref_dd) == q_ds.shape[0] - 3
    check(f"f3 DynamicsDataset.from_dict dt={dt_}: oracle == reference (bit), {q_ds.shape[0]} q / {u_ds.shape[0]} u rows", ok)
    dyn_gold[f"q_ds_{dt_}"], dyn_gold[f"u_ds_{dt_}"] = ref_dd.q, ref_dd.u
_, qi_, ui_ = D.interp_grid(td, qd_, ud_)
ref_dd = du.DynamicsDataset(qi_, ui_, 0.08, history=2)
q_ds, u_ds, _ = D.subsample(qi_, ui_, 0.08)
check("f3 DynamicsDataset.__init__ dt=0.08 on gridded arrays: oracle == reference (bit)",
      np.array_equal(ref_dd.q, q_ds) and np.array_equal(ref_dd.u, u_ds))
dyn_gold["q_grid"], dyn_gold["u_grid"], dyn_gold["q_ds_init_0.08"], dyn_gold["u_ds_init_0.08"] = qi_, ui_, ref_dd.q, ref_dd.u
np.savez_compressed(f"{GOLD}/dynamics.npz", **dyn_gold)

# ---------------------------------------------------------------------------- 6. second validation pickle + the sweep (f2)
pid_bags = synth.load_bags("pid_data_clean")
pid = {}
ref_pds, ref_prds = fwd.RobotData(pid_bags), rnn.RobotData(pid_bags, 4)
pf, pl_ = R.mlp_dataset(pid_bags)
pst = R.zscore_stats(pf, pl_)
pfn, pln = R.apply_zscore(pf, pl_, pst)
check("a1 pid pickle: MLP dataset+zscore bit-identical", bit_equal(pfn, ref_pds.featrs.cpu()) and bit_equal(pln, ref_pds.labels.cpu()))
prf, prl = R.rnn_dataset(pid_bags, 4)
prst = R.zscore_stats(prf, prl)
prfn, prln = R.apply_zscore(prf, prl, prst)
check("a3 pid pickle: RNN dataset+zscore bit-identical (T=4)", bit_equal(prfn, ref_prds.featrs.cpu()) and bit_equal(prln, ref_prds.labels.cpu()))
for k, v in pst.items():
    pid[f"mlp_{k}"] = v.numpy()
for k, v in prst.items():
    pid[f"rnn4_{k}"] = v.numpy()


def ref_module(name):
    p = name.split("_")
    if p[0] == "Basic":
        m = fwd.Basic_Model(8, int(p[1]), 6, int(p[2]), False).cpu()
        m.load_state_dict(torch.load(f"{REF}/model_data/fwd_models/{name}.pt", map_location="cpu"))
    else:
        m = (rnn.GRU_Model if p[0] == "GRU" else rnn.LSTM_Model)(8, int(p[1]), 6, int(p[2]), False).cpu()   # evaluated stateless (eval_models.py:81-82)
        m.load_state_dict(torch.load(f"{REF}/model_data/rnn_models/{name}.pt", map_location="cpu"))
    return m.eval()


sweep = {}
HS = (1, 10, 50)
for pk, pbags, (sm_, sr_), (fm_, lm_), (fr_, lr_) in (("mpc", bags, (mst, rst), (mfn, mln), (rfn, rln)),
                                                     ("pid", pid_bags, (pst, prst), (pfn, pln), (prfn, prln))):
    rows32_p = torch.as_tensor(pbags[0], dtype=torch.float64).to(torch.float32)
    for name in names_fwd + names_rnn:
        mlp = name.startswith("Basic")
        ref_m = ref_module(name)
        mine, meta = R.build_from_name(name, ckpt(name))
        st_ = sm_ if mlp else sr_
        got = R.sweep_model(mine, meta, pbags, HS)
        ok = True
        with torch.no_grad():
            if mlp:
                y1 = ref_m(fm_)
                lab = lm_
            else:
                ref_m.h_state = None
                y1 = ref_m(fr_)
                lab = lr_
            if pk == "pid":
                pid[name] = y1.numpy()
            e1 = R.calc_error(y1, lab)
            ok = ok and e1 == got[1]
            sweep[f"{pk}/{name}/H1"] = np.asarray(e1)
            for H_ in HS[1:]:
                init_, U_, tgt_ = R.open_loop_inputs(rows32_p, H_, 0 if mlp else 4)
                tr = ref_loop_mlp(ref_m, st_, init_, U_) if mlp else ref_loop_rnn(ref_m, st_, init_, U_)
                mae_c, rmse_c = R.error_curves(tr, tgt_, st_)
                ok = ok and torch.equal(mae_c, got[H_][0]) and torch.equal(rmse_c, got[H_][1])
                sweep[f"{pk}/{name}/H{H_}_mae"], sweep[f"{pk}/{name}/H{H_}_rmse"] = mae_c.numpy(), rmse_c.numpy()
        check(f"f2 sweep {pk} {name}: oracle curves == reference modules in the python loop (bit), H1 mae {e1[0]:.4f}, "
              f"H50 mae[-1] {float(sweep[f'{pk}/{name}/H50_mae'][-1]):.3f}", ok)
np.savez_compressed(f"{GOLD}/pid.npz", **pid)
np.savez_compressed(f"{GOLD}/sweep.npz", **sweep)

# ---------------------------------------------------------------------------- 7. training step (f4)
# The reference's own Basic_Model class, nn.L1Loss and optim.Adam (basic_model.py:87-110) stepped over explicit batches ==
# oracle/train_ref.py on the restated module, bit for bit; the per-batch losses and the parameters after 25 steps of 16 are
# stored as a golden (tests/golden/train.npz) for the GPU kernel's test.
from oracle import train_ref as TR  # noqa: E402

train_gold = {}
for hidden_, layers_, batch_ in ((16, 4, 16), (64, 2, 32)):
    torch.manual_seed(100 + hidden_)
    ref_t = fwd.Basic_Model(8, hidden_, 6, layers_, False).cpu()
    mine_t = R.RefBasic(8, hidden_, 6, layers_)
    mine_t.load_state_dict(ref_t.state_dict())
    init_flat = torch.cat([p_.detach().reshape(-1) for p_ in ref_t.parameters()]).clone()
    order_ = torch.randperm(mfn.shape[0], generator=torch.Generator().manual_seed(5))[:25 * batch_ - 3]      # last batch short
    opt_ref = torch.optim.Adam(ref_t.parameters(), lr=fwd.LEARN_R)
    crit_ref = torch.nn.L1Loss()
    ref_t.train()
    ref_losses = []
    for b0_ in range(0, len(order_), batch_):                       # the body of basic_model.py:102-110 on explicit batches
        idx_ = order_[b0_:b0_ + batch_]
        opt_ref.zero_grad()
        loss_ = crit_ref(ref_t(ref_ds.featrs.cpu()[idx_]), ref_ds.labels.cpu()[idx_])
        loss_.backward()
        opt_ref.step()
        ref_losses.append(loss_.item())
    my_losses, _ = TR.train_steps(mine_t, mfn, mln, order_, batch_, fwd.LEARN_R)
    flat_ref = torch.cat([p_.detach().reshape(-1) for p_ in ref_t.parameters()])
    flat_mine = torch.cat([p_.detach().reshape(-1) for p_ in mine_t.parameters()])
    check(f"f4 training: {len(ref_losses)} Adam steps of Basic_{hidden_}_{layers_} (batch {batch_}): oracle == reference class + optimizer (bit), "
          f"loss {ref_losses[0]:.4f} -> {ref_losses[-1]:.4f}", torch.equal(flat_ref, flat_mine) and torch.equal(torch.tensor(ref_losses), my_losses))
    tag = f"Basic_{hidden_}_{layers_}_{batch_}"
    train_gold[f"{tag}/init"], train_gold[f"{tag}/order"] = init_flat.numpy(), order_.numpy()
    train_gold[f"{tag}/losses"], train_gold[f"{tag}/final"] = np.asarray(ref_losses, dtype=np.float32), flat_ref.numpy()
vl_ref = 0.0
with torch.no_grad():
    ref_t.eval()
    nb_ = 0
    for b0_ in range(0, 800, 32):                                   # basic_model.py:112-122 on the first 800 samples
        vl_ref += crit_ref(ref_t(ref_ds.featrs.cpu()[b0_:b0_ + 32]), ref_ds.labels.cpu()[b0_:b0_ + 32]).item()
        nb_ += 1
check("f4 validation loss: oracle == reference loop (bit)", vl_ref / nb_ == TR.validation_loss(mine_t, mfn[:800], mln[:800], 32))
train_gold["Basic_64_2_32/valid800"] = np.asarray(vl_ref / nb_)
np.savez_compressed(f"{GOLD}/train.npz", **train_gold)

# ---------------------------------------------------------------------------- 7b. training step of the recurrent models (f4)
# The reference's GRU_Model / LSTM_Model classes (deep_models.py:44-92), nn.L1Loss and optim.Adam stepped the way
# deep_models.py:114-127 does (init_hidden before the batch when STATE_F) == oracle/train_ref.train_steps_rnn on the restated
# modules, bit for bit, for state_f False and True; losses + parameters after 20 steps stored in tests/golden/train_rnn.npz.
rf8, rl8 = R.rnn_dataset(bags, 8)
rfn8, rln8 = R.apply_zscore(rf8, rl8, R.zscore_stats(rf8, rl8))
rnn_gold = {}
for kind_, hidden_, layers_, ft_, batch_, state_f_ in (("GRU", 32, 2, 4, 16, False), ("LSTM", 16, 4, 8, 8, True), ("GRU", 16, 1, 4, 16, True)):
    feats_, labs_ = (rfn, rln) if ft_ == 4 else (rfn8, rln8)
    torch.manual_seed(200 + hidden_)
    ref_t = getattr(rnn, f"{kind_}_Model")(8, hidden_, 6, layers_, state_f_).cpu()
    mine_t = (R.RefGRU if kind_ == "GRU" else R.RefLSTM)(8, hidden_, 6, layers_)
    mine_t.load_state_dict(ref_t.state_dict())
    init_flat = torch.cat([p_.detach().reshape(-1) for p_ in ref_t.parameters()]).clone()
    order_ = torch.randperm(feats_.shape[0], generator=torch.Generator().manual_seed(6))[:20 * batch_ - 5]    # last batch short
    opt_ref = torch.optim.Adam(ref_t.parameters(), lr=rnn.LEARN_R)
    crit_ref = torch.nn.L1Loss()
    ref_t.train()
    ref_losses = []
    for b0_ in range(0, len(order_), batch_):                       # the body of deep_models.py:115-127 on explicit batches
        idx_ = order_[b0_:b0_ + batch_]
        if state_f_:
            ref_t.init_hidden(len(idx_))
        opt_ref.zero_grad()
        loss_ = crit_ref(ref_t(feats_[idx_]), labs_[idx_])
        loss_.backward()
        opt_ref.step()
        ref_losses.append(loss_.item())
    my_losses, _ = TR.train_steps_rnn(mine_t, feats_, labs_, order_, batch_, rnn.LEARN_R)
    flat_ref = torch.cat([p_.detach().reshape(-1) for p_ in ref_t.parameters()])
    flat_mine = torch.cat([p_.detach().reshape(-1) for p_ in mine_t.parameters()])
    check(f"f4 training: {len(ref_losses)} Adam steps of {kind_}_{hidden_}_{layers_}_{ft_} (batch {batch_}, state_f {state_f_}): oracle == reference "
          f"class + optimizer (bit), loss {ref_losses[0]:.4f} -> {ref_losses[-1]:.4f}",
          torch.equal(flat_ref, flat_mine) and torch.equal(torch.tensor(ref_losses), my_losses))
    if kind_ == "GRU" and hidden_ == 16:
        continue                                                    # the stateful GRU is a check only
    tag = f"{kind_}_{hidden_}_{layers_}_{ft_}_{batch_}"
    rnn_gold[f"{tag}/init"], rnn_gold[f"{tag}/order"] = init_flat.numpy(), order_.numpy()
    rnn_gold[f"{tag}/losses"], rnn_gold[f"{tag}/final"] = np.asarray(ref_losses, dtype=np.float32), flat_ref.numpy()
vl_ref = 0.0
with torch.no_grad():
    ref_t = rnn.LSTM_Model(8, 16, 6, 4, True).cpu()
    torch.nn.utils.vector_to_parameters(torch.from_numpy(rnn_gold["LSTM_16_4_8_8/final"]).clone(), ref_t.parameters())
    mine_t = R.RefLSTM(8, 16, 6, 4)
    mine_t.load_state_dict(ref_t.state_dict())
    ref_t.eval()
    nb_ = 0
    for b0_ in range(0, 400, 8):                                    # deep_models.py:130-146 on the first 400 windows
        ref_t.init_hidden(len(rfn8[b0_:b0_ + 8]))
        vl_ref += crit_ref(ref_t(rfn8[b0_:b0_ + 8]), rln8[b0_:b0_ + 8]).item()
        nb_ += 1
check("f4 validation loss of a recurrent model: oracle == reference loop (bit)", vl_ref / nb_ == TR.validation_loss_rnn(mine_t, rfn8[:400], rln8[:400], 8))
rnn_gold["LSTM_16_4_8_8/valid400"] = np.asarray(vl_ref / nb_)
np.savez_compressed(f"{GOLD}/train_rnn.npz", **rnn_gold)

# ---------------------------------------------------------------------------- 8. rosbag2 CDR decode (f4) -- fixture + consistency
# The deserialiser of the reference (rclpy + mpclab_msgs) is absent, so nothing here is a parity pin (oracle/cdr_ref.py
# says so); this section exports one shipped bag as a fixture and checks what CAN be checked against the reference's own
# files: message sizes, header.stamp == t, the 20 ms period, metadata.yaml's message count, setup_data's cut.
import sqlite3  # noqa: E402
import yaml  # noqa: E402
from oracle import cdr_ref as CDR  # noqa: E402

bag_dir = f"{REF}/data/rosbags/DAgger_827"
con = sqlite3.connect(f"file:{bag_dir}/DAgger_827_GPS_0.db3?mode=ro", uri=True)
raw = con.execute("SELECT TIMESTAMP, DATA FROM MESSAGES ORDER BY ID").fetchall()
meta_y = yaml.safe_load(open(f"{bag_dir}/metadata.yaml"))["rosbag2_bagfile_information"]
check("f4 cdr: message count == metadata.yaml, type VehicleStateMsg, serialization cdr",
      len(raw) == meta_y["message_count"] and meta_y["topics_with_message_count"][0]["topic_metadata"]["type"] == "mpclab_msgs/msg/VehicleStateMsg"
      and meta_y["topics_with_message_count"][0]["topic_metadata"]["serialization_format"] == "cdr")
msgs_d = [CDR.decode(d) for _, d in raw]
ts = np.array([m_["t"] for m_ in msgs_d])
stamp = np.array([m_["header"]["stamp"]["sec"] + 1e-9 * m_["header"]["stamp"]["nanosec"] for m_ in msgs_d])
check("f4 cdr: every message 468 bytes, t == header.stamp to 1 us, period 20 ms (median within 1 ms, jitter < 15 ms), recorder timestamps follow t",
      all(len(d) == 468 for _, d in raw) and np.abs(ts - stamp).max() < 1e-6 and abs(np.median(np.diff(ts)) - 0.02) < 1e-3 and np.abs(np.diff(ts) - 0.02).max() < 15e-3
      and np.abs(np.array([t_ for t_, _ in raw]) * 1e-9 - ts).max() < 0.1)
rows9 = np.array([CDR.row9(m_) for m_ in msgs_d])
check("f4 cdr: decoded fields inside the ranges of the validation pickles (BARC track, speeds, controls)",
      np.abs(rows9[:, 1:3]).max() < 10 and np.abs(rows9[:, 5]).max() < 6 and np.abs(rows9[:, 7]).max() <= 2.0 and np.abs(rows9[:, 8]).max() <= 0.5)
vals = [float(i) + 0.25 for i in range(34)]
enc = CDR.encode(vals, ("", "", ""))
check("f4 cdr: the independent CDR writer with empty frame_ids reproduces the shipped layout (468 bytes, same offsets)",
      len(enc) == 468 and CDR.row9(CDR.decode(enc)) == [vals[i] for i in CDR.ROW_OF_VALS])
clean = CDR.setup_data({"DAgger_827": dict(enumerate(msgs_d))})
check("f4 setup_data restatement: sorted by t, last second dropped", len(clean[0]) == int((ts <= ts.max() - 1.0).sum()) and clean[0][0] == rows9[np.argsort(ts, kind="stable")[0], 1:].tolist())
np.savez_compressed(f"{GOLD}/cdr_bag.npz", blobs=np.frombuffer(b"".join(d for _, d in raw), dtype=np.uint8).reshape(len(raw), 468),
                    timestamps=np.array([t_ for t_, _ in raw], dtype=np.int64), rows9=rows9, clean=np.asarray(clean[0]))

bad = [n for n, ok in checks if not ok]
print(f"{len(checks) - len(bad)}/{len(checks)} checks passed")
print("PINNED" if not bad else "NOT PINNED: " + "; ".join(bad))
sys.exit(0 if not bad else 1)
