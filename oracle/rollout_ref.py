"""CPU restatement of the reference's learned-dynamics path (TEST INFRASTRUCTURE, see __init__).

Every function cites the reference file:line it follows (paths relative to /root/reference).
Arithmetic is delegated to the same third-party library the reference calls (PyTorch
``nn.Linear`` / ``nn.GRU`` / ``nn.LSTM`` / ``softplus``; unpinned in the reference, 2.11.0 here),
so that on identical weights and inputs the restatement is bit-identical to the reference
modules -- ``oracle/pin_reference.py`` checks exactly that.

Conventions (DESIGN.md "Rollout semantics"):
  row    = [x, y, psi, w_psi, v_long, v_tran, u_a, u_steer]      scripts/setup_data.py:70-80
  state  = row[:6], control = row[6:8]
  q      = [v_long, v_tran, w_psi, x, y, psi] = state[Q_PERM]    model_scripts/noise_model.py:108
"""
from __future__ import annotations

import copy
from collections import OrderedDict

import torch
import torch.nn as nn
import torch.nn.functional as F

STATE, CTRL, FRAME = 6, 2, 8
Q_PERM = [4, 5, 3, 0, 1, 2]   # q[i] = state[Q_PERM[i]]
Q_INV = [3, 4, 5, 2, 0, 1]    # state[j] = q[Q_INV[j]]


# --------------------------------------------------------------------------------------
# a1 / a3  dataset build + z-score statistics
# --------------------------------------------------------------------------------------
def mlp_dataset(bags):
    """features = bag[i], labels = bag[i+1][:6]      model_scripts/basic_model.py:13-26"""
    feats, labs = [], []
    for bag in bags:
        for i in range(len(bag) - 1):
            feats.append(bag[i])
            labs.append(bag[i + 1][:6])
    return (torch.tensor(feats, dtype=torch.float32),
            torch.tensor(labs, dtype=torch.float32))


def rnn_dataset(bags, size):
    """windows bag[i:i+size], label bag[i+size][:6]   model_scripts/deep_models.py:13-27"""
    feats, labs = [], []
    for bag in bags:
        for i in range(len(bag) - size - 1):
            feats.append(bag[i:i + size])
            labs.append(bag[i + size][:6])
    return (torch.tensor(feats, dtype=torch.float32),
            torch.tensor(labs, dtype=torch.float32))


def zscore_stats(feats, labs):
    """mean / UNBIASED std over dim 0              basic_model.py:28-32, deep_models.py:29-33

    Returns dict(f_mean, f_std, l_mean, l_std); for the RNN dataset f_* are [T, 8] (one set
    of statistics per window position)."""
    return dict(f_mean=torch.mean(feats, dim=0), f_std=torch.std(feats, dim=0),
                l_mean=torch.mean(labs, dim=0), l_std=torch.std(labs, dim=0))


def apply_zscore(feats, labs, st):
    """subtract THEN divide                        basic_model.py:34-35, deep_models.py:35-36"""
    return (feats - st["f_mean"]) / st["f_std"], (labs - st["l_mean"]) / st["l_std"]


# --------------------------------------------------------------------------------------
# a2 / a4 / a5 / a6  the modules (same parameter names -> the shipped state_dicts load)
# --------------------------------------------------------------------------------------
class RefBasic(nn.Module):
    """ReLU MLP inputs->hidden->(hidden->hidden)*layers->output.  basic_model.py:43-61"""

    def __init__(self, inputs, hidden, output, layers, do_drop=False):
        super().__init__()
        mods = [nn.Linear(inputs, hidden)]
        for _ in range(layers):
            mods.append(nn.Linear(hidden, hidden))
            if do_drop:
                mods.append(nn.Dropout(0.1))      # shifts the Linear indices, basic_model.py:53-54
        mods.append(nn.Linear(hidden, output))
        self.layers = nn.ModuleList(mods)

    def forward(self, x):
        for m in self.layers[:-1]:
            x = torch.relu(m(x))                  # relu also wraps Dropout (identity in eval)
        return self.layers[-1](x)


class RefGRU(nn.Module):
    """nn.GRU(batch_first) + Linear on the last frame.          deep_models.py:44-66"""

    def __init__(self, inputs, hidden, output, layers):
        super().__init__()
        self.hidden, self.nlayers = hidden, layers
        self.gru = nn.GRU(inputs, hidden, layers, batch_first=True)
        self.full_c = nn.Linear(hidden, output)

    def zero_state(self, batch, dtype=torch.float32):
        return torch.zeros(self.nlayers, batch, self.hidden, dtype=dtype)   # deep_models.py:57-58

    def forward(self, x, h):
        out, h = self.gru(x, h)
        return self.full_c(out[:, -1, :]), h


class RefLSTM(nn.Module):
    """nn.LSTM(batch_first) + Linear on the last frame.         deep_models.py:68-92"""

    def __init__(self, inputs, hidden, output, layers):
        super().__init__()
        self.hidden, self.nlayers = hidden, layers
        self.lstm = nn.LSTM(inputs, hidden, layers, batch_first=True)
        self.full_c = nn.Linear(hidden, output)

    def zero_state(self, batch, dtype=torch.float32):
        z = torch.zeros(self.nlayers, batch, self.hidden, dtype=dtype)      # deep_models.py:81-83
        return (z, z.clone())

    def forward(self, x, hc):
        out, hc = self.lstm(x, hc)
        return self.full_c(out[:, -1, :]), hc


_ACT = {"relu": nn.ReLU, "tanh": nn.Tanh, "leaky_relu": nn.LeakyReLU, "sigmoid": nn.Sigmoid,
        "selu": nn.SELU, "softplus": nn.Softplus, "identity": nn.Identity}


class RefNoise(nn.Module):
    """n_layers*[Linear, act] + Linear (+out act) over concat(q_hist, u_hist) -> (mu, p);
    sample = mu + softplus(p) * eps.       noise_model.py:19-40, utils/pytorch_utils.py:19-57

    ``model`` is an ``nn.Sequential`` with the same indices as ``build_mlp`` so a reference
    ``noise_model.pkl`` (a state_dict of ``self.model``, noise_model.py:47-52) would load."""

    def __init__(self, state_size=6, action_size=2, history=3, n_layers=2, size=32,
                 activation="tanh", output_activation="identity"):
        super().__init__()
        self.state_size, self.action_size, self.history = state_size, action_size, history
        seq, d = [], (state_size + action_size) * history
        for _ in range(n_layers):
            seq += [nn.Linear(d, size), _ACT[activation]()]
            d = size
        seq += [nn.Linear(d, 2 * state_size), _ACT[output_activation]()]
        self.model = nn.Sequential(*seq)

    def mu_scale(self, q, u):
        z = self.model(torch.concat([q, u], dim=-1))
        mu, p = z.split(self.state_size, dim=1)
        return mu, F.softplus(p)                 # beta=1, threshold=20; noise_model.py:39

    def forward(self, q, u, eps):
        mu, scale = self.mu_scale(q, u)
        return mu + eps * scale                  # Normal(mu, scale).rsample(), noise_model.py:39-40


def load_checkpoint_into(model, sd):
    model.load_state_dict(OrderedDict((k, torch.as_tensor(v)) for k, v in sd.items()))
    return model.eval()


def build_from_name(name, sd=None):
    """File-name schema is the config.  eval_models.py:19-25, 43-52
    Basic_{h}_{L}_{batch} | {GRU,LSTM}_{h}_{L}_{ft}_{batch}_{state_f}"""
    p = name.replace(".pt", "").replace(".npz", "").split("_")
    kind, hidden, layers = p[0], int(p[1]), int(p[2])
    if kind == "Basic":
        m, meta = RefBasic(FRAME, hidden, STATE, layers), dict(kind="mlp")
    else:
        cls = RefGRU if kind == "GRU" else RefLSTM
        m = cls(FRAME, hidden, STATE, layers)
        meta = dict(kind=kind.lower(), ft=int(p[3]), state_f=(p[-1] == "True"))
    meta.update(hidden=hidden, layers=layers)
    if sd is not None:
        load_checkpoint_into(m, sd)
    return m.eval(), meta


# --------------------------------------------------------------------------------------
# noise history (a7)  noise_model.py:86-103 / 114-125
# --------------------------------------------------------------------------------------
class _NoiseHistory:
    """Batched deque(maxlen=history) of (q, u); the first append left-pads by repetition."""

    def __init__(self, history):
        self.h, self.q, self.u = history, [], []

    def push(self, q, u):
        self.q.append(q)
        self.u.append(u)
        while len(self.q) < self.h:               # noise_model.py:97-99
            self.q.append(q)
            self.u.append(u)
        self.q, self.u = self.q[-self.h:], self.u[-self.h:]
        return torch.cat(self.q, dim=-1), torch.cat(self.u, dim=-1)   # oldest first, :101-102


def _noise_term(noise, hist, s, u, eps_t):
    """residual in A.1 state order for current (state s, control u)."""
    qh, uh = hist.push(s[:, Q_PERM], u)
    return noise(qh, uh, eps_t)[:, Q_INV]


# --------------------------------------------------------------------------------------
# cost (new; MPPI-style quadratic tracking cost, DESIGN.md "Cost")
# --------------------------------------------------------------------------------------
def step_cost(s_next, u, ref_t, qw, rw):
    d = s_next - ref_t
    return (d * d * qw).sum(-1) + (u * u * rw).sum(-1)


# --------------------------------------------------------------------------------------
# the closed loop (a8 generalised): reference modules inside a python ``for t in range(H)``
# --------------------------------------------------------------------------------------
@torch.no_grad()
def rollout_mlp(model, st, s0, U, noise=None, eps=None, cost=None):
    """x=([s_t,u_t]-f_mean)/f_std ; y=model(x) ; s_{t+1}=y*l_std+l_mean (+noise)   App. A.2

    s0 [N,6], U [N,H,2] -> states [N,H,6] (states[:,t] = s_{t+1}), costs [N] or None.
    cost = (ref [H,6], qw [6], rw [2])."""
    dt = s0.dtype
    f_mean, f_std, l_mean, l_std = (st[k].to(dt) for k in ("f_mean", "f_std", "l_mean", "l_std"))
    hist = _NoiseHistory(noise.history) if noise is not None else None
    s, out = s0, []
    costs = torch.zeros(s0.shape[0], dtype=dt) if cost is not None else None
    for t in range(U.shape[1]):
        u = U[:, t]
        x = (torch.cat([s, u], dim=-1) - f_mean) / f_std
        nxt = model(x) * l_std + l_mean
        if noise is not None:
            nxt = nxt + _noise_term(noise, hist, s, u, eps[:, t])       # noise_model.py:103
        if cost is not None:
            costs = costs + step_cost(nxt, u, cost[0][t].to(dt), cost[1].to(dt), cost[2].to(dt))
        out.append(nxt)
        s = nxt
    return torch.stack(out, dim=1), costs


@torch.no_grad()
def rollout_rnn(model, st, win0, U, stateful=False, state0=None, noise=None, eps=None, cost=None):
    """X=(Win_t-f_mean)/f_std [N,T,8]; y,state=model(X,state); s_{t+1}=y*l_std+l_mean (+noise);
    Win_{t+1}=concat(Win_t[:,1:], [s_{t+1}, U[:,t]]).                          App. A.4

    stateless: state is zero at every step (deep_models.py:61-62); stateful: the state left by
    step t is the initial state of step t+1 (deep_models.py:65), starting from ``state0`` or 0.
    The noise model sees (q,u) of the newest window frame (App. A.6).
    Returns states [N,H,6], costs [N]|None, final state."""
    dt = win0.dtype
    f_mean, f_std, l_mean, l_std = (st[k].to(dt) for k in ("f_mean", "f_std", "l_mean", "l_std"))
    N = win0.shape[0]
    hist = _NoiseHistory(noise.history) if noise is not None else None
    win, out = win0, []
    state = state0 if state0 is not None else model.zero_state(N, dt)
    costs = torch.zeros(N, dtype=dt) if cost is not None else None
    for t in range(U.shape[1]):
        u = U[:, t]
        if not stateful:
            state = model.zero_state(N, dt)
        y, state = model((win - f_mean) / f_std, state)
        nxt = y * l_std + l_mean
        if noise is not None:
            nxt = nxt + _noise_term(noise, hist, win[:, -1, :STATE], win[:, -1, STATE:], eps[:, t])
        if cost is not None:
            costs = costs + step_cost(nxt, u, cost[0][t].to(dt), cost[1].to(dt), cost[2].to(dt))
        out.append(nxt)
        win = torch.cat([win[:, 1:], torch.cat([nxt, u], dim=-1).unsqueeze(1)], dim=1)
    return torch.stack(out, dim=1), costs, state


def as_double(model):
    """FP64 twin of a module (the reference's init_hidden hard-codes FP32, deep_models.py:58,
    so the FP64 loop calls the submodules with explicit states)."""
    return copy.deepcopy(model).double().eval()


# --------------------------------------------------------------------------------------
# a9  calc_error, batched restatement              eval_models.py:67-96
# --------------------------------------------------------------------------------------
@torch.no_grad()
def calc_error(outputs, labels):
    """per-sample mean|o-l| and sqrt(mean (o-l)^2), averaged over samples (eval_models.py:91-96)."""
    d = outputs - labels
    return d.abs().mean(dim=1).mean().item(), d.pow(2).mean(dim=1).sqrt().mean().item()


def calc_error_sampled(outputs, labels, epochs, rng):
    """SAMPLE=True branch (eval_models.py:70-75): `epochs` draws of random.randint(0, len-1) from `rng`
    (a random.Random or the random module), sums accumulated in Python floats in draw order (:91-96).
    outputs / labels: the model output and label of EVERY dataset sample [B,6]."""
    d = outputs - labels
    mae, rmse = d.abs().mean(dim=1), d.pow(2).mean(dim=1).sqrt()
    diff = rm = 0.0
    for _ in range(epochs):
        i = rng.randint(0, len(outputs) - 1)
        diff += mae[i].item()
        rm += rmse[i].item()
    return diff / epochs, rm / epochs


# --------------------------------------------------------------------------------------
# f2  validation sweep: teacher-forced one-step + open-loop error curves    eval_models.py:14-96
# --------------------------------------------------------------------------------------
def open_loop_inputs(rows32, H, window):
    """Start states / windows, recorded controls and recorded states of every stretch of ONE bag that leaves H
    recorded frames ahead (DESIGN.md "Rollout semantics"):
      MLP (window 0): s0 = row i [:6], U[i,t] = row i+t [6:8], target[i,t] = row i+t+1 [:6]          n = len - H
      RNN (window T): win0 = rows i..i+T-1, U[i,t] = row i+T+t [6:8], target[i,t] = row i+T+t [:6]    n = len - T - H + 1"""
    L = rows32.shape[0]
    if window == 0:
        n = L - H
        i = torch.arange(n)[:, None] + torch.arange(H)[None, :]
        return rows32[:n, :6].contiguous(), rows32[i][:, :, 6:8].contiguous(), rows32[i + 1][:, :, :6].contiguous()
    n = L - window - H + 1
    w = torch.arange(n)[:, None] + torch.arange(window)[None, :]
    i = torch.arange(n)[:, None] + window + torch.arange(H)[None, :]
    return rows32[w].contiguous(), rows32[i][:, :, 6:8].contiguous(), rows32[i][:, :, :6].contiguous()


@torch.no_grad()
def error_curves(pred, target, st):
    """calc_error's per-sample metrics (eval_models.py:91-92) on label-normalised values, per horizon step, averaged
    over the trajectories: pred / target [n,H,6] raw states -> (mae [H], rmse [H])."""
    d = (pred - st["l_mean"]) / st["l_std"] - (target - st["l_mean"]) / st["l_std"]
    return d.abs().mean(dim=-1).double().mean(dim=0), d.pow(2).mean(dim=-1).sqrt().double().mean(dim=0)


@torch.no_grad()
def sweep_model(model, meta, bags, horizons=(1, 10, 50)):
    """One checkpoint over one validation pickle: the statistics are the pickle's own (RobotData.norm); H = 1 is the
    teacher-forced calc_error over EVERY dataset sample (eval_models.py SAMPLE=False), H > 1 the open-loop curves.
    -> {1: (mae, rmse), H: (mae [H], rmse [H])}"""
    mlp = meta["kind"] == "mlp"
    f, l = mlp_dataset(bags) if mlp else rnn_dataset(bags, meta["ft"])
    st = zscore_stats(f, l)
    fn, ln = apply_zscore(f, l, st)
    out = {}
    for H in horizons:
        if H == 1:
            y = model(fn) if mlp else model(fn, model.zero_state(len(fn)))[0]
            out[1] = calc_error(y, ln)
            continue
        rows32 = torch.as_tensor(bags[0], dtype=torch.float64).to(torch.float32)
        init, U, target = open_loop_inputs(rows32, H, 0 if mlp else meta["ft"])
        pred = rollout_mlp(model, st, init, U)[0] if mlp else rollout_rnn(model, st, init, U, stateful=False)[0]
        out[H] = error_curves(pred, target, st)
    return out


# --------------------------------------------------------------------------------------
# a7 / a8  DynamicsNN closed loop            noise_model.py:113-125, scripts/test_model.py:67-87
# --------------------------------------------------------------------------------------
@torch.no_grad()
def rollout_dynnn(noise, q0, U, eps):
    """q_{t+1} = noise(history of q, history of u): DynamicsNN.get_prediction fed back on itself (the
    open_loop=True branch of run_dynamics_model with the noise net as the whole dynamics), batched.
    q0 [N,6] in q order, U [N,H,2] (U[:,t] = the control passed with q_t), eps [N,H,6] -> [N,H,6]."""
    hist, q, out = _NoiseHistory(noise.history), q0, []
    for t in range(U.shape[1]):
        qh, uh = hist.push(q, U[:, t])
        q = noise(qh, uh, eps[:, t])
        out.append(q)
    return torch.stack(out, 1)


@torch.no_grad()
def run_dynamics_model(noise, q_data, u_data, eps, open_loop):
    """The `with noise` trajectory of run_dynamics_model (test_model.py:67-87) for a DynamicsNN model and a
    dataset of history 1: q_noised[0] = q_data[0]; q_noised[k] = get_prediction(q_noised[k-1] if open_loop else
    q_data[k], u_data[k]) for k >= 1 (note the control index: the CURRENT sample's u, :83).
    q_data [T,6] (q order), u_data [T,2], eps [T-1,6] -> [T,6]."""
    hist, out = _NoiseHistory(noise.history), [q_data[0]]
    for k in range(1, len(q_data)):
        q = out[-1] if open_loop else q_data[k]
        qh, uh = hist.push(q[None], u_data[k][None])
        out.append(noise(qh, uh, eps[k - 1][None])[0])
    return torch.stack(out)


# --------------------------------------------------------------------------------------
# error metric used by every tolerance in tests/ and DESIGN.md
# --------------------------------------------------------------------------------------
def rel_err(a, b):
    """max_d |a-b| / max_{n,t} |b_d|  per trajectory-step -> [N,H] (b is the reference)."""
    a, b = a.double(), b.double()
    scale = b.abs().amax(dim=tuple(range(b.dim() - 1))).clamp_min(1e-30)
    return ((a - b).abs() / scale).amax(dim=-1)
