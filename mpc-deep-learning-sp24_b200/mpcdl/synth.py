"""Seeded synthetic rollout inputs of the BARC rosbag shape.

Inputs are cut from a recorded bag (rows of [x, y, psi, w_psi, v_long, v_tran, u_a, u_steer],
20 ms sample period, column order of the reference's scripts/setup_data.py:70-80): initial states /
windows are real rows, controls are the recorded controls plus 0.25 sigma Gaussian exploration noise
clamped to the recorded range -- what an MPPI sampler would hand to the rollout.
Everything is generated on the CPU with an explicit generator, so tests, bench and the oracle see
identical tensors on every machine.
"""
from __future__ import annotations

import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURES = os.path.normpath(os.path.join(_HERE, "..", "..", "tests", "golden"))


def load_rows(name="mpc_data_clean"):
    """Recorded validation bag as float64 [rows, 8] (fixture made by oracle/pin_reference.py)."""
    return np.load(os.path.join(FIXTURES, "data", f"{name}.npy"))


def load_bags(name="mpc_data_clean"):
    """Same data in the reference's pickle shape: list[bag][row][8 python floats]."""
    return [load_rows(name).tolist()]


def make_inputs(rows, n, horizon, window=0, seed=0, explore=0.25):
    """-> (init, U): init = s0 [n,6] if window == 0 else win0 [n,window,8]; U [n,horizon,2] (FP32)."""
    rows_t = torch.as_tensor(np.asarray(rows), dtype=torch.float32)
    nrows = rows_t.shape[0]
    g = torch.Generator().manual_seed(seed)
    span = max(window, 1)
    i0 = torch.randint(0, max(nrows - span - 1, 1), (n,), generator=g)
    if window == 0:
        init = rows_t[i0, :6].contiguous()
        first_u = i0
    else:
        idx = i0[:, None] + torch.arange(window)[None, :]
        init = rows_t[idx].contiguous()
        first_u = i0 + window
    u_rows = rows_t[:, 6:8]
    u_std, u_lo, u_hi = u_rows.std(dim=0), u_rows.amin(dim=0), u_rows.amax(dim=0)
    t_idx = (first_u[:, None] + torch.arange(horizon)[None, :]) % nrows
    U = u_rows[t_idx] + explore * u_std * torch.randn(n, horizon, 2, generator=g)
    U = torch.minimum(torch.maximum(U, u_lo), u_hi).contiguous()
    return init, U


def make_eps(n, horizon, seed=1):
    return torch.randn(n, horizon, 6, generator=torch.Generator().manual_seed(seed))


def make_cost(rows, horizon, seed=2):
    """Reference path [H,6] (a recorded stretch of the bag) and diagonal weights qw [6], rw [2]."""
    rows_t = torch.as_tensor(np.asarray(rows), dtype=torch.float32)
    g = torch.Generator().manual_seed(seed)
    i0 = int(torch.randint(0, rows_t.shape[0], (1,), generator=g))
    idx = (i0 + torch.arange(horizon)) % rows_t.shape[0]
    ref = rows_t[idx, :6].contiguous()
    qw = torch.tensor([1.0, 1.0, 0.5, 0.1, 0.25, 0.25])
    rw = torch.tensor([0.05, 0.1])
    return ref, qw, rw
