"""CPU restatement of the reference's resampling dataset (TEST INFRASTRUCTURE, see oracle/__init__).

    DynamicsDataset.__init__      utils/data_utils.py:21-37   q[::step], u mean-pooled over `step` samples
    DynamicsDataset.from_dict     utils/data_utils.py:39-75   skip idle controls, np.interp onto the 20 ms grid
    DynamicsDataset.__getitem__   utils/data_utils.py:94-96   (q[i:i+h], u[i:i+h], q[i+h])

The arithmetic is numpy's (float64 ``np.arange`` / ``np.interp`` / ``np.mean``), exactly the calls the reference
makes; ``oracle/pin_reference.py`` checks bit-identity against the reference class itself and writes
``tests/golden/dynamics.npz``.
"""
from __future__ import annotations

import numpy as np


def arrays_from_dict(data):
    """The dict walk of from_dict (utils/data_utils.py:44-58): nested per-message dicts keyed by the time stamp in ns
    -> t [M] seconds, q [M,6] = [v_long, v_tran, w_psi, x, y, psi], u [M,2]; messages whose control is (1, 0) or 0 are
    dropped (the car is idle / under manual control)."""
    qs, us, ts = [], [], []
    for t, state in data.items():
        q = np.array([state['v']['v_long'], state['v']['v_tran'], state['w']['w_psi'],
                      state['x']['x'], state['x']['y'], state['e']['psi']])
        u = np.array([state['u']['u_a'], state['u']['u_steer']])
        if np.allclose(u, np.array([1, 0])) or np.allclose(u, 0):
            continue
        qs.append(q)
        us.append(u)
        ts.append(t)
    return np.asarray(ts) / 1e9, np.asarray(qs), np.asarray(us)


def interp_grid(t, q, u, dataset_dt=0.02):
    """utils/data_utils.py:61-72: every channel interpolated onto np.arange(t[0], t[-1], dataset_dt)."""
    t_interp = np.arange(t[0], t[-1], dataset_dt)
    q_i = np.empty((t_interp.shape[0], q.shape[1]))
    for i in range(q.shape[1]):
        q_i[:, i] = np.interp(t_interp, t, q[:, i])
    u_i = np.empty((t_interp.shape[0], u.shape[1]))
    for i in range(u.shape[1]):
        u_i[:, i] = np.interp(t_interp, t, u[:, i])
    return t_interp, q_i, u_i


def subsample(q, u, dt, dataset_dt=0.02):
    """utils/data_utils.py:21-34: step = int(dt // dataset_dt); q[::step]; u truncated to a multiple of step and averaged."""
    step = int(dt // dataset_dt)
    assert np.isclose(step * dataset_dt, dt), "dt must be a multiplier of DATASET_DT. "
    qs = q[::step]
    u = u[:(u.shape[0] - (u.shape[0] % step))]
    return qs, np.mean(u.reshape(-1, step, u.shape[1]), axis=1), step


def resample_ref(t, q, u, dt, dataset_dt=0.02):
    """from_dict's numeric part followed by __init__: -> (q_ds, u_ds)."""
    _, q_i, u_i = interp_grid(t, q, u, dataset_dt)
    qs, us, _ = subsample(q_i, u_i, dt, dataset_dt)
    return qs, us


def get_item(q_ds, u_ds, i, history):
    """utils/data_utils.py:94-96"""
    return q_ds[i:i + history], u_ds[i:i + history], q_ds[i + history]
