"""Drop-in for the trajectory test of scripts/test_model.py: ``run_dynamics_model(model, dataset, open_loop)``.

``run_dynamics_model`` keeps the reference's loop (test_model.py:67-87) for any model with
``get_nominal_prediction`` / ``get_prediction`` -- one kernel launch per step at batch 1, as the reference does.
``run_dynamics_model_fused`` is the same 'with noise' trajectory for an open-loop run in ONE launch (the whole
horizon fused), and batched over many start states / control sequences when asked.  Building the CasADi bicycle,
the track and the resampled ``DynamicsDataset`` (prepare_components / prepare_dataset, :17-64) needs
``mpclab_common`` and raw-dict pickles that are not in the tree: out of scope."""
from __future__ import annotations

import numpy as np
import torch

from .models import CasadiDynamicCLBicycleNoise, DynamicsNN, FeedforwardNoiseModel  # noqa: F401  (test_model.py:10)


def run_dynamics_model(model, dynamics_dataset, open_loop: bool, with_nominal=True):
    """test_model.py:67-87.  `dynamics_dataset` yields (q [hist,6], u [hist,2], _)."""
    q_data, q_nominal, q_noised = [], [], []
    labels = ["data", "nominal", "with noise"]
    for step, (q, u, _) in enumerate(dynamics_dataset):
        try:
            q_data.append(q[-1])
            if with_nominal:
                if len(q_nominal) == 0:
                    q_nominal.append(q[-1])
                else:
                    q_nominal.append(model.get_nominal_prediction(q_nominal[-1] if open_loop else q[-1], u[-1]))
            if len(q_noised) == 0:
                q_noised.append(q[-1])
            else:
                q_noised.append(model.get_prediction(q_noised[-1] if open_loop else q[-1], u[-1]).squeeze(0))
        except ValueError as e:
            print(f"'{e}' at step {step}.")
            break
    return (q_data, q_nominal, q_noised), labels


def run_dynamics_model_fused(model, q_data, u_data, eps=None, seed=0):
    """Open-loop 'with noise' trajectory of run_dynamics_model in one launch.

    q_data [T,6] / u_data [T,2] (one recorded run) or [N,T,6] / [N,T,2] (a batch of runs); returns the same
    shape as q_data: row 0 is the data's first state, row k >= 1 the prediction fed back k times, each step
    using the CURRENT sample's control u_data[k] (test_model.py:83).  eps [.., T-1, 6] or None -> Philox(seed)."""
    single = q_data.ndim == 2
    dev = next(model.noise_model.model.parameters()).device
    as_t = lambda a: None if a is None else torch.as_tensor(np.asarray(a), dtype=torch.float32)
    q, u, e = as_t(q_data), as_t(u_data), as_t(eps)
    if single:
        q, u, e = q[None], u[None], None if e is None else e[None]
    q0, U = q[:, 0].contiguous().to(dev), u[:, 1:].contiguous().to(dev)
    e = None if e is None else e.contiguous().to(dev)
    pred = model.rollout(q0, U, eps=e, seed=seed)     # DynamicsNN or CasadiDynamicCLBicycleNoise (learned nominal)
    out = torch.cat([q[:, :1].to(dev), pred], 1)
    return out[0] if single else out
