"""Drop-in for the reference's resampling dataset, with the arithmetic on the GPU.

    DynamicsDataset(q, u, dt, DATASET_DT=0.02, history=1)      utils/data_utils.py:19-37
    DynamicsDataset.from_dict(data, dt, DATASET_DT, history)   utils/data_utils.py:39-75
    DynamicsDataset.from_pickle(file_name, dt, ...)            utils/data_utils.py:77-89
    len(ds), ds[i] -> (q[i:i+h], u[i:i+h], q[i+h])             utils/data_utils.py:91-96

What stays on the host is the walk over the nested per-message dicts (Python objects); the interpolation onto the
20 ms grid, the sub-sampling of q and the mean-pooling of u run in ONE launch of ``resample_kernel``
(csrc/dataset_kernels.cu) in float64, bit-identical to numpy's ``np.arange`` / ``np.interp`` / ``np.mean``.
``q`` / ``u`` are float64 CUDA tensors; items are views of them.  Plotting (``plot=True``) has no display stack here.
"""
from __future__ import annotations

import math
import pickle as pkl

import numpy as np
import torch
from torch.utils.data import ConcatDataset, Dataset

q_labels = ['v.v_long', 'v.v_tran', 'w.w_psi', 'x.x', 'x.y', 'e.psi']      # utils/data_utils.py:100
u_labels = ['u.u_a', 'u.u_steer']                                          # utils/data_utils.py:101


def _step_of(dt, DATASET_DT):
    step = int(dt // DATASET_DT)
    assert np.isclose(step * DATASET_DT, dt), "dt must be a multiplier of DATASET_DT. "      # data_utils.py:24-25
    return step


class DynamicsDataset(Dataset):
    def __init__(self, q, u, dt, DATASET_DT=0.02, history=1, device="cuda", _resampled=False):
        if _resampled:                                   # from_dict: q / u already sub-sampled on the device
            self.q, self.u = q, u
        else:
            assert q.shape[0] == u.shape[0], "q and u doesn't match in n_entries. "
            step = _step_of(dt, DATASET_DT)
            n = int(q.shape[0])
            # rows already sit on the grid: t_i = i * DATASET_DT makes the kernel's interpolation an exact lookup
            t = torch.arange(n, dtype=torch.float64) * DATASET_DT
            self.q, self.u = _resample(t, q, u, DATASET_DT, step, n, device)
        self.history = history
        assert self.q.shape[0] >= self.history, (f"Dataset is too short. "
            f"Found {self.q.shape[0]} entries, while history is {self.history}. ")

    @classmethod
    def from_dict(cls, data, dt, DATASET_DT=0.02, history=1, plot=False, device="cuda"):
        qs, us, ts = [], [], []
        for t, state in data.items():                    # data_utils.py:46-58 (host: python objects)
            q = (state['v']['v_long'], state['v']['v_tran'], state['w']['w_psi'],
                 state['x']['x'], state['x']['y'], state['e']['psi'])
            u = np.array([state['u']['u_a'], state['u']['u_steer']])
            if np.allclose(u, np.array([1, 0])) or np.allclose(u, 0):
                continue
            qs.append(q)
            us.append(u)
            ts.append(t)
        t, q, u = np.asarray(ts) / 1e9, np.asarray(qs, dtype=np.float64), np.asarray(us, dtype=np.float64)
        step = _step_of(dt, DATASET_DT)
        n_interp = max(int(math.ceil((t[-1] - t[0]) / DATASET_DT)), 0)        # len(np.arange(t[0], t[-1], DATASET_DT))
        qd, ud = _resample(t, q, u, DATASET_DT, step, n_interp, device)
        return cls(qd, ud, dt, DATASET_DT, history, device, _resampled=True)

    @classmethod
    def from_pickle(cls, file_name, dt, DATASET_DT=0.02, history=1, plot=False, device="cuda"):
        with open(file_name, 'rb') as f:
            full_data = pkl.load(f)
            size = len(full_data)
        sets = [cls.from_dict(data, dt, DATASET_DT, history, plot, device) for data in full_data.values()]
        return sets[0] if size == 1 else ConcatDataset(sets)

    def __len__(self):
        return self.q.shape[0] - self.history

    def __getitem__(self, i):
        """Returns q, u, next_q"""
        return self.q[i:i + self.history], self.u[i:i + self.history], self.q[i + self.history]


def _resample(t, q, u, DATASET_DT, step, n_interp, device):
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("mpcdl: DynamicsDataset resamples on a CUDA device (there is no CPU path)")
    to = lambda a: torch.as_tensor(np.asarray(a) if not isinstance(a, torch.Tensor) else a, dtype=torch.float64).to(dev).contiguous()  # noqa: E731
    return torch.ops.mpcdl.resample(to(t), to(q), to(u), float(DATASET_DT), int(step), int(n_interp))
