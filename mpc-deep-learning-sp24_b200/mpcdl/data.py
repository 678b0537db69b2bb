"""Dataset build + z-score statistics of the reference (host logic, one-off, vectorised).

    basic_model.RobotData   model_scripts/basic_model.py:11-41   features bag[i], labels bag[i+1][:6]
    deep_models.RobotData   model_scripts/deep_models.py:11-42   windows bag[i:i+size], labels bag[i+size][:6]

The reference keeps the statistics in local variables and throws them away; the
rollout needs them (de-normalise the prediction, re-normalise the fed-back window), so this
RobotData keeps them.  RNN statistics are per window position: f_mean / f_std are [size, 8].
"""
from __future__ import annotations

import numpy as np
import torch
from torch.utils.data import Dataset


def _bag_array(bag):
    return torch.as_tensor(np.asarray(bag, dtype=np.float64)).to(torch.float32)   # float64 list -> fp32, as torch.tensor(list)


def mlp_dataset(bags):
    feats = torch.cat([_bag_array(b)[:-1] for b in bags if len(b) > 1])
    labels = torch.cat([_bag_array(b)[1:, :6] for b in bags if len(b) > 1])
    return feats.contiguous(), labels.contiguous()


def rnn_dataset(bags, size):
    feats, labels = [], []
    for b in bags:
        n = len(b) - size - 1                       # deep_models.py:18
        if n <= 0:
            continue
        a = _bag_array(b)
        feats.append(a.unfold(0, size, 1)[:n].permute(0, 2, 1))      # [n, size, 8]
        labels.append(a[size:size + n, :6])
    return torch.cat(feats).contiguous(), torch.cat(labels).contiguous()


def zscore_stats(feats, labels):
    """mean / UNBIASED std over dim 0 (basic_model.py:29-32, deep_models.py:30-33)."""
    return dict(f_mean=torch.mean(feats, dim=0), f_std=torch.std(feats, dim=0),
                l_mean=torch.mean(labels, dim=0), l_std=torch.std(labels, dim=0))


class RobotData(Dataset):
    """RobotData(data) == basic_model.RobotData ; RobotData(data, size) == deep_models.RobotData.
    ``featrs`` / ``labels`` are z-scored exactly like the reference; ``stats`` keeps the statistics."""

    def __init__(self, data, size=None, device="cpu"):
        self.size = size
        f, l = mlp_dataset(data) if size is None else rnn_dataset(data, size)
        self.stats = zscore_stats(f, l)
        self.featrs = ((f - self.stats["f_mean"]) / self.stats["f_std"]).to(device)     # subtract THEN divide
        self.labels = ((l - self.stats["l_mean"]) / self.stats["l_std"]).to(device)

    def __len__(self):
        return len(self.featrs)

    def __getitem__(self, i):
        return self.featrs[i], self.labels[i]
