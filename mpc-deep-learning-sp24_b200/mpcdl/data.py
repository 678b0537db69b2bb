"""Dataset build + z-score statistics of the reference: host restatement (vectorised torch) and the device path
(csrc/dataset_kernels.cu).

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


# ------------------------------------------------------------------------------------------------ device path (f3)
class DeviceBags:
    """The recorded rows of all bags on the GPU ([R,8] FP32: the pickle's doubles rounded to nearest, as
    ``torch.tensor(list)`` does) plus the bag lengths; what the dataset kernels (csrc/dataset_kernels.cu) index."""

    def __init__(self, bags, device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("mpcdl: the dataset kernels run on a CUDA device (there is no CPU path); "
                               "use RobotData(data, size) without device= for the host restatement")
        self.lens = [len(b) for b in bags]
        rows = np.concatenate([np.asarray(b, dtype=np.float64).reshape(-1, 8) for b in bags]) if bags else np.zeros((0, 8))
        self.rows = torch.as_tensor(rows).to(torch.float32).to(self.device).contiguous()

    def offsets(self, per_bag):
        """per_bag(len) -> samples of a bag; returns (bag_row_off, samp_off) device int64 [n_bags] and n.
        Bags that yield no sample are dropped (the reference's ``range(len(bag) - ...)`` is empty for them)."""
        row_off, samp_off, r, n = [], [], 0, 0
        for ln in self.lens:
            k = max(per_bag(ln), 0)
            if k > 0:
                row_off.append(r)
                samp_off.append(n)
                n += k
            r += ln
        if n == 0:
            raise RuntimeError("mpcdl: the bags are too short for this window / horizon")
        return (torch.tensor(row_off, dtype=torch.int64, device=self.device),
                torch.tensor(samp_off, dtype=torch.int64, device=self.device), n)


def stats_from_vector(vec, size):
    """the kernels' `norm` vector [2*T*8+12] -> dict(f_mean, f_std [T,8] | [8], l_mean, l_std [6]) (views)"""
    T = size if size else 1
    shape = (T, 8) if size else (8,)
    return dict(f_mean=vec[:T * 8].view(shape), f_std=vec[T * 8:2 * T * 8].view(shape),
                l_mean=vec[2 * T * 8:2 * T * 8 + 6], l_std=vec[2 * T * 8 + 6:])


def device_dataset(bags, size=None, device="cuda", normalise=True, stats_vec=None):
    """RobotData.__init__ + norm on the GPU: -> (featrs, labels, stats_vec).  Two launches: the statistics (unless
    given) and the gather + z-score.  Statistics are within one FP32 rounding of the FP64 ones (two-pass double accumulation;
    against torch's float32 cascade sum: std <= 1e-6 relative, mean <= 1e-6 of max(|mean|, std)); given the same statistics the z-scored values are bit-identical."""
    db = bags if isinstance(bags, DeviceBags) else DeviceBags(bags, device)
    w = int(size) if size else 0
    row_off, samp_off, n = db.offsets((lambda ln: ln - 1) if w == 0 else (lambda ln: ln - w - 1))   # basic_model.py:19 / deep_models.py:18
    if stats_vec is None:
        stats_vec = torch.ops.mpcdl.dataset_stats(db.rows, row_off, samp_off, w, n)
    featrs, labels = torch.ops.mpcdl.dataset_build(db.rows, row_off, samp_off, w, n, stats_vec if normalise else None)
    return featrs, labels, stats_vec


def open_loop_inputs(bags, horizon, size=None, device="cuda"):
    """Start states / windows, recorded controls [n,H,2] and recorded states [n,H,6] of every stretch that leaves
    `horizon` recorded frames ahead (what an open-loop validation sweep feeds the rollout): one launch."""
    db = bags if isinstance(bags, DeviceBags) else DeviceBags(bags, device)
    w = int(size) if size else 0
    row_off, samp_off, n = db.offsets((lambda ln: ln - horizon) if w == 0 else (lambda ln: ln - w - horizon + 1))
    return torch.ops.mpcdl.rollout_inputs(db.rows, row_off, samp_off, w, n, int(horizon))


def save_stats(path, stats):
    """Persist the z-score statistics (the reference throws them away).  Convention: next to a checkpoint
    ``X.pt`` they live in ``X.stats.npz`` and ``mpcdl.load_model`` picks them up."""
    np.savez(path, **{k: torch.as_tensor(v).detach().cpu().numpy() for k, v in stats.items()})


def load_stats(path):
    z = np.load(path)
    return {k: torch.from_numpy(z[k]) for k in ("f_mean", "f_std", "l_mean", "l_std")}


def stats_path(ckpt_path):
    base = ckpt_path
    for ext in (".pt", ".npz"):
        if base.endswith(ext):
            base = base[:-len(ext)]
    return base + ".stats.npz"


class RobotData(Dataset):
    """RobotData(data) == basic_model.RobotData ; RobotData(data, size) == deep_models.RobotData.
    ``featrs`` / ``labels`` are z-scored exactly like the reference; ``stats`` keeps the statistics.

    device="cpu" (default) is the host restatement on torch (bit-identical statistics to the reference);
    a CUDA device builds the windows, the statistics and the z-score with the dataset kernels and leaves everything
    on the GPU (``stats_vec`` is then directly the rollout's `norm` vector)."""

    def __init__(self, data, size=None, device="cpu"):
        self.size = size
        if torch.device(device).type == "cuda":
            self.featrs, self.labels, self.stats_vec = device_dataset(data, size, device)
            self.stats = stats_from_vector(self.stats_vec, size)
            return
        f, l = mlp_dataset(data) if size is None else rnn_dataset(data, size)
        self.stats = zscore_stats(f, l)
        self.featrs = ((f - self.stats["f_mean"]) / self.stats["f_std"]).to(device)     # subtract THEN divide
        self.labels = ((l - self.stats["l_mean"]) / self.stats["l_std"]).to(device)

    def __len__(self):
        return len(self.featrs)

    def __getitem__(self, i):
        return self.featrs[i], self.labels[i]
