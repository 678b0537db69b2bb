"""state_dict -> packed weight blob (host side of mpcdl_pack_*, include/mpcdl.h).

Key names are the reference checkpoints' (SURVEY.md App. B): ``layers.{i}.weight/bias`` for
Basic_Model (model_scripts/basic_model.py:48-56), ``gru|lstm.weight_ih_l{l}`` ... ``full_c.*`` for
the RNN models (model_scripts/deep_models.py:54-55, 78-79), ``{2i}.weight/bias`` for the
``nn.Sequential`` built by utils/pytorch_utils.py:52-57.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

_FP = C.POINTER(C.c_float)


def _np(t):
    if isinstance(t, torch.Tensor):
        t = t.detach().cpu().numpy()
    return np.ascontiguousarray(t, dtype=np.float32)


def _ptrs(arrs):
    return (_FP * len(arrs))(*[a.ctypes.data_as(_FP) for a in arrs])


def _blob(desc):
    n = _lib.lib.mpcdl_blob_floats(C.byref(desc))
    if n == 0:
        _lib.check(-2)
    return np.empty(n, dtype=np.float32)


def pack_mlp(sd, hidden: int, layers: int, tensor_core: bool = False) -> torch.Tensor:
    """Basic_Model state_dict -> blob.  Dropout entries (do_drop=True) hold no parameters, so the
    Linear entries are simply the sorted ``layers.{i}`` indices present (basic_model.py:53-54)."""
    idx = sorted({int(k.split(".")[1]) for k in sd if k.startswith("layers.")})
    if len(idx) != layers + 2:
        raise RuntimeError(f"mpcdl: expected {layers + 2} Linear entries, state_dict has {len(idx)}")
    W = [_np(sd[f"layers.{i}.weight"]) for i in idx]
    b = [_np(sd[f"layers.{i}.bias"]) for i in idx]
    shapes = [(hidden, 8)] + [(hidden, hidden)] * layers + [(6, hidden)]
    for w, s in zip(W, shapes):
        if w.shape != s:
            raise RuntimeError(f"mpcdl: weight shape {w.shape} != {s}")
    desc = _lib.ModelDesc(_lib.MLP, hidden, layers, 0)
    if tensor_core:
        n = _lib.lib.mpcdl_blob_floats_tc(C.byref(desc))
        if n == 0:
            _lib.check(-2)
        blob = np.empty(n, dtype=np.float32)
        _lib.check(_lib.lib.mpcdl_pack_mlp_tc(C.byref(desc), _ptrs(W), _ptrs(b), blob.ctypes.data_as(_FP)))
        return torch.from_numpy(blob)
    blob = _blob(desc)
    _lib.check(_lib.lib.mpcdl_pack_mlp(C.byref(desc), _ptrs(W), _ptrs(b), blob.ctypes.data_as(_FP)))
    return torch.from_numpy(blob)


def pack_rnn(sd, kind: int, hidden: int, layers: int, window: int, tensor_core: bool = False) -> torch.Tensor:
    """GRU_Model / LSTM_Model state_dict -> blob of the FP32 kernels, or (tensor_core=True) the tf32 hi/lo
    UMMA-layout blob of the tcgen05 kernel (mpcdl_pack_rnn_tc)."""
    pre = "gru" if kind == _lib.GRU else "lstm"
    G = 3 if kind == _lib.GRU else 4
    w_ih = [_np(sd[f"{pre}.weight_ih_l{l}"]) for l in range(layers)]
    w_hh = [_np(sd[f"{pre}.weight_hh_l{l}"]) for l in range(layers)]
    b_ih = [_np(sd[f"{pre}.bias_ih_l{l}"]) for l in range(layers)]
    b_hh = [_np(sd[f"{pre}.bias_hh_l{l}"]) for l in range(layers)]
    for l in range(layers):
        if w_ih[l].shape != (G * hidden, 8 if l == 0 else hidden) or w_hh[l].shape != (G * hidden, hidden):
            raise RuntimeError(f"mpcdl: layer {l} weight shapes {w_ih[l].shape}/{w_hh[l].shape} do not match hidden={hidden}")
    fc_w, fc_b = _np(sd["full_c.weight"]), _np(sd["full_c.bias"])
    desc = _lib.ModelDesc(kind, hidden, layers, window)
    if tensor_core:
        n = _lib.lib.mpcdl_blob_floats_tc(C.byref(desc))
        if n == 0:
            _lib.check(-2)
        blob = np.empty(n, dtype=np.float32)
        _lib.check(_lib.lib.mpcdl_pack_rnn_tc(C.byref(desc), _ptrs(w_ih), _ptrs(w_hh), _ptrs(b_ih), _ptrs(b_hh),
                                              fc_w.ctypes.data_as(_FP), fc_b.ctypes.data_as(_FP), blob.ctypes.data_as(_FP)))
        return torch.from_numpy(blob)
    blob = _blob(desc)
    _lib.check(_lib.lib.mpcdl_pack_rnn(C.byref(desc), _ptrs(w_ih), _ptrs(w_hh), _ptrs(b_ih), _ptrs(b_hh),
                                       fc_w.ctypes.data_as(_FP), fc_b.ctypes.data_as(_FP), blob.ctypes.data_as(_FP)))
    return torch.from_numpy(blob)


def pack_noise(sd, history: int, size: int, n_layers: int) -> torch.Tensor:
    """state_dict of the build_mlp nn.Sequential (Linear at even indices)."""
    W = [_np(sd[f"{2 * i}.weight"]) for i in range(n_layers + 1)]
    b = [_np(sd[f"{2 * i}.bias"]) for i in range(n_layers + 1)]
    K = 8 * history
    for i, w in enumerate(W):
        want = (12 if i == n_layers else size, K)
        if w.shape != want:
            raise RuntimeError(f"mpcdl: noise layer {i} weight {w.shape} != {want}")
        K = size
    blob = np.empty(_lib.lib.mpcdl_noise_blob_floats(history, size, n_layers), dtype=np.float32)
    _lib.check(_lib.lib.mpcdl_pack_noise(history, size, n_layers, _ptrs(W), _ptrs(b), blob.ctypes.data_as(_FP)))
    return torch.from_numpy(blob)


def pack_norm(stats) -> torch.Tensor:
    """dict(f_mean, f_std [8] or [T,8]; l_mean, l_std [6]) -> flat [2*T*8+12] in the kernel's order."""
    return torch.cat([torch.as_tensor(stats[k], dtype=torch.float32).reshape(-1)
                      for k in ("f_mean", "f_std", "l_mean", "l_std")]).contiguous()


def identity_norm(window: int) -> torch.Tensor:
    """mean 0 / std 1: (x-0)/1 and y*1+0 are exact, so rollout(H=1) == Module.forward bit for bit."""
    t = max(window, 1)
    return torch.cat([torch.zeros(t * 8), torch.ones(t * 8), torch.zeros(6), torch.ones(6)])
