"""torch.ops.mpcdl.* -- the operator boundary of the fused rollout path.

The operators are COMPILED: ``csrc/torch_ops.cpp`` registers them with ``TORCH_LIBRARY(mpcdl, ...)`` /
``TORCH_LIBRARY_IMPL(mpcdl, CUDA, ...)`` (built into ``libmpcdl_torch.so`` next to this file); each checks its
tensors, fills the plain-C argument struct of include/mpcdl.h and calls ``mpcdl_rollout`` & co. on
``at::cuda::getCurrentCUDAStream()``.  They exist for the CUDA dispatch key only, so CPU tensors fail in the
dispatcher ("no kernel for CPU"): there is no fallback, and importing this module raises if the library is missing.
The ctypes binding (``_lib``) remains for hosts without torch and serves the packers and the launch planner here.

Reference surface replaced: ``Basic_Model.forward`` model_scripts/basic_model.py:58-61,
``GRU_Model.forward`` / ``LSTM_Model.forward`` model_scripts/deep_models.py:60-66 / 85-92,
``FeedforwardNoiseModel.forward`` model_scripts/noise_model.py:34-40 and the feedback loop of
``run_dynamics_model`` scripts/test_model.py:67-87, batched over N trajectories and fused over H steps.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib

TORCH_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libmpcdl_torch.so")
if not os.path.exists(TORCH_LIB_PATH):
    raise ImportError(f"{TORCH_LIB_PATH} is missing: build it with `make -C mpc-deep-learning-sp24_b200/csrc` "
                      "(or __graft_entry__.build()); mpcdl has no fallback path")
torch.ops.load_library(TORCH_LIB_PATH)

OPS = {_lib.MLP: torch.ops.mpcdl.rollout_mlp, _lib.GRU: torch.ops.mpcdl.rollout_gru, _lib.LSTM: torch.ops.mpcdl.rollout_lstm}


def plan(kind, hidden, layers, window, N, H, stateful=False, noise_cfg=None, tune=(), precision=0):
    """Launch plan the library would use (works without a GPU): tile size, grid, smem, ring mode."""
    a = _lib.RolloutArgs()
    a.model = _lib.ModelDesc(kind, hidden, layers, window if kind != _lib.MLP else 0)
    a.stateful = int(stateful)
    for f in ("wblob", "wblob_tc", "norm", "init", "U", "states"):
        setattr(a, f, 16)           # never dereferenced by the planner
    a.precision = _lib.PRECISION[precision] if isinstance(precision, str) else int(precision)
    if stateful:
        a.carry = 16
    if noise_cfg is not None:
        a.noise_wblob = 16
        a.noise_history, a.noise_size, a.noise_layers, a.noise_act = noise_cfg
    a.N, a.H = N, H
    if len(tune) == 4:
        a.tune_rt, a.tune_ntg, a.tune_nslots, a.tune_kc = tune
    a.scratch, a.scratch_bytes = 16, _lib.lib.mpcdl_rollout_scratch_bytes(C.byref(a))
    p = _lib.Plan()
    _lib.check(_lib.lib.mpcdl_rollout_plan(C.byref(a), C.byref(p)))
    return p.as_dict()


def ffma_peak_tflops(device=None, iters=20000, reps=5, both=False):
    """Measured FP32 FMA peak of the current GPU (TFLOP/s), CUDA events, best of `reps`: the better of the scalar FFMA and
    the packed FFMA2 micro-kernel (the rollout kernels issue FFMA2).  both=True -> (ffma, ffma2)."""
    dev = torch.device(device or "cuda")
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    grid = sms * 2
    out = torch.empty(grid * 256, device=dev, dtype=torch.float32)
    st = torch.cuda.current_stream(dev).cuda_stream
    res = []
    for fn in (_lib.lib.mpcdl_ffma_peak, _lib.lib.mpcdl_ffma2_peak):
        best = float("inf")
        for r in range(reps + 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.check(fn(C.c_void_p(out.data_ptr()), iters, grid, C.c_void_p(st)))
            e1.record()
            e1.synchronize()
            if r:
                best = min(best, e0.elapsed_time(e1))
        res.append(2.0 * grid * 256 * iters * 64 / (best * 1e-3) / 1e12)
    return tuple(res) if both else max(res)
