"""torch.ops.mpcdl.* -- the operator boundary of the fused rollout path.

``torch.ops.mpcdl.rollout_{mlp,gru,lstm}`` are registered for the CUDA dispatch key only and call
``mpcdl_rollout`` (include/mpcdl.h) through ctypes with raw device pointers on the current CUDA
stream.  CPU tensors therefore fail in the dispatcher ("no kernel for CPU"): there is no fallback.

Reference surface replaced: ``Basic_Model.forward`` model_scripts/basic_model.py:58-61,
``GRU_Model.forward`` / ``LSTM_Model.forward`` model_scripts/deep_models.py:60-66 / 85-92,
``FeedforwardNoiseModel.forward`` model_scripts/noise_model.py:34-40 and the feedback loop of
``run_dynamics_model`` scripts/test_model.py:67-87, batched over N trajectories and fused over H steps.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Tuple

import torch

from . import _lib

_SCHEMA = ("(Tensor wblob, Tensor norm, Tensor init, Tensor U, int hidden, int layers, int window, bool stateful, "
           "Tensor(a!)? carry, Tensor? noise_wblob, int[] noise_cfg, Tensor? eps, int seed, int traj_offset, "
           "Tensor? cost_ref, float[] cost_w, bool return_states, bool return_costs, int[] tune, int precision, "
           "Tensor? wblob_tc, float[] ctrl, int ctrl_seed, int n_samples, bool shared_init) -> (Tensor, Tensor)")

_libdef = torch.library.Library("mpcdl", "DEF")
for _name in ("rollout_mlp", "rollout_gru", "rollout_lstm"):
    _libdef.define(_name + _SCHEMA)


def _f32c(t: Optional[torch.Tensor], name: str, dev=None) -> int:
    if t is None:
        return 0
    if t.dtype != torch.float32:
        raise RuntimeError(f"mpcdl: {name} must be float32, got {t.dtype}")
    if not t.is_cuda:
        raise RuntimeError(f"mpcdl: {name} must be a CUDA tensor (no CPU fallback), got {t.device}")
    if dev is not None and t.device != dev:
        raise RuntimeError(f"mpcdl: {name} is on {t.device}, expected {dev}")
    if not t.is_contiguous():
        raise RuntimeError(f"mpcdl: {name} must be contiguous")
    return t.data_ptr()


_BLOB_FLOATS = {}     # (tensor_core, kind, hidden, layers, window) -> floats: a pure function of the description


def _blob_floats(desc, tc: bool) -> int:
    key = (tc, desc.kind, desc.hidden, desc.layers, desc.window)
    n = _BLOB_FLOATS.get(key)
    if n is None:
        n = (_lib.lib.mpcdl_blob_floats_tc if tc else _lib.lib.mpcdl_blob_floats)(C.byref(desc))
        if n:
            _BLOB_FLOATS[key] = n
    return n


def _rollout_impl(kind: int, wblob, norm, init, U, hidden, layers, window, stateful, carry, noise_wblob, noise_cfg,
                  eps, seed, traj_offset, cost_ref, cost_w, return_states, return_costs, tune, precision=0, wblob_tc=None,
                  ctrl=(), ctrl_seed=0, n_samples=0, shared_init=False):
    dev = U.device
    T = window if kind != _lib.MLP else 0
    if len(ctrl):
        # MPPI call: U is the nominal sequence [H,2]; the N sampled sequences exist only in registers
        if len(ctrl) != 6 or U.dim() != 2 or U.shape[1] != 2 or n_samples < 1:
            raise RuntimeError(f"mpcdl: sampled controls need U [H,2], ctrl = sigma[2]+lo[2]+hi[2] and n_samples >= 1; got U {tuple(U.shape)}")
        N, H = int(n_samples), int(U.shape[0])
    else:
        if U.dim() != 3 or U.shape[2] != 2:
            raise RuntimeError(f"mpcdl: U must be [N,H,2], got {tuple(U.shape)}")
        N, H = int(U.shape[0]), int(U.shape[1])
    one = (6,) if kind == _lib.MLP else (T, 8)
    want = one if shared_init else (N,) + one
    if tuple(init.shape) != want:
        raise RuntimeError(f"mpcdl: init must be {want}, got {tuple(init.shape)}")
    nt = max(T, 1)
    if norm.numel() != 2 * nt * 8 + 12:
        raise RuntimeError(f"mpcdl: norm must hold {2 * nt * 8 + 12} floats (f_mean,f_std [{nt},8]; l_mean,l_std [6])")
    a = _lib.RolloutArgs()
    a.model = _lib.ModelDesc(kind, hidden, layers, T)
    a.precision = int(precision)
    if a.precision != 0:
        need = _blob_floats(a.model, True)
        if need == 0:
            _lib.check(-2)
        if wblob_tc is None or wblob_tc.numel() != need:
            raise RuntimeError(f"mpcdl: tensor-core precision needs wblob_tc with {need} floats (pack_rnn(..., tensor_core=True))")
        a.wblob_tc = _f32c(wblob_tc, "wblob_tc", dev)
    else:
        need = _blob_floats(a.model, False)
        if need == 0:
            _lib.check(-2)
        if wblob.numel() != need:
            raise RuntimeError(f"mpcdl: wblob has {wblob.numel()} floats, model needs {need}")
        a.wblob = _f32c(wblob, "wblob", dev)
    a.stateful = int(bool(stateful) and kind != _lib.MLP)
    a.norm = _f32c(norm, "norm", dev)
    a.init = _f32c(init, "init", dev)
    a.U = _f32c(U, "U", dev)
    states = torch.empty((N, H, 6), device=dev, dtype=torch.float32) if return_states else torch.empty(0, device=dev)
    costs = torch.empty((N,), device=dev, dtype=torch.float32) if return_costs else torch.empty(0, device=dev)
    a.states = states.data_ptr() if return_states else 0
    a.costs = costs.data_ptr() if return_costs else 0
    if cost_ref is not None:
        if tuple(cost_ref.shape) != (H, 6) or len(cost_w) != 8:
            raise RuntimeError("mpcdl: cost_ref must be [H,6] and cost_w = qw[6] + rw[2]")
        a.cost_ref = _f32c(cost_ref, "cost_ref", dev)
    if len(cost_w) == 8:
        a.cost_qw = (C.c_float * 6)(*cost_w[:6])
        a.cost_rw = (C.c_float * 2)(*cost_w[6:])
    if noise_wblob is not None:
        hist, size, nl, act = noise_cfg
        if noise_wblob.numel() != _lib.lib.mpcdl_noise_blob_floats(hist, size, nl):
            raise RuntimeError("mpcdl: noise_wblob size does not match noise_cfg")
        a.noise_wblob = _f32c(noise_wblob, "noise_wblob", dev)
        a.noise_history, a.noise_size, a.noise_layers, a.noise_act = hist, size, nl, act
        if eps is not None:
            if tuple(eps.shape) != (N, H, 6):
                raise RuntimeError(f"mpcdl: eps must be [N,H,6], got {tuple(eps.shape)}")
            a.eps = _f32c(eps, "eps", dev)
    a.seed = seed & 0xFFFFFFFFFFFFFFFF
    a.traj_offset = traj_offset
    if a.stateful:
        planes = layers * (2 if kind == _lib.LSTM else 1)
        if carry is None or tuple(carry.shape) != (planes, N, hidden):
            raise RuntimeError(f"mpcdl: stateful rollout needs carry [{planes},{N},{hidden}]")
        a.carry = _f32c(carry, "carry", dev)
    a.N, a.H = N, H
    if len(tune) == 4:
        a.tune_rt, a.tune_ntg, a.tune_nslots, a.tune_kc = tune
    if len(ctrl):
        a.sample_controls = 1
        a.ctrl_sigma, a.ctrl_lo, a.ctrl_hi = (C.c_float * 2)(*ctrl[0:2]), (C.c_float * 2)(*ctrl[2:4]), (C.c_float * 2)(*ctrl[4:6])
        a.ctrl_seed = ctrl_seed & 0xFFFFFFFFFFFFFFFF
    a.shared_init = int(bool(shared_init))
    nscratch = _lib.lib.mpcdl_rollout_scratch_bytes(C.byref(a)) if a.precision == _lib.PRECISION["bf16"] else 0
    if nscratch:
        # per-CTA layer output sequences of the BF16 LSTM-256 kernel; from the caching allocator, so the buffer is
        # tied to the current stream like every other tensor of the call
        scratch = torch.empty(nscratch, device=dev, dtype=torch.uint8)
        a.scratch, a.scratch_bytes = scratch.data_ptr(), nscratch
    stream = torch.cuda.current_stream(dev).cuda_stream
    if torch.cuda.current_device() == dev.index:         # the launch goes to the CURRENT device: switch only when needed
        _lib.check(_lib.lib.mpcdl_rollout(C.byref(a), C.c_void_p(stream)))
    else:
        with torch.cuda.device(dev):
            _lib.check(_lib.lib.mpcdl_rollout(C.byref(a), C.c_void_p(stream)))
    return states, costs


def _make(kind):
    def impl(wblob, norm, init, U, hidden, layers, window, stateful, carry, noise_wblob, noise_cfg, eps, seed,
             traj_offset, cost_ref, cost_w, return_states, return_costs, tune, precision, wblob_tc, ctrl, ctrl_seed,
             n_samples, shared_init):
        return _rollout_impl(kind, wblob, norm, init, U, hidden, layers, window, stateful, carry, noise_wblob,
                             noise_cfg, eps, seed, traj_offset, cost_ref, cost_w, return_states, return_costs, tune,
                             precision, wblob_tc, ctrl, ctrl_seed, n_samples, shared_init)
    return impl


_libimpl = torch.library.Library("mpcdl", "IMPL", "CUDA")
_libimpl.impl("rollout_mlp", _make(_lib.MLP))
_libimpl.impl("rollout_gru", _make(_lib.GRU))
_libimpl.impl("rollout_lstm", _make(_lib.LSTM))



# ---------------------------------------------------------------------------- noise net / error metrics
_libdef.define("noise_forward(Tensor noise_wblob, int[] noise_cfg, Tensor q, Tensor u, Tensor? eps, int seed, "
               "int traj_offset, int step0, bool return_params) -> (Tensor, Tensor, Tensor)")
_libdef.define("dynamics_nn_rollout(Tensor noise_wblob, int[] noise_cfg, Tensor q0, Tensor U, Tensor? eps, int seed, "
               "int traj_offset) -> Tensor")
_libdef.define("error_metrics(Tensor output, Tensor labels) -> (Tensor, Tensor)")


def _noise_args(noise_wblob, noise_cfg, dev):
    hist, size, nl, act = noise_cfg
    if noise_wblob.numel() != _lib.lib.mpcdl_noise_blob_floats(hist, size, nl):
        raise RuntimeError("mpcdl: noise_wblob size does not match noise_cfg")
    a = _lib.NoiseArgs()
    a.noise_wblob = _f32c(noise_wblob, "noise_wblob", dev)
    a.history, a.size, a.n_layers, a.act = hist, size, nl, act
    return a


def _launch_noise(a, dev):
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib.mpcdl_noise_forward(C.byref(a), C.c_void_p(stream)))


def _noise_forward_impl(noise_wblob, noise_cfg, q, u, eps, seed, traj_offset, step0, return_params):
    """FeedforwardNoiseModel.forward (model_scripts/noise_model.py:34-40): q [B,6*hist], u [B,2*hist] -> [B,6]."""
    dev = q.device
    a = _noise_args(noise_wblob, noise_cfg, dev)
    hist = a.history
    if q.dim() != 2 or q.shape[1] != 6 * hist or tuple(u.shape) != (q.shape[0], 2 * hist):
        raise RuntimeError(f"mpcdl: noise forward needs q [B,{6 * hist}] and u [B,{2 * hist}], got {tuple(q.shape)} {tuple(u.shape)}")
    B = int(q.shape[0])
    a.q, a.u = _f32c(q, "q", dev), _f32c(u, "u", dev)
    if eps is not None:
        if tuple(eps.shape) != (B, 6):
            raise RuntimeError(f"mpcdl: eps must be [B,6], got {tuple(eps.shape)}")
        a.eps = _f32c(eps, "eps", dev)
    a.seed, a.traj_offset, a.step0 = seed & 0xFFFFFFFFFFFFFFFF, traj_offset, step0
    out = torch.empty((B, 6), device=dev, dtype=torch.float32)
    mu = torch.empty((B, 6) if return_params else (0,), device=dev, dtype=torch.float32)
    scale = torch.empty((B, 6) if return_params else (0,), device=dev, dtype=torch.float32)
    a.out = out.data_ptr()
    if return_params:
        a.mu, a.scale = mu.data_ptr(), scale.data_ptr()
    a.N, a.H = B, 1
    _launch_noise(a, dev)
    return out, mu, scale


def _dynnn_rollout_impl(noise_wblob, noise_cfg, q0, U, eps, seed, traj_offset):
    """DynamicsNN.get_prediction fed back on itself (noise_model.py:113-125, test_model.py:67-87), fused over H."""
    dev = U.device
    a = _noise_args(noise_wblob, noise_cfg, dev)
    if U.dim() != 3 or U.shape[2] != 2 or tuple(q0.shape) != (U.shape[0], 6):
        raise RuntimeError(f"mpcdl: DynamicsNN rollout needs q0 [N,6] and U [N,H,2], got {tuple(q0.shape)} {tuple(U.shape)}")
    N, H = int(U.shape[0]), int(U.shape[1])
    a.closed_loop = 1
    a.q, a.u = _f32c(q0, "q0", dev), _f32c(U, "U", dev)
    if eps is not None:
        if tuple(eps.shape) != (N, H, 6):
            raise RuntimeError(f"mpcdl: eps must be [N,H,6], got {tuple(eps.shape)}")
        a.eps = _f32c(eps, "eps", dev)
    a.seed, a.traj_offset = seed & 0xFFFFFFFFFFFFFFFF, traj_offset
    out = torch.empty((N, H, 6), device=dev, dtype=torch.float32)
    a.out = out.data_ptr()
    a.N, a.H = N, H
    _launch_noise(a, dev)
    return out


def _error_metrics_impl(output, labels):
    """Per-sample MAE / RMSE of eval_models.calc_error (model_scripts/eval_models.py:91-92)."""
    dev = output.device
    if output.dim() != 2 or output.shape[1] != 6 or output.shape != labels.shape:
        raise RuntimeError(f"mpcdl: error_metrics needs output and labels [B,6], got {tuple(output.shape)} {tuple(labels.shape)}")
    B = int(output.shape[0])
    mae = torch.empty(B, device=dev, dtype=torch.float32)
    rmse = torch.empty(B, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib.mpcdl_error_metrics(C.c_void_p(_f32c(output, "output", dev)), C.c_void_p(_f32c(labels, "labels", dev)),
                                               B, C.c_void_p(mae.data_ptr()), C.c_void_p(rmse.data_ptr()), C.c_void_p(stream)))
    return mae, rmse


_libdef.define("mppi_reduce(Tensor costs, Tensor U, float[] ctrl, int ctrl_seed, int traj_offset, float lam, Tensor? cmin) -> Tensor")
_libdef.define("mppi_min(Tensor costs) -> Tensor")


def _mppi_min_impl(costs):
    dev = costs.device
    ws = torch.empty(_lib.lib.mpcdl_mppi_workspace_floats(1), device=dev, dtype=torch.float32)
    out = torch.empty(1, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib.mpcdl_mppi_min(C.c_void_p(_f32c(costs, "costs", dev)), costs.numel(), C.c_void_p(ws.data_ptr()),
                                          C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
    return out


def _mppi_reduce_impl(costs, U, ctrl, ctrl_seed, traj_offset, lam, cmin):
    """sums [2H+2] = [sum_n w_n (u_n - U) per (t, channel), sum_n w_n, cmin]; include/mpcdl.h mpcdl_mppi_reduce."""
    dev = costs.device
    if U.dim() != 2 or U.shape[1] != 2 or len(ctrl) != 6 or costs.dim() != 1:
        raise RuntimeError("mpcdl: mppi_reduce needs costs [N], U [H,2], ctrl = sigma[2]+lo[2]+hi[2]")
    H = int(U.shape[0])
    a = _lib.MppiArgs()
    a.costs, a.N, a.traj_offset, a.H = _f32c(costs, "costs", dev), costs.numel(), traj_offset, H
    a.U = _f32c(U, "U", dev)
    a.ctrl_sigma, a.ctrl_lo, a.ctrl_hi = (C.c_float * 2)(*ctrl[0:2]), (C.c_float * 2)(*ctrl[2:4]), (C.c_float * 2)(*ctrl[4:6])
    a.ctrl_seed, a.lam = ctrl_seed & 0xFFFFFFFFFFFFFFFF, lam
    if cmin is not None:
        a.cmin = _f32c(cmin, "cmin", dev)
    ws = torch.empty(_lib.lib.mpcdl_mppi_workspace_floats(H), device=dev, dtype=torch.float32)
    sums = torch.empty(2 * H + 2, device=dev, dtype=torch.float32)
    a.workspace, a.sums = ws.data_ptr(), sums.data_ptr()
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib.mpcdl_mppi_reduce(C.byref(a), C.c_void_p(stream)))
    return sums


_libimpl.impl("mppi_min", _mppi_min_impl)
_libimpl.impl("mppi_reduce", _mppi_reduce_impl)
_libimpl.impl("noise_forward", _noise_forward_impl)
_libimpl.impl("dynamics_nn_rollout", _dynnn_rollout_impl)
_libimpl.impl("error_metrics", _error_metrics_impl)

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


def ffma_peak_tflops(device=None, iters=20000, reps=5) -> float:
    """Measured FP32 FFMA peak of the current GPU (TFLOP/s), CUDA events, best of `reps`."""
    dev = torch.device(device or "cuda")
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    grid = sms * 2
    out = torch.empty(grid * 256, device=dev, dtype=torch.float32)
    st = torch.cuda.current_stream(dev).cuda_stream
    best = float("inf")
    for r in range(reps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(_lib.lib.mpcdl_ffma_peak(C.c_void_p(out.data_ptr()), iters, grid, C.c_void_p(st)))
        e1.record()
        e1.synchronize()
        if r:
            best = min(best, e0.elapsed_time(e1))
    return 2.0 * grid * 256 * iters * 64 / (best * 1e-3) / 1e12
