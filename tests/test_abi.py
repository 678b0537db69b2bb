"""CPU: libmpcdl.so loads and exports every entry point include/mpcdl.h declares; argument checks
that need no device return the documented status codes (no compute calls without a GPU)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "mpcdl.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mpcdl_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    from mpcdl import _lib
    names = _declared()
    assert len(names) >= 10
    for n in names:
        assert hasattr(_lib.lib, n), f"{n} declared in include/mpcdl.h but not exported by libmpcdl.so"
    assert set(_lib.EXPORTS) <= set(names)
    assert _lib.lib.mpcdl_version() >= 100


def test_struct_layout_matches_header():
    from mpcdl import _lib
    # mpcdl_rollout_args: 4 ints model + stateful + 7 pointers ... the ctypes mirror must agree with the
    # compiler's layout; the library reports its own sizeof through the planner accepting our struct.
    assert C.sizeof(_lib.ModelDesc) == 16
    assert C.sizeof(_lib.Plan) == 52
    assert C.sizeof(_lib.RolloutArgs) % 8 == 0


def test_status_codes_without_device():
    from mpcdl import _lib
    lib = _lib.lib
    bad = _lib.ModelDesc(_lib.MLP, 48, 2, 0)
    assert lib.mpcdl_blob_floats(C.byref(bad)) == 0
    assert b"hidden=48" in lib.mpcdl_last_error()
    good = _lib.ModelDesc(_lib.MLP, 64, 4, 0)
    assert lib.mpcdl_blob_floats(C.byref(good)) == (8 * 64 + 64) + 4 * (64 * 64 + 64) + (64 * 8 + 8)
    a = _lib.RolloutArgs()
    a.model = good
    a.N, a.H = 0, 5
    p = _lib.Plan()
    assert lib.mpcdl_rollout_plan(C.byref(a), C.byref(p)) == -1          # MPCDL_ERR_ARG
    a.N = 10
    assert lib.mpcdl_rollout_plan(C.byref(a), C.byref(p)) == -1          # null pointers
    with pytest.raises(RuntimeError):
        _lib.check(-1)
    assert lib.mpcdl_rollout(None, None) == -1


def test_noise_entry_points_without_device():
    """mpcdl_noise_forward / mpcdl_error_metrics: argument checks come before any device work."""
    from mpcdl import _lib
    lib = _lib.lib
    assert C.sizeof(_lib.NoiseArgs) == 120 and _lib.NoiseArgs.q.offset == 32 and _lib.NoiseArgs.N.offset == 104
    assert lib.mpcdl_noise_forward(None, None) == -1
    a = _lib.NoiseArgs()
    assert lib.mpcdl_noise_forward(C.byref(a), None) == -1               # null pointers
    for f in ("noise_wblob", "q", "u", "out"):
        setattr(a, f, 16)
    a.history, a.size, a.n_layers, a.act, a.N, a.H = 3, 32, 2, 2, 8, 2
    assert lib.mpcdl_noise_forward(C.byref(a), None) == -1               # one-step forward with H = 2
    assert b"H=2" in lib.mpcdl_last_error()
    a.H, a.size = 1, 30
    assert lib.mpcdl_noise_forward(C.byref(a), None) == -2               # size not a multiple of 4
    a.size, a.history = 32, 9
    assert lib.mpcdl_noise_forward(C.byref(a), None) == -2
    a.history, a.closed_loop, a.mu = 3, 1, 16
    assert lib.mpcdl_noise_forward(C.byref(a), None) == -1               # mu/scale exist for the forward only
    assert lib.mpcdl_error_metrics(None, None, 4, None, None, None) == -1


def test_plans_fit_the_machine():
    from mpcdl import _lib, ops
    for kind, hid, layers, win, n, h in [(0, 64, 4, 0, 4096, 50), (0, 16, 1, 0, 1000, 50), (0, 256, 4, 0, 65536, 10),
                                         (1, 64, 4, 4, 65536, 100), (2, 64, 4, 4, 65536, 100), (2, 256, 4, 4, 65536, 100),
                                         (1, 128, 2, 8, 100, 3), (2, 16, 1, 1, 1, 1)]:
        for st in ((False, True) if kind else (False,)):
            for nz in (None, [3, 32, 2, 2]):
                p = ops.plan(kind, hid, layers, win, n, h, stateful=st, noise_cfg=nz)
                assert p["smem_bytes"] <= 232448 and p["threads"] <= 512 and p["threads"] % 32 == 0
                assert 1 <= p["grid"] <= 148 and p["ntiles"] * p["tile_traj"] >= n
                assert p["tile_traj"] == p["rt"] * p["n_tg"]
    # 1 M trajectories of the hidden-64 MLP: two independent 128-trajectory groups per CTA; few tiles: one group, more SMs
    p = ops.plan(0, 64, 4, 0, 1 << 20, 50)
    assert p["variant"] == 2 and p["tile_traj"] == 128 and p["threads"] == 256 and p["ntiles"] == 8192 and p["grid"] == 148
    p = ops.plan(0, 64, 4, 0, 16384, 50)
    assert p["variant"] == 2 and p["tile_traj"] == 128 and p["threads"] == 128 and p["grid"] == 128
    assert ops.plan(0, 64, 4, 0, 4096, 50)["tile_traj"] == 32


def test_tensor_core_plans_and_errors():
    """tcgen05 path: planner geometry (one 128-trajectory tile per CTA, ring fits the SM) and the
    'no silent fallback' rule for unsupported shapes."""
    from mpcdl import _lib, ops
    for prec in ("tf32x3", "tf32"):
        for kind in (1, 2):
            for st in (False, True):
                for nz in (None, [3, 32, 2, 2]):
                    p = ops.plan(kind, 64, 4, 4, 65536, 100, stateful=st, noise_cfg=nz, precision=prec)
                    assert p["variant"] == 4 and p["tile_traj"] == 128 and p["threads"] == 384
                    assert 120 * 1024 <= p["smem_bytes"] <= 232448 and p["nslots"] >= 2
                    assert p["ksplit"] in (1, 2, 4) and p["grid"] == 148 and p["ntiles"] == 512
                    if p["resident"]:
                        assert p["nslots"] * p["ksplit"] >= 8          # a layer's two 64-row matrices stay in the ring
    for kind in (1, 2):                                                    # window 8: same kernel, the sequence through a per-CTA scratch
        p = ops.plan(kind, 64, 4, 8, 65536, 100, noise_cfg=[3, 32, 2, 2], precision="tf32x3")
        assert p["variant"] == 4 and p["grid"] == 148 and 120 * 1024 <= p["smem_bytes"] <= 232448
        a = _lib.RolloutArgs()
        a.model = _lib.ModelDesc(kind, 64, 4, 8)
        a.precision, a.N, a.H = _lib.PRECISION["tf32x3"], 65536, 100
        assert _lib.lib.mpcdl_rollout_scratch_bytes(C.byref(a)) == 148 * 8 * 64 * 128 * 4
        a.N = 1000
        assert _lib.lib.mpcdl_rollout_scratch_bytes(C.byref(a)) == 8 * 8 * 64 * 128 * 4
        a.model = _lib.ModelDesc(kind, 64, 4, 4)
        assert _lib.lib.mpcdl_rollout_scratch_bytes(C.byref(a)) == 0
    for args in [(1, 128, 2, 4), (1, 64, 4, 6), (0, 16, 4, 0)]:           # hidden 128 / window 6 / MLP hidden 16: FP32 kernels only
        with pytest.raises(RuntimeError):
            ops.plan(args[0], args[1], args[2], args[3], 100, 5, precision="tf32x3")
    # tcgen05 MLP + residual-noise net: per-tile noise buffers; two tiles per CTA only where both sets fit beside the weights
    p = ops.plan(0, 64, 4, 0, 1 << 20, 50, noise_cfg=[3, 32, 2, 2], precision="tf32")
    assert p["variant"] == 5 and p["ksplit"] == 2 and p["smem_bytes"] <= 232448
    p = ops.plan(0, 64, 4, 0, 1 << 20, 50, noise_cfg=[3, 32, 2, 2], precision="tf32x3")
    assert p["variant"] == 5 and p["ksplit"] in (1, 2) and p["grid"] == 148 and p["smem_bytes"] <= 232448
    # tcgen05 MLP (rollout_mlp_tc.cu): 128-trajectory tiles, two ping-pong tiles per CTA once there are more tiles than SMs
    for prec in ("tf32x3", "tf32"):
        for layers in (1, 2, 4):
            p = ops.plan(0, 64, layers, 0, 4096, 50, precision=prec)
            assert p["variant"] == 5 and p["tile_traj"] == 128 and p["ntiles"] == 32 and p["grid"] == 32 and p["ksplit"] == 1
            p = ops.plan(0, 64, layers, 0, 1 << 20, 50, precision=prec)
            # one-pass TF32 without the noise net and >= 8 tiles per SM: four tile slots of 128 TMEM columns, four epilogue warps each
            assert p["grid"] == 148 and p["ntiles"] == 8192 and (p["ksplit"], p["threads"]) == ((4, 640) if prec == "tf32" else (2, 576))
            assert 120 * 1024 <= p["smem_bytes"] <= 232448
            p = ops.plan(0, 64, layers, 0, 65536, 50, precision=prec)
            assert p["ksplit"] == 2 and p["threads"] == 576 and p["grid"] == 148
    d = _lib.ModelDesc(_lib.GRU, 32, 2, 4)
    assert _lib.lib.mpcdl_blob_floats_tc(C.byref(d)) == 0
    # BF16 tcgen05 kernel of the hidden-256 LSTM (rollout_lstm_tc.cu): 128-trajectory tiles, 16 epilogue warps + one
    # control warpgroup, CTA pairs (even grid), a per-CTA scratch of 2 x window operand tiles; everything else refuses the precision
    p = ops.plan(2, 256, 4, 4, 65536, 100, precision="bf16")
    assert p["variant"] == 6 and p["tile_traj"] == 128 and p["threads"] == 640 and p["grid"] == 148 and p["ntiles"] == 512
    assert p["smem_bytes"] <= 232448
    assert ops.plan(2, 256, 4, 4, 129, 5, precision="bf16")["grid"] == 2 and ops.plan(2, 256, 4, 4, 3 * 128, 5, precision="bf16")["grid"] == 4
    a = _lib.RolloutArgs()
    a.model, a.precision, a.N, a.H = _lib.ModelDesc(2, 256, 4, 4), 3, 1000, 5
    assert _lib.lib.mpcdl_rollout_scratch_bytes(C.byref(a)) == 8 * (4 * 128 * 256 * 2 + 4 * 8 * 128 * 4)
    a.precision = 0
    assert _lib.lib.mpcdl_rollout_scratch_bytes(C.byref(a)) == 0
    # with the residual-noise net the scratch also holds its history [8 * history][128] FP32 per CTA
    a.precision, a.noise_wblob, a.noise_history, a.noise_size, a.noise_layers, a.noise_act = 3, 16, 3, 32, 2, 2
    assert _lib.lib.mpcdl_rollout_scratch_bytes(C.byref(a)) == 8 * (4 * 128 * 256 * 2 + 4 * 8 * 128 * 4 + 24 * 128 * 4)
    # stateful models and the default noise net run on the BF16 kernel (round 2); other families / sizes and a noise net
    # whose buffers exceed the 64 KB the kernel can lend it are refused
    assert ops.plan(2, 256, 4, 4, 100, 5, precision="bf16", stateful=True)["variant"] == 6
    assert ops.plan(2, 256, 4, 4, 100, 5, precision="bf16", noise_cfg=[3, 32, 2, 2])["variant"] == 6
    for args, kw in [((2, 64, 4, 4), {}), ((1, 256, 4, 4), {}), ((2, 256, 4, 4), dict(noise_cfg=[8, 64, 2, 2]))]:
        with pytest.raises(RuntimeError):
            ops.plan(*args, 100, 5, precision="bf16", **kw)


def test_training_entry_points_without_device():
    """mpcdl_train_mlp_epoch / mpcdl_l1_batches: argument checks come before any device work"""
    from mpcdl import _lib
    lib = _lib.lib
    assert lib.mpcdl_train_param_count(64, 4) == 17606 and lib.mpcdl_train_param_count(16, 1) == 518
    assert lib.mpcdl_train_mlp_epoch(None, None) == -1
    a = _lib.TrainArgs()
    assert lib.mpcdl_train_mlp_epoch(C.byref(a), None) == -1 and b"non-NULL" in lib.mpcdl_last_error()
    for f in ("theta", "adam_m", "adam_v", "featrs", "labels", "perm"):
        setattr(a, f, 16)
    a.n_models, a.n_samples, a.batch, a.n_rows, a.hidden, a.layers = 1, 100, 16, 100, 48, 2
    a.lr, a.beta1, a.beta2, a.eps = 1e-3, 0.9, 0.999, 1e-8
    assert lib.mpcdl_train_mlp_epoch(C.byref(a), None) == -2 and b"hidden=48" in lib.mpcdl_last_error()   # MPCDL_ERR_UNSUPPORTED
    a.hidden, a.lr = 64, 0.0
    assert lib.mpcdl_train_mlp_epoch(C.byref(a), None) == -1 and b"Adam" in lib.mpcdl_last_error()
    assert lib.mpcdl_l1_batches(None, None, 10, 16, None, None) == -1


def test_recurrent_training_entry_points_without_device():
    """mpcdl_train_rnn_epoch: parameter / scratch counts and the argument checks, which come before any device work"""
    from mpcdl import _lib
    lib = _lib.lib
    # GRU_64_4 89 478 and LSTM_64_4 119 174 parameters
    assert lib.mpcdl_train_rnn_param_count(_lib.GRU, 64, 4) == 89478 and lib.mpcdl_train_rnn_param_count(_lib.LSTM, 64, 4) == 119174
    assert lib.mpcdl_train_rnn_param_count(0, 64, 4) == 0 and lib.mpcdl_train_rnn_work_floats(3, 64, 4, 4) == 0
    # scratch: k-major weight copies + [L][T][16][h] outputs + [L][T][16][4h | 5h] saved gates + 2 [T][16][Gh] gate gradients + dY + the
    # micro-batch + the gradient sums of a batch of more than 16 rows
    assert lib.mpcdl_train_rnn_work_floats(_lib.GRU, 64, 4, 4) == 89480 + 16384 + 65536 + 2 * 12288 + 4096 + 512 + 89480
    assert lib.mpcdl_train_rnn_work_floats(_lib.LSTM, 16, 2, 8) == 3944 + 4096 + 20480 + 2 * 8192 + 2048 + 1024 + 3944
    assert lib.mpcdl_train_rnn_epoch(None, None) == -1
    a = _lib.TrainRnnArgs()
    assert lib.mpcdl_train_rnn_epoch(C.byref(a), None) == -1 and b"non-NULL" in lib.mpcdl_last_error()
    for f in ("theta", "adam_m", "adam_v", "featrs", "labels", "perm", "work"):
        setattr(a, f, 16)
    a.n_models, a.n_samples, a.batch, a.n_rows, a.kind, a.hidden, a.layers, a.window = 1, 100, 16, 100, _lib.GRU, 64, 4, 4
    a.lr, a.beta1, a.beta2, a.eps, a.param_stride = 1e-3, 0.9, 0.999, 1e-8, 89480
    for field, bad in (("hidden", 48), ("window", 9), ("kind", 0)):
        good = getattr(a, field)
        setattr(a, field, bad)
        assert lib.mpcdl_train_rnn_epoch(C.byref(a), None) == -2 and f"{field}={bad}".encode() in lib.mpcdl_last_error()   # MPCDL_ERR_UNSUPPORTED
        setattr(a, field, good)
    a.param_stride = 89478
    assert lib.mpcdl_train_rnn_epoch(C.byref(a), None) == -1 and b"param_stride" in lib.mpcdl_last_error()
    a.param_stride, a.theta = 89480, 20
    assert lib.mpcdl_train_rnn_epoch(C.byref(a), None) == -1 and b"16-byte" in lib.mpcdl_last_error()
    a.theta, a.lr = 16, -1.0
    assert lib.mpcdl_train_rnn_epoch(C.byref(a), None) == -1 and b"Adam" in lib.mpcdl_last_error()


def test_mppi_entry_points_without_device():
    from mpcdl import _lib
    lib = _lib.lib
    assert lib.mpcdl_mppi_workspace_floats(50) >= 102 and lib.mpcdl_mppi_workspace_floats(0) == 0
    assert lib.mpcdl_mppi_reduce(None, None) == -1
    a = _lib.MppiArgs()
    for f in ("costs", "U", "workspace", "sums"):
        setattr(a, f, 16)
    a.N, a.H, a.lam = 10, 513, 1.0
    assert lib.mpcdl_mppi_reduce(C.byref(a), None) == -1                 # H beyond MPPI_MAX_H
    a.H, a.lam = 50, 0.0
    assert lib.mpcdl_mppi_reduce(C.byref(a), None) == -1 and b"lambda" in lib.mpcdl_last_error()
    assert lib.mpcdl_mppi_min(None, 5, None, None, None) == -1
    # sampled-control rollouts validate their bounds before any device work
    r = _lib.RolloutArgs()
    r.model = _lib.ModelDesc(_lib.MLP, 64, 4, 0)
    for f in ("wblob", "norm", "init", "U", "costs"):
        setattr(r, f, 16)
    r.N, r.H, r.sample_controls = 100, 10, 1
    r.ctrl_lo, r.ctrl_hi = (C.c_float * 2)(1.0, 0.0), (C.c_float * 2)(0.0, 1.0)
    p = _lib.Plan()
    assert lib.mpcdl_rollout_plan(C.byref(r), C.byref(p)) == -1 and b"ctrl_lo" in lib.mpcdl_last_error()


def test_ctypes_mirror_matches_the_compiled_header(tmp_path):
    """The ctypes structures must have the layout a C compiler gives include/mpcdl.h: compile a probe with gcc that
    prints sizeof / offsetof of every structure the binding mirrors and compare field by field."""
    import shutil
    import subprocess
    from mpcdl import _lib
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    structs = {"mpcdl_rollout_args": _lib.RolloutArgs, "mpcdl_noise_args": _lib.NoiseArgs, "mpcdl_mppi_args": _lib.MppiArgs,
               "mpcdl_plan": _lib.Plan, "mpcdl_model_desc": _lib.ModelDesc, "mpcdl_dataset_desc": _lib.DatasetDesc,
               "mpcdl_resample_args": _lib.ResampleArgs, "mpcdl_train_args": _lib.TrainArgs,
               "mpcdl_train_rnn_args": _lib.TrainRnnArgs}
    cfield = {"lam": "lambda"}                       # `lambda` is a Python keyword: the mirror calls it lam
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "mpcdl.h"', 'int main(void) {']
    for cname, cls in structs.items():
        lines.append(f'  printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'  printf("{cname}.{fname} %zu\\n", offsetof({cname}, {cfield.get(fname, fname)}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "probe.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-I", os.path.join(root, "include"), str(src), "-o", str(exe)], check=True)
    out = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for cname, cls in structs.items():
        assert int(out[cname]) == C.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert int(out[f"{cname}.{fname}"]) == getattr(cls, fname).offset, f"{cname}.{fname}"
