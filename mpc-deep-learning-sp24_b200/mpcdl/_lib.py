"""ctypes binding of the C ABI in include/mpcdl.h (libmpcdl.so, built by csrc/Makefile).

The library is the product: there is no Python / PyTorch / CPU fallback.  If the shared object is
missing or fails to load, importing this module raises immediately.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmpcdl.so")

MLP, GRU, LSTM = 0, 1, 2
PRECISION = {"fp32": 0, "tf32x3": 1, "tf32": 2, "bf16": 3}
ACT = {"identity": 0, "relu": 1, "tanh": 2, "sigmoid": 3}


class ModelDesc(C.Structure):
    _fields_ = [("kind", C.c_int32), ("hidden", C.c_int32), ("layers", C.c_int32), ("window", C.c_int32)]


class RolloutArgs(C.Structure):
    _fields_ = [
        ("model", ModelDesc),
        ("stateful", C.c_int32),
        ("wblob", C.c_void_p),
        ("norm", C.c_void_p),
        ("init", C.c_void_p),
        ("U", C.c_void_p),
        ("states", C.c_void_p),
        ("costs", C.c_void_p),
        ("cost_ref", C.c_void_p),
        ("cost_qw", C.c_float * 6),
        ("cost_rw", C.c_float * 2),
        ("noise_wblob", C.c_void_p),
        ("noise_history", C.c_int32),
        ("noise_size", C.c_int32),
        ("noise_layers", C.c_int32),
        ("noise_act", C.c_int32),
        ("eps", C.c_void_p),
        ("seed", C.c_uint64),
        ("traj_offset", C.c_int64),
        ("carry", C.c_void_p),
        ("N", C.c_int64),
        ("H", C.c_int32),
        ("tune_rt", C.c_int32),
        ("tune_ntg", C.c_int32),
        ("tune_nslots", C.c_int32),
        ("tune_kc", C.c_int32),
        ("precision", C.c_int32),
        ("wblob_tc", C.c_void_p),
        ("sample_controls", C.c_int32),
        ("shared_init", C.c_int32),
        ("ctrl_sigma", C.c_float * 2),
        ("ctrl_lo", C.c_float * 2),
        ("ctrl_hi", C.c_float * 2),
        ("ctrl_seed", C.c_uint64),
        ("scratch", C.c_void_p),
        ("scratch_bytes", C.c_uint64),
        ("n_models", C.c_int32),
        ("share_inputs", C.c_int32),
    ]


class DatasetDesc(C.Structure):
    _fields_ = [("rows", C.c_void_p), ("bag_row_off", C.c_void_p), ("samp_off", C.c_void_p),
                ("n_bags", C.c_int32), ("window", C.c_int32), ("n", C.c_int64)]


class ResampleArgs(C.Structure):
    _fields_ = [("t", C.c_void_p), ("q", C.c_void_p), ("u", C.c_void_p), ("M", C.c_int64), ("dataset_dt", C.c_double),
                ("step", C.c_int32), ("n_interp", C.c_int64), ("q_out", C.c_void_p), ("u_out", C.c_void_p)]


class TrainArgs(C.Structure):
    _fields_ = [("theta", C.c_void_p), ("adam_m", C.c_void_p), ("adam_v", C.c_void_p), ("featrs", C.c_void_p), ("labels", C.c_void_p),
                ("perm", C.c_void_p), ("losses", C.c_void_p), ("n_rows", C.c_int64), ("n_models", C.c_int32), ("n_samples", C.c_int32),
                ("batch", C.c_int32), ("hidden", C.c_int32), ("layers", C.c_int32), ("step0", C.c_int64),
                ("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double)]


class TrainRnnArgs(C.Structure):
    _fields_ = [("theta", C.c_void_p), ("adam_m", C.c_void_p), ("adam_v", C.c_void_p), ("featrs", C.c_void_p), ("labels", C.c_void_p),
                ("perm", C.c_void_p), ("losses", C.c_void_p), ("work", C.c_void_p), ("n_rows", C.c_int64), ("param_stride", C.c_int64),
                ("n_models", C.c_int32), ("n_samples", C.c_int32), ("batch", C.c_int32), ("kind", C.c_int32), ("hidden", C.c_int32),
                ("layers", C.c_int32), ("window", C.c_int32), ("step0", C.c_int64),
                ("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double)]


class MppiArgs(C.Structure):
    _fields_ = [
        ("costs", C.c_void_p),
        ("N", C.c_int64),
        ("traj_offset", C.c_int64),
        ("H", C.c_int32),
        ("U", C.c_void_p),
        ("ctrl_sigma", C.c_float * 2),
        ("ctrl_lo", C.c_float * 2),
        ("ctrl_hi", C.c_float * 2),
        ("ctrl_seed", C.c_uint64),
        ("lam", C.c_float),
        ("cmin", C.c_void_p),
        ("workspace", C.c_void_p),
        ("sums", C.c_void_p),
    ]


class NoiseArgs(C.Structure):
    _fields_ = [
        ("noise_wblob", C.c_void_p),
        ("history", C.c_int32), ("size", C.c_int32), ("n_layers", C.c_int32), ("act", C.c_int32),
        ("closed_loop", C.c_int32),
        ("q", C.c_void_p), ("u", C.c_void_p), ("eps", C.c_void_p),
        ("seed", C.c_uint64),
        ("traj_offset", C.c_int64),
        ("step0", C.c_int32),
        ("out", C.c_void_p), ("mu", C.c_void_p), ("scale", C.c_void_p),
        ("N", C.c_int64),
        ("H", C.c_int32),
    ]


class Plan(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("rt", "rj", "n_tg", "tile_traj", "threads", "grid", "smem_bytes",
                                         "resident", "nslots", "slot_floats", "ntiles", "variant", "ksplit")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


EXPORTS = ("mpcdl_blob_floats", "mpcdl_pack_mlp", "mpcdl_pack_rnn", "mpcdl_blob_floats_tc", "mpcdl_pack_rnn_tc", "mpcdl_pack_mlp_tc", "mpcdl_noise_blob_floats", "mpcdl_pack_noise",
           "mpcdl_rollout", "mpcdl_rollout_static_scratch", "mpcdl_rollout_scratch_bytes", "mpcdl_rollout_plan", "mpcdl_noise_forward", "mpcdl_error_metrics", "mpcdl_mppi_workspace_floats", "mpcdl_mppi_min", "mpcdl_mppi_reduce", "mpcdl_ffma_peak", "mpcdl_ffma2_peak", "mpcdl_last_error", "mpcdl_version",
           "mpcdl_dataset_stats", "mpcdl_dataset_build", "mpcdl_rollout_inputs", "mpcdl_curve_metrics", "mpcdl_resample",
           "mpcdl_train_param_count", "mpcdl_train_mlp_epoch", "mpcdl_train_rnn_param_count", "mpcdl_train_rnn_work_floats",
           "mpcdl_train_rnn_epoch", "mpcdl_l1_batches", "mpcdl_cdr_decode",
           "mpcdl_debug_poison_smem")


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C mpc-deep-learning-sp24_b200/csrc` "
                          "(or __graft_entry__.build()); mpcdl has no fallback path")
    lib = C.CDLL(LIB_PATH)
    fpp = C.POINTER(C.POINTER(C.c_float))
    fp = C.POINTER(C.c_float)
    lib.mpcdl_blob_floats.restype = C.c_size_t
    lib.mpcdl_blob_floats.argtypes = [C.POINTER(ModelDesc)]
    lib.mpcdl_pack_mlp.restype = C.c_int
    lib.mpcdl_pack_mlp.argtypes = [C.POINTER(ModelDesc), fpp, fpp, fp]
    lib.mpcdl_pack_rnn.restype = C.c_int
    lib.mpcdl_pack_rnn.argtypes = [C.POINTER(ModelDesc), fpp, fpp, fpp, fpp, fp, fp, fp]
    lib.mpcdl_blob_floats_tc.restype = C.c_size_t
    lib.mpcdl_blob_floats_tc.argtypes = [C.POINTER(ModelDesc)]
    lib.mpcdl_pack_rnn_tc.restype = C.c_int
    lib.mpcdl_pack_rnn_tc.argtypes = [C.POINTER(ModelDesc), fpp, fpp, fpp, fpp, fp, fp, fp]
    lib.mpcdl_pack_mlp_tc.restype = C.c_int
    lib.mpcdl_pack_mlp_tc.argtypes = [C.POINTER(ModelDesc), fpp, fpp, fp]
    lib.mpcdl_noise_blob_floats.restype = C.c_size_t
    lib.mpcdl_noise_blob_floats.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.mpcdl_pack_noise.restype = C.c_int
    lib.mpcdl_pack_noise.argtypes = [C.c_int, C.c_int, C.c_int, fpp, fpp, fp]
    lib.mpcdl_rollout.restype = C.c_int
    lib.mpcdl_rollout.argtypes = [C.POINTER(RolloutArgs), C.c_void_p]
    lib.mpcdl_rollout_static_scratch.restype = C.c_int
    lib.mpcdl_rollout_static_scratch.argtypes = [C.POINTER(RolloutArgs), C.c_void_p]
    lib.mpcdl_rollout_scratch_bytes.restype = C.c_size_t
    lib.mpcdl_rollout_scratch_bytes.argtypes = [C.POINTER(RolloutArgs)]
    lib.mpcdl_rollout_plan.restype = C.c_int
    lib.mpcdl_rollout_plan.argtypes = [C.POINTER(RolloutArgs), C.POINTER(Plan)]
    lib.mpcdl_noise_forward.restype = C.c_int
    lib.mpcdl_noise_forward.argtypes = [C.POINTER(NoiseArgs), C.c_void_p]
    lib.mpcdl_mppi_workspace_floats.restype = C.c_size_t
    lib.mpcdl_mppi_workspace_floats.argtypes = [C.c_int]
    lib.mpcdl_mppi_min.restype = C.c_int
    lib.mpcdl_mppi_min.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpcdl_mppi_reduce.restype = C.c_int
    lib.mpcdl_mppi_reduce.argtypes = [C.POINTER(MppiArgs), C.c_void_p]
    lib.mpcdl_error_metrics.restype = C.c_int
    lib.mpcdl_error_metrics.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpcdl_ffma_peak.restype = C.c_int
    lib.mpcdl_ffma_peak.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.mpcdl_train_param_count.restype = C.c_int64
    lib.mpcdl_train_param_count.argtypes = [C.c_int32, C.c_int32]
    lib.mpcdl_train_rnn_param_count.restype = C.c_int64
    lib.mpcdl_train_rnn_param_count.argtypes = [C.c_int32, C.c_int32, C.c_int32]
    lib.mpcdl_train_rnn_work_floats.restype = C.c_int64
    lib.mpcdl_train_rnn_work_floats.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    lib.mpcdl_train_rnn_epoch.restype = C.c_int
    lib.mpcdl_train_rnn_epoch.argtypes = [C.POINTER(TrainRnnArgs), C.c_void_p]
    lib.mpcdl_train_mlp_epoch.restype = C.c_int
    lib.mpcdl_train_mlp_epoch.argtypes = [C.POINTER(TrainArgs), C.c_void_p]
    lib.mpcdl_l1_batches.restype = C.c_int
    lib.mpcdl_l1_batches.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]
    lib.mpcdl_cdr_decode.restype = C.c_int
    lib.mpcdl_cdr_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpcdl_ffma2_peak.restype = C.c_int
    lib.mpcdl_ffma2_peak.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.mpcdl_dataset_stats.restype = C.c_int
    lib.mpcdl_dataset_stats.argtypes = [C.POINTER(DatasetDesc), C.c_void_p, C.c_void_p]
    lib.mpcdl_dataset_build.restype = C.c_int
    lib.mpcdl_dataset_build.argtypes = [C.POINTER(DatasetDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpcdl_rollout_inputs.restype = C.c_int
    lib.mpcdl_rollout_inputs.argtypes = [C.POINTER(DatasetDesc), C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpcdl_curve_metrics.restype = C.c_int
    lib.mpcdl_curve_metrics.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.mpcdl_resample.restype = C.c_int
    lib.mpcdl_resample.argtypes = [C.POINTER(ResampleArgs), C.c_void_p]
    lib.mpcdl_debug_poison_smem.restype = C.c_int
    lib.mpcdl_debug_poison_smem.argtypes = [C.c_void_p]
    lib.mpcdl_last_error.restype = C.c_char_p
    lib.mpcdl_version.restype = C.c_int
    return lib


lib = _load()


def check(rc: int):
    """Status -> exception, mirroring the reference's error behaviour (torch raises RuntimeError on
    shape/device mismatch)."""
    if rc != 0:
        raise RuntimeError(f"mpcdl error {rc}: {lib.mpcdl_last_error().decode()}")
