"""mpcdl -- fused B200 (sm_100a) rollout kernels behind the reference's nn.Module call surface.

Importing the package loads libmpcdl.so (ctypes) and registers torch.ops.mpcdl.*; it raises if the
library has not been built (no fallback).
"""
from . import _lib, ops, pack, data, data_utils, synth, parallel, train, rosbags  # noqa: F401
from .models import (Basic_Model, GRU_Model, LSTM_Model, FeedforwardNoiseModel, build_mlp,   # noqa: F401
                     CasadiDynamicCLBicycleNoise, DynamicsNN, from_checkpoint_name, load_model)
from .data import RobotData, zscore_stats              # noqa: F401

INPUTS, OUTPUT = 8, 6                                  # basic_model.py:63-64, deep_models.py:94-95
__all__ = ["Basic_Model", "GRU_Model", "LSTM_Model", "FeedforwardNoiseModel", "CasadiDynamicCLBicycleNoise", "DynamicsNN", "RobotData", "zscore_stats",
           "from_checkpoint_name", "load_model", "ops", "pack", "data", "synth", "parallel"]
