"""Drop-in for ``import basic_model as fwd`` (model_scripts/basic_model.py): the names eval_models.py uses --
``Basic_Model``, ``RobotData``, ``INPUTS``, ``OUTPUT``, ``DATA``, ``dataset`` (eval_models.py:16,22-23) -- backed
by the fused kernels.  ``DATA`` / ``dataset`` are loaded on first access (see _datafile.py); the training grid of
the reference's ``__main__`` (basic_model.py:75-144) is out of scope."""
from __future__ import annotations

from . import _datafile
from .data import RobotData as _RobotData
from .models import Basic_Model  # noqa: F401

INPUTS = 8      # [pos.(2), orient.(1), vel.(3), acceleration (2)]   basic_model.py:63
OUTPUT = 6      # [positions (2), orientation (1), velocities (3)]   basic_model.py:64
HIDDEN, LAYERS, BATCH_S, LEARN_R, DO_DROP, EPOCHS = 16, 4, 16, 0.001, False, 100   # basic_model.py:65-71


class RobotData(_RobotData):
    """basic_model.RobotData(data): features bag[i], labels bag[i+1][:6], z-scored (basic_model.py:11-41)."""

    def __init__(self, data):
        super().__init__(data, None)


def __getattr__(name):
    if name == "DATA":
        globals()["DATA"] = _datafile.load_data()
        return globals()["DATA"]
    if name == "dataset":
        globals()["dataset"] = RobotData(__getattr__("DATA") if "DATA" not in globals() else globals()["DATA"])
        return globals()["dataset"]
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
