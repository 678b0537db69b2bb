"""Drop-in for ``import deep_models as rnn`` (model_scripts/deep_models.py): ``GRU_Model``, ``LSTM_Model``,
``RobotData(data, size)``, ``INPUTS``, ``OUTPUT``, ``FT_SIZE``, ``DATA``, ``dataset`` (eval_models.py:40,46-47,62)
backed by the fused kernels; ``DATA`` / ``dataset`` load on first access.  Training (deep_models.py:106-189) is out
of scope."""
from __future__ import annotations

from . import _datafile
from .data import RobotData as _RobotData
from .models import GRU_Model, LSTM_Model  # noqa: F401

INPUTS = 8      # deep_models.py:94
OUTPUT = 6      # deep_models.py:95
HIDDEN, LAYERS, FT_SIZE, BATCH_S, LEARN_R, STATE_F = 16, 4, 4, 16, 0.001, False   # deep_models.py:96-102


class RobotData(_RobotData):
    """deep_models.RobotData(data, size): windows bag[i:i+size], labels bag[i+size][:6], per-position z-score
    (deep_models.py:11-42)."""

    def __init__(self, data, size):
        super().__init__(data, size)


def __getattr__(name):
    if name == "DATA":
        globals()["DATA"] = _datafile.load_data()
        return globals()["DATA"]
    if name == "dataset":
        data = globals()["DATA"] if "DATA" in globals() else __getattr__("DATA")
        globals()["dataset"] = RobotData(data, FT_SIZE)
        return globals()["dataset"]
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
