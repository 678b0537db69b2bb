"""Where the module-level ``DATA`` of basic_model / deep_models comes from.

The reference unpickles ``../mpc_data_clean.pkl`` at import time (basic_model.py:9, deep_models.py:9), which
ties importing to the working directory.  Here the load is lazy (first access of ``DATA`` / ``dataset``) and
looks, in order, at ``$MPCDL_DATA``, ``../mpc_data_clean.pkl`` (the reference's location) and the bundled
validation bag (tests/golden/data/mpc_data_clean.npy, the same rows exported as an array)."""
from __future__ import annotations

import os
import pickle


def load_data():
    cand = [os.environ.get("MPCDL_DATA"), "../mpc_data_clean.pkl"]
    for p in cand:
        if p and os.path.exists(p):
            if p.endswith(".npy"):
                import numpy as np
                return [np.load(p).tolist()]
            with open(p, "rb") as f:
                return pickle.load(f)
    from . import synth
    return synth.load_bags("mpc_data_clean")
