"""Drop-in for model_scripts/eval_models.py: the one-step evaluation harness on the fused kernels.

Same names and arguments as the reference (``eval_basic``, ``eval_deep``, ``calc_error``, ``build_test``,
``SAMPLE``, ``EPOCHS``, ``params``, ``result``); what changes is the execution: the reference runs
``model(featrs)`` 10 000 times at batch 1 (eval_models.py:72-92), here the sampled rows are gathered
once, ONE forward launch evaluates them (``Module.forward`` = the rollout kernel with H = 1) and ONE
launch of ``mpcdl_error_metrics`` produces the per-sample MAE / RMSE, which are then accumulated on the host in
the reference's order (Python floats, sample order).  The sample indices come from the same
``random.randint`` calls, so a seeded run evaluates the same rows as the reference.

Differences kept deliberately: ``calc_error`` also RETURNS ``(mae, rmse)`` (the reference only prints);
``view_test`` writes the collected series to ``eval_test.npz`` (no display stack / matplotlib here).
"""
from __future__ import annotations

import os
import random

import torch

from . import basic_model as fwd
from . import deep_models as rnn

SAMPLE = True
EPOCHS = 10000
device = torch.device("cuda" if torch.cuda.is_available() else "cpu")   # the reference evaluates on CPU (:11)

FWD_DIR = "../model_data/fwd_models"      # eval_models.py:18 (listing) / :28 (loading, relative to model_data/)
RNN_DIR = "../model_data/rnn_models"


def _load_state(path):
    if path.endswith(".npz"):
        import numpy as np
        return {k: torch.from_numpy(v) for k, v in np.load(path).items()}
    return torch.load(path, map_location="cpu")


def _files(listing, loading):
    d = listing if os.path.isdir(listing) else loading
    return d, sorted(f for f in os.listdir(d) if f.endswith((".pt", ".npz")))


def eval_basic(one_model=True, model_dir=None):
    """eval_models.py:14-35.  Returns {file: (mae, rmse)} (the reference returns None)."""
    curr_h, out = 0, {}
    dataset = fwd.RobotData(fwd.DATA)
    d, files = _files(model_dir or FWD_DIR, "fwd_models")
    for file in files:
        name = file.split("_")
        model_type = getattr(fwd, f"{name[0]}_Model")
        hidden, layers = int(name[1]), int(name[2])
        model = model_type(fwd.INPUTS, hidden, fwd.OUTPUT, layers, False)
        model.load_state_dict(_load_state(os.path.join(d, file)))
        model.to(device)
        if curr_h != hidden:
            print("\n=======================")
            curr_h = hidden
        out[file] = calc_error(dataset, model, file)
        if one_model:
            return out
    return out


def eval_deep(one_model=True, model_dir=None):
    """eval_models.py:37-65 (the dataset is rebuilt only when ft_size really changes)."""
    curr_h, curr_ft, out, dataset = 0, 0, {}, None
    d, files = _files(model_dir or RNN_DIR, "rnn_models")
    for file in files:
        name = file.split("_")
        model_type = getattr(rnn, f"{name[0]}_Model")
        hidden, layers, ft_size = int(name[1]), int(name[2]), int(name[3])
        state_f = name[-1].rsplit(".", 1)[0] == "True"
        model = model_type(rnn.INPUTS, hidden, rnn.OUTPUT, layers, state_f)
        model.load_state_dict(_load_state(os.path.join(d, file)))
        model.to(device)
        if curr_h != hidden:
            print("\n=======================")
            curr_h = hidden
        if curr_ft != ft_size:
            dataset = rnn.RobotData(rnn.DATA, ft_size)
            curr_ft = ft_size
        out[file] = calc_error(dataset, model, file, state_f)
        if one_model:
            return out
    return out


def calc_error(dataset, model, file, state_f=False):
    """eval_models.py:67-96, batched: two kernel launches instead of `epochs` batch-1 forwards."""
    epochs = EPOCHS if SAMPLE else len(dataset)
    idx = [random.randint(0, len(dataset) - 1) for _ in range(epochs)] if SAMPLE else list(range(epochs))
    dev = next(model.parameters()).device
    sel = torch.as_tensor(idx, dtype=torch.long)
    featrs = dataset.featrs.cpu()[sel].to(dev)
    labels = dataset.labels.cpu()[sel].to(dev)
    if state_f:
        # the reference calls init_hidden(1) before EVERY sample (:81-82): each sample starts from a zero state
        model.init_hidden(featrs.size(0), dev)
    model.eval()
    with torch.no_grad():
        output = model(featrs)
    mae, rmse = torch.ops.mpcdl.error_metrics(output.contiguous(), labels.contiguous())
    lab, outp, mae, rmse = labels.cpu(), output.cpu(), mae.cpu().tolist(), rmse.cpu().tolist()
    for key, col in (("x_pos", 0), ("y_pos", 1), ("v_long", 4), ("v_tran", 5)):   # build_test, :109-120
        params[key].extend(lab[:, col].tolist())
        result[key].extend(outp[:, col].tolist())
    diff = rm = 0.0
    for a, b in zip(mae, rmse):            # same accumulation order and type as :91-92
        diff += a
        rm += b
    print(f"\n{file}")
    print(f"Avg. difference: {(diff / epochs):.4f}")
    print(f"Avg. rmsq error: {(rm / epochs):.4f}")
    return diff / epochs, rm / epochs


# ------------------------------------------------------------------------------------------------ validation sweep (f2)
def _sweep_checkpoints(model_dirs):
    """{name: state_dict} of every checkpoint under the given directories (default: the reference's two, :18 / :41)."""
    out = {}
    for d in model_dirs:
        d2, files = _files(d, os.path.basename(d))
        for f in files:
            out[f.rsplit(".", 1)[0]] = _load_state(os.path.join(d2, f))
    return out


def _pad_mlp_state(sd, hidden, layers, to_layers):
    """Basic_Model state_dict with `layers` hidden layers -> an equivalent one with `to_layers`: the extra layers are
    identity matrices with zero bias placed before the output layer.  Their inputs are ReLU outputs (>= 0), so
    relu(I x + 0) = x exactly (every added FMA term is x*1 or x*0 onto 0): the padded model computes the same bits."""
    if layers == to_layers:
        return {k: torch.as_tensor(v) for k, v in sd.items()}
    out = {f"layers.{i}.{p}": torch.as_tensor(sd[f"layers.{i}.{p}"]) for i in range(layers + 1) for p in ("weight", "bias")}
    for i in range(layers + 1, to_layers + 1):
        out[f"layers.{i}.weight"], out[f"layers.{i}.bias"] = torch.eye(hidden), torch.zeros(hidden)
    out[f"layers.{to_layers + 1}.weight"] = torch.as_tensor(sd[f"layers.{layers + 1}.weight"])
    out[f"layers.{to_layers + 1}.bias"] = torch.as_tensor(sd[f"layers.{layers + 1}.bias"])
    return out


def sweep(checkpoints=None, pickles=("mpc", "pid"), horizons=(1, 10, 50), device=None, model_dirs=(FWD_DIR, RNN_DIR)):
    """The whole validation sweep as ONE call: every checkpoint x every validation pickle x every
    horizon.  The reference runs ``calc_error`` per checkpoint at batch 1 and only one step deep
    (model_scripts/eval_models.py:14-96); here, per pickle and per model family (MLP hidden 16, MLP hidden 64, GRU, LSTM)
    the checkpoints' packed weights are STACKED and one fused-rollout launch evaluates all of them on all samples:

      H = 1   teacher forced: the model's prediction on every dataset sample against its label (``calc_error`` with
              SAMPLE = False) -> (mae, rmse) in the pickle's own z-score units
      H > 1   open loop from every stretch of the bag that leaves H recorded frames ahead, fed with the recorded
              controls -> per-step curves (mae [H], rmse [H]) of the predicted against the recorded states, same metric

    checkpoints: {name: state_dict} (names follow the file-name schema, eval_models.py:19-25 / 43-52) or None = every
    file under `model_dirs`.  pickles: names resolved like ``basic_model.DATA`` ("mpc" -> mpc_data_clean) or a
    {label: bags} dict.  GRU / LSTM checkpoints are evaluated stateless, as ``calc_error`` does (init_hidden before
    every sample, :81-82).  Dataset build, statistics, evaluation inputs, rollout and metrics all run on the GPU.
    Returns {pickle: {name: {H: (mae, rmse)}}} with numpy values."""
    from . import _lib, data, ops, pack, synth
    dev = torch.device(device if device is not None else "cuda")
    ckpts = checkpoints if checkpoints is not None else _sweep_checkpoints(model_dirs)
    if not isinstance(pickles, dict):
        pickles = {p: synth.load_bags(f"{p}_data_clean") for p in pickles}
    groups = {}                                                        # (kind, hidden, window) -> [(name, layers, sd)]
    for name, sd in ckpts.items():
        p = name.split("_")
        kind = {"Basic": _lib.MLP, "GRU": _lib.GRU, "LSTM": _lib.LSTM}[p[0]]
        groups.setdefault((kind, int(p[1]), 0 if kind == _lib.MLP else int(p[3])), []).append((name, int(p[2]), sd))
    out = {}
    for label, bags in pickles.items():
        db = data.DeviceBags(bags, dev)
        res = out.setdefault(label, {})
        for (kind, hidden, window), members in groups.items():
            lay = max(m[1] for m in members) if kind == _lib.MLP else members[0][1]
            if kind != _lib.MLP and any(m[1] != lay for m in members):
                raise RuntimeError("mpcdl: stacked GRU / LSTM checkpoints must share the layer count")
            blobs = [pack.pack_mlp(_pad_mlp_state(sd, hidden, L, lay), hidden, lay) if kind == _lib.MLP
                     else pack.pack_rnn({k: torch.as_tensor(v) for k, v in sd.items()}, kind, hidden, lay, window) for _, L, sd in members]
            wblob = torch.cat(blobs).to(dev)
            M = len(members)
            featrs, labels, svec = data.device_dataset(db, window or None, dev)
            ident = pack.identity_norm(window).to(dev)

            def launch(norm, init, U):
                return ops.OPS[kind](wblob, norm, init, U, hidden, lay, window, False, None, None, [], None, 0, 0, None, [],
                                     True, False, [], 0, None, [], 0, 0, False, None, None, M, True)[0]
            for H in horizons:
                if H == 1:
                    if kind == _lib.MLP:
                        init, U = featrs[:, :6].contiguous(), featrs[:, None, 6:8].contiguous()
                    else:
                        init, U = featrs, torch.zeros(featrs.shape[0], 1, 2, device=dev)
                    pred, target, lstat, norm = launch(ident, init, U), labels[:, None, :].contiguous(), ident[-12:], None
                else:
                    init, U, target = data.open_loop_inputs(db, H, window or None, dev)
                    pred, lstat = launch(svec, init, U), svec[-12:]
                mae, rmse = torch.ops.mpcdl.curve_metrics(pred.view(M, -1, H, 6), target, lstat.contiguous())
                mae, rmse = mae.cpu().numpy(), rmse.cpu().numpy()
                for i, (name, _, _) in enumerate(members):
                    res.setdefault(name, {})[H] = (float(mae[i, 0]), float(rmse[i, 0])) if H == 1 else (mae[i], rmse[i])
    return out


params = {"x_pos": [], "y_pos": [], "v_long": [], "v_tran": []}
result = {"x_pos": [], "y_pos": [], "v_long": [], "v_tran": []}


def build_test(in_arr, out_dict):
    """eval_models.py:109-120."""
    out_dict["x_pos"].append(in_arr[0])
    out_dict["y_pos"].append(in_arr[1])
    out_dict["v_long"].append(in_arr[4])
    out_dict["v_tran"].append(in_arr[5])


def view_test(path="eval_test.npz"):
    """eval_models.py:123-133 plots prediction vs actual; here the same series are written to `path`."""
    import numpy as np
    np.savez(path, **{f"prediction_{k}": np.asarray(v, dtype=np.float32) for k, v in result.items()},
             **{f"actual_{k}": np.asarray(v, dtype=np.float32) for k, v in params.items()})
    return path


if __name__ == "__main__":
    import sys
    if len(sys.argv) < 2 or sys.argv[1] not in ("eval_basic", "eval_deep"):
        print("Usage: python -m mpcdl.eval_models eval_basic|eval_deep [one_model True|False]")
    else:
        one = len(sys.argv) < 3 or sys.argv[2] == "True"
        (eval_basic if sys.argv[1] == "eval_basic" else eval_deep)(one)
        view_test()
