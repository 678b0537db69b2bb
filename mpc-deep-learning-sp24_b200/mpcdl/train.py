"""Training of the forward-dynamics MLP on the GPU: ``train_model`` of model_scripts/basic_model.py:76-139
with every EPOCH as one kernel launch (csrc/train_kernels.cu: forward, nn.L1Loss, backward and the torch.optim.Adam step of
all batches fused, parameters and gradients in shared memory; one CTA per model, so an ensemble or the reference's
hidden x layers grid of one family trains side by side).

    model = mpcdl.Basic_Model(8, 16, 6, 4, False).cuda()
    hist = mpcdl.train.train_model(model, dataset, epochs=100, batch_s=16, learn_r=1e-3, seed=0)

keeps the reference's loop: 80 / 20 random_split, a fresh shuffle of the training samples every epoch, Adam(lr), L1 loss,
validation loss = mean of the per-batch L1 losses, ReduceLROnPlateau(default settings) on it, best state kept, the
"Early stop triggered." / "Bad loss" messages.  What differs is where the random numbers come from: the sample orders are
drawn with an explicit torch.Generator (the reference uses the global RNG through DataLoader), so a run is reproducible
and the oracle (oracle/train_ref.py) can be handed the same orders.  Dropout (do_drop=True) is not implemented in the
fused step and raises.

Recurrent models (round 2): ``train_model`` of model_scripts/deep_models.py:106-146 for GRU_Model / LSTM_Model runs on
csrc/train_rnn_kernels.cu (forward over the window, L1 loss, back-propagation through time, Adam; one CTA per model).  The
reference's loop differs from the MLP's in two places, both kept: the DataLoader does NOT shuffle (``shuffle=False``, :112),
so every epoch walks the random_split subset in the same order; and ``init_hidden`` runs before every batch when STATE_F
(:118-119), so a training batch always starts from h = 0 whether the model is stateful or not.  Batches of more than 16
windows (the shipped checkpoints: BATCH_S 16 and 64) run as micro-batches of 16 whose gradients are summed before the Adam step."""
from __future__ import annotations

import torch

from . import _lib  # noqa: F401  (loads libmpcdl.so; raises if it is missing)


def _flat(model):
    ps = [p for p in model.parameters()]
    return torch.cat([p.detach().reshape(-1) for p in ps]).float().contiguous(), ps


def _unflat(vec, ps):
    with torch.no_grad():
        o = 0
        for p in ps:
            n = p.numel()
            p.copy_(vec[o:o + n].view_as(p))
            o += n


class AdamState:
    """exp_avg / exp_avg_sq / step of torch.optim.Adam for the flattened parameter vectors of n_models models"""

    def __init__(self, n_models, n_params, device):
        self.n_params = n_params
        self.m = torch.zeros(n_models, n_params, device=device)
        self.v = torch.zeros(n_models, n_params, device=device)
        self.step = 0
        self._work = None                         # scratch of the recurrent kernel, kept between epochs


def _check_model(model):
    """-> ("mlp", hidden, layers) | ("gru" | "lstm", hidden, layers)"""
    from .models import Basic_Model, GRU_Model, LSTM_Model
    if isinstance(model, GRU_Model):
        return "gru", model.hidden, model.layers
    if isinstance(model, LSTM_Model):
        return "lstm", model.hidden, model.layers
    if not isinstance(model, Basic_Model):
        raise RuntimeError("mpcdl.train: the fused training step exists for Basic_Model (basic_model.py) and GRU_Model / LSTM_Model "
                           "(deep_models.py); got " + type(model).__name__)
    if any(isinstance(l, torch.nn.Dropout) for l in model.layers):
        raise RuntimeError("mpcdl.train: do_drop=True (Dropout between the layers) is not implemented in the fused training step")
    hidden, layers, _ = model._desc()
    return "mlp", hidden, layers


def param_count(model):
    kind, hidden, layers = _check_model(model)
    if kind == "mlp":
        return _lib.lib.mpcdl_train_param_count(hidden, layers)
    return _lib.lib.mpcdl_train_rnn_param_count(_lib.GRU if kind == "gru" else _lib.LSTM, hidden, layers)


def train_epoch(models, state, featrs, labels, perm, batch_s=16, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
    """One epoch for one model or a list of same-shaped models in ONE launch.  featrs [n,8] / labels [n,6] CUDA tensors
    (RobotData.featrs / .labels), perm int32 [n_models, n_samples] (or [n_samples]): row index of every training sample in
    the order the batches take them.  Updates the modules' parameters and `state` in place; returns the per-batch L1 losses
    [n_models, n_batches] (what ``loss.item()`` would give, basic_model.py:106)."""
    models = list(models) if isinstance(models, (list, tuple)) else [models]
    kind, hidden, layers = _check_model(models[0])
    for m in models[1:]:
        if _check_model(m) != (kind, hidden, layers):
            raise RuntimeError("mpcdl.train: the models of one launch must share family, hidden and layers")
    dev = featrs.device
    flats = [_flat(m) for m in models]
    perm = perm.to(device=dev, dtype=torch.int32)
    if perm.dim() == 1:
        perm = perm[None].expand(len(models), -1)
    perm = perm.contiguous()
    if kind == "mlp":
        theta = torch.stack([f[0].to(dev) for f in flats]).contiguous()
        losses = torch.ops.mpcdl.train_mlp_epoch(theta, state.m, state.v, featrs.contiguous(), labels.contiguous(), perm, hidden, layers,
                                                 int(batch_s), int(state.step), float(lr), float(betas[0]), float(betas[1]), float(eps))
    else:
        if featrs.dim() != 3:
            raise RuntimeError("mpcdl.train: a recurrent model trains on windows [n, ft_size, 8] (deep_models.RobotData.featrs)")
        k = _lib.GRU if kind == "gru" else _lib.LSTM
        n_params = state.n_params
        stride = (n_params + 3) // 4 * 4          # the kernel reads a model's vectors as float4: one model every 4k floats
        theta, am, av = (torch.zeros(len(models), stride, device=dev) for _ in range(3))
        theta[:, :n_params] = torch.stack([f[0].to(dev) for f in flats])
        am[:, :n_params], av[:, :n_params] = state.m, state.v
        wf = _lib.lib.mpcdl_train_rnn_work_floats(k, hidden, layers, int(featrs.shape[1]))
        if state._work is None or state._work.numel() < len(models) * wf:
            state._work = torch.empty(len(models) * wf, device=dev)
        losses = torch.ops.mpcdl.train_rnn_epoch(theta, am, av, featrs.contiguous(), labels.contiguous(), perm, state._work, k,
                                                 hidden, layers, int(batch_s), int(state.step), float(lr), float(betas[0]), float(betas[1]),
                                                 float(eps))
        theta = theta[:, :n_params]
        state.m.copy_(am[:, :n_params]); state.v.copy_(av[:, :n_params])
    state.step += (perm.shape[1] + batch_s - 1) // batch_s
    for (_, ps), row in zip(flats, theta):
        _unflat(row, ps)
    return losses


def validation_loss(model, featrs, labels, batch_s=16):
    """v_loss of basic_model.py:112-122: mean over the batches of consecutive samples of the per-batch L1 loss"""
    with torch.no_grad():
        if getattr(model, "state_f", False):
            model.init_hidden(featrs.shape[0], featrs.device)       # deep_models.py:136-137: zeros before every validation batch
        out = model(featrs)
    return torch.ops.mpcdl.l1_batches(out.contiguous(), labels.contiguous(), int(batch_s)).mean().item()


class _Plateau:
    """torch.optim.lr_scheduler.ReduceLROnPlateau with its default settings (mode min, factor 0.1, patience 10, rel
    threshold 1e-4, cooldown 0, min_lr 0, eps 1e-8), on a plain float learning rate"""

    def __init__(self, lr):
        self.lr, self.best, self.bad = lr, float("inf"), 0

    def step(self, metric):
        if metric < self.best * (1.0 - 1e-4):
            self.best, self.bad = metric, 0
        else:
            self.bad += 1
        if self.bad > 10:
            new = self.lr * 0.1
            if self.lr - new > 1e-8:
                self.lr = new
            self.bad = 0
        return self.lr


def split_indices(n, generator):
    """random_split(dataset, [int(0.8 n), n - int(0.8 n)]) with an explicit generator (basic_model.py:82-85)"""
    t_size = int(n * 0.8)
    p = torch.randperm(n, generator=generator)
    return p[:t_size], p[t_size:]


def train_model(model, dataset, epochs=100, batch_s=16, learn_r=1e-3, seed=0, path=None, verbose=True, bad_loss=0.1):
    """basic_model.train_model (model_scripts/basic_model.py:76-139) on the GPU.  `dataset` = mpcdl.RobotData (or anything
    with .featrs / .labels).  Returns dict(valid_loss=[...], train_loss=[...], lr=[...], best=state_dict, code=None | 1);
    code 1 = the reference's `return 1` (an epoch ended with valid_loss > bad_loss = 0.1; its caller re-initialises and
    tries again: `while train_model(Basic_Model): continue`, basic_model.py:143)."""
    dev = next(model.parameters()).device
    if dev.type != "cuda":
        raise RuntimeError("mpcdl.train: the model must live on a CUDA device (no CPU fallback)")
    featrs, labels = dataset.featrs.to(dev).float().contiguous(), dataset.labels.to(dev).float().contiguous()
    g = torch.Generator().manual_seed(seed)
    t_idx, v_idx = split_indices(featrs.shape[0], g)
    vf, vl = featrs[v_idx.to(dev)].contiguous(), labels[v_idx.to(dev)].contiguous()
    kind, hidden, layers = _check_model(model)
    state = AdamState(1, param_count(model), dev)
    sched = _Plateau(learn_r)
    lr, min_loss, patience = learn_r, float("inf"), 0
    hist = dict(valid_loss=[], train_loss=[], lr=[], best=None, code=None)
    for i in range(epochs):
        # basic_model.py:84 DataLoader(shuffle=True): a new order every epoch; deep_models.py:112 shuffle=False: the subset's own order
        order = t_idx[torch.randperm(t_idx.numel(), generator=g)] if kind == "mlp" else t_idx
        losses = train_epoch(model, state, featrs, labels, order, batch_s, lr)
        v_loss = validation_loss(model, vf, vl, batch_s)
        lr = sched.step(v_loss)
        hist["valid_loss"].append(v_loss); hist["train_loss"].append(losses.mean().item()); hist["lr"].append(lr)
        if verbose:
            print(f"Epoch {i + 1}: valid_loss = {v_loss:.4f}")
        if v_loss < min_loss:
            min_loss, patience = v_loss, 0
            hist["best"] = {k: v.detach().clone() for k, v in model.state_dict().items()}
            if path is not None:
                torch.save(hist["best"], path)
        else:
            patience += 1
        if patience > 10 and verbose:
            print("Early stop triggered.")                   # the reference prints and trains on (basic_model.py:133-134)
        if v_loss > bad_loss:                                # ... and gives up at ANY epoch whose validation loss is above 0.1 (:135-137)
            if verbose:
                print("Bad loss. Restarting:")
            hist["code"] = 1
            return hist
    return hist
