"""CPU restatement of the reference's training step for the forward-dynamics MLP -- TEST INFRASTRUCTURE ONLY (imported by
tests/ and oracle/pin_reference.py, never by the mpcdl package).

    reference: train_model  model_scripts/basic_model.py:76-110
        optimizer = optim.Adam(model.parameters(), lr=LEARN_R); criterion = nn.L1Loss()
        for featrs, labels in t_load:     # DataLoader(batch_size=BATCH_S, shuffle=True)
            optimizer.zero_grad(); output = model(featrs); loss = criterion(output, labels); loss.backward(); optimizer.step()
    validation loss: :112-122 (sum of loss.item() over v_load / len(v_load))

The arithmetic lives in torch (nn.Linear / relu autograd, nn.L1Loss, torch.optim.Adam: an unpinned third-party dependency of
the reference; installed torch is the oracle).  The only thing restated is WHICH samples form a batch: the
reference lets DataLoader draw the order from the global RNG, here the order is an explicit index vector so that the GPU
kernel can be given the same one.  Pinned by oracle/pin_reference.py section 7: the reference's own Basic_Model class,
optimizer and criterion stepped over the same batches give bit-identical parameters.

Recurrent models (train_steps_rnn):
    reference: train_model  model_scripts/deep_models.py:106-146
        t_load = DataLoader(t_data, batch_size=BATCH_S, shuffle=False)
        for featrs, labels in t_load:
            if STATE_F: model.init_hidden(batch_s)          # zeros before EVERY batch
            optimizer.zero_grad(); output = model(featrs); loss = criterion(output, labels); loss.backward(); optimizer.step()
    model(featrs) with a zero hidden state = GRU_Model / LSTM_Model.forward (deep_models.py:60-66, 85-92) for both values of
    state_f.  Pinned by oracle/pin_reference.py section 7b against the reference's own classes (state_f False and True)."""
import torch
from torch import nn, optim


def train_steps(model, featrs, labels, order, batch_s=16, lr=1e-3, optimizer=None):
    """Run the batches of `order` (row indices, in training order; the last batch may be short like DataLoader's) through
    zero_grad / forward / L1Loss / backward / Adam.step.  Returns (per-batch losses, optimizer)."""
    opt = optimizer if optimizer is not None else optim.Adam(model.parameters(), lr=lr)
    for gp in opt.param_groups:
        gp["lr"] = lr
    crit = nn.L1Loss()
    model.train()
    losses = []
    for b0 in range(0, len(order), batch_s):
        idx = order[b0:b0 + batch_s]
        opt.zero_grad()
        loss = crit(model(featrs[idx]), labels[idx])
        loss.backward()
        opt.step()
        losses.append(loss.item())
    return torch.tensor(losses), opt


def train_steps_rnn(model, featrs, labels, order, batch_s=16, lr=1e-3, optimizer=None):
    """The same loop for rollout_ref.RefGRU / RefLSTM (forward(x, state) -> (output, state)): every batch starts from the zero
    state (deep_models.py:118-119 + :60-61).  featrs [n, ft_size, 8]."""
    opt = optimizer if optimizer is not None else optim.Adam(model.parameters(), lr=lr)
    for gp in opt.param_groups:
        gp["lr"] = lr
    crit = nn.L1Loss()
    model.train()
    losses = []
    for b0 in range(0, len(order), batch_s):
        idx = order[b0:b0 + batch_s]
        opt.zero_grad()
        loss = crit(model(featrs[idx], model.zero_state(len(idx)))[0], labels[idx])
        loss.backward()
        opt.step()
        losses.append(loss.item())
    return torch.tensor(losses), opt


@torch.no_grad()
def validation_loss_rnn(model, featrs, labels, batch_s=16):
    """deep_models.py:130-146"""
    crit = nn.L1Loss()
    model.eval()
    tot, nb = 0.0, 0
    for b0 in range(0, featrs.shape[0], batch_s):
        x = featrs[b0:b0 + batch_s]
        tot += crit(model(x, model.zero_state(x.shape[0]))[0], labels[b0:b0 + batch_s]).item()
        nb += 1
    return tot / nb


@torch.no_grad()
def validation_loss(model, featrs, labels, batch_s=16):
    """basic_model.py:112-122"""
    crit = nn.L1Loss()
    model.eval()
    tot, nb = 0.0, 0
    for b0 in range(0, featrs.shape[0], batch_s):
        tot += crit(model(featrs[b0:b0 + batch_s]), labels[b0:b0 + batch_s]).item()
        nb += 1
    return tot / nb
