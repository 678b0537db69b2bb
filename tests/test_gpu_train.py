"""GPU parity of the fused training epoch (csrc/train_kernels.cu) against the oracle (oracle/train_ref.py:
the reference's loop on torch's autograd, L1Loss and Adam with the same batch order).

Bounds.  A step is FP32 arithmetic in another summation order than torch's, so single steps agree to rounding; Adam then
divides by sqrt(v): a gradient component that is ZERO up to rounding (sums of +-1/(6B) that cancel) gets an update of
+-lr whatever its size, so a handful of parameters may differ by up to 2 lr after a step while the loss does not notice.
  per-batch loss, first 20 steps        <= 1e-5 relative
  parameters after 1 / 20 steps         >= 99 % within 1e-5 (absolute, parameters are O(0.1..1)); max <= 2.5 lr
  one full epoch (228 steps of 16)      per-batch losses <= 2 % relative (median <= 1e-3), validation loss <= 1 % relative
  repeat launches                       bit-identical (fixed-order sums)
"""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _data(bags):
    import mpcdl
    ds = mpcdl.RobotData(bags)                       # z-scored features / labels, as basic_model.py:69 builds them
    return ds.featrs.float(), ds.labels.float()


def _models(hidden, layers, seed):
    import mpcdl
    from oracle import rollout_ref as R
    torch.manual_seed(seed)
    ref = R.RefBasic(8, hidden, 6, layers)
    m = mpcdl.Basic_Model(8, hidden, 6, layers, False)
    m.load_state_dict(ref.state_dict())
    return m.to(DEV), ref


def _flat(model):
    return torch.cat([p.detach().reshape(-1).cpu() for p in model.parameters()])


@pytest.mark.parametrize("hidden,layers,batch", [(16, 4, 16), (64, 4, 16), (64, 1, 32), (16, 2, 64), (32, 2, 16)])
def test_training_steps_match_the_oracle(hidden, layers, batch, bags):
    from mpcdl import train
    from oracle import train_ref as T
    f, l = _data(bags)
    m, ref = _models(hidden, layers, seed=hidden + layers)
    g = torch.Generator().manual_seed(1)
    order = torch.randperm(f.shape[0], generator=g)
    fd, ld = f.to(DEV), l.to(DEV)
    state = train.AdamState(1, _flat(ref).numel(), DEV)
    lr = 1e-3
    # one step
    want1, opt = T.train_steps(ref, f, l, order[:batch], batch, lr)
    got1 = train.train_epoch(m, state, fd, ld, order[:batch], batch, lr).cpu()[0]
    assert abs(got1[0] - want1[0]).item() <= 1e-5 * abs(want1[0]).item()
    d = (_flat(m) - _flat(ref)).abs()
    assert (d <= 1e-5).float().mean().item() >= 0.99 and d.max().item() <= 2.5 * lr, (d.max().item(), (d <= 1e-5).float().mean().item())
    # 19 more steps, continuing both optimizers' state
    nxt = order[batch:20 * batch]
    want, opt = T.train_steps(ref, f, l, nxt, batch, lr, optimizer=opt)
    got = train.train_epoch(m, state, fd, ld, nxt, batch, lr).cpu()[0]
    assert state.step == 20
    assert ((got - want).abs() / want.abs()).max().item() <= 1e-4, ((got - want).abs() / want.abs()).max().item()
    d = (_flat(m) - _flat(ref)).abs()
    assert (d <= 2e-5).float().mean().item() >= 0.98 and d.max().item() <= 5 * lr, (d.max().item(), (d <= 2e-5).float().mean().item())


@pytest.mark.parametrize("tag", ["Basic_16_4_16", "Basic_64_2_32"])
def test_training_kernel_against_the_reference_made_golden(tag, bags):
    """25 Adam steps (the last batch short) against tests/golden/train.npz -- losses and parameters produced by the REFERENCE's
    Basic_Model class, nn.L1Loss and optim.Adam (oracle/pin_reference.py section 7)."""
    import os
    import numpy as np
    import mpcdl
    from conftest import GOLD
    from mpcdl import train
    g = np.load(os.path.join(GOLD, "train.npz"))
    hidden, layers, batch = (int(x) for x in tag.split("_")[1:])
    f, l = _data(bags)
    m = mpcdl.Basic_Model(8, hidden, 6, layers, False)
    torch.nn.utils.vector_to_parameters(torch.from_numpy(g[f"{tag}/init"]).clone(), m.parameters())
    m.to(DEV)
    state = train.AdamState(1, g[f"{tag}/init"].size, DEV)
    got = train.train_epoch(m, state, f.to(DEV), l.to(DEV), torch.from_numpy(g[f"{tag}/order"]), batch, 1e-3).cpu()[0]
    want = torch.from_numpy(g[f"{tag}/losses"])
    assert got.shape == want.shape and ((got - want).abs() / want.abs()).max().item() <= 1e-4
    d = (_flat(m) - torch.from_numpy(g[f"{tag}/final"])).abs()
    assert (d <= 2e-5).float().mean().item() >= 0.98 and d.max().item() <= 5e-3, (d.max().item(), (d <= 2e-5).float().mean().item())
    if tag == "Basic_64_2_32":
        v = train.validation_loss(m, f[:800].to(DEV), l[:800].to(DEV), 32)
        assert abs(v - float(g[f"{tag}/valid800"])) <= 1e-4 * v


def test_full_epoch_and_validation_loss(bags):
    """one epoch of the reference's default configuration (hidden 16 x 4 layers, batch 16, 80 / 20 split)"""
    from mpcdl import train
    from oracle import train_ref as T
    f, l = _data(bags)
    m, ref = _models(16, 4, seed=3)
    g = torch.Generator().manual_seed(7)
    t_idx, v_idx = train.split_indices(f.shape[0], g)
    order = t_idx[torch.randperm(t_idx.numel(), generator=g)]
    assert t_idx.numel() == int(f.shape[0] * 0.8) and order.numel() % 16 != 0          # the last batch is short (DataLoader keeps it)
    fd, ld = f.to(DEV), l.to(DEV)
    state = train.AdamState(1, _flat(ref).numel(), DEV)
    want, _ = T.train_steps(ref, f, l, order, 16, 1e-3)
    got = train.train_epoch(m, state, fd, ld, order, 16, 1e-3).cpu()[0]
    assert got.shape == want.shape == ((order.numel() + 15) // 16,)
    rel = (got - want).abs() / want.abs()
    assert rel.max().item() <= 2e-2 and rel.median().item() <= 1e-3, (rel.max().item(), rel.median().item())
    v_want = T.validation_loss(ref, f[v_idx], l[v_idx], 16)
    v_got = train.validation_loss(m, fd[v_idx.to(DEV)].contiguous(), ld[v_idx.to(DEV)].contiguous(), 16)
    assert abs(v_got - v_want) <= 1e-2 * v_want, (v_got, v_want)
    # the validation kernel alone, on identical model outputs: FP32 sums of 96 terms
    m2, ref2 = _models(16, 4, seed=3)
    v2 = train.validation_loss(m2, fd[v_idx.to(DEV)].contiguous(), ld[v_idx.to(DEV)].contiguous(), 16)
    assert abs(v2 - T.validation_loss(ref2, f[v_idx], l[v_idx], 16)) <= 1e-6 * v2 + 1e-7


def test_ensemble_launch_equals_single_launches_bitwise(bags):
    """n_models CTAs in one launch == one launch per model, bit for bit; a repeated launch from the same state too"""
    from mpcdl import train
    f, l = _data(bags)
    fd, ld = f.to(DEV), l.to(DEV)
    g = torch.Generator().manual_seed(2)
    orders = torch.stack([torch.randperm(f.shape[0], generator=g)[:400] for _ in range(3)])
    ms = [_models(64, 2, seed=s)[0] for s in (1, 2, 3)]
    singles = [copy.deepcopy(m) for m in ms]
    st = train.AdamState(3, _flat(ms[0]).numel(), DEV)
    loss = train.train_epoch(ms, st, fd, ld, orders, 16, 1e-3)
    assert loss.shape == (3, 25) and st.step == 25
    for i, m1 in enumerate(singles):
        s1 = train.AdamState(1, _flat(m1).numel(), DEV)
        l1 = train.train_epoch(m1, s1, fd, ld, orders[i], 16, 1e-3)
        assert torch.equal(l1[0], loss[i]) and torch.equal(_flat(m1), _flat(ms[i]))
        assert torch.equal(s1.m[0], st.m[i]) and torch.equal(s1.v[0], st.v[i])


def test_train_model_learns_and_follows_the_reference_protocol(bags):
    """train_model end to end on the validation bag: the loss falls, the best state is kept, a run is reproducible, and the
    reference's `v_loss > 0.1 -> return 1` exit fires for an untrained net with a tiny learning rate."""
    import mpcdl
    from mpcdl import train
    ds = mpcdl.RobotData(bags)
    torch.manual_seed(0)
    m = mpcdl.Basic_Model(8, 16, 6, 4, False).to(DEV)
    h = train.train_model(m, ds, epochs=12, batch_s=16, learn_r=1e-3, seed=0, verbose=False, bad_loss=1e9)
    assert h["code"] is None and len(h["valid_loss"]) == 12 and h["valid_loss"][-1] < 0.6 * h["valid_loss"][0], h["valid_loss"]
    assert h["best"] is not None and min(h["valid_loss"]) == h["valid_loss"][[i for i, v in enumerate(h["valid_loss"]) if v == min(h["valid_loss"])][0]]
    torch.manual_seed(0)
    m2 = mpcdl.Basic_Model(8, 16, 6, 4, False).to(DEV)
    h2 = train.train_model(m2, ds, epochs=12, batch_s=16, learn_r=1e-3, seed=0, verbose=False, bad_loss=1e9)
    assert h2["valid_loss"] == h["valid_loss"]
    torch.manual_seed(0)
    m3 = mpcdl.Basic_Model(8, 16, 6, 4, False).to(DEV)
    h3 = train.train_model(m3, ds, epochs=5, batch_s=16, learn_r=1e-9, seed=0, verbose=False)
    assert h3["code"] == 1 and len(h3["valid_loss"]) == 1
    with pytest.raises(RuntimeError):
        train.train_model(mpcdl.Basic_Model(8, 16, 6, 2, True).to(DEV), ds, epochs=1, verbose=False)
    with pytest.raises(RuntimeError):
        train.train_model(mpcdl.GRU_Model(8, 64, 6, 2, False).to(DEV), ds, epochs=1, verbose=False)
