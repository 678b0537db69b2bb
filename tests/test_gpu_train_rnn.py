"""GPU parity of the fused training epoch of the recurrent models (csrc/train_rnn_kernels.cu) against the oracle
(oracle/train_ref.train_steps_rnn: the reference's loop, deep_models.py:114-127, on torch's nn.GRU / nn.LSTM autograd, L1Loss and
Adam with the same batches; pinned bit for bit against the reference's GRU_Model / LSTM_Model classes, pin_reference.py 7b).

Bounds (the MLP's, tests/test_gpu_train.py).  A step is FP32 arithmetic in another summation order than torch's, so single
steps agree to rounding; Adam then divides by sqrt(v): a gradient component that is ZERO up to rounding gets an update of
+-lr whatever its size, so a handful of parameters may differ by a few lr while the loss does not notice.
  per-batch loss, first 20 steps        <= 1e-4 relative (first step 1e-5)
  parameters after 1 / 20 steps         >= 99 % / 98 % within 1e-5 / 2e-5 absolute; max <= 2.5 lr / 5 lr
  repeat launches, ensembles            bit-identical (fixed-order sums)
"""
import copy
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _data(bags, window):
    import mpcdl
    ds = mpcdl.RobotData(bags, window)               # z-scored windows / labels, as deep_models.py:104 builds them
    return ds.featrs.float(), ds.labels.float()


def _models(kind, hidden, layers, state_f, seed):
    import mpcdl
    from oracle import rollout_ref as R
    torch.manual_seed(seed)
    ref = (R.RefGRU if kind == "gru" else R.RefLSTM)(8, hidden, 6, layers)
    m = (mpcdl.GRU_Model if kind == "gru" else mpcdl.LSTM_Model)(8, hidden, 6, layers, state_f)
    m.load_state_dict(ref.state_dict())
    return m.to(DEV), ref


def _flat(model):
    return torch.cat([p.detach().reshape(-1).cpu() for p in model.parameters()])


# the reference's grid: hidden {64, 128} x layers {2, 4} x ft_size {4, 8} x batch {8, 16} (deep_models.py:182-186), its defaults
# (hidden 16, 4 layers, window 4, batch 16: :96-101), and odd shapes (one layer, window 1, a batch of 5)
@pytest.mark.parametrize("kind,hidden,layers,window,batch,state_f", [
    ("gru", 64, 4, 4, 16, False), ("lstm", 64, 4, 4, 16, False), ("gru", 128, 2, 8, 8, True), ("lstm", 128, 2, 8, 8, False),
    ("gru", 16, 4, 4, 16, False), ("lstm", 16, 4, 4, 16, True), ("gru", 32, 1, 1, 5, False), ("lstm", 32, 1, 2, 5, False),
    ("gru", 64, 2, 8, 16, False), ("lstm", 64, 2, 8, 8, True),
    # batches above 16 rows run as micro-batches of 16 with summed gradients: 64 = the BATCH_S of half the shipped checkpoints
    ("gru", 64, 4, 4, 64, True), ("lstm", 64, 4, 4, 64, False), ("gru", 32, 2, 4, 40, False), ("lstm", 16, 2, 8, 17, False)])
def test_rnn_training_steps_match_the_oracle(kind, hidden, layers, window, batch, state_f, bags):
    from mpcdl import train
    from oracle import train_ref as T
    f, l = _data(bags, window)
    m, ref = _models(kind, hidden, layers, state_f, seed=hidden + layers + window)
    assert train.param_count(m) == _flat(ref).numel()
    g = torch.Generator().manual_seed(1)
    order = torch.randperm(f.shape[0], generator=g)
    fd, ld = f.to(DEV), l.to(DEV)
    state = train.AdamState(1, _flat(ref).numel(), DEV)
    lr = 1e-3
    # one step
    want1, opt = T.train_steps_rnn(ref, f, l, order[:batch], batch, lr)
    got1 = train.train_epoch(m, state, fd, ld, order[:batch], batch, lr).cpu()[0]
    assert abs(got1[0] - want1[0]).item() <= 1e-5 * abs(want1[0]).item(), (got1, want1)
    d = (_flat(m) - _flat(ref)).abs()
    assert (d <= 1e-5).float().mean().item() >= 0.99 and d.max().item() <= 2.5 * lr, (d.max().item(), (d <= 1e-5).float().mean().item())
    # 19 more steps (the last batch short), continuing both optimizers' state
    nxt = order[batch:20 * batch - 3]
    want, opt = T.train_steps_rnn(ref, f, l, nxt, batch, lr, optimizer=opt)
    got = train.train_epoch(m, state, fd, ld, nxt, batch, lr).cpu()[0]
    assert state.step == 20 and got.shape == want.shape
    assert ((got - want).abs() / want.abs()).max().item() <= 1e-4, ((got - want).abs() / want.abs()).max().item()
    d = (_flat(m) - _flat(ref)).abs()
    assert (d <= 2e-5).float().mean().item() >= 0.98 and d.max().item() <= 5 * lr, (d.max().item(), (d <= 2e-5).float().mean().item())
    # the Adam moments are the optimizer's
    em = torch.cat([opt.state[p]["exp_avg"].reshape(-1) for p in ref.parameters()])
    ev = torch.cat([opt.state[p]["exp_avg_sq"].reshape(-1) for p in ref.parameters()])
    assert (state.m[0].cpu() - em).abs().max().item() <= 1e-4 * em.abs().max().item() + 1e-7
    assert (state.v[0].cpu() - ev).abs().max().item() <= 1e-4 * ev.abs().max().item() + 1e-9


@pytest.mark.parametrize("tag", ["GRU_32_2_4_16", "LSTM_16_4_8_8"])
def test_rnn_training_kernel_against_the_reference_made_golden(tag, bags):
    """20 Adam steps (the last batch short) against tests/golden/train_rnn.npz -- losses and parameters produced by the
    REFERENCE's GRU_Model / LSTM_Model classes, nn.L1Loss and optim.Adam (oracle/pin_reference.py section 7b)."""
    import mpcdl
    from conftest import GOLD
    from mpcdl import train
    g = np.load(os.path.join(GOLD, "train_rnn.npz"))
    kind, hidden, layers, window, batch = tag.split("_")[0], *(int(x) for x in tag.split("_")[1:])
    f, l = _data(bags, window)
    m = (mpcdl.GRU_Model if kind == "GRU" else mpcdl.LSTM_Model)(8, hidden, 6, layers, kind == "LSTM")
    torch.nn.utils.vector_to_parameters(torch.from_numpy(g[f"{tag}/init"]).clone(), m.parameters())
    m.to(DEV)
    state = train.AdamState(1, g[f"{tag}/init"].size, DEV)
    got = train.train_epoch(m, state, f.to(DEV), l.to(DEV), torch.from_numpy(g[f"{tag}/order"]), batch, 1e-3).cpu()[0]
    want = torch.from_numpy(g[f"{tag}/losses"])
    assert got.shape == want.shape and ((got - want).abs() / want.abs()).max().item() <= 1e-4
    d = (_flat(m) - torch.from_numpy(g[f"{tag}/final"])).abs()
    assert (d <= 2e-5).float().mean().item() >= 0.98 and d.max().item() <= 5e-3, (d.max().item(), (d <= 2e-5).float().mean().item())
    if kind == "LSTM":
        m2 = mpcdl.LSTM_Model(8, hidden, 6, layers, True)
        torch.nn.utils.vector_to_parameters(torch.from_numpy(g[f"{tag}/final"]).clone(), m2.parameters())
        m2.to(DEV)
        v = train.validation_loss(m2, f[:400].to(DEV), l[:400].to(DEV), 8)
        assert abs(v - float(g[f"{tag}/valid400"])) <= 1e-5 * v


def test_rnn_ensemble_launch_equals_single_launches_bitwise(bags):
    """n_models CTAs in one launch == one launch per model, bit for bit; a repeated launch from the same state too"""
    from mpcdl import train
    f, l = _data(bags, 4)
    fd, ld = f.to(DEV), l.to(DEV)
    g = torch.Generator().manual_seed(2)
    orders = torch.stack([torch.randperm(f.shape[0], generator=g)[:150] for _ in range(3)])
    for kind in ("gru", "lstm"):
        ms = [_models(kind, 64, 2, False, seed=s)[0] for s in (1, 2, 3)]
        singles = [copy.deepcopy(m) for m in ms]
        again = [copy.deepcopy(m) for m in ms]
        st = train.AdamState(3, _flat(ms[0]).numel(), DEV)
        loss = train.train_epoch(ms, st, fd, ld, orders, 16, 1e-3)
        assert loss.shape == (3, 10) and st.step == 10 and torch.isfinite(loss).all()
        for i, m1 in enumerate(singles):
            s1 = train.AdamState(1, _flat(m1).numel(), DEV)
            l1 = train.train_epoch(m1, s1, fd, ld, orders[i], 16, 1e-3)
            assert torch.equal(l1[0], loss[i]) and torch.equal(_flat(m1), _flat(ms[i]))
            assert torch.equal(s1.m[0], st.m[i]) and torch.equal(s1.v[0], st.v[i])
        # in place of initcheck (compute-sanitizer is closed on this pool): the repeat runs with NaN in every word of the scratch
        # buffer and NaN-poisoned shared memory on every SM -- a read of anything the launch did not write would show
        from mpcdl import _lib
        st2 = train.AdamState(3, _flat(again[0]).numel(), DEV)
        k = _lib.GRU if kind == "gru" else _lib.LSTM
        st2._work = torch.full((3 * _lib.lib.mpcdl_train_rnn_work_floats(k, 64, 2, 4),), float("nan"), device=DEV)
        torch.ops.mpcdl.debug_poison_smem(fd)
        loss2 = train.train_epoch(again, st2, fd, ld, orders, 16, 1e-3)
        assert torch.equal(loss2, loss) and all(torch.equal(_flat(a), _flat(b)) for a, b in zip(again, ms))


def test_rnn_full_epoch_follows_the_oracle(bags):
    """one epoch of the reference's default configuration (GRU hidden 16 x 4 layers, window 4, batch 16) over the 80 % split
    in the subset's own order (DataLoader(shuffle=False), deep_models.py:112), then the validation loss of :130-146"""
    from mpcdl import train
    from oracle import train_ref as T
    f, l = _data(bags, 4)
    m, ref = _models("gru", 16, 4, False, seed=3)
    g = torch.Generator().manual_seed(7)
    t_idx, v_idx = train.split_indices(f.shape[0], g)
    fd, ld = f.to(DEV), l.to(DEV)
    state = train.AdamState(1, _flat(ref).numel(), DEV)
    want, _ = T.train_steps_rnn(ref, f, l, t_idx, 16, 1e-3)
    got = train.train_epoch(m, state, fd, ld, t_idx, 16, 1e-3).cpu()[0]
    assert got.shape == want.shape == ((t_idx.numel() + 15) // 16,)
    rel = (got - want).abs() / want.abs()
    assert rel.max().item() <= 2e-2 and rel.median().item() <= 1e-3, (rel.max().item(), rel.median().item())
    v_want = T.validation_loss_rnn(ref, f[v_idx], l[v_idx], 16)
    v_got = train.validation_loss(m, fd[v_idx.to(DEV)].contiguous(), ld[v_idx.to(DEV)].contiguous(), 16)
    assert abs(v_got - v_want) <= 1e-2 * v_want, (v_got, v_want)


def test_rnn_train_model_learns_and_follows_the_reference_protocol(bags):
    """train_model end to end on the validation bag for both families: the loss falls, the best state is kept, a run is
    reproducible, a stateful model trains like a stateless one (init_hidden before every batch), the bad-loss exit fires"""
    import mpcdl
    from mpcdl import train
    ds = mpcdl.RobotData(bags, 4)
    hists = {}
    for kind, cls in (("gru", mpcdl.GRU_Model), ("lstm", mpcdl.LSTM_Model)):
        for state_f in (False, True):
            torch.manual_seed(0)
            m = cls(8, 16, 6, 2, state_f).to(DEV)
            h = train.train_model(m, ds, epochs=6, batch_s=16, learn_r=1e-3, seed=0, verbose=False, bad_loss=1e9)
            assert h["code"] is None and len(h["valid_loss"]) == 6 and h["valid_loss"][-1] < 0.7 * h["valid_loss"][0], h["valid_loss"]
            assert h["best"] is not None
            hists[kind, state_f] = h["valid_loss"]
        assert hists[kind, False] == hists[kind, True]
    torch.manual_seed(0)
    m3 = mpcdl.GRU_Model(8, 16, 6, 2, False).to(DEV)
    h3 = train.train_model(m3, ds, epochs=5, batch_s=16, learn_r=1e-9, seed=0, verbose=False)
    assert h3["code"] == 1 and len(h3["valid_loss"]) == 1
    with pytest.raises(RuntimeError):                 # an MLP dataset (rows, not windows)
        train.train_model(mpcdl.GRU_Model(8, 16, 6, 2, False).to(DEV), mpcdl.RobotData(bags), epochs=1, verbose=False)


def test_rnn_more_models_than_sms_run_two_per_sm_with_identical_results(bags):
    """above one model per SM the launcher picks the 128-register instantiation (two CTAs per SM): same arithmetic, so a
    model's result must not depend on which instantiation trained it"""
    from mpcdl import train
    f, l = _data(bags, 4)
    fd, ld = f.to(DEV), l.to(DEV)
    n_big = torch.cuda.get_device_properties(0).multi_processor_count + 12
    order = torch.randperm(f.shape[0], generator=torch.Generator().manual_seed(4))[:5 * 16 - 2]
    for kind in ("gru", "lstm"):
        base = [_models(kind, 32, 2, False, seed=s)[0] for s in (5, 6)]
        many = [copy.deepcopy(base[i % 2]) for i in range(n_big)]
        st = train.AdamState(n_big, _flat(base[0]).numel(), DEV)
        loss = train.train_epoch(many, st, fd, ld, order, 16, 1e-3)
        assert loss.shape == (n_big, 5)
        for i in (0, 1):
            s1 = train.AdamState(1, _flat(base[i]).numel(), DEV)
            l1 = train.train_epoch(base[i], s1, fd, ld, order, 16, 1e-3)
            for j in (i, i + 2, n_big - 1 - ((n_big - 1 - i) % 2)):       # copies of model i sit at the indices of its parity
                assert j % 2 == i and j < n_big
                assert torch.equal(l1[0], loss[j]) and torch.equal(_flat(base[i]), _flat(many[j])), (kind, i, j)
