"""Every compiled FP32 template instance runs on the GPU at least once (round-1 verdict missing #7, next #8).

The shipped checkpoints are hidden 16/64 MLPs and hidden-64 / window-4 GRU / LSTM; the reference's own training grid is
hidden {64,128} x layers {2,4} x ft_size {4,8} (deep_models.py:182-184) and hidden {16,64} x layers {1,2,4} for the MLP
(basic_model.py:141-147), and launch.h instantiates hidden 16..256.  Here: seeded random-init modules of every
(family, hidden, register tile) in MPCDL_MLP_CONFIGS / MPCDL_MLP2_CONFIGS / MPCDL_RNN_CONFIGS, ragged N, H = 1 and short
horizons, window 1 / 4 / 8, stateless + stateful, do_drop = True, with and without the residual-noise net, against
oracle/rollout_ref.py.  Tolerance: T1 (one-step max rel err <= 1e-5) and <= 2e-5 over the short horizon (random-init
networks are contractive, so the max norm is meaningful here; the chaotic shipped checkpoints use quantiles, T2).
A hypothesis property test then draws configurations at random from the same space."""
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as hst

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _stats(window, gold_stats):
    mlp = gold_stats[0]
    if window == 0:
        return mlp
    if window == 4:
        return gold_stats[1]
    # per-position statistics for other windows: the MLP's row statistics, perturbed per position so that a kernel mixing up
    # positions shows
    scale = 1.0 + 0.01 * torch.arange(window, dtype=torch.float32)[:, None]
    return dict(f_mean=mlp["f_mean"][None, :] * scale, f_std=mlp["f_std"][None, :] * scale, l_mean=mlp["l_mean"], l_std=mlp["l_std"])


def _modules(kind, hidden, layers, do_drop, seed):
    """(mpcdl module on the GPU, oracle module) sharing one seeded random state_dict"""
    import mpcdl
    from oracle import rollout_ref as R
    torch.manual_seed(seed)
    if kind == "mlp":
        ref = R.RefBasic(8, hidden, 6, layers, do_drop).eval()
        m = mpcdl.Basic_Model(8, hidden, 6, layers, do_drop)
    elif kind == "gru":
        ref = R.RefGRU(8, hidden, 6, layers).eval()
        m = mpcdl.GRU_Model(8, hidden, 6, layers, False)
    else:
        ref = R.RefLSTM(8, hidden, 6, layers).eval()
        m = mpcdl.LSTM_Model(8, hidden, 6, layers, False)
    with torch.no_grad():                                       # weights of O(1) so that gates leave the linear regime
        for p in ref.parameters():
            p.mul_(2.0)
    m.load_state_dict(ref.state_dict())
    return m.eval().to(DEV), ref


def _noise(seed):
    import mpcdl
    from oracle import rollout_ref as R
    torch.manual_seed(seed)
    ref = R.RefNoise(6, 2, 3, 2, 32, "tanh", "identity").eval()
    nm = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    nm.model.load_state_dict(ref.model.state_dict())
    return nm.to(DEV), ref


def _check(kind, hidden, layers, window, n, h, tune, stateful, noise, do_drop, gold_stats, seed=0):
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, ref = _modules(kind, hidden, layers, do_drop, seed)
    st = _stats(window, gold_stats)
    rows = synth.load_rows()
    init, U = synth.make_inputs(rows, n, h, window=window, seed=seed + 1)
    cost = synth.make_cost(rows, h)
    nm = nref = eps = None
    if noise:
        nm, nref = _noise(seed + 2)
        eps = synth.make_eps(n, h, seed=seed + 3)
    kw = dict(noise=nm, eps=eps.to(DEV) if noise else None, cost=tuple(c.to(DEV) for c in cost), tune=tune)
    with torch.no_grad():
        if kind == "mlp":
            got = m.rollout(init.to(DEV), U.to(DEV), st, **kw)
            want = R.rollout_mlp(ref, st, init, U, noise=nref, eps=eps, cost=cost)
        else:
            got = m.rollout(init.to(DEV), U.to(DEV), st, stateful=stateful, **kw)
            want = R.rollout_rnn(ref, st, init, U, stateful=stateful, noise=nref, eps=eps, cost=cost)
    # R.rel_err scales by max |ref_d| over the batch; a one-trajectory batch whose v_tran is 1e-3 would turn an FP32 rounding
    # into 1e-5 "relative", so the scale is never taken below the spread of that state component in the data
    scale = torch.maximum(want[0].abs().amax(dim=(0, 1)), st["l_mean"].abs() + st["l_std"])
    err = ((got[0].cpu() - want[0]).abs() / scale).amax(dim=-1)
    assert torch.isfinite(got[0]).all()
    assert err[:, 0].max().item() <= 1e-5, (kind, hidden, layers, window, n, h, tune, err[:, 0].max().item())
    assert err.max().item() <= 2e-5, (kind, hidden, layers, window, n, h, tune, err.max().item())
    cerr = ((got[1].cpu() - want[1]).abs() / want[1].abs().clamp_min(1e-3)).max().item()
    assert cerr <= 1e-4, cerr
    if stateful and kind != "mlp":                              # the carried state after H steps (deep_models.py:65)
        carry = got[2].cpu()
        final = want[2] if kind == "gru" else torch.cat(list(want[2]), 0)
        assert carry.shape == final.shape
        assert ((carry - final).abs().max() / final.abs().max()).item() <= 2e-5


# (hidden, RT, RJ) of MPCDL_MLP_CONFIGS: the generic ring-fed kernel (tune kc = -1 keeps the weight-resident kernel out)
MLP_GENERIC = [(16, 2), (16, 8), (32, 2), (32, 8), (64, 2), (64, 4), (64, 8), (128, 2), (128, 8), (256, 2), (256, 8)]
# (hidden, RT, KSPLIT) of MPCDL_MLP2_CONFIGS
MLP_RESIDENT = [(64, 4, 2), (64, 8, 2), (64, 8, 1), (32, 4, 2), (32, 8, 1), (16, 4, 2), (16, 8, 1)]
# (hidden, RT) of MPCDL_RNN_CONFIGS
RNN_TILES = [(16, 2), (16, 4), (32, 2), (32, 4), (64, 2), (64, 4), (64, 8), (128, 2), (128, 4), (256, 2), (256, 4)]


@pytest.mark.parametrize("hidden,rt", MLP_GENERIC)
@pytest.mark.parametrize("noise", [False, True])
def test_generic_mlp_kernel_instances(hidden, rt, noise, gold_stats):
    layers = {16: 1, 32: 2, 64: 4, 128: 2, 256: 1}[hidden]
    _check("mlp", hidden, layers, 0, 301, 5, (rt, 0, 0, -1), False, noise, do_drop=hidden == 32, gold_stats=gold_stats, seed=hidden + rt)


@pytest.mark.parametrize("hidden,rt,ks", MLP_RESIDENT)
def test_resident_mlp_kernel_instances(hidden, rt, ks, gold_stats):
    from mpcdl import _lib, ops
    n = 517 if rt == 8 else 77
    tune = (rt, 0, 0, ks)
    assert ops.plan(_lib.MLP, hidden, 2, 0, n, 4, tune=tune)["variant"] == 2
    _check("mlp", hidden, 2, 0, n, 4, tune, False, False, do_drop=True, gold_stats=gold_stats, seed=3 * hidden + rt + ks)


@pytest.mark.parametrize("n,h", [(1, 1), (127, 3), (128, 2), (129, 5), (1000, 7), (40000, 3)])
def test_two_group_mlp_kernel_bitwise_equal(n, h, gold_stats):
    """rollout_mlp2_kernel<64,8,8,1,2> (two independent 128-trajectory groups per CTA, tune kc = 4) == the one-group kernel
    (kc = 1) == the generic kernel, bit for bit: states, costs, cost-only launch."""
    import mpcdl
    from conftest import load_ckpt
    from mpcdl import _lib, ops, synth
    m, _ = mpcdl.from_checkpoint_name("Basic_64_4_32", load_ckpt("Basic_64_4_32"), torch.device(DEV))
    rows = synth.load_rows()
    init, U = synth.make_inputs(rows, n, h, window=0, seed=31)
    cost = tuple(c.to(DEV) for c in synth.make_cost(rows, h))
    init, U, st = init.to(DEV), U.to(DEV), gold_stats[0]
    p = ops.plan(_lib.MLP, 64, 4, 0, n, h, tune=(8, 0, 0, 4))
    assert p["variant"] == 2 and p["threads"] == 256 and p["tile_traj"] == 128 and p["grid"] == min(p["ntiles"], 148)
    s2, c2 = m.rollout(init, U, st, cost=cost, tune=(8, 0, 0, 4))
    s1, c1 = m.rollout(init, U, st, cost=cost, tune=(8, 0, 0, 1))
    sg, cg = m.rollout(init, U, st, cost=cost, tune=(0, 0, 0, -1))
    assert torch.isfinite(s2).all() and torch.equal(s2, s1) and torch.equal(c2, c1) and torch.equal(s2, sg) and torch.equal(c2, cg)
    assert torch.equal(m.rollout(init, U, st, cost=cost, return_states=False, tune=(8, 0, 0, 4))[1], c2)


@pytest.mark.parametrize("kind", ["gru", "lstm"])
@pytest.mark.parametrize("hidden,rt", RNN_TILES)
def test_rnn_kernel_instances(kind, hidden, rt, gold_stats):
    layers, window = (1, 8) if hidden <= 32 else (2, 4) if hidden <= 128 else (1, 1)
    for stateful in (False, True):
        _check(kind, hidden, layers, window, 131, 3, (rt, 0, 0, 0), stateful, noise=hidden == 128, do_drop=False,
               gold_stats=gold_stats, seed=hidden + rt + stateful)


@pytest.mark.parametrize("kind,hidden,layers,window", [("gru", 128, 2, 8), ("lstm", 128, 4, 4), ("gru", 64, 2, 8), ("lstm", 64, 4, 8),
                                                       ("gru", 128, 4, 4), ("lstm", 128, 2, 8)])
def test_reference_training_grid_shapes(kind, hidden, layers, window, gold_stats):
    """deep_models.py:182-184 trains hidden {64,128} x layers {2,4} x ft_size {4,8}: the launch the planner picks on its own"""
    _check(kind, hidden, layers, window, 1000, 4, (), False, noise=False, do_drop=False, gold_stats=gold_stats, seed=window + layers)


@pytest.mark.parametrize("kind", ["gru", "lstm"])
@pytest.mark.parametrize("layers,n,h,stateful,noise", [(1, 131, 3, False, False), (2, 1000, 4, True, True), (4, 20000, 2, False, True),
                                                       (2, 300, 5, True, False)])
@pytest.mark.parametrize("precision,one_step,horizon", [("tf32x3", 2e-5, 6e-5), ("tf32", 5e-3, 2e-2)])
def test_tensor_core_window_8(kind, layers, n, h, stateful, noise, precision, one_step, horizon, gold_stats):
    """ft_size 8 of deep_models.py:184 on the tcgen05 kernel (round-1 verdict missing #5): the layer output sequence goes through
    the per-CTA scratch instead of registers (rollout_rnn_tc_epi.inl); 20 000 trajectories = more tiles than CTAs, so the
    scratch is reused.  Bounds = the stated T3 one-step bounds of the window-4 kernel."""
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, ref = _modules(kind, 64, layers, False, seed=layers + n)
    st = _stats(8, gold_stats)
    rows = synth.load_rows()
    init, U = synth.make_inputs(rows, n, h, window=8, seed=n)
    cost = synth.make_cost(rows, h)
    nm = nref = eps = None
    if noise:
        nm, nref = _noise(n + 2)
        eps = synth.make_eps(n, h, seed=n + 3)
    with torch.no_grad():
        got = m.rollout(init.to(DEV), U.to(DEV), st, stateful=stateful, noise=nm, eps=eps.to(DEV) if noise else None,
                        cost=tuple(c.to(DEV) for c in cost), precision=precision)
        want = R.rollout_rnn(ref, st, init, U, stateful=stateful, noise=nref, eps=eps, cost=cost)
    scale = torch.maximum(want[0].abs().amax(dim=(0, 1)), st["l_mean"].abs() + st["l_std"])
    err = ((got[0].cpu() - want[0]).abs() / scale).amax(dim=-1)
    assert torch.isfinite(got[0]).all()
    assert err[:, 0].max().item() <= one_step, (kind, layers, precision, err[:, 0].max().item())
    assert err.max().item() <= horizon, (kind, layers, precision, err.max().item())
    if stateful:
        carry = got[2].cpu()
        final = want[2] if kind == "gru" else torch.cat(list(want[2]), 0)
        assert ((carry - final).abs().max() / final.abs().max()).item() <= 10 * horizon


@pytest.mark.parametrize("layers,n,h", [(1, 300, 4), (4, 40000, 3), (2, 19001, 5)])
@pytest.mark.parametrize("precision,one_step,horizon", [("tf32x3", 2e-5, 6e-5), ("tf32", 5e-3, 2e-2)])
def test_tensor_core_mlp_with_noise_net(layers, n, h, precision, one_step, horizon, gold_stats):
    """tcgen05 MLP kernel + FeedforwardNoiseModel residual (round-1 verdict missing #5: "the MLP tcgen05 kernel has no noise
    path either"): per-tile noise buffers, the net evaluated by the tile's epilogue warps under the input layer's MMAs.
    40 000 / 19 001 trajectories = more tiles than SMs (two tiles per CTA where both buffer sets fit, one for 4-layer tf32x3)."""
    from mpcdl import _lib, ops, synth
    from oracle import rollout_ref as R
    m, ref = _modules("mlp", 64, layers, False, seed=layers + n)
    nm, nref = _noise(n + 2)
    st = gold_stats[0]
    rows = synth.load_rows()
    init, U = synth.make_inputs(rows, n, h, window=0, seed=n)
    eps = synth.make_eps(n, h, seed=n + 3)
    cost = synth.make_cost(rows, h)
    p = ops.plan(_lib.MLP, 64, layers, 0, n, h, noise_cfg=nm.cfg(), precision=precision)
    assert p["variant"] == 5 and p["ksplit"] == (1 if n <= 148 * 128 or (layers == 4 and precision == "tf32x3") else 2)
    with torch.no_grad():
        got = m.rollout(init.to(DEV), U.to(DEV), st, noise=nm, eps=eps.to(DEV), cost=tuple(c.to(DEV) for c in cost), precision=precision)
        ns = min(n, 3000)                                       # the oracle runs on a slice that ends at the ragged tail
        sl = slice(n - ns, n)
        want = R.rollout_mlp(ref, st, init[sl], U[sl], noise=nref, eps=eps[sl], cost=cost)
    assert torch.isfinite(got[0]).all() and torch.isfinite(got[1]).all()
    scale = torch.maximum(want[0].abs().amax(dim=(0, 1)), st["l_mean"].abs() + st["l_std"])
    err = ((got[0][sl].cpu() - want[0]).abs() / scale).amax(dim=-1)
    assert err[:, 0].max().item() <= one_step, (layers, n, precision, err[:, 0].max().item())
    assert err.max().item() <= horizon, (layers, n, precision, err.max().item())
    cerr = ((got[1][sl].cpu() - want[1]).abs() / want[1].abs().clamp_min(1e-3)).max().item()
    assert cerr <= 20 * horizon, cerr
    # without the net the same launch differs: the residual is really applied
    plain = m.rollout(init.to(DEV), U.to(DEV), st, precision=precision)[0]
    assert (plain - got[0]).abs().max().item() > 1e-4


@pytest.mark.parametrize("kind", ["mlp", "gru", "lstm"])
def test_single_trajectory_single_step(kind, gold_stats):
    _check(kind, 64, 1, 0 if kind == "mlp" else 1, 1, 1, (), kind != "mlp", noise=True, do_drop=False, gold_stats=gold_stats, seed=5)


@settings(max_examples=20, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(kind=hst.sampled_from(["mlp", "gru", "lstm"]), hidden=hst.sampled_from([16, 32, 64, 128, 256]), layers=hst.sampled_from([1, 2, 4]),
       window=hst.sampled_from([1, 4, 8]), n=hst.integers(1, 700), h=hst.integers(1, 6), stateful=hst.booleans(), noise=hst.booleans(),
       do_drop=hst.booleans())
def test_property_random_configurations(kind, hidden, layers, window, n, h, stateful, noise, do_drop, gold_stats):
    if hidden == 256 and kind != "mlp":
        layers, window = min(layers, 2), min(window, 4)         # keeps the CPU oracle of an example under a second
    _check(kind, hidden, layers, 0 if kind == "mlp" else window, n, h, (), stateful, noise, do_drop, gold_stats,
           seed=hidden + layers + window + n)
