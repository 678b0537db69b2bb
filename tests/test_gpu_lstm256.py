"""GPU parity tests of the BF16 tcgen05 kernel for the wide LSTM (rollout_lstm_tc.cu; BASELINE.json config 4:
"LSTM rnn_model rollout, hidden 256, BF16 tensor-core path vs FP32 tolerance study").

The reference ships no hidden-256 checkpoint: the weights are LSTM_Model(8, 256, 6, L, False) under
torch.manual_seed(0), the oracle is the reference module definition in the closed loop
(oracle/rollout_ref.py), run in FP32 and FP64 on the same inputs.

T3 bounds stated for this path (rel err = max_d |dy_d| / max|ref_d| per trajectory-step; measured on B200:
one-step max 5.7e-4, H = 20 p50 2.1e-4 / p90 3.8e-4 / p99 5.3e-4 / max 6.7e-4, profiles/r01f/lt_check.log):
  one-step (vs torch FP32)      <= 2e-3
  H = 20 vs the FP64 oracle     p50 <= 6e-4, p90 <= 1.2e-3, p99 <= 2e-3
Other depths / windows of the same family (random init, some output dimensions nearly zero, so the metric's
denominator max|ref_d| is small; measured: 3 layers one-step max 4.0e-3, absolute difference 4e-5):
  one-step <= 1e-2;  p50 <= 3e-3, p90 <= 6e-3, p99 <= 1e-2
The FP32 FFMA kernel on the same model stays inside T1 / T2 (one-step <= 1e-5).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu
ONE_STEP, QUANT = 2e-3, (6e-4, 1.2e-3, 2e-3)            # config 4: 4 layers, window 4
ONE_STEP_ANY, QUANT_ANY = 1e-2, (3e-3, 6e-3, 1e-2)      # any other depth / window


def _dev():
    return torch.device("cuda:0")


def _sd(layers, hidden=256):
    torch.manual_seed(0)
    ref = torch.nn.ModuleDict(dict(lstm=torch.nn.LSTM(8, hidden, layers, batch_first=True), full_c=torch.nn.Linear(hidden, 6)))
    return {k: v.detach().clone() for k, v in ref.state_dict().items()}


def _pair(layers, window=4):
    import mpcdl
    from oracle import rollout_ref as R
    name = f"LSTM_256_{layers}_{window}_0_False"
    sd = _sd(layers)
    m, _ = mpcdl.from_checkpoint_name(name, sd, _dev())
    ref, _ = R.build_from_name(name, sd)
    return m, ref


def _quantiles(e):
    e = e.flatten().double()
    return [torch.quantile(e, x).item() for x in (0.5, 0.9, 0.99)]


def _stats(gold_stats, window):
    st = gold_stats[1]
    if window == 4:
        return st
    if window > 4:      # the shipped statistics are per position of a 4-frame window: repeat them for longer windows
        rep = lambda a: torch.cat([a, a])[:window].contiguous()
        return dict(f_mean=rep(st["f_mean"]), f_std=rep(st["f_std"]), l_mean=st["l_mean"], l_std=st["l_std"])
    return dict(f_mean=st["f_mean"][:window].contiguous(), f_std=st["f_std"][:window].contiguous(), l_mean=st["l_mean"], l_std=st["l_std"])


@pytest.mark.parametrize("layers,window,N,H", [(4, 4, 300, 20), (2, 4, 129, 6), (1, 4, 128, 3), (3, 3, 77, 5), (2, 1, 40, 4), (2, 8, 130, 3)])
def test_t3_bf16_lstm256_horizon(layers, window, N, H, gold_stats):
    """Ragged N (partial last tile), several depths / windows: one-step against torch FP32, quantiles against FP64."""
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, ref = _pair(layers, window)
    st = _stats(gold_stats, window)
    init, U = synth.make_inputs(synth.load_rows(), N, H, window=window, seed=3)
    y32 = R.rollout_rnn(ref, st, init, U, stateful=False)[0]
    y64 = R.rollout_rnn(R.as_double(ref), {k: v.double() for k, v in st.items()}, init.double(), U.double(), stateful=False)[0]
    y = m.rollout(init.to(_dev()), U.to(_dev()), st, precision="bf16")[0].cpu()
    assert y.shape == (N, H, 6) and not torch.isnan(y).any()
    one, quant = (ONE_STEP, QUANT) if (layers, window) == (4, 4) else (ONE_STEP_ANY, QUANT_ANY)
    assert R.rel_err(y, y32)[:, 0].max().item() <= one
    q = _quantiles(R.rel_err(y, y64))
    for a, b in zip(q, quant):
        assert a <= b, (q, quant)
    # the FP32 FFMA kernel of the same model is the parity path: T1 one-step
    yk = m.rollout(init.to(_dev()), U.to(_dev()), st, precision="fp32")[0].cpu()
    assert R.rel_err(yk, y32)[:, 0].max().item() <= 1e-5


def test_bf16_lstm256_multi_tile_cost_and_shards(gold_stats):
    """More tiles than SMs (CTAs loop over tiles), cost-only mode, bitwise repeatability, shard invariance."""
    from mpcdl import synth
    m, _ = _pair(2)
    st = gold_stats[1]
    N, H = 148 * 128 + 77, 3
    rows = synth.load_rows()
    init, U = synth.make_inputs(rows, N, H, window=4, seed=5)
    init, U = init.to(_dev()), U.to(_dev())
    cost = tuple(c.to(_dev()) for c in synth.make_cost(rows, H))
    s, c = m.rollout(init, U, st, precision="bf16", cost=cost)
    s2, c2 = m.rollout(init, U, st, precision="bf16", cost=cost)
    assert torch.equal(s, s2) and torch.equal(c, c2)
    _, c3 = m.rollout(init, U, st, precision="bf16", cost=cost, return_states=False)
    assert torch.equal(c3, c)
    s32, c32 = m.rollout(init, U, st, precision="fp32", cost=cost)
    assert ((s - s32).abs().amax((0, 1)) / s32.abs().amax((0, 1))).max().item() <= ONE_STEP_ANY
    assert ((c - c32).abs() / c32.abs()).max().item() <= 1e-3
    # a tile-aligned shard gives the same bits as the same trajectories inside the full batch
    k = 100 * 128
    part = m.rollout(init[k:].contiguous(), U[k:].contiguous(), st, precision="bf16")[0]
    assert torch.equal(part, s[k:])


def test_bf16_lstm256_sampled_controls_shared_init(gold_stats):
    """MPPI call form: controls drawn in-kernel around a nominal sequence, one shared initial window; the same
    Philox stream as the FP32 kernel, so the two precisions roll out the SAME control samples."""
    from mpcdl import synth
    m, _ = _pair(2)
    st = gold_stats[1]
    rows = synth.load_rows()
    H = 4
    init, U = synth.make_inputs(rows, 1, H, window=4, seed=9)
    cost = tuple(c.to(_dev()) for c in synth.make_cost(rows, H))
    sample = dict(n=1000, sigma=(0.2, 0.1), lo=(-0.875, -0.285), hi=(0.237, 0.279), seed=11)
    kw = dict(cost=cost, sample=sample, shared_init=True)
    sb, cb = m.rollout(init[0].to(_dev()), U[0].to(_dev()), st, precision="bf16", **kw)
    sf, cf = m.rollout(init[0].to(_dev()), U[0].to(_dev()), st, precision="fp32", **kw)
    assert sb.shape == (1000, H, 6)
    assert ((sb - sf).abs().amax((0, 1)) / sf.abs().amax((0, 1))).max().item() <= ONE_STEP_ANY
    assert ((cb - cf).abs() / cf.abs()).max().item() <= 1e-3
    half = m.rollout(init[0].to(_dev()), U[0].to(_dev()), st, precision="bf16", traj_offset=512,
                     cost=cost, sample=dict(sample, n=488), shared_init=True)[0]
    assert torch.equal(half, sb[512:])


def test_bf16_rejects_unsupported(gold_stats):
    """No silent fallback: BF16 exists for the hidden-256 LSTM only; a noise net too large for the shared memory the kernel
    can lend it (the A_h tile, 64 KB) is refused."""
    import mpcdl
    m, _ = _pair(1)
    win, U = torch.zeros(8, 4, 8, device=_dev()), torch.zeros(8, 2, 2, device=_dev())
    noise = mpcdl.FeedforwardNoiseModel(6, 2, 8, n_layers=2, size=64, activation="tanh", output_activation="identity").to(_dev())
    with pytest.raises(RuntimeError):
        m.rollout(win, U, None, precision="bf16", noise=noise)
    with pytest.raises(RuntimeError):           # hidden 256 has no TF32 kernel
        m.rollout(win, U, None, precision="tf32")
    with pytest.raises(RuntimeError):           # CPU tensors: no fallback
        m.rollout(win.cpu(), U.cpu(), None, precision="bf16")


def test_bf16_lstm256_full_size_properties(gold_stats):
    """BASELINE config 4 at its full size (65 536 trajectories x horizon 100, 512 tiles on 74 CTA pairs):
    size-independent properties instead of an oracle run -- (1) a trajectory's result does not depend on which tile /
    CTA / wave it lands in: rolling out a permutation of the batch gives the permuted result BIT FOR BIT;
    (2) a 256-trajectory sample agrees with the FP32 FFMA kernel on the same trajectories within the T3 bound over
    the whole horizon; (3) costs equal the cost recomputed from the returned states."""
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, _ = _pair(4)
    st = gold_stats[1]
    N, H = 65536, 100
    rows = synth.load_rows()
    init, U = synth.make_inputs(rows, N, H, window=4, seed=0)
    init, U = init.to(_dev()), U.to(_dev())
    cost = tuple(c.to(_dev()) for c in synth.make_cost(rows, H))
    s, c = m.rollout(init, U, st, precision="bf16", cost=cost)
    assert not torch.isnan(s).any()
    perm = torch.randperm(N, generator=torch.Generator().manual_seed(1)).to(_dev())
    sp, cp = m.rollout(init[perm].contiguous(), U[perm].contiguous(), st, precision="bf16", cost=cost)
    assert torch.equal(sp, s[perm]) and torch.equal(cp, c[perm])
    del sp, cp
    idx = perm[:256]
    s32 = m.rollout(init[idx].contiguous(), U[idx].contiguous(), st, precision="fp32")[0]
    q = _quantiles(R.rel_err(s[idx].cpu(), s32.cpu().double()))
    for a, b in zip(q, QUANT):
        assert a <= b, (q, QUANT)
    ref, qw, rw = cost
    want = (((s[idx] - ref[None]) ** 2) * qw).sum((1, 2)) + ((U[idx] ** 2) * rw).sum((1, 2))
    assert ((c[idx] - want).abs() / want.abs()).max().item() <= 1e-4


def test_c4_tolerance_study_h100_against_fp64_oracle(gold_stats):
    """BASELINE.json config 4 as a TEST (it used to be the tool output profiles/r01f/c4_tolerance_study.md): 256
    trajectories x the full horizon 100 of the 4-layer hidden-256 LSTM against the FP64 ORACLE, per-step quantiles.
    Measured (r01f): torch FP32 p50 6.7e-8 / max 3.7e-7; FP32 kernel p50 3.7e-7 / max 2.3e-6; BF16 kernel p50 2.2e-4 /
    max 7.9e-4, flat over the horizon.  Bounds: FP32 kernel p50 <= 1e-6, p90 <= 3e-6, p99 <= 5e-6 (its 256-deep sums run k-ascending
    in one FMA chain where torch's sgemm blocks them, so it sits ~5x above torch FP32's own error and 2x torch's is not the
    bound here, unlike the hidden-64 models of T2); BF16 kernel p50 <= 6e-4, p90 <= 1.2e-3, p99 <= 2e-3 at EVERY checked step (1, 10, 50, 100) and over all steps."""
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, ref = _pair(4)
    st = gold_stats[1]
    N, H = 256, 100
    init, U = synth.make_inputs(synth.load_rows(), N, H, window=4, seed=3)
    y64 = R.rollout_rnn(R.as_double(ref), {k: v.double() for k, v in st.items()}, init.double(), U.double(), stateful=False)[0]
    e_torch = R.rel_err(R.rollout_rnn(ref, st, init, U, stateful=False)[0], y64)
    e_fp32 = R.rel_err(m.rollout(init.to(_dev()), U.to(_dev()), st, precision="fp32")[0].cpu(), y64)
    e_bf16 = R.rel_err(m.rollout(init.to(_dev()), U.to(_dev()), st, precision="bf16")[0].cpu(), y64)
    for a, bound in zip(_quantiles(e_fp32), (1e-6, 3e-6, 5e-6)):
        assert a <= bound, (_quantiles(e_fp32), _quantiles(e_torch))
    assert e_fp32.max().item() <= 2e-5
    for sl in (slice(0, 1), slice(9, 10), slice(49, 50), slice(99, 100), slice(0, H)):
        for got, bound in zip(_quantiles(e_bf16[:, sl]), QUANT):
            assert got <= bound, (sl, got, bound)
    # flat over the horizon: the closed loop does not amplify the BF16 operand rounding of this model
    assert _quantiles(e_bf16[:, 99:100])[0] <= 3.0 * _quantiles(e_bf16[:, 0:1])[0] + 1e-4


@pytest.mark.parametrize("layers,window,N,H", [(4, 4, 300, 12), (2, 4, 129, 6), (2, 1, 40, 5), (1, 8, 130, 3)])
def test_bf16_lstm256_stateful_carry(layers, window, N, H, gold_stats):
    """Stateful hidden-256 LSTM on the BF16 kernel (deep_models.py:88-91: (h, c) carried from call to call): states within
    the T3 bounds of the FP64 oracle, the returned carry against the oracle's final (h, c), and H steps in one launch ==
    two launches of H1 + H2 steps chained through the carry, bit for bit."""
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, ref = _pair(layers, window)
    st = _stats(gold_stats, window)
    init, U = synth.make_inputs(synth.load_rows(), N, H, window=window, seed=9)
    di, dU = init.to(_dev()), U.to(_dev())
    want, _, fin = R.rollout_rnn(R.as_double(ref), {k: v.double() for k, v in st.items()}, init.double(), U.double(), stateful=True)
    s, _, carry = m.rollout(di, dU, st, precision="bf16", stateful=True)
    assert torch.isfinite(s).all() and carry.shape == (2 * layers, N, 256)
    one, quant = (ONE_STEP, QUANT) if (layers, window) == (4, 4) else (ONE_STEP_ANY, QUANT_ANY)
    err = R.rel_err(s.cpu(), want)
    assert err[:, 0].max().item() <= one
    for got, bound in zip(_quantiles(err), quant):
        assert got <= bound, (_quantiles(err), quant)
    fin = torch.cat(list(fin), 0)
    assert ((carry.cpu().double() - fin).abs().max() / fin.abs().max()).item() <= 2e-2
    # the FP32 kernel's stateful result is the tight reference for the states
    s32 = m.rollout(di, dU, st, precision="fp32", stateful=True)[0]
    assert R.rel_err(s.cpu(), s32.cpu()).max().item() <= 5 * quant[2]
    # chaining: H1 steps, then H - H1 steps from the returned carry (the window the second call starts from is rebuilt from
    # the first call's states and controls, App. A.4)
    if window > 1 and H >= window + 2:
        H1 = window + 1
        s_a, _, c_a = m.rollout(di, dU[:, :H1].contiguous(), st, precision="bf16", stateful=True)
        frames = torch.cat([s_a, dU[:, :H1]], dim=-1)                      # appended frames [s_{t+1}, u_t]
        win_b = frames[:, H1 - window:H1].contiguous()
        s_b = m.rollout(win_b, dU[:, H1:].contiguous(), st, precision="bf16", stateful=True, carry=c_a)[0]
        assert torch.equal(torch.cat([s_a, s_b], 1), s)


@pytest.mark.parametrize("stateful", [False, True])
def test_bf16_lstm256_residual_noise(stateful, gold_stats):
    """Residual-noise net fused into the BF16 kernel (noise_model.py:34-40, history 3, 32 x 2): supplied eps against the
    oracle with the same draws; in-kernel Philox draws are shard invariant (traj_offset) and repeatable."""
    import mpcdl
    from conftest import load_ckpt
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, ref = _pair(2)
    st = gold_stats[1]
    N, H = 300, 10
    init, U = synth.make_inputs(synth.load_rows(), N, H, window=4, seed=10)
    eps = synth.make_eps(N, H, seed=4)
    noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    noise.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    noise.to(_dev())
    rn = R.RefNoise(6, 2, 3, 2, 32, "tanh", "identity")
    rn.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    rn.eval()
    di, dU = init.to(_dev()), U.to(_dev())
    want = R.rollout_rnn(R.as_double(ref), {k: v.double() for k, v in st.items()}, init.double(), U.double(), stateful=stateful,
                         noise=R.as_double(rn), eps=eps.double())[0]
    got = m.rollout(di, dU, st, precision="bf16", noise=noise, eps=eps.to(_dev()), stateful=stateful)[0]
    err = R.rel_err(got.cpu(), want)
    assert err[:, 0].max().item() <= ONE_STEP_ANY
    for g_, bound in zip(_quantiles(err), QUANT_ANY):
        assert g_ <= bound, (_quantiles(err), QUANT_ANY)
    # the noise term itself is FP32 on the CUDA cores: against the FP32 kernel with the same draws the gap stays BF16-sized
    g32 = m.rollout(di, dU, st, precision="fp32", noise=noise, eps=eps.to(_dev()), stateful=stateful)[0]
    assert R.rel_err(got.cpu(), g32.cpu()).max().item() <= 5 * QUANT_ANY[2]
    a = m.rollout(di, dU, st, precision="bf16", noise=noise, seed=3, stateful=stateful)[0]
    assert torch.equal(a, m.rollout(di, dU, st, precision="bf16", noise=noise, seed=3, stateful=stateful)[0])
    h = 128                                      # a tile boundary and a ragged tail
    b = m.rollout(di[h:].contiguous(), dU[h:].contiguous(), st, precision="bf16", noise=noise, seed=3, traj_offset=h, stateful=stateful)[0]
    assert torch.equal(b, a[h:])
    assert not torch.equal(a, got)               # different draws
