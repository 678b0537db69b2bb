"""GPU parity tests: the CUDA path (through torch.ops.mpcdl -> C ABI) against the golden vectors made
from the reference modules (oracle/pin_reference.py) and against the live oracle.

Tolerances (DESIGN.md "Tolerances"):
  T1 one-step      max|y_kernel - y_ref| / max|y_ref| <= 1e-5       (teacher forced, every row/window)
  T2 full horizon  vs an FP64 run of the same module: kernel error quantiles (p50,p90,p99) <= 2x the
                   quantiles of torch-FP32's own error + 1e-6, and kernel-vs-torch-FP32 p90 <= 1e-5
"""
import numpy as np
import pytest
import torch

from conftest import FWD, RNN, load_ckpt

pytestmark = pytest.mark.gpu
T1 = 1e-5


def _dev():
    return torch.device("cuda:0")


def _t1(y, ref):
    ref = torch.as_tensor(ref)
    return ((y.cpu() - ref).abs().max() / ref.abs().max()).item()


def _model(name):
    import mpcdl
    m, meta = mpcdl.from_checkpoint_name(name, load_ckpt(name), _dev())
    return m, meta


@pytest.mark.parametrize("name", FWD)
def test_t1_onestep_mlp(name, bags, gold_onestep):
    import mpcdl
    ds = mpcdl.RobotData(bags)
    m, _ = _model(name)
    y = m(ds.featrs.to(_dev()))
    assert y.shape == (4550, 6)
    assert _t1(y, gold_onestep[name]) <= T1


@pytest.mark.parametrize("name", RNN)
def test_t1_onestep_rnn(name, bags, gold_onestep):
    import mpcdl
    ds = mpcdl.RobotData(bags, 4)
    m, meta = _model(name)
    x = ds.featrs.to(_dev())
    y = m(x)
    assert y.shape == (4546, 6)
    assert _t1(y, gold_onestep[name]) <= T1
    if meta["state_f"]:      # second call carries h_state (deep_models.py:61-65)
        y2 = m(x)
        assert _t1(y2, gold_onestep[name + "_second"]) <= T1
        with pytest.raises(RuntimeError):          # fixed batch size of the carried state
            m(x[:7])


def _quantiles(e):
    e = e.flatten().double()
    return [torch.quantile(e, q).item() for q in (0.5, 0.9, 0.99)]


@pytest.mark.parametrize("name", FWD + RNN)
def test_golden_rollout(name, gold_rollouts, gold_stats):
    """N=64, H=24 closed loop produced by the REFERENCE modules in a python loop."""
    from oracle import rollout_ref as R
    m, meta = _model(name)
    mlp = meta["kind"] == "mlp"
    st = gold_stats[0] if mlp else gold_stats[1]
    init = torch.from_numpy(gold_rollouts["s0" if mlp else "win0"]).to(_dev())
    U = torch.from_numpy(gold_rollouts["U" if mlp else "Uw"]).to(_dev())
    out = m.rollout(init, U, st)
    ref = torch.from_numpy(gold_rollouts[name])
    err = R.rel_err(out[0].cpu(), ref)
    assert err[:, 0].max().item() <= T1                       # first step is a one-step prediction
    q = _quantiles(err)
    assert q[1] <= 1e-5, q                                    # p90 over the 24-step horizon
    assert err.max().item() <= 1e-3, err.max().item()


def test_golden_noise_cost(gold_rollouts, gold_stats):
    import mpcdl
    from oracle import rollout_ref as R
    noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    noise.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    noise.to(_dev())
    g = gold_rollouts
    cost = tuple(torch.from_numpy(g[k]) for k in ("cost_ref", "cost_qw", "cost_rw"))
    eps = torch.from_numpy(g["eps"]).to(_dev())
    for name, init_k, u_k, st in (("Basic_64_4_32", "s0", "U", gold_stats[0]), ("GRU_64_4_4_16_False", "win0", "Uw", gold_stats[1])):
        m, _ = _model(name)
        states, costs = m.rollout(torch.from_numpy(g[init_k]).to(_dev()), torch.from_numpy(g[u_k]).to(_dev()), st,
                                  noise=noise, eps=eps, cost=cost)
        err = R.rel_err(states.cpu(), torch.from_numpy(g[name + "_noise"]))
        assert err[:, 0].max().item() <= T1
        assert _quantiles(err)[1] <= 1e-5
        cref = torch.from_numpy(g[name + "_noise_cost"])
        assert ((costs.cpu() - cref).abs() / cref.abs()).max().item() <= 1e-4
    # batch-1 golden made through the reference's own DynamicsNN.get_prediction wrapper (noise_model.py:114-125)
    m, _ = _model("Basic_64_4_32")
    states, _ = m.rollout(torch.from_numpy(g["s0"][:1]).to(_dev()), torch.from_numpy(g["U"][:1]).to(_dev()), gold_stats[0],
                          noise=noise, eps=torch.from_numpy(g["eps_b1"]).to(_dev()))
    err = R.rel_err(states.cpu(), torch.from_numpy(g["Basic_64_4_32_noise_b1"]))
    assert err.max().item() <= 1e-4


@pytest.mark.parametrize("name,N,H", [("Basic_64_4_32", 4096, 50), ("Basic_16_1_32", 1000, 50),
                                      ("GRU_64_4_4_16_False", 300, 50), ("GRU_64_4_4_64_True", 300, 50),
                                      ("LSTM_64_4_4_16_False", 300, 50), ("LSTM_64_4_4_64_True", 300, 50)])
def test_t2_full_horizon(name, N, H, gold_stats):
    """T2: error quantiles against the FP64 oracle are no worse than 2x torch-FP32's own."""
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, meta = _model(name)
    mlp = meta["kind"] == "mlp"
    st = gold_stats[0] if mlp else gold_stats[1]
    init, U = synth.make_inputs(synth.load_rows(), N, H, window=meta["window"], seed=3)
    ref_m, _ = R.build_from_name(name, load_ckpt(name))
    ref64 = R.as_double(ref_m)
    st64 = {k: v.double() for k, v in st.items()}
    if mlp:
        y32 = R.rollout_mlp(ref_m, st, init, U)[0]
        y64 = R.rollout_mlp(ref64, st64, init.double(), U.double())[0]
    else:
        y32 = R.rollout_rnn(ref_m, st, init, U, stateful=meta["state_f"])[0]
        y64 = R.rollout_rnn(ref64, st64, init.double(), U.double(), stateful=meta["state_f"])[0]
    yk = m.rollout(init.to(_dev()), U.to(_dev()), st)[0].cpu()
    qk, qt = _quantiles(R.rel_err(yk, y64)), _quantiles(R.rel_err(y32, y64))
    for a, b in zip(qk, qt):
        assert a <= 2 * b + 1e-6, (qk, qt)
    assert _quantiles(R.rel_err(yk, y32))[1] <= 1e-5


# ------------------------------------------------------------------------------ T3: tensor-core precisions
# Stated bounds of the tcgen05 path, rel err = max_d |dy_d| / max|ref_d| per trajectory-step:
#   tf32x3  one-step <= 2e-5  (vs the reference modules' FP32 output);  H = 50 vs FP64: p50 <= 4e-5, p90 <= 1e-4, p99 <= 3e-4
#   tf32    one-step <= 5e-3;                                           H = 50 vs FP64: p50 <= 3e-3, p90 <= 8e-3, p99 <= 3e-2
TC_BOUNDS = {"tf32x3": (2e-5, (4e-5, 1e-4, 3e-4)), "tf32": (5e-3, (3e-3, 8e-3, 3e-2))}


@pytest.mark.parametrize("prec", ["tf32x3", "tf32"])
@pytest.mark.parametrize("name", RNN)
def test_t3_tensor_core_onestep(name, prec, bags, gold_onestep):
    """Module.forward on every validation window through the tcgen05 kernel (stateful: carried h_state)."""
    import mpcdl
    ds = mpcdl.RobotData(bags, 4)
    m, meta = _model(name)
    m.precision = prec
    x = ds.featrs.to(_dev())
    y = m(x)
    assert _t1(y, gold_onestep[name]) <= TC_BOUNDS[prec][0]
    if meta["state_f"]:
        y2 = m(x)
        assert _t1(y2, gold_onestep[name + "_second"]) <= 2 * TC_BOUNDS[prec][0]


@pytest.mark.parametrize("prec", ["tf32x3", "tf32"])
@pytest.mark.parametrize("name", ["GRU_64_4_4_16_False", "GRU_64_4_4_64_True", "LSTM_64_4_4_16_False", "LSTM_64_4_4_64_True"])
def test_t3_tensor_core_horizon(name, prec, gold_stats):
    """Ragged N (300 = 2 tiles + 44), H = 50: error quantiles against the FP64 oracle within the stated bound."""
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, meta = _model(name)
    st = gold_stats[1]
    init, U = synth.make_inputs(synth.load_rows(), 300, 50, window=4, seed=3)
    ref64 = R.as_double(R.build_from_name(name, load_ckpt(name))[0])
    y64 = R.rollout_rnn(ref64, {k: v.double() for k, v in st.items()}, init.double(), U.double(), stateful=meta["state_f"])[0]
    out = m.rollout(init.to(_dev()), U.to(_dev()), st, precision=prec)
    q = _quantiles(R.rel_err(out[0].cpu(), y64))
    for a, b in zip(q, TC_BOUNDS[prec][1]):
        assert a <= b, (q, TC_BOUNDS[prec][1])
    if meta["state_f"]:                       # the returned carry continues the rollout exactly like one long rollout
        a = m.rollout(init.to(_dev()), U[:, :20].contiguous().to(_dev()), st, precision=prec)
        win = torch.cat([init.to(_dev()), torch.cat([a[0], U[:, :20].to(_dev())], 2)], 1)[:, -4:].contiguous()
        b = m.rollout(win, U[:, 20:].contiguous().to(_dev()), st, precision=prec, carry=a[2])
        assert torch.equal(torch.cat([a[0], b[0]], 1), out[0])


@pytest.mark.parametrize("prec", ["tf32x3", "tf32"])
def test_t3_tensor_core_noise_cost(prec, gold_rollouts, gold_stats):
    """Residual-noise injection with supplied eps + tracking cost on the tcgen05 path vs the golden rollout
    made by the reference modules; Philox draws are invariant to sharding (traj_offset)."""
    import mpcdl
    from oracle import rollout_ref as R
    noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    noise.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    noise.to(_dev())
    g = gold_rollouts
    cost = tuple(torch.from_numpy(g[k]) for k in ("cost_ref", "cost_qw", "cost_rw"))
    eps = torch.from_numpy(g["eps"]).to(_dev())
    name = "GRU_64_4_4_16_False"
    m, _ = _model(name)
    win0, Uw = torch.from_numpy(g["win0"]).to(_dev()), torch.from_numpy(g["Uw"]).to(_dev())
    states, costs = m.rollout(win0, Uw, gold_stats[1], noise=noise, eps=eps, cost=cost, precision=prec)
    err = R.rel_err(states.cpu(), torch.from_numpy(g[name + "_noise"]))
    assert err[:, 0].max().item() <= TC_BOUNDS[prec][0]
    assert _quantiles(err)[1] <= TC_BOUNDS[prec][1][1]
    cref = torch.from_numpy(g[name + "_noise_cost"])
    assert ((costs.cpu() - cref).abs() / cref.abs()).max().item() <= (1e-3 if prec == "tf32x3" else 0.2)
    # cost-only mode returns the same costs; Philox noise keyed by the global trajectory index
    _, c2 = m.rollout(win0, Uw, gold_stats[1], noise=noise, eps=eps, cost=cost, precision=prec, return_states=False)
    assert torch.equal(c2, costs)
    full = m.rollout(win0, Uw, gold_stats[1], noise=noise, seed=7, precision=prec)[0]
    part = m.rollout(win0[40:].contiguous(), Uw[40:].contiguous(), gold_stats[1], noise=noise, seed=7, traj_offset=40, precision=prec)[0]
    assert torch.equal(full[40:], part)


# tcgen05 MLP (rollout_mlp_tc.cu), hidden-64 checkpoints.  Stated bounds (profiles/r01c/mc_check_v2.log):
#   tf32x3  one-step <= 5e-6 (FP32-class: measured 4e-7..9e-7);  H = 50 vs FP64: p50 <= 1e-5, p90 <= 3e-5, p99 <= 1e-4
#   tf32    one-step <= 3e-3;                                     H = 50 vs FP64: p50 <= 4e-3, p90 <= 1.2e-2, p99 <= 4e-2
MC_BOUNDS = {"tf32x3": (5e-6, (1e-5, 3e-5, 1e-4)), "tf32": (3e-3, (4e-3, 1.2e-2, 4e-2))}
FWD64 = [n for n in FWD if n.startswith("Basic_64")]


@pytest.mark.parametrize("prec", ["tf32x3", "tf32"])
@pytest.mark.parametrize("name", FWD64)
def test_t3_mlp_tensor_core_onestep(name, prec, bags, gold_onestep):
    """Basic_Model.forward on every validation row through the tcgen05 MLP kernel (4550 rows = 35 tiles + 70)."""
    import mpcdl
    ds = mpcdl.RobotData(bags)
    m, _ = _model(name)
    m.precision = prec
    y = m(ds.featrs.to(_dev()))
    assert y.shape == (4550, 6)
    assert _t1(y, gold_onestep[name]) <= MC_BOUNDS[prec][0]


@pytest.mark.parametrize("prec", ["tf32x3", "tf32"])
@pytest.mark.parametrize("n", [300, 40000])
@pytest.mark.parametrize("name", ["Basic_64_4_32", "Basic_64_1_32"])
def test_t3_mlp_tensor_core_horizon(name, n, prec, gold_stats):
    """H = 50, ragged N; 40000 > 148 tiles exercises the two-tiles-per-CTA ping-pong and the persistent loop.
    Quantiles vs the FP64 oracle within the stated bound; cost-only mode returns identical costs."""
    from mpcdl import synth
    from oracle import rollout_ref as R
    m, _ = _model(name)
    st = gold_stats[0]
    init, U = synth.make_inputs(synth.load_rows(), n, 50, window=0, seed=5)
    ref64 = R.as_double(R.build_from_name(name, load_ckpt(name))[0])
    ns = min(n, 2000)                                     # the FP64 oracle runs on a slice that spans tile boundaries
    sl = slice(n - ns, n)
    y64 = R.rollout_mlp(ref64, {k: v.double() for k, v in st.items()}, init[sl].double(), U[sl].double())[0]
    cost = (torch.zeros(50, 6), torch.ones(6), torch.ones(2))
    states, costs = m.rollout(init.to(_dev()), U.to(_dev()), st, precision=prec, cost=cost)
    assert states.shape == (n, 50, 6) and torch.isfinite(states).all()
    q = _quantiles(R.rel_err(states[sl].cpu(), y64))
    for a, b in zip(q, MC_BOUNDS[prec][1]):
        assert a <= b, (q, MC_BOUNDS[prec][1])
    _, c2 = m.rollout(init.to(_dev()), U.to(_dev()), st, precision=prec, cost=cost, return_states=False)
    assert torch.equal(c2, costs)
    want = (states ** 2).sum((1, 2)) + (U.to(_dev()) ** 2).sum((1, 2))
    assert ((costs - want).abs() / want).max().item() <= 1e-4
    # a shard of the batch gives the same bits as the same trajectories inside the full batch
    part = m.rollout(init[128:].contiguous().to(_dev()), U[128:].contiguous().to(_dev()), st, precision=prec)[0]
    assert torch.equal(part, states[128:])


@pytest.mark.parametrize("name", ["Basic_64_4_32", "Basic_64_1_32"])
def test_t3_mlp_tensor_core_four_tiles_per_cta(name, gold_stats):
    """One-pass TF32 with FOUR tile slots per CTA (128 TMEM columns each, one epilogue thread per trajectory): forced with
    tune=(0, 4) on a ragged batch (97 tiles -> 25 CTAs, the last with one tile) and picked by the planner at >= 8 tiles per SM.
    Same MMAs as the two-tile form; the FP32 output layer sums its 64 units in one thread instead of two halves, so the two
    forms agree to FP32 rounding of that sum (<= 2e-6 of the state scale per step, fed back), not bit for bit.  Bounds: T3."""
    from mpcdl import ops, synth
    from oracle import rollout_ref as R
    m, _ = _model(name)
    st = gold_stats[0]
    n = 96 * 128 + 70
    init, U = synth.make_inputs(synth.load_rows(), n, 50, window=0, seed=6)
    d_init, d_U = init.to(_dev()), U.to(_dev())
    cost = (torch.zeros(50, 6), torch.ones(6), torch.ones(2))
    s4, c4 = m.rollout(d_init, d_U, st, precision="tf32", cost=cost, tune=(0, 4, 0, 0))
    s2, c2 = m.rollout(d_init, d_U, st, precision="tf32", cost=cost, tune=(0, 2, 0, 0))
    assert torch.isfinite(s4).all() and not torch.equal(s4, torch.zeros_like(s4))
    one = R.rel_err(s4[:, :1].cpu(), s2[:, :1].cpu()).max().item()
    assert one <= 2e-6, one                                # the first step differs only by the output layer's summation order
    ns = 1500
    sl = slice(n - ns, n)
    ref64 = R.as_double(R.build_from_name(name, load_ckpt(name))[0])
    y64 = R.rollout_mlp(ref64, {k: v.double() for k, v in st.items()}, init[sl].double(), U[sl].double())[0]
    q = _quantiles(R.rel_err(s4[sl].cpu(), y64))
    for a, b in zip(q, MC_BOUNDS["tf32"][1]):
        assert a <= b, (q, MC_BOUNDS["tf32"][1])
    want = (s4 ** 2).sum((1, 2)) + (d_U ** 2).sum((1, 2))
    assert ((c4 - want).abs() / want).max().item() <= 1e-4
    # shard invariance and repeatability, bit for bit, in the four-tile form
    part = m.rollout(d_init[640:].contiguous(), d_U[640:].contiguous(), st, precision="tf32", tune=(0, 4, 0, 0))[0]
    assert torch.equal(part, s4[640:])
    assert torch.equal(m.rollout(d_init, d_U, st, precision="tf32", cost=cost, tune=(0, 4, 0, 0))[0], s4)
    # the planner's pick at >= 8 tiles per SM, cost-only (1200 tiles per ... 160 000 trajectories): equals the forced form
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert ops.plan(0, 64, 4, 0, 8 * sms * 128, 50, precision="tf32")["ksplit"] == 4
    with pytest.raises(RuntimeError):                      # tf32x3 needs 192 columns per tile: no four-tile form
        m.rollout(d_init, d_U, st, precision="tf32x3", tune=(0, 4, 0, 0))


def test_tensor_core_rejects_unsupported():
    m, _ = _model("Basic_16_4_32")
    with pytest.raises(RuntimeError):           # hidden 16: FP32 kernels only, no silent fallback
        m.rollout(torch.zeros(8, 6, device=_dev()), torch.zeros(8, 3, 2, device=_dev()), None, precision="tf32x3")
    g, _ = _model("GRU_64_4_4_16_False")
    with pytest.raises(RuntimeError):           # window 3: the tcgen05 kernel is built for FT_SIZE = 4 only
        g.rollout(torch.zeros(8, 3, 8, device=_dev()), torch.zeros(8, 3, 2, device=_dev()), None, precision="tf32")
    with pytest.raises(RuntimeError):
        g.rollout(torch.zeros(8, 4, 8, device=_dev()), torch.zeros(8, 3, 2, device=_dev()), None, precision="bf16")


@pytest.mark.parametrize("prec", ["tf32x3", "tf32"])
@pytest.mark.parametrize("n,h", [(1, 1), (129, 3), (300, 9)])
def test_tensor_core_noise_buffers_inside_the_ring(prec, n, h, gold_stats):
    """GRU + residual-noise net on the tcgen05 kernel: with the noise net's buffers aliased into the weight ring (what makes a
    tf32x3 layer resident, api.cu make_tc_plan) the result equals the streaming launch (tune nslots = -1: own noise region,
    history pushed step by step) bit for bit -- the history rebuilt from the window buffer is the deque of
    noise_model.py:95-99, including the left-padding of the first steps."""
    import mpcdl
    from mpcdl import _lib, ops, synth
    m, _ = mpcdl.from_checkpoint_name("GRU_64_4_4_16_False", load_ckpt("GRU_64_4_4_16_False"), _dev())
    noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    noise.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    noise.to(_dev())
    st = gold_stats[1]
    init, U = synth.make_inputs(synth.load_rows(), n, h, window=4, seed=17)
    eps = synth.make_eps(n, h, seed=6).to(_dev())
    cost = tuple(c.to(_dev()) for c in synth.make_cost(synth.load_rows(), h))
    init, U = init.to(_dev()), U.to(_dev())
    pa = ops.plan(_lib.GRU, 64, 4, 4, n, h, noise_cfg=noise.cfg(), precision=prec)
    ps = ops.plan(_lib.GRU, 64, 4, 4, n, h, noise_cfg=noise.cfg(), precision=prec, tune=(0, 0, -1, 0))
    if prec == "tf32x3":
        assert pa["resident"] == 1 and ps["resident"] == 0 and pa["smem_bytes"] <= 232448
    for kw in (dict(eps=eps), dict(seed=9)):
        a = m.rollout(init, U, st, noise=noise, cost=cost, precision=prec, **kw)
        b = m.rollout(init, U, st, noise=noise, cost=cost, precision=prec, tune=(0, 0, -1, 0), **kw)
        assert torch.isfinite(a[0]).all() and torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
