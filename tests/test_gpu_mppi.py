"""GPU tests of the MPPI planner call: in-kernel control sampling, shared initial state, and the
softmin update, against the CPU restatement in oracle/mppi_ref.py (Philox4x32-10 pinned to the Random123 known-answer
vectors; the MPPI update itself has no reference counterpart -- parity unpinned, see the oracle's header).
Bounds: sampled controls agree with the oracle's to 2e-6 absolute (GPU sincosf/logf vs numpy); a rollout on
in-kernel samples agrees with the same kernel fed the oracle's samples to p90 <= 1e-5 (T2 metric); the update sums
agree to 1e-4 relative."""
import numpy as np
import pytest
import torch

from conftest import load_ckpt

pytestmark = pytest.mark.gpu
SIG, LO, HI = (0.06, 0.03), (-0.875, -0.285), (0.237, 0.279)


def _dev():
    return torch.device("cuda:0")


def _setup(name, gold_stats, H=30):
    import mpcdl
    from mpcdl import synth
    m, meta = mpcdl.from_checkpoint_name(name, load_ckpt(name), _dev())
    st = gold_stats[0] if meta["kind"] == "mlp" else gold_stats[1]
    rows = synth.load_rows()
    init, U = synth.make_inputs(rows, 1, H, window=meta["window"], seed=2)
    cost = tuple(c.to(_dev()) for c in synth.make_cost(rows, H))
    return m, meta, st, init[0].contiguous().to(_dev()), U[0].contiguous().to(_dev()), cost


@pytest.mark.parametrize("name,prec,noise_on", [("Basic_64_4_32", "fp32", False), ("Basic_64_4_32", "fp32", True),
                                                 ("Basic_16_2_32", "fp32", False), ("Basic_64_4_32", "tf32x3", False),
                                                 ("GRU_64_4_4_16_False", "fp32", False), ("GRU_64_4_4_16_False", "tf32x3", True),
                                                 ("LSTM_64_4_4_64_True", "fp32", False)])
def test_sampled_rollout_matches_explicit_controls(name, prec, noise_on, gold_stats):
    import mpcdl
    from oracle import mppi_ref as M
    from oracle import rollout_ref as R
    m, meta, st, init, Ub, cost = _setup(name, gold_stats)
    noise = None
    if noise_on:
        noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
        noise.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
        noise.to(_dev())
    N, off, seed = 700, 1000, 11
    sample = dict(n=N, sigma=SIG, lo=LO, hi=HI, seed=seed)
    kw = dict(noise=noise, seed=3, traj_offset=off, cost=cost, precision=prec)
    s_a, c_a = m.rollout(init, Ub, st, sample=sample, shared_init=True, **kw)[:2]
    u = torch.from_numpy(M.sample_controls(Ub.cpu().numpy(), seed, np.arange(off, off + N), SIG, LO, HI)).to(_dev())
    assert LO[0] <= u[..., 0].min() and u[..., 0].max() <= HI[0] and (u[..., 0] == HI[0]).any()      # the clamp is exercised
    init_n = init.expand(N, *init.shape).contiguous()
    s_b, c_b = m.rollout(init_n, u, st, **kw)[:2]
    err = R.rel_err(s_a.cpu(), s_b.cpu())
    q = torch.quantile(err.flatten(), torch.tensor([0.5, 0.9], dtype=torch.float64))
    bound = 1e-5 if prec == "fp32" else 1e-4
    assert err[:, 0].max().item() <= bound and q[1].item() <= bound, q
    assert ((c_a - c_b).abs() / c_b.abs()).max().item() <= (1e-4 if prec == "fp32" else 1e-3)
    # shared_init is the same as N copies of the state, bit for bit; cost-only mode returns the same costs
    s_c, c_c = m.rollout(init_n, Ub, st, sample=sample, **kw)[:2]
    assert torch.equal(s_a, s_c) and torch.equal(c_a, c_c)
    c_d = m.rollout(init, Ub, st, sample=sample, shared_init=True, return_states=False, **kw)[1]
    assert torch.equal(c_a, c_d)
    # shard invariance: the samples are keyed by the global trajectory index
    sample2 = dict(sample, n=N - 300)
    c_e = m.rollout(init, Ub, st, sample=sample2, shared_init=True, return_states=False, **dict(kw, traj_offset=off + 300))[1]
    assert torch.equal(c_e, c_a[300:])


def test_sampled_controls_against_oracle_philox(gold_stats):
    """cost = 1 * u_a^2 over H = 1 reads the sampled control straight out of the kernel."""
    from oracle import mppi_ref as M
    m, meta, st, init, Ub, _ = _setup("Basic_16_1_32", gold_stats, H=1)
    N, seed = 5000, 0xABCDEF0123
    for ch in (0, 1):
        rw = torch.zeros(2)
        rw[ch] = 1.0
        cost = (torch.zeros(1, 6).to(_dev()), torch.zeros(6).to(_dev()), rw.to(_dev()))
        big = dict(n=N, sigma=SIG, lo=(-10, -10), hi=(10, 10), seed=seed)
        c = m.rollout(init, Ub, st, sample=big, shared_init=True, cost=cost, return_states=False, traj_offset=(1 << 33) + 5)[1]
        u = M.sample_controls(Ub.cpu().numpy(), seed, np.arange(N, dtype=np.uint64) + np.uint64((1 << 33) + 5), SIG, (-10, -10), (10, 10))
        assert np.abs(np.sqrt(c.cpu().numpy()) - np.abs(u[:, 0, ch])).max() <= 2e-6


def test_mppi_reduce_and_step_against_oracle(gold_stats):
    import mpcdl
    from mpcdl.mppi import MPPI
    from oracle import mppi_ref as M
    m, meta, st, init, Ub, cost = _setup("Basic_64_4_32", gold_stats, H=30)
    N, lam = 40000, 5.0
    plan = MPPI(m, st, N, SIG, LO, HI, cost, lam=lam, seed=21)
    U1, info = plan.step(init, Ub)
    costs = info["costs"].cpu().numpy()
    u = M.sample_controls(Ub.cpu().numpy(), 21, np.arange(N), SIG, LO, HI)
    U_ref, S, wsum, cmin = M.mppi_update(Ub.cpu().numpy(), u, costs, lam, LO, HI)
    assert abs(info["cmin"].item() - cmin) <= 1e-6 * abs(cmin)
    assert abs(info["wsum"].item() - wsum) <= 1e-4 * wsum
    assert np.abs(U1.cpu().numpy() - U_ref).max() <= 1e-5
    # two shards reduced with the global minimum add up to the whole (what the multi-GPU path all-reduces)
    ctrl = list(SIG) + list(LO) + list(HI)
    cm = torch.ops.mpcdl.mppi_min(info["costs"])
    full = torch.ops.mpcdl.mppi_reduce(info["costs"], Ub, ctrl, 21, 0, lam, None)
    a = torch.ops.mpcdl.mppi_reduce(info["costs"][:15000].contiguous(), Ub, ctrl, 21, 0, lam, cm)
    b = torch.ops.mpcdl.mppi_reduce(info["costs"][15000:].contiguous(), Ub, ctrl, 21, 15000, lam, cm)
    tot = (a + b)[:-1]
    assert ((tot - full[:-1]).abs() / full[:-1].abs().clamp_min(1e-3)).max().item() <= 1e-4
    assert torch.equal(torch.ops.mpcdl.mppi_reduce(info["costs"], Ub, ctrl, 21, 0, lam, None), full)      # deterministic
    # iterating the planner lowers the mean sampled cost
    means, U = [], Ub
    for _ in range(6):
        U, inf = plan.step(init, U)
        means.append(inf["costs"].mean().item())
    assert means[-1] < means[0]
