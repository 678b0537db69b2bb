"""CPU: the oracle (oracle/rollout_ref.py) against the golden vectors that oracle/pin_reference.py
produced by running the UNMODIFIED reference modules (one-step on every validation row/window for
all 16 checkpoints; N=64 x H=24 closed loops; noise + cost).  Bit-exact: same torch build, same
arithmetic -- if a torch upgrade changes CPU kernels these become <= 1e-6 comparisons."""
import numpy as np
import pytest
import torch

from conftest import FWD, RNN, load_ckpt
from oracle import rollout_ref as R

torch.set_num_threads(1)


def _close(a, b, tol=0.0):
    a, b = torch.as_tensor(a), torch.as_tensor(b)
    if torch.equal(a, b):
        return True
    return tol > 0 and ((a - b).abs().max() / b.abs().max()).item() <= tol


def test_stats_match_golden(bags, gold_stats):
    mst = R.zscore_stats(*R.mlp_dataset(bags))
    rst = R.zscore_stats(*R.rnn_dataset(bags, 4))
    for k in ("f_mean", "f_std", "l_mean", "l_std"):
        assert _close(mst[k], gold_stats[0][k]) and _close(rst[k], gold_stats[1][k])
    assert tuple(rst["f_mean"].shape) == (4, 8)            # per window position


@pytest.mark.parametrize("name", FWD)
def test_onestep_mlp(name, bags, gold_onestep):
    f, l = R.mlp_dataset(bags)
    fn, ln = R.apply_zscore(f, l, R.zscore_stats(f, l))
    m, _ = R.build_from_name(name, load_ckpt(name))
    with torch.no_grad():
        y = m(fn)
    assert _close(y, gold_onestep[name], 1e-6)
    mae, rmse = R.calc_error(y, ln)                        # eval_models.py:91-96
    assert np.allclose([mae, rmse], gold_onestep[name + "_err"], rtol=1e-5)


@pytest.mark.parametrize("name", RNN)
def test_onestep_rnn(name, bags, gold_onestep):
    f, l = R.rnn_dataset(bags, 4)
    fn, _ = R.apply_zscore(f, l, R.zscore_stats(f, l))
    m, meta = R.build_from_name(name, load_ckpt(name))
    with torch.no_grad():
        y, st = m(fn, m.zero_state(len(fn)))
        assert _close(y, gold_onestep[name], 1e-6)
        if meta["state_f"]:
            y2, _ = m(fn, st)
            assert _close(y2, gold_onestep[name + "_second"], 1e-6)


@pytest.mark.parametrize("name", FWD + RNN)
def test_rollout_golden(name, gold_rollouts, gold_stats):
    g = gold_rollouts
    m, meta = R.build_from_name(name, load_ckpt(name))
    if meta["kind"] == "mlp":
        y = R.rollout_mlp(m, gold_stats[0], torch.from_numpy(g["s0"]), torch.from_numpy(g["U"]))[0]
    else:
        y = R.rollout_rnn(m, gold_stats[1], torch.from_numpy(g["win0"]), torch.from_numpy(g["Uw"]), stateful=meta["state_f"])[0]
    assert _close(y, g[name], 1e-6)


def test_noise_cost_golden(gold_rollouts, gold_stats):
    g = gold_rollouts
    noise = R.RefNoise(6, 2, 3, 2, 32, "tanh", "identity")
    noise.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    noise.eval()
    cost = tuple(torch.from_numpy(g[k]) for k in ("cost_ref", "cost_qw", "cost_rw"))
    m, _ = R.build_from_name("Basic_64_4_32", load_ckpt("Basic_64_4_32"))
    y, c = R.rollout_mlp(m, gold_stats[0], torch.from_numpy(g["s0"]), torch.from_numpy(g["U"]), noise=noise,
                         eps=torch.from_numpy(g["eps"]), cost=cost)
    assert _close(y, g["Basic_64_4_32_noise"], 1e-6) and _close(c, g["Basic_64_4_32_noise_cost"], 1e-6)
    y1, _ = R.rollout_mlp(m, gold_stats[0], torch.from_numpy(g["s0"][:1]), torch.from_numpy(g["U"][:1]), noise=noise,
                          eps=torch.from_numpy(g["eps_b1"]))
    assert _close(y1, g["Basic_64_4_32_noise_b1"], 1e-6)     # made through the reference's DynamicsNN wrapper
    gm, _ = R.build_from_name("GRU_64_4_4_16_False", load_ckpt("GRU_64_4_4_16_False"))
    y, c, _ = R.rollout_rnn(gm, gold_stats[1], torch.from_numpy(g["win0"]), torch.from_numpy(g["Uw"]), noise=noise,
                            eps=torch.from_numpy(g["eps"]), cost=cost)
    assert _close(y, g["GRU_64_4_4_16_False_noise"], 1e-6) and _close(c, g["GRU_64_4_4_16_False_noise_cost"], 1e-6)


def test_fp32_vs_fp64_one_step_is_tight(gold_stats):
    """The tolerance story of DESIGN.md: one step of torch-FP32 vs FP64 sits near 1e-7."""
    from mpcdl import synth
    m, _ = R.build_from_name("Basic_64_4_32", load_ckpt("Basic_64_4_32"))
    s0, U = synth.make_inputs(synth.load_rows(), 512, 1, seed=5)
    y32 = R.rollout_mlp(m, gold_stats[0], s0, U)[0]
    st64 = {k: v.double() for k, v in gold_stats[0].items()}
    y64 = R.rollout_mlp(R.as_double(m), st64, s0.double(), U.double())[0]
    assert R.rel_err(y32, y64).max().item() < 2e-6


# ------------------------------------------------------------------ a6 / a7 / a8 / a9 goldens (tests/golden/aux.npz)
def _noise_oracle():
    m = R.RefNoise(6, 2, 3, 2, 32, "tanh", "identity")
    m.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    return m.eval()


def test_noise_forward_matches_reference_golden(gold_aux):
    """a6: the golden output came from the reference FeedforwardNoiseModel.forward under a shared seed."""
    g = gold_aux
    m = _noise_oracle()
    with torch.no_grad():
        out = m(torch.from_numpy(g["noise_q"]), torch.from_numpy(g["noise_u"]), torch.from_numpy(g["noise_eps"]))
    assert torch.equal(out, torch.from_numpy(g["noise_out"]))


@pytest.mark.parametrize("open_loop", [True, False])
def test_run_dynamics_model_matches_reference_golden(open_loop, gold_aux):
    """a8: the golden trajectory came from the reference's own run_dynamics_model loop around its DynamicsNN."""
    g = gold_aux
    tr = R.run_dynamics_model(_noise_oracle(), torch.from_numpy(g["rdm_q"]), torch.from_numpy(g["rdm_u"]),
                              torch.from_numpy(g["rdm_eps"]), open_loop)
    assert torch.equal(tr, torch.from_numpy(g["rdm_open" if open_loop else "rdm_closed"]))


def test_rollout_dynnn_batched_is_the_same_recursion(gold_aux):
    g = gold_aux
    m = _noise_oracle()
    q, u, e = (torch.from_numpy(g[k]) for k in ("rdm_q", "rdm_u", "rdm_eps"))
    tr = R.rollout_dynnn(m, q[:1], u[None, 1:], e[None])
    assert torch.equal(tr[0], torch.from_numpy(g["rdm_open"][1:]))
    out = R.rollout_dynnn(m, torch.from_numpy(g["dyn_q0"]), torch.from_numpy(g["dyn_U"]), torch.from_numpy(g["dyn_eps"]))
    assert torch.equal(out, torch.from_numpy(g["dyn_out"]))
    # a row of the batch is independent of the rest of the batch
    one = R.rollout_dynnn(m, torch.from_numpy(g["dyn_q0"][7:8]), torch.from_numpy(g["dyn_U"][7:8]), torch.from_numpy(g["dyn_eps"][7:8]))
    assert torch.allclose(one[0], out[7], rtol=0, atol=1e-6)


def test_calc_error_sampled_matches_reference_golden(gold_aux, gold_onestep, bags):
    """a9 SAMPLE=True: same random.randint stream as the reference -> the values its calc_error printed."""
    import random
    from mpcdl import RobotData
    for name, size in (("Basic_64_4_32", None), ("GRU_64_4_4_16_True", 4)):
        ds = RobotData(bags, size)
        random.seed(5)
        mae, rmse = R.calc_error_sampled(torch.from_numpy(gold_onestep[name]), ds.labels, 2000, random)
        assert (mae, rmse) == tuple(gold_aux[name + "_err_sampled"])


def test_philox_known_answers_and_mppi_update():
    """oracle/mppi_ref.py: Philox4x32-10 against the Random123 known-answer vectors; Box-Muller moments; the update
    is a convex combination of the samples when nothing is clamped."""
    from oracle import mppi_ref as M
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for c, k, want in kat:
        assert tuple(M.philox4x32(np.array(c, dtype=np.uint32), np.array(k, dtype=np.uint32))) == want
    z = M.control_normals(5, np.arange(20000), 8)
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1) < 0.01
    assert not np.array_equal(M.control_normals(5, [7], 8), M.control_normals(6, [7], 8))
    U = np.zeros((8, 2), np.float32)
    u = M.sample_controls(U, 5, np.arange(100), (0.1, 0.1), (-9, -9), (9, 9))
    costs = (u ** 2).sum((1, 2))
    Un, S, wsum, cmin = M.mppi_update(U, u, costs, 0.05, (-9, -9), (9, 9))
    assert cmin == costs.astype(np.float64).min() and wsum >= 1.0
    assert (np.abs(Un) <= np.abs(u).max(0) + 1e-7).all()
    assert (Un ** 2).sum() < np.median(costs)                         # pulled towards the cheap samples
