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


# ------------------------------------------------------------------------------ f2 / f3 goldens (round 2)
def _npz(name):
    import os
    from conftest import GOLD
    return np.load(os.path.join(GOLD, name))


def test_dynamics_dataset_oracle_matches_reference_golden():
    """oracle/dataset_ref.py against tests/golden/dynamics.npz (made by the reference's DynamicsDataset): bit-exact."""
    from oracle import dataset_ref as D
    g = _npz("dynamics.npz")
    for dt in (0.1, 0.04, 0.02):
        q, u = D.resample_ref(g["t"], g["q"], g["u"], dt)
        assert np.array_equal(q, g[f"q_ds_{dt}"]) and np.array_equal(u, g[f"u_ds_{dt}"])
    q, u, step = D.subsample(g["q_grid"], g["u_grid"], 0.08)
    assert step == 4 and np.array_equal(q, g["q_ds_init_0.08"]) and np.array_equal(u, g["u_ds_init_0.08"])
    with pytest.raises(AssertionError):          # 0.06 // 0.02 == 2.0 in floating point: the reference rejects this dt
        D.subsample(g["q_grid"], g["u_grid"], 0.06)


def test_dynamics_dataset_host_logic():
    """The host side of mpcdl.data_utils.DynamicsDataset: the step rule (incl. the reference's 0.06 quirk) and the grid
    length handed to the kernel equal numpy's."""
    import math
    from mpcdl import data_utils as du
    assert du._step_of(0.1, 0.02) == 5 and du._step_of(0.08, 0.02) == 4
    with pytest.raises(AssertionError):
        du._step_of(0.06, 0.02)
    rng = np.random.default_rng(1)
    for _ in range(300):
        t0 = float(rng.uniform(0, 2e9))
        span = float(rng.uniform(0.01, 50))
        n = len(np.arange(t0, t0 + span, 0.02))
        assert n == max(int(math.ceil((t0 + span - t0) / 0.02)), 0)
    with pytest.raises(RuntimeError):            # no CPU path for the arithmetic
        du.DynamicsDataset(np.zeros((10, 6)), np.zeros((10, 2)), 0.1, device="cpu")


@pytest.mark.parametrize("name", ["Basic_16_1_32", "Basic_64_4_32", "GRU_64_4_4_16_False"])
def test_sweep_oracle_matches_golden(name):
    """oracle sweep (teacher-forced H = 1 + open-loop curves) against tests/golden/sweep.npz, whose curves were computed
    from the REFERENCE modules driven in the python loop (oracle/pin_reference.py section 6): bit-exact."""
    from conftest import load_ckpt
    from mpcdl import synth
    g = _npz("sweep.npz")
    m, meta = R.build_from_name(name, load_ckpt(name))
    hs = (1, 10, 50) if name.startswith("Basic") else (1, 10)
    for pk in ("mpc", "pid"):
        got = R.sweep_model(m, meta, synth.load_bags(f"{pk}_data_clean"), hs)
        assert np.array_equal(np.asarray(got[1]), g[f"{pk}/{name}/H1"])
        for H in hs[1:]:
            assert np.array_equal(got[H][0].numpy(), g[f"{pk}/{name}/H{H}_mae"])
            assert np.array_equal(got[H][1].numpy(), g[f"{pk}/{name}/H{H}_rmse"])


def test_pid_pickle_onestep_golden():
    """second validation pickle (pid_data_clean): statistics and one-step outputs of the reference modules"""
    from conftest import load_ckpt
    from mpcdl import synth
    g = _npz("pid.npz")
    bags = synth.load_bags("pid_data_clean")
    f, l = R.mlp_dataset(bags)
    st = R.zscore_stats(f, l)
    assert all(np.array_equal(st[k].numpy(), g[f"mlp_{k}"]) for k in st)
    fn, _ = R.apply_zscore(f, l, st)
    for name in ("Basic_16_2_16", "Basic_64_2_64"):
        m, _ = R.build_from_name(name, load_ckpt(name))
        with torch.no_grad():
            assert np.array_equal(m(fn).numpy(), g[name])


def test_identity_padding_of_shallow_mlps_is_exact():
    """sweep() stacks MLP checkpoints of different depth by appending identity layers: relu(I x + 0) == x bit for bit."""
    from conftest import load_ckpt
    from mpcdl.eval_models import _pad_mlp_state
    x = torch.randn(500, 8, generator=torch.Generator().manual_seed(0)) * 2
    for name, hid, lay in (("Basic_16_1_32", 16, 1), ("Basic_64_2_32", 64, 2)):
        sd = load_ckpt(name)
        a, _ = R.build_from_name(name, sd)
        b = R.RefBasic(8, hid, 6, 4)
        b.load_state_dict(_pad_mlp_state(sd, hid, lay, 4))
        with torch.no_grad():
            assert torch.equal(a(x), b.eval()(x))


@pytest.mark.parametrize("tag", ["Basic_16_4_16", "Basic_64_2_32"])
def test_training_oracle_reproduces_the_reference_made_golden(tag):
    """oracle/train_ref.py (the reference's loop: L1Loss + Adam over explicit batches) against tests/golden/train.npz, which
    oracle/pin_reference.py section 7 made with the reference's own Basic_Model class and optimizer: bit for bit."""
    import os
    import numpy as np
    import torch
    from conftest import GOLD
    from oracle import rollout_ref as R, train_ref as T
    g = np.load(os.path.join(GOLD, "train.npz"))
    hidden, layers, batch = (int(x) for x in tag.split("_")[1:])
    z = np.load(os.path.join(GOLD, "stats.npz"))
    from mpcdl import synth
    f, l = R.mlp_dataset(synth.load_bags())
    st = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("mlp_")}
    fn, ln = R.apply_zscore(f, l, st)
    m = R.RefBasic(8, hidden, 6, layers)
    torch.nn.utils.vector_to_parameters(torch.from_numpy(g[f"{tag}/init"]).clone(), m.parameters())
    torch.set_num_threads(1)
    losses, _ = T.train_steps(m, fn, ln, torch.from_numpy(g[f"{tag}/order"]), batch, 1e-3)
    final = torch.cat([p.detach().reshape(-1) for p in m.parameters()])
    assert torch.equal(losses, torch.from_numpy(g[f"{tag}/losses"])) and torch.equal(final, torch.from_numpy(g[f"{tag}/final"]))
    if tag == "Basic_64_2_32":
        assert T.validation_loss(m, fn[:800], ln[:800], 32) == float(g[f"{tag}/valid800"])
