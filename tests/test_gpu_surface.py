"""GPU parity of the call surface around the rollout kernels: the stand-alone noise net,
the DynamicsNN / CasadiDynamicCLBicycleNoise wrappers, run_dynamics_model and eval_models.calc_error, against
golden vectors produced by the reference's own code (tests/golden/aux.npz, oracle/pin_reference.py 3c/3d).
Tolerance T1 = 1e-5 (max |dy| / max |ref|) one-step; closed loops: one-step T1 and p90 over the horizon <= 1e-5."""
import random

import numpy as np
import pytest
import torch

from conftest import FWD, RNN, load_ckpt

pytestmark = pytest.mark.gpu
T1 = 1e-5


def _dev():
    return torch.device("cuda:0")


def _noise():
    import mpcdl
    n = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    n.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    return n.to(_dev())


def _rel(a, b):
    a, b = torch.as_tensor(a).cpu().double(), torch.as_tensor(b).cpu().double()
    return ((a - b).abs().max() / b.abs().max()).item()


def test_noise_forward_golden(gold_aux):
    """a6: FeedforwardNoiseModel.forward with supplied eps vs the reference module's output (512 samples)."""
    g, n = gold_aux, _noise()
    q, u, eps = (torch.from_numpy(g[k]).to(_dev()) for k in ("noise_q", "noise_u", "noise_eps"))
    out, mu, scale = n(q, u, eps, return_params=True)
    assert out.shape == (512, 6)
    assert _rel(out, g["noise_out"]) <= T1
    assert _rel(mu, g["noise_mu"]) <= T1 and _rel(scale, g["noise_scale"]) <= T1
    # ragged batch (1 sample, the reference's get_prediction shape) and numpy in / numpy out
    one = n.get_prediction(g["noise_q"][:1], g["noise_u"][:1], g["noise_eps"][:1])
    assert isinstance(one, np.ndarray) and one.shape == (1, 6) and _rel(one, g["noise_out"][:1]) <= T1
    # without eps the draws come from Philox keyed by torch's seed: reproducible, and distributed like the Normal
    torch.manual_seed(3)
    n._calls = 0
    a = n(q, u)
    torch.manual_seed(3)
    n._calls = 0
    b = n(q, u)
    c = n(q, u)
    assert torch.equal(a, b) and not torch.equal(b, c)
    z = ((a - mu) / scale).flatten()
    assert abs(z.mean().item()) < 0.1 and abs(z.std().item() - 1.0) < 0.1
    with pytest.raises(RuntimeError):
        n(q[:, :17].contiguous(), u)


def test_dynamics_nn_rollout_golden(gold_aux):
    """a7/a8: the noise net as the whole dynamics, fused over H = 40 for N = 300 (2 tiles + 44)."""
    import mpcdl
    g, n = gold_aux, _noise()
    dyn = mpcdl.DynamicsNN(None, n, False, history=3)
    q0, U, eps = (torch.from_numpy(g[k]).to(_dev()) for k in ("dyn_q0", "dyn_U", "dyn_eps"))
    out = dyn.rollout(q0, U, eps=eps)
    ref = torch.from_numpy(g["dyn_out"])
    assert out.shape == ref.shape
    err = ((out.cpu() - ref).abs() / ref.abs().amax((0, 1))).amax(-1)
    assert err[:, 0].max().item() <= T1
    assert torch.quantile(err.flatten(), 0.9).item() <= T1, torch.quantile(err.flatten(), torch.tensor([0.5, 0.9, 0.99, 1.0]))
    # Philox: shard invariance through traj_offset
    full = dyn.rollout(q0, U, seed=9)
    part = dyn.rollout(q0[130:].contiguous(), U[130:].contiguous(), seed=9, traj_offset=130)
    assert torch.equal(full[130:], part)
    with pytest.raises(RuntimeError):
        mpcdl.DynamicsNN(None, n, False, history=2).rollout(q0, U)


@pytest.mark.parametrize("open_loop", [True, False])
def test_run_dynamics_model_golden(open_loop, gold_aux):
    """a8: the reference loop (batch 1, deques on the host, one launch per step) and its fused form."""
    import mpcdl
    from mpcdl import test_model as tm
    g, n = gold_aux, _noise()
    q, u, eps = g["rdm_q"], g["rdm_u"], g["rdm_eps"]
    ref = g["rdm_open" if open_loop else "rdm_closed"]

    class _Eps:                                    # feeds the golden draws to get_prediction call by call
        def __init__(self, model):
            self.m, self.k = model, 0

        def get_prediction(self, qq, uu):
            self.k += 1
            return self.m.get_prediction(qq, uu, eps=eps[self.k - 1][None])

    dyn = mpcdl.DynamicsNN(None, n, False, history=3)
    dataset = [(q[k][None], u[k][None], None) for k in range(len(q))]
    (qd, qn, qz), labels = tm.run_dynamics_model(_Eps(dyn), dataset, open_loop, with_nominal=False)
    assert labels == ["data", "nominal", "with noise"] and len(qz) == len(q)
    got = np.stack([np.asarray(x, dtype=np.float32) for x in qz])
    err = np.abs(got - ref).max(-1) / np.abs(ref).max()
    assert err[1] <= T1 and np.quantile(err, 0.9) <= T1, err
    if open_loop:
        fused = tm.run_dynamics_model_fused(mpcdl.DynamicsNN(None, n, False, history=3), q, u, eps=eps)
        errf = (fused.cpu().numpy() - ref).__abs__().max(-1) / np.abs(ref).max()
        assert errf[1] <= T1 and np.quantile(errf, 0.9) <= T1, errf


def test_bicycle_noise_wrapper_with_learned_nominal(gold_rollouts, gold_stats):
    """a7: CasadiDynamicCLBicycleNoise with the learned MLP as the nominal step (north_star's substitution):
    batch-1 get_prediction calls and the fused rollout both reproduce the golden made with the reference's wrapper."""
    import mpcdl
    from conftest import load_ckpt as ck
    g, n = gold_rollouts, _noise()
    m, _ = mpcdl.from_checkpoint_name("Basic_64_4_32", ck("Basic_64_4_32"), _dev())
    Q = mpcdl.models.Q_PERM
    w = mpcdl.CasadiDynamicCLBicycleNoise(m, n, False, history=3, stats=gold_stats[0])
    ref = g["Basic_64_4_32_noise_b1"][0][:, Q]               # [H,6] in q order
    q = g["s0"][0][Q]
    got = []
    for t in range(8):
        q = w.get_prediction(q, g["U"][0, t], eps=g["eps_b1"][0, t][None])[0]
        got.append(q)
    err = np.abs(np.stack(got) - ref[:8]).max(-1) / np.abs(ref).max()
    assert err[0] <= T1 and err.max() <= 1e-4, err
    w2 = mpcdl.CasadiDynamicCLBicycleNoise(m, n, False, history=3, stats=gold_stats[0])
    s0 = torch.from_numpy(g["s0"][:, Q]).to(_dev())
    out = w2.rollout(s0, torch.from_numpy(g["U"]).to(_dev()), eps=torch.from_numpy(g["eps"]).to(_dev()))
    refb = torch.from_numpy(g["Basic_64_4_32_noise"])[:, :, Q]
    e = ((out.cpu() - refb).abs() / refb.abs().amax((0, 1))).amax(-1)
    assert e[:, 0].max().item() <= T1 and torch.quantile(e.flatten(), 0.9).item() <= T1
    with pytest.raises(RuntimeError):
        mpcdl.CasadiDynamicCLBicycleNoise(m, n, True)         # Frenet frame needs mpclab_common's track


@pytest.mark.parametrize("name", FWD + RNN)
def test_calc_error_all_samples(name, bags, gold_onestep):
    """a9 SAMPLE=False: whole dataset, every checkpoint; golden = oracle value pinned to the reference's printout."""
    from mpcdl import eval_models as ev, basic_model as fwd, deep_models as rnn
    import mpcdl
    m, meta = mpcdl.from_checkpoint_name(name, load_ckpt(name), _dev())
    ds = fwd.RobotData(bags) if meta["kind"] == "mlp" else rnn.RobotData(bags, 4)
    ev.SAMPLE = False
    mae, rmse = ev.calc_error(ds, m, name, meta["state_f"])
    want = gold_onestep[name + "_err"]
    assert abs(mae - want[0]) <= 2e-6 and abs(rmse - want[1]) <= 2e-6, (mae, rmse, want)


def test_calc_error_sampled_and_eval_entry_points(bags, gold_aux, tmp_path, monkeypatch):
    """a9 SAMPLE=True under random.seed(5) evaluates the rows the reference evaluates; eval_basic / eval_deep walk a
    model directory with the reference's file-name schema."""
    from mpcdl import eval_models as ev, basic_model as fwd, deep_models as rnn
    import mpcdl
    ev.SAMPLE, ev.EPOCHS = True, 2000
    for name, ds in (("Basic_64_4_32", fwd.RobotData(bags)), ("GRU_64_4_4_16_True", rnn.RobotData(bags, 4))):
        m, meta = mpcdl.from_checkpoint_name(name, load_ckpt(name), _dev())
        random.seed(5)
        mae, rmse = ev.calc_error(ds, m, name, meta["state_f"])
        want = gold_aux[name + "_err_sampled"]
        assert abs(mae - want[0]) <= 2e-6 and abs(rmse - want[1]) <= 2e-6
    import os, shutil
    from conftest import GOLD
    for sub, names in (("fwd_models", ["Basic_16_1_32", "Basic_64_4_32"]), ("rnn_models", ["GRU_64_4_4_16_False", "LSTM_64_4_4_64_True"])):
        os.makedirs(tmp_path / sub)
        for nme in names:
            shutil.copy(os.path.join(GOLD, "checkpoints", nme + ".npz"), tmp_path / sub / (nme + ".npz"))
    ev.SAMPLE = False
    monkeypatch.setattr(fwd, "DATA", bags, raising=False)
    monkeypatch.setattr(rnn, "DATA", bags, raising=False)
    rb = ev.eval_basic(one_model=False, model_dir=str(tmp_path / "fwd_models"))
    rd = ev.eval_deep(one_model=False, model_dir=str(tmp_path / "rnn_models"))
    assert len(rb) == 2 and len(rd) == 2 and len(ev.eval_basic(True, str(tmp_path / "fwd_models"))) == 1
    assert all(0 < v[0] < 1 and v[0] <= v[1] for v in list(rb.values()) + list(rd.values()))
    p = ev.view_test(str(tmp_path / "series.npz"))
    z = np.load(p)
    assert len(z["prediction_x_pos"]) == len(z["actual_x_pos"]) > 0


def test_surface_has_no_cpu_fallback(gold_aux):
    import mpcdl
    n = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    with pytest.raises((RuntimeError, NotImplementedError)):
        n(torch.zeros(4, 18), torch.zeros(4, 6), torch.zeros(4, 6))


def test_host_pipeline_matches_synchronous_calls(gold_stats):
    """HostPipeline (overlapped copy-in / kernel / copy-out over pinned host buffers) returns, batch by batch, exactly
    what a synchronous rollout of the same batch returns -- including when the device input slots are reused."""
    import mpcdl
    from mpcdl import parallel, synth
    m, _ = mpcdl.from_checkpoint_name("Basic_64_4_32", load_ckpt("Basic_64_4_32"), _dev())
    st = gold_stats[0]
    rows = synth.load_rows()
    batches = [synth.make_inputs(rows, 1000, 20, window=0, seed=s) for s in range(5)]
    outs = [torch.empty(1000, 20, 6).pin_memory() for _ in batches]
    pipe = parallel.HostPipeline(lambda a, b, out: m.rollout(a, b, st, out=(out, None))[0], depth=2, device=_dev())
    for (init, U), o in zip(batches, outs):
        pipe.submit((init.pin_memory(), U.pin_memory()), o)
    pipe.drain(sync=True)
    for (init, U), o in zip(batches, outs):
        want = m.rollout(init.to(_dev()), U.to(_dev()), st)[0].cpu()
        assert torch.equal(o, want)
