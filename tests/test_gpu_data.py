"""GPU tests of the data side of the path: dataset build + statistics on the device, the
DynamicsDataset resampler, evaluation inputs, stacked-model launches and the validation sweep, against the oracle and the
golden vectors made from the reference (oracle/pin_reference.py sections 5 and 6).

Tolerances
  statistics          two-pass double accumulation on the device, float32 cascade sum in torch: std <= 1e-6 relative to
                      torch.std, mean <= 1e-6 of the column's scale max(|mean|, std) against torch.mean (a column whose
                      mean is 1e-3 of its spread -- u_a -- carries torch's own FP32 summation error of 1.6e-6 relative
                      to the mean), and both within one FP32 rounding (1.2e-7 relative) of the FP64 statistics
  z-scored features   bit-identical when given the reference's statistics; <= 5e-6 absolute with the device's own
  resampling (f64)    bit-identical to numpy (np.arange / np.interp / np.mean)
  T1 one-step         <= 1e-5 (every window of the second validation pickle, all 16 checkpoints)
  sweep               H = 1 within 1e-5 relative of the reference's calc_error; open-loop curves within 2e-3 relative
                      (means over ~4500 chaotic trajectories of FP32 closed loops, cf. T2)
"""
import os

import numpy as np
import pytest
import torch

from conftest import FWD, GOLD, RNN, load_ckpt

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


def _npz(name):
    return np.load(os.path.join(GOLD, name))


def _rel(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).abs() / b.abs().clamp_min(1e-30)).max().item()


def _stats_ok(got, st, f, l):
    """device statistics vs the oracle's FP32 ones (st) and vs FP64 statistics of the same samples (f, l)"""
    for pre, x in (("f", f), ("l", l)):
        gm, gs = got[f"{pre}_mean"].double().cpu(), got[f"{pre}_std"].double().cpu()
        wm, ws = st[f"{pre}_mean"].double(), st[f"{pre}_std"].double()
        assert gm.shape == wm.shape and gs.shape == ws.shape
        assert ((gm - wm).abs() / torch.maximum(wm.abs(), ws)).max().item() <= 1e-6, pre
        assert ((gs - ws).abs() / ws).max().item() <= 1e-6, pre
        tm, ts = x.double().mean(0), x.double().std(0)
        assert ((gm - tm).abs() / tm.abs().clamp_min(1e-30)).max().item() <= 1.2e-7, pre
        assert ((gs - ts).abs() / ts).max().item() <= 1.2e-7, pre


@pytest.mark.parametrize("pk", ["mpc", "pid"])
@pytest.mark.parametrize("size", [None, 4])
def test_device_dataset_and_statistics(pk, size):
    import mpcdl
    from mpcdl import data, synth
    from oracle import rollout_ref as R
    bags = synth.load_bags(f"{pk}_data_clean")
    f, l = R.mlp_dataset(bags) if size is None else R.rnn_dataset(bags, size)
    st = R.zscore_stats(f, l)
    fn, ln = R.apply_zscore(f, l, st)
    ds = mpcdl.RobotData(bags, size, device=_dev())
    assert ds.featrs.shape == fn.shape and ds.labels.shape == ln.shape and len(ds) == len(fn)
    _stats_ok(ds.stats, st, f, l)
    assert (ds.featrs.cpu() - fn).abs().max().item() <= 5e-6 and (ds.labels.cpu() - ln).abs().max().item() <= 5e-6
    # given the reference's statistics the z-score is the same two roundings: bit-identical
    from mpcdl import pack
    f2, l2, _ = data.device_dataset(bags, size, _dev(), stats_vec=pack.pack_norm(st).to(_dev()))
    assert torch.equal(f2.cpu(), fn) and torch.equal(l2.cpu(), ln)
    raw_f, raw_l, _ = data.device_dataset(bags, size, _dev(), normalise=False)
    assert torch.equal(raw_f.cpu(), f) and torch.equal(raw_l.cpu(), l)


def test_device_dataset_multi_bag_and_short_bags():
    """windows never cross a bag boundary; bags too short for a sample are skipped like the reference's empty range()"""
    from mpcdl import data, synth
    from oracle import rollout_ref as R
    rows = synth.load_rows().tolist()
    bags = [rows[:700], rows[700:703], rows[703:2000], rows[2000:2001], rows[2001:4551]]
    for size in (None, 4, 8):
        f, l = R.mlp_dataset(bags) if size is None else R.rnn_dataset([b for b in bags if len(b) - (size or 0) - 1 > 0], size)
        rf, rl, sv = data.device_dataset(bags, size, _dev(), normalise=False)
        assert torch.equal(rf.cpu(), f) and torch.equal(rl.cpu(), l)
        st = R.zscore_stats(f, l)
        got = data.stats_from_vector(sv, size)
        _stats_ok(got, st, f, l)


@pytest.mark.parametrize("window", [0, 4])
def test_open_loop_inputs_match_oracle(window):
    from mpcdl import data, synth
    from oracle import rollout_ref as R
    bags = synth.load_bags("pid_data_clean")
    rows32 = torch.as_tensor(bags[0], dtype=torch.float64).to(torch.float32)
    for H in (1, 10, 50):
        want = R.open_loop_inputs(rows32, H, window)
        got = data.open_loop_inputs(bags, H, window or None, _dev())
        assert all(torch.equal(g.cpu(), w) for g, w in zip(got, want))


def test_dynamics_dataset_resampling_bit_exact():
    """mpcdl.data_utils.DynamicsDataset against the reference's own outputs (tests/golden/dynamics.npz), float64, bit for bit"""
    from collections import OrderedDict
    from mpcdl.data_utils import DynamicsDataset
    g = _npz("dynamics.npz")
    raw = OrderedDict()
    for t, q, u in zip(g["t"], g["q"], g["u"]):
        raw[int(round(t * 1e9))] = {"v": {"v_long": q[0], "v_tran": q[1]}, "w": {"w_psi": q[2]}, "x": {"x": q[3], "y": q[4]},
                                    "e": {"psi": q[5]}, "u": {"u_a": u[0], "u_steer": u[1]}}
    assert np.array_equal(np.asarray(list(raw)) / 1e9, g["t"])          # the time stamps survive the ns round trip
    for dt in (0.1, 0.04, 0.02):
        ds = DynamicsDataset.from_dict(raw, dt, history=3, device=_dev())
        assert ds.q.dtype == torch.float64 and np.array_equal(ds.q.cpu().numpy(), g[f"q_ds_{dt}"])
        assert np.array_equal(ds.u.cpu().numpy(), g[f"u_ds_{dt}"])
        assert len(ds) == g[f"q_ds_{dt}"].shape[0] - 3
        q, u, nq = ds[4]
        assert np.array_equal(q.cpu().numpy(), g[f"q_ds_{dt}"][4:7]) and np.array_equal(nq.cpu().numpy(), g[f"q_ds_{dt}"][7])
        assert np.array_equal(u.cpu().numpy(), g[f"u_ds_{dt}"][4:7])
    ds = DynamicsDataset(g["q_grid"], g["u_grid"], 0.08, history=2, device=_dev())
    assert np.array_equal(ds.q.cpu().numpy(), g["q_ds_init_0.08"]) and np.array_equal(ds.u.cpu().numpy(), g["u_ds_init_0.08"])
    with pytest.raises(AssertionError):
        DynamicsDataset(g["q_grid"], g["u_grid"], 0.06, device=_dev())
    with pytest.raises(AssertionError):
        DynamicsDataset(g["q_grid"][:2], g["u_grid"][:2], 0.02, history=3, device=_dev())


@pytest.mark.parametrize("name", FWD + RNN)
def test_t1_onestep_second_validation_pickle(name):
    """T1 on pid_data_clean (SURVEY.md 4a asked for both pickles): dataset built on the device with the reference's
    statistics, Module.forward on every sample, against the reference modules' outputs."""
    import mpcdl
    from mpcdl import data, pack, synth
    g = _npz("pid.npz")
    m, meta = mpcdl.from_checkpoint_name(name, load_ckpt(name), _dev())
    size = meta["window"] or None
    pre = "mlp_" if size is None else "rnn4_"
    st = {k: torch.from_numpy(g[pre + k]) for k in ("f_mean", "f_std", "l_mean", "l_std")}
    f, _, _ = data.device_dataset(synth.load_bags("pid_data_clean"), size, _dev(), stats_vec=pack.pack_norm(st).to(_dev()))
    m.state_f = False                                   # the golden is the stateless (first-call) output
    y = m(f).cpu()
    ref = torch.from_numpy(g[name])
    assert y.shape == ref.shape and ((y - ref).abs().max() / ref.abs().max()).item() <= 1e-5


def test_stacked_models_equal_single_launches():
    """n_models stacked in one launch (shared inputs, model-major outputs) == the same kernel run per checkpoint, bit for bit;
    shallow MLPs padded with identity layers included."""
    import mpcdl
    from mpcdl import _lib, ops, pack, synth
    from mpcdl.eval_models import _pad_mlp_state
    dev = _dev()
    z = np.load(os.path.join(GOLD, "stats.npz"))
    rows = synth.load_rows()
    for names, kind, hidden, lay, window, pre in ((["Basic_64_1_32", "Basic_64_2_64", "Basic_64_4_32"], _lib.MLP, 64, 4, 0, "mlp_"),
                                                  ([n for n in RNN if n.startswith("GRU")], _lib.GRU, 64, 4, 4, "rnn4_")):
        st = {k[len(pre):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(pre)}
        norm = pack.pack_norm(st).to(dev)
        init, U = synth.make_inputs(rows, 333, 7, window=window, seed=9)
        init, U = init.to(dev), U.to(dev)
        blobs = [pack.pack_mlp(_pad_mlp_state(load_ckpt(n), hidden, int(n.split("_")[2]), lay), hidden, lay) if kind == _lib.MLP
                 else pack.pack_rnn(load_ckpt(n), kind, hidden, lay, window) for n in names]

        def run(wblob, M, tune):
            return ops.OPS[kind](wblob.to(dev), norm, init, U, hidden, lay, window, False, None, None, [], None, 0, 0, None, [],
                                 True, False, tune, 0, None, [], 0, 0, False, None, None, M, True)[0]
        stacked = run(torch.cat(blobs), len(names), []).view(len(names), 333, 7, 6)
        for i, n in enumerate(names):
            single = run(blobs[i], 1, [0, 0, 0, -1] if kind == _lib.MLP else [])      # the generic (ring-fed) kernel
            assert torch.equal(stacked[i], single), n
            m, _ = mpcdl.from_checkpoint_name(n, load_ckpt(n), dev)                    # and the model's own rollout to T1
            m.state_f = False                                                          # the stacked launch above is stateless
            own = m.rollout(init, U, st)[0]
            assert ((own - stacked[i]).abs().amax(dim=(0, 1)) / own.abs().amax(dim=(0, 1))).max().item() <= 1e-5, n
    with pytest.raises(RuntimeError):                   # 333 trajectories do not split evenly over 2 stacked models
        ops.OPS[_lib.GRU](torch.cat(blobs[:2]).to(dev), norm, init, U, 64, 4, 4, False, None, None, [], None, 0, 0, None, [],
                          True, False, [], 0, None, [], 0, 0, False, None, None, 2, False)


def test_validation_sweep_matches_reference_curves():
    """mpcdl.eval_models.sweep: all 16 checkpoints x both validation pickles x H in {1, 10, 50} in one call, against the
    curves computed from the reference modules in the python loop (tests/golden/sweep.npz)."""
    from mpcdl import eval_models
    g = _npz("sweep.npz")
    ck = {n: load_ckpt(n) for n in FWD + RNN}
    res = eval_models.sweep(ck, ("mpc", "pid"), (1, 10, 50), device=_dev())
    worst = {1: 0.0, 10: 0.0, 50: 0.0}
    for pk in ("mpc", "pid"):
        for n in FWD + RNN:
            got = res[pk][n]
            worst[1] = max(worst[1], _rel(np.asarray(got[1]), g[f"{pk}/{n}/H1"]))
            for H in (10, 50):
                assert got[H][0].shape == (H,)
                worst[H] = max(worst[H], _rel(got[H][0], g[f"{pk}/{n}/H{H}_mae"]), _rel(got[H][1], g[f"{pk}/{n}/H{H}_rmse"]))
    assert worst[1] <= 1e-5 and worst[10] <= 2e-3 and worst[50] <= 2e-3, worst


def test_stats_persist_beside_the_checkpoint(tmp_path):
    import mpcdl
    from mpcdl import data, synth
    sd = load_ckpt("GRU_64_4_4_16_False")
    path = str(tmp_path / "GRU_64_4_4_16_False.pt")
    torch.save(sd, path)
    m, meta = mpcdl.load_model(path, _dev())
    assert m.stats is None and meta["window"] == 4
    ds = mpcdl.RobotData(synth.load_bags("mpc_data_clean"), 4, device=_dev())
    data.save_stats(data.stats_path(path), ds.stats)
    m, _ = mpcdl.load_model(path, _dev())
    assert all(torch.equal(m.stats[k], ds.stats[k].cpu()) for k in ds.stats)
    init, U = synth.make_inputs(synth.load_rows(), 50, 5, window=4, seed=1)
    a = m.rollout(init.to(_dev()), U.to(_dev()), m.stats)[0]
    b = m.rollout(init.to(_dev()), U.to(_dev()), ds.stats)[0]
    assert torch.equal(a, b)
