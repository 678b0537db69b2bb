"""Tiny instances of every kernel family, for compute-sanitizer (SURVEY.md 5, round-1 verdict next #1b).

    compute-sanitizer --tool racecheck python tools/sanitize_cases.py rnn
    families: mlp2 mlp rnn rnn_tc mlp_tc lstm_tc aux | all

Every case runs a ragged N (so a partially filled tile exists) and a short horizon; results are only checked for
finiteness here -- parity is what tests/ is for."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
GOLD = os.path.join(ROOT, "tests", "golden")


def ckpt(name):
    return {k: torch.from_numpy(v) for k, v in np.load(os.path.join(GOLD, "checkpoints", f"{name}.npz")).items()}


def main(fams):
    import mpcdl
    from mpcdl import synth
    dev = torch.device("cuda:0")
    z = np.load(os.path.join(GOLD, "stats.npz"))
    st_mlp = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("mlp_")}
    st_rnn = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("rnn4_")}
    rows = synth.load_rows()
    noise = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    noise.model.load_state_dict(ckpt("noise_seed0_h3_s32_l2"))
    noise.to(dev)
    cost = tuple(c.to(dev) for c in synth.make_cost(rows, 3))

    def roll(name, n, h=3, label="", **kw):
        m, meta = mpcdl.from_checkpoint_name(name, ckpt(name) if not name.startswith("LSTM_256") else None, dev)
        init, U = synth.make_inputs(rows, n, h, window=meta["window"], seed=4)
        st = st_mlp if meta["kind"] == "mlp" else st_rnn
        out = m.rollout(init.to(dev), U.to(dev), st, **kw)
        torch.cuda.synchronize()
        ok = all(torch.isfinite(o).all().item() for o in out if o is not None)
        print(f"case {name} N={n} H={h} {label} {kw.get('precision', 'fp32')} tune={kw.get('tune', ())}: finite={ok}", flush=True)
        assert ok

    if "mlp2" in fams:
        for name in ("Basic_64_4_32", "Basic_16_1_32", "Basic_64_1_32"):
            roll(name, 70)
        roll("Basic_64_4_32", 300, tune=(8, 0, 0, 0), cost=cost)
        roll("Basic_64_4_32", 300, tune=(8, 0, 0, 1))
    if "mlp" in fams:
        roll("Basic_64_4_32", 70, noise=noise, seed=5, cost=cost, label="noise")
        roll("Basic_16_2_32", 70, noise=noise, eps=synth.make_eps(70, 3).to(dev), label="noise+eps")
        roll("Basic_64_4_32", 70, tune=(2, 0, 3, 0), label="ring")
    if "rnn" in fams:
        for name in ("GRU_64_4_4_16_False", "LSTM_64_4_4_16_False"):
            for rt in (2, 4, 8):
                roll(name, 70, tune=(rt, 0, 0, 0))
        roll("GRU_64_4_4_64_True", 70, label="stateful")
        roll("LSTM_64_4_4_64_True", 70, label="stateful")
        roll("GRU_64_4_4_16_False", 70, noise=noise, seed=5, cost=cost, label="noise")
    if "rnn_tc" in fams:
        for prec in ("tf32x3", "tf32"):
            roll("GRU_64_4_4_16_False", 200, precision=prec)
            roll("LSTM_64_4_4_64_True", 200, precision=prec, label="stateful")
            roll("GRU_64_4_4_16_False", 200, precision=prec, noise=noise, seed=5, cost=cost, label="noise")
    if "mlp_tc" in fams:
        for prec in ("tf32x3", "tf32"):
            roll("Basic_64_4_32", 300, precision=prec, cost=cost)
    if "lstm_tc" in fams:
        torch.manual_seed(0)
        name = "LSTM_256_2_4_0_False"
        mod = torch.nn.ModuleDict(dict(lstm=torch.nn.LSTM(8, 256, 2, batch_first=True), full_c=torch.nn.Linear(256, 6)))
        m, _ = mpcdl.from_checkpoint_name(name, {k: v.detach().clone() for k, v in mod.state_dict().items()}, dev)
        init, U = synth.make_inputs(rows, 300, 2, window=4, seed=4)
        out = m.rollout(init.to(dev), U.to(dev), st_rnn, precision="bf16")[0]
        torch.cuda.synchronize()
        print(f"case {name} bf16 N=300 H=2: finite={torch.isfinite(out).all().item()}", flush=True)
    if "aux" in fams:
        q, u = torch.randn(200, 18, device=dev), torch.randn(200, 6, device=dev)
        o = noise(q, u, torch.randn(200, 6, device=dev))
        o2 = noise(q, u)
        dn = mpcdl.DynamicsNN(None, noise, False, history=3)
        o3 = dn.rollout(torch.randn(70, 6, device=dev), torch.randn(70, 3, 2, device=dev), seed=3)
        mae, rmse = torch.ops.mpcdl.error_metrics(torch.randn(999, 6, device=dev), torch.randn(999, 6, device=dev))
        from mpcdl.mppi import MPPI
        pl = MPPI(mpcdl.from_checkpoint_name("Basic_64_4_32", ckpt("Basic_64_4_32"), dev)[0], st_mlp, 700,
                  (0.06, 0.03), (-0.875, -0.285), (0.237, 0.279), cost, lam=1.0)
        init, U = synth.make_inputs(rows, 1, 3, window=0, seed=2)
        pl.step(init[0].to(dev), U[0].to(dev))
        torch.cuda.synchronize()
        print("case aux: finite=", all(torch.isfinite(t).all().item() for t in (o, o2, o3, mae, rmse)), flush=True)
    print("sanitize cases done", flush=True)


if __name__ == "__main__":
    f = sys.argv[1:] or ["all"]
    if "all" in f:
        f = ["mlp2", "mlp", "rnn", "rnn_tc", "mlp_tc", "lstm_tc", "aux"]
    main(f)
