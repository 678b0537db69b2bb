"""Determinism of every kernel family under perturbation (round-1 verdict, weak #1).

compute-sanitizer is closed on the GPU pool ("runs under it have left GPUs needing a reset",
profiles/r02/gru_repro/compute_sanitizer_closed.txt), so initcheck / racecheck are replaced by what can be observed from
outside: every variant runs 30 times back to back and all outputs must be torch.equal to the first, while
  * before each launch the shared memory of every SM is filled with NaN patterns (mpcdl_debug_poison_smem): a read of
    shared memory the kernel has not written turns into NaN / changed bits instead of quietly reusing the previous
    launch's identical contents;
  * the result buffers are pre-filled with NaN (out=): an element the kernel does not write shows up;
  * a second stream keeps thrashing L2 (256 MiB memsets) and re-poisoning shared memory of idle SMs while the rollout
    runs, so the CTAs see different memory timing on every repeat.
The FP32 tile variants (RT = 2 / 4 / 8) must moreover agree with EACH OTHER bit for bit (same arithmetic order per
trajectory; profiles/r02/gru_repro/cmp.log shows it held across plain and ncu runs)."""
import numpy as np
import pytest
import torch

from conftest import load_ckpt

pytestmark = pytest.mark.gpu
REPEATS = 30


def _dev():
    return torch.device("cuda:0")


class _Perturb:
    def __init__(self):
        self.side = torch.cuda.Stream(_dev())
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=_dev())
        self.tok = torch.zeros(1, device=_dev())

    def before(self):
        torch.ops.mpcdl.debug_poison_smem(self.tok)                 # same stream: every SM's shared memory is NaN now
        self.side.wait_stream(torch.cuda.current_stream(_dev()))
        with torch.cuda.stream(self.side):                          # concurrent with the rollout that follows
            self.flush.zero_()
            torch.ops.mpcdl.debug_poison_smem(self.tok)
            self.flush.zero_()

    def after(self):
        torch.cuda.current_stream(_dev()).wait_stream(self.side)


def _repeat(fn, check_extra=None):
    pt = _Perturb()
    first = [t.clone() for t in fn(None)]
    assert all(torch.isfinite(t).all() for t in first)
    for _ in range(REPEATS):
        bufs = [torch.full_like(t, float("nan")) for t in first]
        pt.before()
        out = fn(bufs)
        pt.after()
        for a, b in zip(first, out):
            assert torch.equal(a, b)
    return first


def _stats(kind, gold_stats):
    return gold_stats[0] if kind == "mlp" else gold_stats[1]


def _noise():
    import mpcdl
    n = mpcdl.FeedforwardNoiseModel(6, 2, 3, n_layers=2, size=32, activation="tanh", output_activation="identity")
    n.model.load_state_dict(load_ckpt("noise_seed0_h3_s32_l2"))
    return n.to(_dev())


CASES = [
    # name, N, H, precision, tune, stateful, noise
    ("Basic_64_4_32", 300, 6, "fp32", (), False, False),               # weight-resident MLP kernel, K split over 2 thread sets
    ("Basic_64_4_32", 5000, 4, "fp32", (8, 0, 0, 0), False, False),    # 8x8 tiles, persistent loop over several tiles
    ("Basic_16_2_32", 300, 6, "fp32", (), False, False),
    ("Basic_64_4_32", 300, 6, "fp32", (0, 0, 0, -1), False, False),    # generic ring-fed MLP kernel
    ("Basic_64_4_32", 300, 6, "fp32", (), False, True),                # generic kernel + residual-noise net
    ("GRU_64_4_4_16_False", 300, 5, "fp32", (2, 0, 0, 0), False, False),
    ("GRU_64_4_4_16_False", 300, 5, "fp32", (4, 0, 0, 0), False, False),
    ("GRU_64_4_4_16_False", 300, 5, "fp32", (8, 0, 0, 0), False, False),
    ("GRU_64_4_4_16_False", 300, 5, "fp32", (), False, True),
    ("GRU_64_4_4_64_True", 300, 5, "fp32", (), True, False),
    ("LSTM_64_4_4_16_False", 300, 5, "fp32", (2, 0, 0, 0), False, False),
    ("LSTM_64_4_4_16_False", 300, 5, "fp32", (4, 0, 0, 0), False, False),
    ("LSTM_64_4_4_64_True", 300, 5, "fp32", (8, 0, 0, 0), True, False),
    ("GRU_64_4_4_16_False", 300, 5, "tf32x3", (), False, False),
    ("GRU_64_4_4_16_False", 300, 5, "tf32x3", (), False, True),
    ("GRU_64_4_4_64_True", 300, 5, "tf32", (), True, False),
    ("LSTM_64_4_4_16_False", 300, 5, "tf32x3", (), False, False),
    ("LSTM_64_4_4_64_True", 300, 5, "tf32", (), True, True),
    ("Basic_64_4_32", 700, 6, "tf32x3", (), False, False),
    ("Basic_64_4_32", 40000, 3, "tf32", (), False, False),              # two tiles per CTA ping-pong
]


@pytest.mark.parametrize("name,N,H,prec,tune,stateful,noise_on", CASES)
def test_repeatable_under_perturbation(name, N, H, prec, tune, stateful, noise_on, gold_stats):
    import mpcdl
    from mpcdl import synth
    m, meta = mpcdl.from_checkpoint_name(name, load_ckpt(name), _dev())
    st = _stats(meta["kind"], gold_stats)
    rows = synth.load_rows()
    init, U = synth.make_inputs(rows, N, H, window=meta["window"], seed=21)
    init, U = init.to(_dev()), U.to(_dev())
    cost = tuple(c.to(_dev()) for c in synth.make_cost(rows, H))
    noise = _noise() if noise_on else None
    carry0 = torch.randn(meta["layers"] * (2 if meta["kind"] == "lstm" else 1), N, meta["hidden"], device=_dev()) * 0.3 if stateful else None

    def fn(bufs):
        kw = dict(noise=noise, seed=5, cost=cost, precision=prec, tune=tune, out=bufs[:2] if bufs else None)
        if stateful:
            c = carry0.clone()
            s, k, c = m.rollout(init, U, st, stateful=True, carry=c, **kw)
            return [s, k, c]
        s, k = m.rollout(init, U, st, stateful=False, **kw)[:2]
        return [s, k]
    _repeat(fn)


def test_fp32_tile_variants_agree_bitwise(gold_stats):
    """a trajectory's FP32 result does not depend on the register tile / CTA size that processed it"""
    import mpcdl
    from mpcdl import synth
    rows = synth.load_rows()
    for name in ("GRU_64_4_4_16_False", "LSTM_64_4_4_16_False"):
        m, meta = mpcdl.from_checkpoint_name(name, load_ckpt(name), _dev())
        init, U = synth.make_inputs(rows, 500, 8, window=4, seed=22)
        outs = [m.rollout(init.to(_dev()), U.to(_dev()), gold_stats[1], tune=(rt, 0, 0, 0))[0] for rt in (2, 4, 8)]
        assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


def test_bf16_lstm256_repeatable_under_perturbation(gold_stats):
    import mpcdl
    from mpcdl import synth
    torch.manual_seed(0)
    mod = torch.nn.ModuleDict(dict(lstm=torch.nn.LSTM(8, 256, 2, batch_first=True), full_c=torch.nn.Linear(256, 6)))
    m, _ = mpcdl.from_checkpoint_name("LSTM_256_2_4_0_False", {k: v.detach().clone() for k, v in mod.state_dict().items()}, _dev())
    init, U = synth.make_inputs(synth.load_rows(), 600, 4, window=4, seed=23)
    init, U = init.to(_dev()), U.to(_dev())
    _repeat(lambda bufs: [m.rollout(init, U, gold_stats[1], precision="bf16", out=(bufs[0], None) if bufs else None)[0]])


def test_aux_kernels_repeatable_under_perturbation(gold_stats):
    import mpcdl
    from mpcdl import data, synth
    noise = _noise()
    g = torch.Generator().manual_seed(3)
    q, u, e = (torch.randn(500, k, generator=g).to(_dev()) for k in (18, 6, 6))
    _repeat(lambda bufs: [noise(q, u, e)])
    dn = mpcdl.DynamicsNN(None, noise, False, history=3)
    q0, U = torch.randn(300, 6, generator=g).to(_dev()), torch.randn(300, 5, 2, generator=g).to(_dev())
    _repeat(lambda bufs: [dn.rollout(q0, U, seed=3)])
    bags = synth.load_bags("pid_data_clean")
    _repeat(lambda bufs: list(data.device_dataset(bags, 4, _dev())))
    m, _ = mpcdl.from_checkpoint_name("Basic_64_4_32", load_ckpt("Basic_64_4_32"), _dev())
    from mpcdl.mppi import MPPI
    rows = synth.load_rows()
    cost = tuple(c.to(_dev()) for c in synth.make_cost(rows, 6))
    init, Ub = synth.make_inputs(rows, 1, 6, window=0, seed=2)

    def plan(bufs):
        pl = MPPI(m, gold_stats[0], 3000, (0.06, 0.03), (-0.875, -0.285), (0.237, 0.279), cost, lam=1.0, seed=4)
        return [pl.step(init[0].to(_dev()), Ub[0].to(_dev()))[0]]
    _repeat(plan)
