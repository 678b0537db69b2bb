"""BASELINE.json config 4: "LSTM rnn_model rollout, hidden 256, BF16 tensor-core path vs FP32 tolerance study".
Error of the closed-loop rollout against the FP64 oracle as a function of the step, for torch FP32 (the reference
arithmetic), the FP32 FFMA kernel and the BF16 tcgen05 kernel.   python tools/c4_tolerance.py [N] [H] > study.md"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import torch
import mpcdl, bench
from mpcdl import synth
from oracle import rollout_ref as R

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
H = int(sys.argv[2]) if len(sys.argv) > 2 else 100
dev = torch.device("cuda:0")
name = "LSTM_256_4_4_0_False"
sd = bench.load_ckpt(name)
st = bench.load_stats(4)
m, _ = mpcdl.from_checkpoint_name(name, sd, dev)
ref, _ = R.build_from_name(name, sd)
init, U = synth.make_inputs(synth.load_rows(), N, H, window=4, seed=3)
y64 = R.rollout_rnn(R.as_double(ref), {k: v.double() for k, v in st.items()}, init.double(), U.double(), stateful=False)[0]
runs = {
    "torch FP32 (reference arithmetic)": R.rollout_rnn(ref, st, init, U, stateful=False)[0],
    "kernel FP32 (FFMA)": m.rollout(init.to(dev), U.to(dev), st, precision="fp32")[0].cpu(),
    "kernel BF16 (tcgen05, CTA pairs)": m.rollout(init.to(dev), U.to(dev), st, precision="bf16")[0].cpu(),
}
steps = [s for s in (1, 2, 5, 10, 25, 50, 100, 200) if s <= H]
print(f"# C4 tolerance study: {name} (seeded random init), N = {N} trajectories, horizon {H}, synthetic BARC inputs\n")
print("Error metric (SURVEY.md 8c): max_d |y_d - y64_d| / max_{n,t} |y64_d| per trajectory-step, against the FP64 run of the same module.\n")
for label, y in runs.items():
    e = R.rel_err(y, y64)
    print(f"## {label}\n\n| step | p50 | p90 | p99 | max |\n|---|---|---|---|---|")
    for s in steps:
        c = e[:, s - 1].double()
        print(f"| {s} | " + " | ".join(f"{torch.quantile(c, q).item():.2e}" for q in (0.5, 0.9, 0.99)) + f" | {c.max().item():.2e} |")
    a = e.flatten().double()
    print(f"| all {H} | " + " | ".join(f"{torch.quantile(a, q).item():.2e}" for q in (0.5, 0.9, 0.99)) + f" | {a.max().item():.2e} |\n")
d = (runs["kernel BF16 (tcgen05, CTA pairs)"] - y64).abs().amax((0, 1))
print("Largest absolute BF16 error per state dimension [x, y, psi, w_psi, v_long, v_tran]: " + ", ".join(f"{v:.2e}" for v in d.tolist()))
print("Largest |state| per dimension in the FP64 run: " + ", ".join(f"{v:.2e}" for v in y64.abs().amax((0, 1)).tolist()))
