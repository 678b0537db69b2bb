"""Round-1 verdict, weak #1: the seeded smoke() GRU case printed 5.37e-07 plain and 7.70e-06 under ncu.

    python tools/gru_repro.py TAG          # writes gpurun_out/repro/TAG.npz (GPU outputs AND the CPU oracle outputs)
    python tools/gru_repro.py --cmp A B    # which side moved between two runs?

Each run records, for the smoke cases (seed 11) of the GRU and the LSTM: the inputs, the GPU result of every FP32
tile variant (tune rt = 2/4/8; 20 repeats each, all compared bit for bit), the CPU oracle result, and the thread /
ISA settings torch's CPU kernels ran with."""
import hashlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
OUT = os.path.join(ROOT, "gpurun_out", "repro")


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()[:12]


def run(tag):
    import mpcdl
    from oracle import rollout_ref as R
    os.makedirs(OUT, exist_ok=True)
    dev = torch.device("cuda:0")
    gold = os.path.join(ROOT, "tests", "golden")
    z = np.load(os.path.join(gold, "stats.npz"))
    st = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("rnn4_")}
    rows = mpcdl.synth.load_rows()
    rec = {}
    print(f"[{tag}] torch threads {torch.get_num_threads()} interop {torch.get_num_interop_threads()} "
          f"cpu {torch.backends.cpu.get_cpu_capability()} mkldnn {torch.backends.mkldnn.is_available()} "
          f"OMP_NUM_THREADS={os.environ.get('OMP_NUM_THREADS')} affinity {len(os.sched_getaffinity(0))}")
    for name, n, h in (("GRU_64_4_4_16_False", 128, 6), ("LSTM_64_4_4_16_False", 128, 6)):
        sd = {k: torch.from_numpy(v) for k, v in np.load(os.path.join(gold, "checkpoints", f"{name}.npz")).items()}
        m, meta = mpcdl.from_checkpoint_name(name, sd, dev)
        ref, _ = R.build_from_name(name, sd)
        init, U = mpcdl.synth.make_inputs(rows, n, h, window=4, seed=11)
        rec[f"{name}_init"], rec[f"{name}_U"] = init.numpy(), U.numpy()
        want = R.rollout_rnn(ref, st, init, U)[0]
        want2 = R.rollout_rnn(ref, st, init, U)[0]
        torch.set_num_threads(1)
        want1 = R.rollout_rnn(ref, st, init, U)[0]
        torch.set_num_threads(os.cpu_count())
        rec[f"{name}_want"], rec[f"{name}_want1"] = want.numpy(), want1.numpy()
        print(f"[{tag}] {name} oracle sha {sha(want.numpy())} repeat-equal {torch.equal(want, want2)} "
              f"1-thread-equal {torch.equal(want, want1)} (max diff {(want - want1).abs().max().item():.2e})")
        di, dU = init.to(dev), U.to(dev)
        for tune in ((), (2, 0, 0, 0), (4, 0, 0, 0), (8, 0, 0, 0)):
            outs = [m.rollout(di, dU, st, tune=tune)[0].cpu() for _ in range(20)]
            same = all(torch.equal(outs[0], o) for o in outs[1:])
            err = R.rel_err(outs[0], want)
            key = f"{name}_got_rt{tune[0] if tune else 0}"
            rec[key] = outs[0].numpy()
            print(f"[{tag}] {name} tune={tune or 'auto'} sha {sha(outs[0].numpy())} repeats-equal {same} "
                  f"one-step {err[:, 0].max().item():.2e} p90 {torch.quantile(err.flatten(), 0.9).item():.2e}")
    np.savez(os.path.join(OUT, f"{tag}.npz"), **rec)


def cmp(a, b):
    A, B = np.load(os.path.join(OUT, f"{a}.npz")), np.load(os.path.join(OUT, f"{b}.npz"))
    for k in A.files:
        eq = np.array_equal(A[k], B[k])
        d = float(np.abs(A[k].astype(np.float64) - B[k].astype(np.float64)).max())
        print(f"{k:40s} {'IDENTICAL' if eq else 'DIFFERENT'}  max|diff| {d:.3e}")


if __name__ == "__main__":
    if sys.argv[1] == "--cmp":
        cmp(sys.argv[2], sys.argv[3])
    else:
        run(sys.argv[1])
