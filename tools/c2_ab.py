"""Tile variants of the weight-resident FP32 MLP kernel (rollout_mlp2, tune = (rt, n_tg, nslots, ksplit)) on the C2 shape and
around it.  CUDA events, L2 flushed between launches.  (profiles/r02/c2_experiment/ holds the run that also had the
two-group experiment kernel, tune kc = 3, which was dropped: 0.364 ms against 0.287 ms.)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import mpcdl
    from conftest import load_ckpt
    from mpcdl import ops, synth, _lib
    import numpy as np
    dev = torch.device("cuda:0")
    z = np.load(os.path.join(ROOT, "tests", "golden", "stats.npz"))
    st = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("mlp_")}
    rows = synth.load_rows()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    pk = ops.ffma_peak_tflops(dev, both=True)
    peak = max(pk)
    print(f"fp32 peak: ffma {pk[0]:.1f}  ffma2 {pk[1]:.1f} TFLOP/s")
    for name in ("Basic_64_4_32", "Basic_64_2_64"):
        m, meta = mpcdl.from_checkpoint_name(name, load_ckpt(name), dev)
        flop = 2 * (8 * 64 + meta["layers"] * 64 * 64 + 6 * 64)
        for n, h, states in ((4096, 50, True), (2048, 50, True), (4736, 50, True), (8192, 50, True), (9472, 50, True), (16384, 50, True),
                             (65536, 50, False), (1 << 20, 50, False)):
            if name != "Basic_64_4_32" and n != 4096:
                continue
            init, U = synth.make_inputs(rows, min(n, 65536), h, window=0, seed=0)
            idx = torch.arange(n) % init.shape[0]
            init, U = init[idx].contiguous().to(dev), U[idx].contiguous().to(dev)
            cost = tuple(c.to(dev) for c in synth.make_cost(rows, h))
            for tune in ((), (4, 8, 0, 2), (8, 0, 0, 2), (8, 0, 0, 1), (8, 0, 0, 4)):
                try:
                    p = ops.plan(_lib.MLP, 64, meta["layers"], 0, n, h, tune=tune)
                    fn = lambda: m.rollout(init, U, st, cost=cost, return_states=states, tune=tune)   # noqa: E731
                    for _ in range(3):
                        fn()
                    reps = 30 if n <= 16384 else 5
                    ms = []
                    for _ in range(reps):
                        flush.zero_()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(); fn(); e1.record()
                        torch.cuda.synchronize()
                        ms.append(e0.elapsed_time(e1))
                    ms.sort()
                    med = ms[len(ms) // 2]
                    print(f"{name} N={n:8d} H={h} tune={str(tune):14s} variant {p['variant']} tile {p['tile_traj']:3d} grid {p['grid']:3d} "
                          f"ks {p['ksplit']}  median {med:8.4f} ms  min {ms[0]:8.4f}  {n * h / (med * 1e-3):10.3e} traj-steps/s  "
                          f"frac {n * h * flop / (med * 1e-3) / 1e12 / peak:.3f}")
                except Exception as e:
                    print(f"{name} N={n} tune={tune}: {type(e).__name__}: {str(e)[:100]}")


if __name__ == "__main__":
    main()
