"""Per-rank shard sizes of the strong-scaled default workload (1 M trajectories over 1 / 2 / 4 / 8 ranks) timed on ONE GPU:
kernel time of a shard against the ideal (time at 1 M divided by the rank count) = what wave quantisation costs."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import mpcdl
    from conftest import load_ckpt
    from mpcdl import ops, synth, _lib
    dev = torch.device("cuda:0")
    z = np.load(os.path.join(ROOT, "tests", "golden", "stats.npz"))
    st = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("mlp_")}
    rows = synth.load_rows()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    m, _ = mpcdl.from_checkpoint_name("Basic_64_4_32", load_ckpt("Basic_64_4_32"), dev)
    H = 50
    init, U = synth.make_inputs(rows, 65536, H, window=0, seed=0)
    cost = tuple(c.to(dev) for c in synth.make_cost(rows, H))
    base = None
    for ranks in (1, 2, 4, 8):
        n = (1 << 20) // ranks
        idx = torch.arange(n) % 65536
        i_d, U_d = init[idx].contiguous().to(dev), U[idx].contiguous().to(dev)
        for tune in ((), (8, 0, 0, 1), (8, 0, 0, 4)):
            p = ops.plan(_lib.MLP, 64, 4, 0, n, H, tune=tune)
            fn = lambda: m.rollout(i_d, U_d, st, cost=cost, return_states=False, tune=tune)   # noqa: E731
            for _ in range(3):
                fn()
            ms = []
            for _ in range(7):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record()
                torch.cuda.synchronize()
                ms.append(e0.elapsed_time(e1))
            med = sorted(ms)[3]
            if base is None:
                base = med
            print(f"ranks {ranks} shard {n:8d} tune={str(tune):13s} tile {p['tile_traj']:3d} threads {p['threads']} grid {p['grid']} ntiles {p['ntiles']:5d}: "
                  f"{med:8.3f} ms  ideal {base / ranks:7.3f}  efficiency {base / ranks / med:.3f}")


if __name__ == "__main__":
    main()
