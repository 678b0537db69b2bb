"""Development aid: one tiny BF16 LSTM-256 rollout (for compute-sanitizer)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import torch, mpcdl, bench
from mpcdl import synth
dev = torch.device("cuda:0")
N, H, L = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
W = int(sys.argv[4]) if len(sys.argv) > 4 else 4
name = f"LSTM_256_{L}_{W}_0_False"
m, meta = mpcdl.from_checkpoint_name(name, bench.load_ckpt(name), dev)
ST = bench.load_stats(4)
ST = dict(f_mean=ST['f_mean'][:W].contiguous(), f_std=ST['f_std'][:W].contiguous(), l_mean=ST['l_mean'], l_std=ST['l_std'])
dbg = torch.zeros(64, dtype=torch.int32).pin_memory()      # MPCDL_LT_DEBUG builds: progress words + timed-out waits
os.environ["MPCDL_LT_DBG"] = hex(dbg.data_ptr())
NAMES = ["producer:empty", "mma:full", "mma:d_empty", "mma:ax_full", "mma:ah_full", "epi:d_full"]
def report():
    d = dbg.numpy().view("uint32")
    print(f"progress: epilogue chunks {d[0]}  mma chunks {d[1]}  slots produced {d[2]}")
    for r in range(2):
        w = d[48 + 8 * r:48 + 8 * r + 7]
        print(f"  MMA warp of CTA {r}: kcycles waiting full {w[1]} d_empty {w[2]} ax_full {w[3]} ah_full {w[4]} | total {w[6]}")
    for r in range(2):
        for i, n in enumerate(NAMES):
            if d[16 + 2 * (i + 8 * r)]:
                print(f"  TIMEOUT rank {r} {n}: cta {d[16 + 2 * (i + 8 * r)] & 0xFFFF} count {d[17 + 2 * (i + 8 * r)]}")
init, U = synth.make_inputs(synth.load_rows(), N, H, window=W, seed=3)
try:
    y = m.rollout(init.to(dev), U.to(dev), ST, precision="bf16")[0]
    torch.cuda.synchronize()
except Exception as ex:
    print("FAILED:", str(ex).splitlines()[0]); report(); sys.exit(1)
report()
y32 = m.rollout(init.to(dev), U.to(dev), ST, precision="fp32")[0]
from oracle import rollout_ref as R
e = R.rel_err(y.cpu(), y32.cpu())
print("rel err vs fp32 kernel per step (max over trajectories):", [f"{v:.1e}" for v in e.max(0).values.tolist()], "median step1", f"{e[:,0].median().item():.1e}")
print("ok", N, H, L, W, "max abs diff vs fp32 kernel", (y - y32).abs().max().item(), "nan", int(torch.isnan(y).sum()), flush=True)
