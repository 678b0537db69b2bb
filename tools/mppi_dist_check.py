"""torchrun --nproc-per-node 2 tools/mppi_dist_check.py : an MPPI iteration sharded over the ranks equals the same
iteration on one GPU (samples keyed by the global trajectory index; all-gather of costs + all-reduce of 2H+1 sums)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import numpy as np, torch, torch.distributed as dist
import mpcdl
from mpcdl import synth
from mpcdl.mppi import MPPI, apply_update

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
GOLD = os.path.join(ROOT, "tests", "golden")
ck = lambda n: {k: torch.from_numpy(v) for k, v in np.load(f"{GOLD}/checkpoints/{n}.npz").items()}
z = np.load(f"{GOLD}/stats.npz")
SIG, LO, HI = (0.06, 0.03), (-0.875, -0.285), (0.237, 0.279)
ok = True
for name, prec, N in (("Basic_64_4_32", "fp32", 100001), ("GRU_64_4_4_16_False", "tf32x3", 30000)):
    m, meta = mpcdl.from_checkpoint_name(name, ck(name), dev)
    pre = "mlp_" if meta["window"] == 0 else "rnn4_"
    st = {k[len(pre):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(pre)}
    rows = synth.load_rows()
    H = 30
    init, U = synth.make_inputs(rows, 1, H, window=meta["window"], seed=2)
    init, U = init[0].contiguous().to(dev), U[0].contiguous().to(dev)
    cost = tuple(c.to(dev) for c in synth.make_cost(rows, H))
    for gather in (True, False):
        plan = MPPI(m, st, N, SIG, LO, HI, cost, lam=5.0, precision=prec, seed=4, gather_costs=gather)
        U_new, info = plan.step(init, U)
        # single-GPU evaluation of the same iteration
        c = m.rollout(init, U, st, cost=cost, return_states=False, precision=prec, seed=4, stateful=False,
                      sample=dict(n=N, sigma=SIG, lo=LO, hi=HI, seed=4), shared_init=True)[1]
        sums = torch.ops.mpcdl.mppi_reduce(c, U, list(SIG) + list(LO) + list(HI), 4, 0, 5.0, None)
        U_ref = apply_update(U, sums, LO, HI)
        err = (U_new - U_ref).abs().max().item()
        same_costs = torch.equal(info["costs"], c[info["lo"]:info["hi"]])
        gathered = info["costs_all"] is None or torch.equal(info["costs_all"], c)
        print(f"rank {rank} {name} {prec} gather={gather}: |U_new - U_ref| max {err:.2e}, shard costs identical {same_costs}, gathered identical {gathered}", flush=True)
        ok = ok and err <= 1e-5 and same_costs and gathered
dist.barrier()
dist.destroy_process_group()
print("MPPI DIST OK" if ok else "MPPI DIST FAIL", flush=True)
sys.exit(0 if ok else 1)
