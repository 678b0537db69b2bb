import os, sys, cProfile, pstats
ROOT = "/root/repo"
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import torch, mpcdl, bench
from mpcdl import synth
dev = torch.device("cuda:0")
m, meta = mpcdl.from_checkpoint_name("Basic_64_4_32", bench.load_ckpt("Basic_64_4_32"), dev)
st = bench.load_stats(0)
init, U = synth.make_inputs(synth.load_rows(), 128, 2, window=0, seed=0)
idv, Ud = init.to(dev), U.to(dev)
cost = tuple(c.to(dev) for c in synth.make_cost(synth.load_rows(), 2))
for _ in range(200): m.rollout(idv, Ud, st, cost=cost)
pr = cProfile.Profile(); pr.enable()
for _ in range(3000): m.rollout(idv, Ud, st, cost=cost)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
