"""Development aid: host-side cost of one rollout() / HostPipeline.submit() call (tiny kernel, wall clock per call)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200"))
import torch, mpcdl, bench
from mpcdl import synth, parallel
dev = torch.device("cuda:0")
name = "Basic_64_4_32"
m, meta = mpcdl.from_checkpoint_name(name, bench.load_ckpt(name), dev)
st = bench.load_stats(0)
init, U = synth.make_inputs(synth.load_rows(), 128, 2, window=0, seed=0)
ip, Up = init.pin_memory(), U.pin_memory()
idv, Ud = ip.to(dev), Up.to(dev)
cost = tuple(c.to(dev) for c in synth.make_cost(synth.load_rows(), 2))
def t(fn, n=3000):
    for _ in range(200): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    t1 = time.perf_counter(); torch.cuda.synchronize()
    return (t1 - t0) / n * 1e6
print(f"rollout(states)            : {t(lambda: m.rollout(idv, Ud, st)):.1f} us per call (host)")
print(f"rollout(states + cost)     : {t(lambda: m.rollout(idv, Ud, st, cost=cost)):.1f} us per call (host)")
res = torch.empty(128, 2, 6).pin_memory()
pipe = parallel.HostPipeline(lambda a, b, out: m.rollout(a, b, st, cost=cost, out=(out, None))[0], depth=2, device=dev)
print(f"HostPipeline.submit        : {t(lambda: pipe.submit((ip, Up), res)):.1f} us per call (host)")
pipe.drain()
e = torch.empty(1 << 10, device=dev)
print(f"torch.empty                : {t(lambda: torch.empty(128, 2, 6, device=dev)):.1f} us")
