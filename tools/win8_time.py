"""Device time of the window-8 GRU / LSTM (hidden 64, 4 layers: ft_size 8 of deep_models.py:184) at C3's size, FP32 kernel vs
the tcgen05 kernel with the sequence in the per-CTA scratch.  Random-init weights, synthetic inputs, CUDA events, L2 flushed."""
import sys, os, json
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mpc-deep-learning-sp24_b200"))
import mpcdl
from mpcdl import synth

dev = torch.device("cuda:0")
N, H = 65536, 100
rows = synth.load_rows()
out = {}
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for kind in ("gru", "lstm"):
    torch.manual_seed(1)
    m = (mpcdl.GRU_Model if kind == "gru" else mpcdl.LSTM_Model)(8, 64, 6, 4, False).eval().to(dev)
    init, U = synth.make_inputs(rows, N, H, window=8, seed=3)
    init, U = init.to(dev), U.to(dev)
    mean = init.mean(0); std = init.std(0)
    if True:
        st = dict(f_mean=mean.cpu(), f_std=std.cpu(), l_mean=mean[-1, :6].cpu(), l_std=std[-1, :6].cpu())
    for prec in ("fp32", "tf32x3", "tf32"):
        ts = []
        for i in range(5):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            s, _ = m.rollout(init, U, st, precision=prec)[:2]
            e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        out[f"{kind}_{prec}"] = {"ms": round(min(ts[2:]), 3), "finite": bool(torch.isfinite(s).all())}
        print(kind, prec, out[f"{kind}_{prec}"], flush=True)
os.makedirs("gpurun_out/win8", exist_ok=True)
json.dump(out, open("gpurun_out/win8/time.json", "w"), indent=1)
