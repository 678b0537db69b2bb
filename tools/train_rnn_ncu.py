"""One short launch of the recurrent training kernel for ncu: python tools/train_rnn_ncu.py [gru|lstm] [hidden] [n_models] [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200")); sys.path.insert(0, ROOT)
import torch, mpcdl
from mpcdl import _lib, synth
kind = sys.argv[1] if len(sys.argv) > 1 else "gru"
hidden = int(sys.argv[2]) if len(sys.argv) > 2 else 64
nm = int(sys.argv[3]) if len(sys.argv) > 3 else 1
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 40
layers, window, batch = 4, 4, 16
ds = mpcdl.RobotData(synth.load_bags(), window)
f, l = ds.featrs.float().cuda(), ds.labels.float().cuda()
torch.manual_seed(0)
m = (mpcdl.GRU_Model if kind == "gru" else mpcdl.LSTM_Model)(8, hidden, 6, layers, False)
k = _lib.GRU if kind == "gru" else _lib.LSTM
npar = _lib.lib.mpcdl_train_rnn_param_count(k, hidden, layers)
stride = (npar + 3) // 4 * 4
wf = _lib.lib.mpcdl_train_rnn_work_floats(k, hidden, layers, window)
flat = torch.cat([p.detach().reshape(-1) for p in m.parameters()]).cuda()
theta = torch.zeros(nm, stride, device="cuda"); theta[:, :npar] = flat
am, av = torch.zeros_like(theta), torch.zeros_like(theta)
work = torch.empty(nm * wf, device="cuda")
g = torch.Generator().manual_seed(0)
orders = torch.stack([torch.randperm(f.shape[0], generator=g)[:steps * batch] for _ in range(nm)]).int().cuda()
for rep in range(2):
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    torch.ops.mpcdl.train_rnn_epoch(theta, am, av, f, l, orders, work, k, hidden, layers, batch, rep * steps, 1e-3, 0.9, 0.999, 1e-8)
    e1.record(); e1.synchronize()
print(f"{kind} hidden {hidden} x {layers}, window {window}, batch {batch}, {nm} models, {steps} steps: {e0.elapsed_time(e1) / steps * 1e3:.1f} us per step")
