"""Epoch time of the fused recurrent training kernel (csrc/train_rnn_kernels.cu) for 1 / 148 / 296 models in one launch, beside
the oracle (the reference's loop on torch CPU, all host threads) -- python tools/train_rnn_time.py  (GPU box)."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200")); sys.path.insert(0, ROOT)
import mpcdl  # noqa: E402
from mpcdl import _lib, synth  # noqa: E402
from oracle import rollout_ref as R, train_ref as T  # noqa: E402

torch.set_num_threads(os.cpu_count())
bags = synth.load_bags()
for kind, hidden, layers, window, batch in (("gru", 16, 4, 4, 16), ("gru", 64, 4, 4, 16), ("lstm", 64, 4, 4, 16), ("gru", 128, 4, 8, 16), ("lstm", 128, 2, 8, 8)):
    ds = mpcdl.RobotData(bags, window)
    f, l = ds.featrs.float(), ds.labels.float()
    n_train = int(f.shape[0] * 0.8)
    torch.manual_seed(0)
    ref = (R.RefGRU if kind == "gru" else R.RefLSTM)(8, hidden, 6, layers)
    g = torch.Generator().manual_seed(0)
    order = torch.randperm(f.shape[0], generator=g)[:n_train]
    T.train_steps_rnn(ref, f, l, order[:10 * batch], batch)
    n_cpu = 60 * batch
    t0 = time.perf_counter(); T.train_steps_rnn(ref, f, l, order[:n_cpu], batch); t_cpu = (time.perf_counter() - t0) / 60
    k = _lib.GRU if kind == "gru" else _lib.LSTM
    npar = _lib.lib.mpcdl_train_rnn_param_count(k, hidden, layers)
    stride = (npar + 3) // 4 * 4
    wf = _lib.lib.mpcdl_train_rnn_work_floats(k, hidden, layers, window)
    fd, ld = f.cuda(), l.cuda()
    flat = torch.cat([p.detach().reshape(-1) for p in ref.parameters()])
    steps = (n_train + batch - 1) // batch
    for nm in (1, 148, 296, 592):
        theta = torch.zeros(nm, stride, device="cuda"); theta[:, :npar] = flat.cuda()
        am, av = torch.zeros_like(theta), torch.zeros_like(theta)
        work = torch.empty(nm * wf, device="cuda")
        orders = torch.stack([torch.randperm(f.shape[0], generator=g)[:n_train] for _ in range(nm)]).int().cuda()
        args = (fd, ld, orders, work, k, hidden, layers, batch)
        torch.ops.mpcdl.train_rnn_epoch(theta, am, av, *args, 0, 1e-3, 0.9, 0.999, 1e-8)
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); torch.ops.mpcdl.train_rnn_epoch(theta, am, av, *args, steps, 1e-3, 0.9, 0.999, 1e-8); e1.record(); e1.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"{kind.upper()}_{hidden}_{layers}_{window}_{batch}: one epoch ({steps} Adam steps): {nm:4d} models in one launch {ms:9.2f} ms = "
              f"{ms / steps * 1e3:8.1f} us per step, {nm * steps / ms * 1e3:10.3e} model-steps/s; torch CPU ({torch.get_num_threads()} threads) "
              f"{t_cpu * 1e3:.2f} ms per step = {1 / t_cpu:.3e} steps/s", flush=True)
