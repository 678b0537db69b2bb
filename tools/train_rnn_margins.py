"""Margins of tests/test_gpu_train_rnn.py::test_rnn_training_steps_match_the_oracle: prints the measured figures beside the bounds."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200")); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import test_gpu_train_rnn as TT
from mpcdl import synth, train
from oracle import train_ref as T
bags = synth.load_bags("mpc_data_clean")
cases = [("gru", 64, 4, 4, 16, False), ("lstm", 64, 4, 4, 16, False), ("gru", 128, 2, 8, 8, True), ("lstm", 128, 2, 8, 8, False),
         ("gru", 16, 4, 4, 16, False), ("lstm", 16, 4, 4, 16, True), ("gru", 32, 1, 1, 5, False), ("lstm", 32, 1, 2, 5, False),
         ("gru", 64, 2, 8, 16, False), ("lstm", 64, 2, 8, 8, True)]
for seed_off in (0, 1000):
    for kind, hidden, layers, window, batch, state_f in cases:
        f, l = TT._data(bags, window)
        m, ref = TT._models(kind, hidden, layers, state_f, seed=hidden + layers + window + seed_off)
        order = torch.randperm(f.shape[0], generator=torch.Generator().manual_seed(1 + seed_off))
        fd, ld = f.cuda(), l.cuda()
        st = train.AdamState(1, TT._flat(ref).numel(), "cuda:0")
        w1, opt = T.train_steps_rnn(ref, f, l, order[:batch], batch, 1e-3)
        g1 = train.train_epoch(m, st, fd, ld, order[:batch], batch, 1e-3).cpu()[0]
        d1 = (TT._flat(m) - TT._flat(ref)).abs()
        nxt = order[batch:20 * batch - 3]
        w, opt = T.train_steps_rnn(ref, f, l, nxt, batch, 1e-3, optimizer=opt)
        g = train.train_epoch(m, st, fd, ld, nxt, batch, 1e-3).cpu()[0]
        d = (TT._flat(m) - TT._flat(ref)).abs()
        print(f"{kind} {hidden} {layers} {window} {batch} seed+{seed_off}: step1 loss rel {abs(g1[0]-w1[0]).item()/w1[0].item():.1e} (1e-5) frac<=1e-5 {(d1<=1e-5).float().mean().item():.4f} (0.99) "
              f"max {d1.max().item():.1e} (2.5e-3) | 20 steps loss rel {((g-w).abs()/w.abs()).max().item():.1e} (1e-4) frac<=2e-5 {(d<=2e-5).float().mean().item():.4f} (0.98) max {d.max().item():.1e} (5e-3)", flush=True)
