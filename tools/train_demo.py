"""End-to-end use of the fused training path (mpcdl.train.train_model, the reference's protocol) on the data the reference
trains on (mpc_data_clean), beside the SHIPPED checkpoint of the same configuration: L1 loss (z-scored units, what
train_model prints) of both over all samples.  python tools/train_demo.py [epochs]      (GPU box)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "mpc-deep-learning-sp24_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
import mpcdl
from mpcdl import synth, train
epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 100
dev = "cuda:0"
bags = synth.load_bags("mpc_data_clean")
ck = lambda n: {k: torch.from_numpy(v) for k, v in np.load(os.path.join(ROOT, "tests", "golden", "checkpoints", f"{n}.npz")).items()}
def l1_all(model, ds, batch):
    return train.validation_loss(model, ds.featrs.float().to(dev).contiguous(), ds.labels.float().to(dev).contiguous(), batch)
for name in ("Basic_64_4_32", "Basic_16_4_32", "GRU_64_4_4_16_False", "LSTM_64_4_4_64_True"):
    shipped, meta = mpcdl.from_checkpoint_name(name, ck(name), dev)
    window = meta["window"] or None
    ds = mpcdl.RobotData(bags, window)
    batch = int(name.split("_")[3 if meta["kind"] == "mlp" else 4])
    torch.manual_seed(0)
    if meta["kind"] == "mlp":
        m = mpcdl.Basic_Model(8, meta["hidden"], 6, meta["layers"], False).to(dev)
    else:
        m = (mpcdl.GRU_Model if meta["kind"] == "gru" else mpcdl.LSTM_Model)(8, meta["hidden"], 6, meta["layers"], meta["state_f"]).to(dev)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    h = train.train_model(m, ds, epochs=epochs, batch_s=batch, learn_r=1e-3, seed=0, verbose=False, bad_loss=1e9)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    m.load_state_dict(h["best"])
    print(f"{name}: {epochs} epochs of batch {batch} in {dt:.1f} s ({dt / epochs * 1e3:.0f} ms per epoch incl. validation); valid_loss first {h['valid_loss'][0]:.4f} "
          f"best {min(h['valid_loss']):.4f}; L1 over all {len(ds)} samples: trained here {l1_all(m, ds, batch):.4f}, shipped checkpoint {l1_all(shipped, ds, batch):.4f}", flush=True)
