"""CPU, world_size 2, gloo: the N>1 host path -- contiguous trajectory shards, no exchange inside
the horizon, ONE all-gather of per-trajectory costs; results identical to a single-rank run.
The rollout itself has no CPU implementation, so the oracle stands in for the kernel here."""
import functools
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import load_ckpt


def test_shard_ranges_cover_and_are_contiguous():
    from mpcdl import parallel
    for n in (1, 2, 7, 64, 65537):
        for world in (1, 2, 3, 8):
            spans = [parallel.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == parallel.shard_sizes(n, world)
    with pytest.raises(ValueError):
        parallel.shard_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, out_dir):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.dirname(here))
    sys.path.insert(0, os.path.join(os.path.dirname(here), "mpc-deep-learning-sp24_b200"))
    from mpcdl import parallel, synth
    from oracle import rollout_ref as R
    torch.set_num_threads(1)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        z = np.load(os.path.join(here, "golden", "stats.npz"))
        st = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("mlp_")}
        rows = synth.load_rows()
        s0, U = synth.make_inputs(rows, n, 12, seed=4)
        eps = synth.make_eps(n, 12, seed=1)
        cost = synth.make_cost(rows, 12)
        sd = {k: torch.from_numpy(v) for k, v in np.load(os.path.join(here, "golden", "checkpoints", "Basic_16_2_32.npz")).items()}
        m, _ = R.build_from_name("Basic_16_2_32", sd)
        noise = R.RefNoise(6, 2, 3, 2, 32, "tanh", "identity")
        noise.model.load_state_dict({k: torch.from_numpy(v) for k, v in
                                     np.load(os.path.join(here, "golden", "checkpoints", "noise_seed0_h3_s32_l2.npz")).items()})

        def rollout_fn(init, u, eps=None, traj_offset=0):
            assert traj_offset == parallel.shard_range(n, rank, world)[0]
            return R.rollout_mlp(m, st, init, u, noise=noise.eval(), eps=eps, cost=cost)

        states, costs, (lo, hi) = parallel.sharded_rollout(rollout_fn, s0, U, eps=eps)
        full_states, full_costs = R.rollout_mlp(m, st, s0, U, noise=noise, eps=eps, cost=cost)
        assert costs.shape == (n,)
        assert torch.equal(states, full_states[lo:hi])            # shard invariance, bit for bit
        assert torch.equal(costs, full_costs)
        w = parallel.mppi_weights(costs, 0.5)
        assert abs(w.sum().item() - 1) < 1e-5
        with pytest.raises(RuntimeError):
            parallel.gather_costs(costs[:1], n)
        torch.save(costs, os.path.join(out_dir, f"costs_{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [64, 37])          # equal shards (all_gather_into_tensor) and ragged shards (padded)
def test_sharded_rollout_world2(n, tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n, str(tmp_path)), nprocs=2, join=True)
    a, b = torch.load(tmp_path / "costs_0.pt"), torch.load(tmp_path / "costs_1.pt")
    assert torch.equal(a, b) and a.shape == (n,)
