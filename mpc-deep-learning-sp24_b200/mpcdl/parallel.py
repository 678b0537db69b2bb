"""Trajectory sharding across the GPUs of one box (one process per GPU, torchrun).

Rollouts are independent per trajectory, so the horizon loop needs no exchange: every rank runs the
fused kernel on a contiguous shard of the N trajectories with replicated weights, and the only
collective of the path is ONE all-gather of the per-trajectory costs afterwards (NCCL over
NVLink/NVSwitch on the GPU box, gloo in the CPU tests).  The Philox noise stream is keyed by the
GLOBAL trajectory index (``traj_offset``), so results do not depend on the number of ranks.

The reference has no distributed code at all; its closest caller is the batch-1
loop ``run_dynamics_model`` (scripts/test_model.py:67-87) that this generalises.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n: int, rank: int, world: int):
    """Contiguous shard [lo, hi) of n trajectories; the first n % world ranks hold one extra."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(n: int, world: int):
    return [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]


def gather_costs(costs_local: torch.Tensor, n: int, group=None) -> torch.Tensor:
    """All-gather of per-trajectory costs -> [n] on every rank (rank order == trajectory order).

    Equal shards use ``all_gather_into_tensor`` (one NCCL kernel); ragged shards are padded to the
    largest shard first and trimmed afterwards."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        if costs_local.numel() != n:
            raise RuntimeError(f"mpcdl: {costs_local.numel()} local costs for n={n} on a single rank")
        return costs_local
    sizes = shard_sizes(n, world)
    rank = dist.get_rank(group)
    if costs_local.numel() != sizes[rank]:
        raise RuntimeError(f"mpcdl: rank {rank} holds {costs_local.numel()} costs, its shard of n={n} is {sizes[rank]}")
    m = max(sizes)
    if min(sizes) == m:
        out = torch.empty(n, dtype=costs_local.dtype, device=costs_local.device)
        dist.all_gather_into_tensor(out, costs_local.contiguous(), group=group)
        return out
    pad = torch.zeros(m, dtype=costs_local.dtype, device=costs_local.device)
    pad[:sizes[rank]] = costs_local
    buf = torch.empty(world * m, dtype=costs_local.dtype, device=costs_local.device)
    dist.all_gather_into_tensor(buf, pad, group=group)
    return torch.cat([buf[r * m:r * m + sizes[r]] for r in range(world)])


def sharded_rollout(rollout_fn, init, U, *, eps=None, group=None, **kw):
    """Run ``rollout_fn(init_shard, U_shard, eps=..., traj_offset=lo, **kw) -> (states|None, costs)``
    on this rank's shard of the GLOBAL tensors and all-gather the costs.

    ``rollout_fn`` is normally ``functools.partial(model.rollout, stats=..., cost=...)``.
    Returns (states_local, costs_global [N], (lo, hi))."""
    n = int(U.shape[0])
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_range(n, rank, world)
    out = rollout_fn(init[lo:hi].contiguous(), U[lo:hi].contiguous(),
                     eps=None if eps is None else eps[lo:hi].contiguous(), traj_offset=lo, **kw)
    states, costs = out[0], out[1]
    if costs is None:
        raise RuntimeError("mpcdl: sharded_rollout needs a cost (the all-gather carries per-trajectory costs)")
    return states, gather_costs(costs, n, group), (lo, hi)


def mppi_weights(costs: torch.Tensor, lam: float = 1.0) -> torch.Tensor:
    """softmin weights of an MPPI update, computed redundantly on every rank from the gathered costs."""
    return torch.softmax(-(costs - costs.min()) / lam, dim=0)
