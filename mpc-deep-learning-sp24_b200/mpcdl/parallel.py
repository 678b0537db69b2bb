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


class HostPipeline:
    """Host-buffer driver of the rollout: copy-in, kernel and copy-out of consecutive batches overlap.

    A planner that keeps its control samples in host memory pays H2D(init, U) + kernel + D2H(result) per call;
    issued on one stream they serialise (C2: 0.42 ms per call against a 0.28 ms kernel).  The pipeline gives every
    in-flight batch its own device input AND result buffers and runs the three stages on three streams (copy-in, the
    caller's current stream for the kernel, copy-out), so batch i+1 is uploaded and batch i-1 downloaded while batch
    i computes.  Every batch still moves all of its bytes; only the waiting is overlapped.

        pipe = HostPipeline(lambda init, U, out: model.rollout(init, U, stats, out=(out, None))[0], depth=2)
        pipe.submit((init_pinned, U_pinned), out_pinned)      # returns at once
        pipe.drain()                                          # current stream (and host, if sync=True) waits

    ``fn(*device_inputs, out) -> Tensor`` runs on the current stream and writes its result into ``out`` (a device
    tensor shaped like the host result, owned by the pipeline slot: nothing is allocated per batch -- with results
    from the caching allocator and ``record_stream`` the first batches of a run each paid a cudaMalloc, which is what
    made the round-1 figure swing 0.29 .. 0.89 ms from box to box).  Host tensors must be pinned for real overlap."""

    def __init__(self, fn, depth: int = 2, device=None):
        self.fn, self.depth = fn, max(1, int(depth))
        self.dev = torch.device(device if device is not None else ("cuda", torch.cuda.current_device()))
        self.s_in, self.s_out = torch.cuda.Stream(self.dev), torch.cuda.Stream(self.dev)
        self.slots = [dict(bufs=None, res=None, ready=torch.cuda.Event(), used=torch.cuda.Event(), done=torch.cuda.Event(),
                           fresh=True) for _ in range(self.depth)]
        self.k = 0

    def submit(self, host_inputs, host_out):
        slot = self.slots[self.k % self.depth]
        self.k += 1
        cur = torch.cuda.current_stream(self.dev)
        if slot["bufs"] is None or any(b.shape != h.shape for b, h in zip(slot["bufs"], host_inputs)):
            slot["bufs"] = [torch.empty(h.shape, dtype=h.dtype, device=self.dev) for h in host_inputs]
        if slot["res"] is None or slot["res"].shape != host_out.shape:
            slot["res"] = torch.empty(host_out.shape, dtype=host_out.dtype, device=self.dev)
        if slot["fresh"]:
            self.s_in.wait_stream(cur)
        else:
            self.s_in.wait_event(slot["used"])          # the kernel that last read these input buffers has finished
            cur.wait_event(slot["done"])                # ... and the last copy-out of this slot's result buffer
        with torch.cuda.stream(self.s_in):
            for b, h in zip(slot["bufs"], host_inputs):
                b.copy_(h, non_blocking=True)
            slot["ready"].record(self.s_in)
        cur.wait_event(slot["ready"])
        res = self.fn(*slot["bufs"], slot["res"])
        slot["used"].record(cur)
        self.s_out.wait_event(slot["used"])
        with torch.cuda.stream(self.s_out):
            host_out.copy_(res, non_blocking=True)
            slot["done"].record(self.s_out)
        slot["fresh"] = False
        return res

    def drain(self, sync: bool = False):
        cur = torch.cuda.current_stream(self.dev)
        cur.wait_stream(self.s_out)
        cur.wait_stream(self.s_in)
        if sync:
            cur.synchronize()
