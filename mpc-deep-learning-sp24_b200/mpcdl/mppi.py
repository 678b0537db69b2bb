"""MPPI planner iteration around the fused rollout.

New relative to the reference, which has no planner: its closest caller is the open-loop feedback of
``run_dynamics_model`` (scripts/test_model.py:67-87) that this batches over N sampled control sequences.  One
iteration is

    u_n,t  = clamp(U[t] + sigma * z_n,t)        z ~ N(0,1), Philox(seed, global trajectory index n, step t), in registers
    cost_n = sum_t (s_t+1 - ref_t)' Q (s_t+1 - ref_t) + u' R u       the fused rollout kernel, cost-only output
    w_n    = exp(-(cost_n - min cost) / lambda)
    U'[t]  = clamp(U[t] + sum_n w_n (u_n,t - U[t]) / sum_n w_n)      mpcdl_mppi_reduce re-generates u_n,t from the keys

so neither the N x H x 2 controls nor the N x H x 6 states ever exist in HBM: per iteration the kernels read the
initial state and the nominal sequence and write N costs.  Sharded over ranks (one process per GPU): every rank
rolls out its slice of the N samples (``traj_offset`` keeps the Philox keys global, so the samples do not depend
on the number of ranks); the collectives are the all-gather of costs (BASELINE.json north_star) -- or, with
``gather_costs=False``, one all-reduce MIN of a scalar -- and one all-reduce SUM of 2H+1 floats.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import parallel


class MPPI:
    def __init__(self, model, stats, n_samples, sigma, u_lo, u_hi, cost, lam=1.0, noise=None, precision=None,
                 seed=0, group=None, gather_costs=True):
        self.model, self.stats, self.n, self.lam = model, stats, int(n_samples), float(lam)
        self.sigma, self.lo, self.hi = tuple(map(float, sigma)), tuple(map(float, u_lo)), tuple(map(float, u_hi))
        self.cost, self.noise, self.precision, self.seed, self.group = cost, noise, precision, int(seed), group
        self.gather = gather_costs
        self.iteration = 0

    def _ctrl(self):
        return list(self.sigma) + list(self.lo) + list(self.hi)

    def step(self, init, U):
        """init: ONE state [6] (MLP) or window [T,8] (GRU/LSTM), U: nominal controls [H,2], both on the GPU.
        Returns (U_new [H,2], info) with info = dict(costs=local costs [n_local], cmin, wsum, lo, hi)."""
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        lo, hi = parallel.shard_range(self.n, rank, world)
        seed = self.seed + self.iteration
        self.iteration += 1
        sample = dict(n=hi - lo, sigma=self.sigma, lo=self.lo, hi=self.hi, seed=seed)
        U = U.contiguous()
        out = self.model.rollout(init.contiguous(), U, self.stats, noise=self.noise, seed=seed, traj_offset=lo, cost=self.cost,
                                 return_states=False, precision=self.precision, sample=sample, shared_init=True,
                                 stateful=False)
        costs = out[1]
        cmin, costs_all = None, None
        if world > 1:
            if self.gather:
                costs_all = parallel.gather_costs(costs, self.n, self.group)
                cmin = torch.ops.mpcdl.mppi_min(costs_all)
            else:
                cmin = torch.ops.mpcdl.mppi_min(costs)
                dist.all_reduce(cmin, op=dist.ReduceOp.MIN, group=self.group)
        sums = torch.ops.mpcdl.mppi_reduce(costs, U, self._ctrl(), seed, lo, self.lam, cmin)
        H = U.shape[0]
        if world > 1:
            dist.all_reduce(sums[:2 * H + 1], op=dist.ReduceOp.SUM, group=self.group)
        U_new = apply_update(U, sums, self.lo, self.hi)
        return U_new, dict(costs=costs, costs_all=costs_all, cmin=sums[2 * H + 1], wsum=sums[2 * H], lo=lo, hi=hi)


def apply_update(U, sums, u_lo, u_hi):
    """U' = clamp(U + S / wsum): 2H floats of host-launched tensor arithmetic on the reduced sums."""
    H = U.shape[0]
    upd = U + sums[:2 * H].view(H, 2) / sums[2 * H]
    lo = torch.tensor(u_lo, device=U.device, dtype=U.dtype)
    hi = torch.tensor(u_hi, device=U.device, dtype=U.dtype)
    return torch.minimum(torch.maximum(upd, lo), hi)
