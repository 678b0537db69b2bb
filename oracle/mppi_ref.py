"""CPU restatement of the MPPI pieces around the rollout (TEST INFRASTRUCTURE; see oracle/__init__).

PARITY UNPINNED for this file: the reference has no planner, no control sampler and no softmin update
(SURVEY.md D1, 8f-1), so there is nothing of the reference to pin these against.  What is restated here is the
published algorithm the kernels implement -- Philox4x32-10 (Salmon et al., SC'11; constants 0xD2511F53 /
0xCD9E8D57, key increments 0x9E3779B9 / 0xBB67AE85), Box-Muller, and the MPPI information-theoretic update
(Williams et al., 2017) -- and it is checked against the Random123 known-answer vectors in tests/test_oracle_golden.py.
The rollout inside an MPPI iteration IS pinned (oracle/rollout_ref.py)."""
from __future__ import annotations

import numpy as np

M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85


def philox4x32(ctr, key, rounds=10):
    """ctr [...,4] uint32, key [...,2] uint32 -> [...,4] uint32."""
    c = [np.asarray(ctr[..., i], dtype=np.uint64) for i in range(4)]
    k = [np.asarray(key[..., i], dtype=np.uint64) for i in range(2)]
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(rounds):
        p0, p1 = np.uint64(M0) * c[0], np.uint64(M1) * c[2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c = [hi1 ^ c[1] ^ k[0], lo1, hi0 ^ c[3] ^ k[1], lo0]
        k = [(k[0] + np.uint64(W0)) & mask, (k[1] + np.uint64(W1)) & mask]
    return np.stack(c, -1).astype(np.uint32)


def box_muller(a, b):
    """two uint32 -> two N(0,1) float32 (csrc/common.cuh box_muller: u = (x + 0.5) * 2^-32)."""
    u1 = (a.astype(np.float32) + np.float32(0.5)) * np.float32(2.3283064365386963e-10)
    u2 = (b.astype(np.float32) + np.float32(0.5)) * np.float32(2.3283064365386963e-10)
    r = np.sqrt(np.float32(-2.0) * np.log(u1), dtype=np.float32)
    ang = np.float32(6.283185307179586) * u2
    return (r * np.cos(ang)).astype(np.float32), (r * np.sin(ang)).astype(np.float32)


def control_normals(seed, traj, H):
    """z [len(traj), H, 2]: counter = (traj lo, traj hi, step, 2), key = (seed lo, seed hi)."""
    traj = np.asarray(traj, dtype=np.uint64)
    n = len(traj)
    ctr = np.zeros((n, H, 4), dtype=np.uint32)
    ctr[..., 0] = (traj & np.uint64(0xFFFFFFFF)).astype(np.uint32)[:, None]
    ctr[..., 1] = (traj >> np.uint64(32)).astype(np.uint32)[:, None]
    ctr[..., 2] = np.arange(H, dtype=np.uint32)[None]
    ctr[..., 3] = 2
    key = np.zeros((n, H, 2), dtype=np.uint32)
    key[..., 0], key[..., 1] = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    r = philox4x32(ctr, key)
    z0, z1 = box_muller(r[..., 0], r[..., 1])
    return np.stack([z0, z1], -1)


def sample_controls(U, seed, traj, sigma, lo, hi):
    """u [n,H,2] = clamp(fma(sigma, z, U), lo, hi) in float32."""
    z = control_normals(seed, traj, U.shape[0]).astype(np.float64)
    u = (np.asarray(sigma, np.float64) * z + U.astype(np.float64)[None]).astype(np.float32)   # fma: one rounding
    return np.minimum(np.maximum(u, np.asarray(lo, np.float32)), np.asarray(hi, np.float32))


def mppi_update(U, u, costs, lam, lo, hi):
    """softmin-weighted update in float64: returns (U_new, S [H,2], wsum, cmin)."""
    c = costs.astype(np.float64)
    cmin = c.min()
    w = np.exp(-(c - cmin) / lam)
    S = (w[:, None, None] * (u.astype(np.float64) - U.astype(np.float64)[None])).sum(0)
    un = np.minimum(np.maximum(U + S / w.sum(), np.asarray(lo, np.float64)), np.asarray(hi, np.float64))
    return un.astype(np.float32), S, w.sum(), cmin
