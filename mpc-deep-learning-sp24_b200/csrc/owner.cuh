// Per-trajectory ("owner" thread) work that brackets the network evaluation of every step:
// residual-noise history, de-normalisation, noise injection, cost accumulation, state store.
// Reference semantics: oracle/rollout_ref.py rollout_mlp / rollout_rnn (pinned against the
// reference modules); noise history = model_scripts/noise_model.py:95-103.
#pragma once
#include "params.h"

namespace mpcdl {

// q[i] = state[Q_PERM[i]]  (model_scripts/noise_model.py:108, utils/data_utils.py:47-49)
__device__ __forceinline__ int q_perm(int i) { return i == 0 ? 4 : i == 1 ? 5 : i == 2 ? 3 : i - 3; }

// push (state s, control u) into the noise input column of trajectory `pos`:
// rows [0,6h) = q history oldest first, rows [6h,8h) = u history; first push left-pads by repetition.
__device__ __forceinline__ void noise_push(float* nin, int TN, int pos, int hist, bool first, const float (&s)[6],
                                           const float (&u)[2]) {
    float q[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) q[i] = s[q_perm(i)];
    if (first) {
        for (int e = 0; e < hist; ++e) {
#pragma unroll
            for (int i = 0; i < 6; ++i) nin[(e * 6 + i) * TN + pos] = q[i];
            nin[(6 * hist + e * 2) * TN + pos] = u[0];
            nin[(6 * hist + e * 2 + 1) * TN + pos] = u[1];
        }
        return;
    }
    for (int e = 0; e + 1 < hist; ++e) {
#pragma unroll
        for (int i = 0; i < 6; ++i) nin[(e * 6 + i) * TN + pos] = nin[((e + 1) * 6 + i) * TN + pos];
        nin[(6 * hist + e * 2) * TN + pos] = nin[(6 * hist + (e + 1) * 2) * TN + pos];
        nin[(6 * hist + e * 2 + 1) * TN + pos] = nin[(6 * hist + (e + 1) * 2 + 1) * TN + pos];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) nin[((hist - 1) * 6 + i) * TN + pos] = q[i];
    nin[(6 * hist + (hist - 1) * 2) * TN + pos] = u[0];
    nin[(6 * hist + (hist - 1) * 2 + 1) * TN + pos] = u[1];
}

// control of trajectory n at step t: read from U [N,H,2], or (MPPI) sampled around the nominal sequence U [H,2]
// (the sampler is kept out of line so the plain [N,H,2] path keeps its registers and instruction footprint)
static __device__ __noinline__ float2 sample_control(const float* U, unsigned long long seed, unsigned long long traj, int t,
                                              float s0, float s1, float lo0, float lo1, float hi0, float hi1) {
    const float2 ub = __ldg(reinterpret_cast<const float2*>(U) + t);
    float z0, z1;
    philox_normal2(seed, traj, static_cast<uint32_t>(t), z0, z1);
    return make_float2(fminf(fmaxf(fmaf(s0, z0, ub.x), lo0), hi0), fminf(fmaxf(fmaf(s1, z1, ub.y), lo1), hi1));
}
template <class IO>
__device__ __forceinline__ float2 load_control(const IO& io, long long n, int t) {
    if (!io.sample) return __ldg(reinterpret_cast<const float2*>(io.U) + n * io.H + t);
    return sample_control(io.U, io.ctrl_seed, static_cast<unsigned long long>(io.traj_offset + n), t, io.sig[0], io.sig[1],
                          io.ulo[0], io.ulo[1], io.uhi[0], io.uhi[1]);
}

// y (normalised network output, column `pos` of ybuf) -> next raw state; optional noise, cost, store.
template <class P>
__device__ __forceinline__ void owner_finish(const P& p, const float* norm_l /* l_mean[6] l_std[6] */,
                                             const float* ybuf, const float* nzb, int TN, int pos, long long n,
                                             bool valid, int t, const float (&u)[2], float (&nxt)[6], float& cost) {
#pragma unroll
    for (int d = 0; d < 6; ++d)   // y * l_std + l_mean with torch's two roundings
        nxt[d] = __fadd_rn(__fmul_rn(ybuf[d * TN + pos], norm_l[6 + d]), norm_l[d]);
    if (p.nz.enabled) {
        float e[6];
        if (p.io.eps) {
            const float* ep = p.io.eps + (n * p.io.H + t) * 6;
#pragma unroll
            for (int i = 0; i < 6; ++i) e[i] = __ldg(ep + i);
        } else {
            philox_normal6(p.io.seed, static_cast<uint64_t>(p.io.traj_offset + n), static_cast<uint32_t>(t), e);
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) {   // mu + eps * softplus(p)   (noise_model.py:39-40), added in q order
            float mu = nzb[i * TN + pos], sc = softplus_f(nzb[(6 + i) * TN + pos]);
            nxt[q_perm(i)] = __fadd_rn(nxt[q_perm(i)], __fadd_rn(mu, __fmul_rn(e[i], sc)));
        }
    }
    if (p.io.costs) {
        float cq = 0.f, cr = 0.f;
        if (p.io.cost_ref) {
#pragma unroll
            for (int d = 0; d < 6; ++d) {
                float df = nxt[d] - __ldg(p.io.cost_ref + t * 6 + d);
                cq += df * df * p.io.qw[d];
            }
        }
        cr = u[0] * u[0] * p.io.rw[0] + u[1] * u[1] * p.io.rw[1];
        cost += cq + cr;
    }
    if (valid && p.io.states) {
        float2* o = reinterpret_cast<float2*>(p.io.states + (n * p.io.H + t) * 6);
        o[0] = make_float2(nxt[0], nxt[1]);
        o[1] = make_float2(nxt[2], nxt[3]);
        o[2] = make_float2(nxt[4], nxt[5]);
    }
}

}  // namespace mpcdl
