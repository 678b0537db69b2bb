// Shared device helpers for the fused rollout kernels (sm_100a only).
//   * mbarrier + cp.async.bulk (TMA 1-D bulk copy, SASS UBLKCP) wrappers for the weight ring
//   * activation functions built on MUFU.EX2 / MUFU.RCP with ~1e-7 absolute error
//   * Philox4x32-10 + Box-Muller for in-kernel residual-noise draws (shard invariant)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mpcdl {

// ------------------------------------------------------------------ mbarrier / bulk copy
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded spin: a protocol bug traps instead of hanging the GPU box.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 26)) __trap();
    }
}
// global -> shared bulk copy completing on an mbarrier (bytes % 16 == 0, both 16 B aligned)
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------ packed FP32 FMA (Blackwell FFMA2)
// fma.rn.f32x2 does two IEEE FP32 fused multiply-adds per lane in ONE issue slot (SASS FFMA2): each half is the same
// round-to-nearest fmaf the scalar instruction computes, so results are bit-identical; what changes is that the FMA pipe,
// not the issue slot, bounds a register-tile GEMM (ncu r02: rollout_mlp2 issued 78 % of the cycles with 67 % FMA).
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void ffma2(uint64_t& acc, uint64_t a, uint64_t b) {      // acc = a * b + acc, per half
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}

// ------------------------------------------------------------------ activations
// sigmoid(x) = 1/(1+2^(-x*log2e)); MUFU.EX2 (2 ulp) + MUFU.RCP (1 ulp): |err| < 1.5e-7 absolute.
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sigmoid_f(float x) {
    return rcp_approx(1.0f + ex2_approx(-1.4426950408889634f * x));
}
// tanh(x) = 1 - 2/(1+e^{2x}); absolute error ~1e-7 (cancellation near 0 stays absolute).
__device__ __forceinline__ float tanh_f(float x) {
    float ax = fabsf(x);
    if (ax < 0.04f) {                       // odd Taylor keeps RELATIVE accuracy for tiny inputs
        float x2 = x * x;
        return x * fmaf(x2, fmaf(x2, 0.13333333f, -0.33333333f), 1.0f);
    }
    float e = ex2_approx(2.8853900817779268f * ax);           // e^{2|x|}
    float r = 1.0f - 2.0f * rcp_approx(1.0f + e);
    return copysignf(r, x);
}
// F.softplus(beta=1, threshold=20)   (reference: model_scripts/noise_model.py:39)
__device__ __forceinline__ float softplus_f(float x) { return x > 20.0f ? x : log1pf(expf(x)); }

enum { ACT_IDENTITY = 0, ACT_RELU = 1, ACT_TANH = 2, ACT_SIGMOID = 3 };
__device__ __forceinline__ float apply_act(int act, float x) {
    switch (act) {
        case ACT_RELU: return fmaxf(x, 0.0f);
        case ACT_TANH: return tanh_f(x);
        case ACT_SIGMOID: return sigmoid_f(x);
        default: return x;
    }
}

// ------------------------------------------------------------------ Philox4x32-10
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u;
        key.y += 0xBB67AE85u;
    }
    return ctr;
}
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& z0, float& z1) {
    float u1 = (static_cast<float>(a) + 0.5f) * 2.3283064365386963e-10f;   // (0,1)
    float u2 = (static_cast<float>(b) + 0.5f) * 2.3283064365386963e-10f;
    float r = sqrtf(-2.0f * logf(u1));
    float s, c;
    sincosf(6.283185307179586f * u2, &s, &c);
    z0 = r * c;
    z1 = r * s;
}
// six N(0,1) draws keyed by (seed, global trajectory index, step): independent of tiling/sharding
__device__ __forceinline__ void philox_normal6(uint64_t seed, uint64_t traj, uint32_t step, float (&z)[6]) {
    uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    uint4 c0 = philox4x32(make_uint4(static_cast<uint32_t>(traj), static_cast<uint32_t>(traj >> 32), step, 0u), key);
    uint4 c1 = philox4x32(make_uint4(static_cast<uint32_t>(traj), static_cast<uint32_t>(traj >> 32), step, 1u), key);
    box_muller(c0.x, c0.y, z[0], z[1]);
    box_muller(c0.z, c0.w, z[2], z[3]);
    box_muller(c1.x, c1.y, z[4], z[5]);
}

// two N(0,1) draws for the sampled control of (trajectory, step): counter word 3 = 2 keeps the stream disjoint
// from the residual-noise draws (words 0 and 1) under the same key
__device__ __forceinline__ void philox_normal2(uint64_t seed, uint64_t traj, uint32_t step, float& z0, float& z1) {
    uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    uint4 c = philox4x32(make_uint4(static_cast<uint32_t>(traj), static_cast<uint32_t>(traj >> 32), step, 2u), key);
    box_muller(c.x, c.y, z0, z1);
}

}  // namespace mpcdl
