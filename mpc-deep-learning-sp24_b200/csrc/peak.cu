// FP32 FFMA peak micro-kernel: 64 independent register accumulators per thread, 8+8 operands,
// no memory traffic in the loop.  bench.py times it to obtain the FP32 roofline denominator on
// the box it runs on (MEASURED_PEAKS.json carries HBM and BF16 peaks only).
#include "launch.h"

namespace mpcdl {

__global__ void __launch_bounds__(256) ffma_peak_kernel(float* out, int iters, float seed) {
    float a[8], w[8], acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = seed + threadIdx.x * 1e-6f + i * 1e-3f;
        w[i] = seed * 0.5f + i * 1e-3f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) s += acc[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// the same 8 x 8 accumulator tile issued as packed FFMA2 (fma.rn.f32x2: two FMAs per lane and issue slot)
__global__ void __launch_bounds__(256) ffma2_peak_kernel(float* out, int iters, float seed) {
    float a[8], w[8];
    uint64_t acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = seed + threadIdx.x * 1e-6f + i * 1e-3f;
        w[i] = seed * 0.5f + i * 1e-3f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = pack2(0.f, 0.f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) ffma2(acc[i][j], pack2(a[i], a[i]), pack2(w[2 * j], w[2 * j + 1]));
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float lo, hi;
            unpack2(acc[i][j], lo, hi);
            s += lo + hi;
        }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

cudaError_t launch_ffma_peak(float* out, int iters, int grid, bool packed, cudaStream_t st) {
    if (packed) ffma2_peak_kernel<<<grid, 256, 0, st>>>(out, iters, 1.0f);
    else ffma_peak_kernel<<<grid, 256, 0, st>>>(out, iters, 1.0f);
    return cudaGetLastError();
}

}  // namespace mpcdl

// Test aid (mpcdl_debug_poison_smem): one CTA per SM claims the maximum dynamic shared memory and fills it with NaN bit
// patterns.  Launched before a rollout in the determinism tests: a kernel that reads shared memory it has not written
// then produces NaN / changed bits instead of silently reusing the previous launch's (usually identical) contents.
namespace mpcdl {

__global__ void __launch_bounds__(256) poison_smem_kernel() {
    extern __shared__ volatile unsigned psm[];                  // volatile: the stores are the whole point
    for (int i = threadIdx.x; i < SMEM_LIMIT / 4; i += blockDim.x) psm[i] = 0x7FC00000u | (unsigned)(i & 0xFFFF);
}

cudaError_t launch_poison_smem(int num_sms, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(poison_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT);
    if (e != cudaSuccess) return e;
    poison_smem_kernel<<<num_sms, 256, SMEM_LIMIT, st>>>();
    return cudaGetLastError();
}

}  // namespace mpcdl
