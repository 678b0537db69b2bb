// Microbenchmark: cost of ONE tcgen05.mma.kind::tf32 (M = 128, K = 8, A from TMEM) as a function of N, alone and with the
// other warps of the CTA hammering TMEM with tcgen05.ld / tcgen05.st (what the epilogue warps of the GRU kernel do).
// Answers why the N-split half cells of rollout_rnn_tc.cu were a loss (profiles/r02/nsplit): is the cost per instruction,
// per N, or contention?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../mpc-deep-learning-sp24_b200/csrc mma_rate.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "common.cuh"
#include "umma.cuh"

using namespace mpcdl;
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

// mode 0: MMAs alone; 1: + 8 warps looping tmem_ld8 on other columns; 2: + 8 warps looping tmem_st8 on other columns
template <int N, bool SS>
__global__ void __launch_bounds__(384, 1) rate_kernel(int nmma, int reps, int mode, long long* cyc) {
    extern __shared__ __align__(1024) float sm[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    __shared__ volatile int stop;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 48 * 1024 / 4; i += blockDim.x) sm[i] = 0.f;
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); stop = 0; }
    if (warp == 8) { umma::tmem_alloc(&slot, 512); umma::tmem_relinquish(); }
    umma::fence_proxy_async_smem();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tbase = slot;
    if (warp == 8) {
        if (tid == 256) {
            constexpr uint32_t LBO = N / 8 * 128;
            const uint32_t idesc = umma::idesc_tf32(128, N);
            const uint64_t bd = umma::smem_desc_kmajor(smem_u32(sm), LBO, 128u);
            const uint64_t ad = umma::smem_desc_kmajor(smem_u32(sm) + 32 * 1024, 128 / 8 * 128, 128u);
            uint32_t par = 0;
            long long best = 1ll << 60;
            for (int r = 0; r < reps; ++r) {
                long long t0 = clock64();
                for (int i = 0; i < nmma; ++i) {
                    if (SS) umma::mma_tf32_ss(tbase, ad, bd, idesc, i > 0);
                    else umma::mma_tf32_ts(tbase, tbase + 256 + (i & 7) * 8, bd, idesc, i > 0);
                }
                umma::mma_commit(&bar);
                mbar_wait(&bar, par);
                par ^= 1u;
                long long t1 = clock64();
                if (t1 - t0 < best) best = t1 - t0;
            }
            cyc[blockIdx.x] = best;
            stop = 1;
        }
    } else if (warp < 8 && mode) {
        const uint32_t tl = tbase + (static_cast<uint32_t>((warp & 3) * 32) << 16) + 384 + (warp >> 2) * 64;
        float v[8];
        uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        float acc = 0.f;
        while (!stop) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                if (mode == 1) { umma::tmem_ld8(tl + c * 8, v); umma::wait_ld(); acc += v[0]; }
                else { umma::tmem_st8(tl + c * 8, z); umma::wait_st(); }
            }
        }
        if (acc == 123.f) cyc[0] = 0;
    }
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 8) umma::tmem_dealloc(tbase, 512);
}

// TMEM read / write rate: `nw` warps (lane quarter = warp % 4) each move `iters` x 32 columns of their 32 lanes
// (4 x tcgen05.ld.32x32b.x8 + wait, or .st): bytes per clock per SM.
__global__ void __launch_bounds__(512, 1) ldst_kernel(int iters, int store, long long* cyc) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) { umma::tmem_alloc(&slot, 512); umma::tmem_relinquish(); }
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tl = slot + (static_cast<uint32_t>((warp & 3) * 32) << 16) + (warp >> 2) * 64;
    float v[4][8];
    uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float acc = 0.f;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (store) {
#pragma unroll
            for (int c = 0; c < 4; ++c) umma::tmem_st8(tl + c * 8 + (i & 1) * 32, z);
            umma::wait_st();
        } else {
#pragma unroll
            for (int c = 0; c < 4; ++c) umma::tmem_ld8(tl + c * 8 + (i & 1) * 32, v[c]);
            umma::wait_ld();
            acc += v[0][0] + v[1][1] + v[2][2] + v[3][3];
        }
    }
    __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    if (acc == 123.f) cyc[1] = 0;
    umma::fence_before_sync();
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(slot, 512);
}
static void run_ldst(int nw, int store, long long* dC) {
    const int iters = 2000;
    long long c[148];
    ldst_kernel<<<148, nw * 32>>>(iters, store, dC);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(c, dC, sizeof(c), cudaMemcpyDeviceToHost));
    double bytes = (double)nw * 32 * 32 * 4 * iters;
    printf("tcgen05.%s x8 : %2d warps : %6.1f B / clk / SM  (%.0f cyc per 32-column pass of a warp)\n", store ? "st" : "ld", nw, bytes / c[0],
           (double)c[0] / iters);
}

template <int N, bool SS>
static void run(int mode, long long* dC) {
    auto k = rate_kernel<N, SS>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    long long c[2][148];
    const int n0 = 16, n1 = 144;
    for (int j = 0; j < 2; ++j) {
        k<<<148, 384, 64 * 1024>>>(j ? n1 : n0, 20, mode, dC);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(c[j], dC, sizeof(c[j]), cudaMemcpyDeviceToHost));
    }
    double per = 0;
    for (int i = 0; i < 148; ++i) per += double(c[1][i] - c[0][i]) / (n1 - n0);
    printf("N %3d A_%s mode %d : %6.1f cyc / MMA  (16 MMAs + commit + wait: %lld cyc)\n", N, SS ? "smem" : "tmem", mode, per / 148, c[0][0]);
}

int main() {
    long long* dC;
    CK(cudaMalloc(&dC, 148 * sizeof(long long)));
    for (int mode = 0; mode < 3; ++mode) {
        run<32, false>(mode, dC); run<64, false>(mode, dC); run<96, false>(mode, dC); run<128, false>(mode, dC);
        run<192, false>(mode, dC); run<256, false>(mode, dC);
    }
    run<64, true>(0, dC); run<96, true>(0, dC); run<192, true>(0, dC); run<256, true>(0, dC);
    for (int st = 0; st < 2; ++st)
        for (int nw = 4; nw <= 16; nw *= 2) run_ldst(nw, st, dC);
    printf("done\n");
    return 0;
}
