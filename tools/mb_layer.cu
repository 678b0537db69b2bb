// Micro-benchmark of ONE dense layer step of the small-N MLP rollout tile (development aid):
//   out[j][n] = relu(b[j] + sum_k W[k][j] * in[k][n]),  K = HID = 64, TN trajectories per CTA,
// repeated `iters` times with ping-pong activations and a CTA barrier per layer, exactly the
// dependency structure of the rollout kernel.  Prints cycles per layer for several mappings.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mb_layer tools/mb_layer.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

constexpr int HID = 64;

template <int RT, int RJ>
__device__ __forceinline__ void fma_tile(float (&acc)[RT][RJ], const float (&a)[RT], const float (&w)[RJ]) {
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int r = 0; r < RJ; ++r) acc[i][r] = fmaf(a[i], w[r], acc[i][r]);
}
template <int N>
__device__ __forceinline__ void ldv(float (&d)[N], const float* p) {
#pragma unroll
    for (int c = 0; c < N / 4; ++c) {
        float4 v = *reinterpret_cast<const float4*>(p + 4 * c);
        d[4 * c] = v.x; d[4 * c + 1] = v.y; d[4 * c + 2] = v.z; d[4 * c + 3] = v.w;
    }
}

template <int RT>
__device__ __forceinline__ void lda(float (&d)[RT], const float* p, int hs) {
#pragma unroll
    for (int c = 0; c < RT / 4; ++c) {
        float4 v = *reinterpret_cast<const float4*>(p + c * hs);
        d[4 * c] = v.x; d[4 * c + 1] = v.y; d[4 * c + 2] = v.z; d[4 * c + 3] = v.w;
    }
}
// MODE 0: plain loop, #pragma unroll UNR.   MODE 1: explicit register double buffer of PF k-steps.
template <int RT, int RJ, int KSPLIT, int MODE, int UNR>
__global__ void __launch_bounds__(512) k_layer(float* out, int iters, int TN, long long* cyc) {
    extern __shared__ __align__(16) float sm[];
    float* W = sm;                       // [HID][HID] + bias[HID]
    float* act0 = W + HID * HID + HID;   // [HID][TN]
    float* act1 = act0 + HID * TN;
    float* red = act1 + HID * TN;        // [KSPLIT-1][HID][TN] partial sums
    for (int i = threadIdx.x; i < HID * HID + HID; i += blockDim.x) W[i] = ((i * 37) % 101 - 50) * 2e-3f;
    for (int i = threadIdx.x; i < 2 * HID * TN; i += blockDim.x) act0[i] = (i % 13) * 0.01f;
    __syncthreads();
    const int n_tg = TN / RT, n_jg = HID / RJ;
    const int per = n_tg * n_jg;                       // threads per K-slice
    const int ks = threadIdx.x / per, t = threadIdx.x % per;
    const int tg = t % n_tg, jg = t / n_tg;
    const bool active = threadIdx.x < per * KSPLIT;
    constexpr int KS = HID / KSPLIT;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const float* in = (it & 1) ? act1 : act0;
        float* o = (it & 1) ? act0 : act1;
        if (active) {
            float acc[RT][RJ];
            {
                float b[RJ];
                ldv<RJ>(b, W + HID * HID + jg * RJ);
#pragma unroll
                for (int i = 0; i < RT; ++i)
#pragma unroll
                    for (int r = 0; r < RJ; ++r) acc[i][r] = ks == 0 ? b[r] : 0.f;
            }
            const int hs = 4 * n_tg;
            const float* ap = in + (ks * KS) * TN + tg * 4;
            const float* wp = W + (ks * KS) * HID + jg * RJ;
            if (MODE == 0) {
#pragma unroll UNR
                for (int k = 0; k < KS; ++k) {
                    float a[RT], w[RJ];
                    lda<RT>(a, ap + k * TN, hs);
                    ldv<RJ>(w, wp + k * HID);
                    fma_tile<RT, RJ>(acc, a, w);
                }
            } else {
                constexpr int PF = UNR;                 // k-steps per register stage
                float a0[PF][RT], w0[PF][RJ], a1[PF][RT], w1[PF][RJ];
#pragma unroll
                for (int u = 0; u < PF; ++u) { lda<RT>(a0[u], ap + u * TN, hs); ldv<RJ>(w0[u], wp + u * HID); }
#pragma unroll 1
                for (int k = 0; k < KS; k += 2 * PF) {
#pragma unroll
                    for (int u = 0; u < PF; ++u) { lda<RT>(a1[u], ap + (k + PF + u) * TN, hs); ldv<RJ>(w1[u], wp + (k + PF + u) * HID); }
#pragma unroll
                    for (int u = 0; u < PF; ++u) fma_tile<RT, RJ>(acc, a0[u], w0[u]);
                    if (k + 2 * PF < KS) {
#pragma unroll
                        for (int u = 0; u < PF; ++u) { lda<RT>(a0[u], ap + (k + 2 * PF + u) * TN, hs); ldv<RJ>(w0[u], wp + (k + 2 * PF + u) * HID); }
                    }
#pragma unroll
                    for (int u = 0; u < PF; ++u) fma_tile<RT, RJ>(acc, a1[u], w1[u]);
                }
            }
            if (KSPLIT > 1) {
                if (ks > 0) {
#pragma unroll
                    for (int r = 0; r < RJ; ++r) {
                        float* d = red + ((ks - 1) * HID + jg * RJ + r) * TN + tg * 4;
#pragma unroll
                        for (int c = 0; c < RT / 4; ++c)
                            *reinterpret_cast<float4*>(d + c * hs) = make_float4(acc[4 * c][r], acc[4 * c + 1][r], acc[4 * c + 2][r], acc[4 * c + 3][r]);
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"r"(per * KSPLIT) : "memory");
                if (ks == 0) {
#pragma unroll
                    for (int s = 0; s < KSPLIT - 1; ++s)
#pragma unroll
                        for (int r = 0; r < RJ; ++r) {
                            float p[RT];
                            lda<RT>(p, red + (s * HID + jg * RJ + r) * TN + tg * 4, hs);
#pragma unroll
                            for (int i = 0; i < RT; ++i) acc[i][r] += p[i];
                        }
                }
            }
            if (ks == 0) {
#pragma unroll
                for (int r = 0; r < RJ; ++r) {
                    float* d = o + (jg * RJ + r) * TN + tg * 4;
#pragma unroll
                    for (int c = 0; c < RT / 4; ++c)
                        *reinterpret_cast<float4*>(d + c * hs) = make_float4(fmaxf(acc[4 * c][r], 0.f), fmaxf(acc[4 * c + 1][r], 0.f),
                                                                            fmaxf(acc[4 * c + 2][r], 0.f), fmaxf(acc[4 * c + 3][r], 0.f));
                }
            }
        }
        __syncthreads();
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    if (threadIdx.x < TN) out[blockIdx.x * TN + threadIdx.x] = act0[threadIdx.x] + act1[threadIdx.x];
}

static float* d_out; static long long* d_cyc; static int n_sm;

template <int RT, int RJ, int KSPLIT, int MODE, int UNR> static void run(int TN) {
    int iters = 2000;
    int threads = (TN / RT) * (HID / RJ) * KSPLIT;
    threads = (threads + 31) / 32 * 32;
    size_t smem = (size_t)(HID * HID + HID + 2 * HID * TN + (KSPLIT - 1) * HID * TN) * 4;
    auto k = k_layer<RT, RJ, KSPLIT, MODE, UNR>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<n_sm, threads, smem>>>(d_out, iters, TN, d_cyc);
    CK(cudaDeviceSynchronize());
    k<<<n_sm, threads, smem>>>(d_out, iters, TN, d_cyc);
    CK(cudaDeviceSynchronize());
    static long long h[1024];
    CK(cudaMemcpy(h, d_cyc, n_sm * sizeof(long long), cudaMemcpyDeviceToHost));
    long long m = 0; for (int i = 0; i < n_sm; ++i) if (h[i] > m) m = h[i];
    double cpl = (double)m / iters;
    printf("TN %3d tile %dx%d ksplit %d mode %d unr %2d thr %3d : %7.1f cyc/layer  %6.1f FMA/clk/SM  (floor %5.0f)\n", TN, RT, RJ, KSPLIT, MODE,
           UNR, threads, cpl, (double)TN * HID * HID / cpl, (double)TN * HID * HID / 128.0);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0)); n_sm = p.multiProcessorCount;
    CK(cudaMalloc(&d_out, 16 << 20)); CK(cudaMalloc(&d_cyc, 1024 * sizeof(long long)));
    // small tile (C2: 4096 trajectories -> 28..32 per SM)
    run<4, 4, 1, 0, 4>(32);  run<4, 4, 1, 0, 8>(32);  run<4, 4, 1, 0, 16>(32); run<4, 4, 1, 0, 64>(32);
    run<4, 4, 1, 1, 2>(32);  run<4, 4, 1, 1, 4>(32);  run<4, 4, 1, 1, 8>(32);
    run<4, 4, 2, 0, 8>(32);  run<4, 4, 2, 0, 32>(32); run<4, 4, 2, 1, 4>(32);
    run<4, 4, 4, 0, 16>(32); run<4, 4, 4, 1, 4>(32);
    run<4, 8, 1, 0, 8>(32);  run<4, 8, 2, 0, 8>(32);  run<4, 8, 4, 0, 16>(32); run<4, 8, 4, 1, 4>(32);
    run<8, 8, 4, 0, 16>(32); run<8, 8, 4, 1, 2>(32);
    run<4, 4, 1, 0, 8>(28);  run<4, 4, 2, 0, 8>(28);
    // larger tiles
    run<4, 4, 1, 0, 8>(64);  run<4, 8, 1, 0, 8>(64);  run<8, 4, 1, 0, 8>(64);  run<8, 8, 1, 0, 8>(64); run<8, 8, 2, 0, 8>(64);
    run<8, 8, 1, 0, 8>(128); run<8, 8, 1, 1, 2>(128); run<8, 8, 2, 0, 8>(128); run<8, 4, 1, 0, 8>(128);
    run<8, 8, 1, 0, 8>(256); run<8, 8, 1, 1, 2>(256); run<8, 8, 1, 0, 16>(256); run<8, 4, 1, 0, 8>(256);
    printf("done\n");
    return 0;
}
