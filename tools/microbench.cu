// Micro-benchmarks that size the FP32 rollout micro-kernel on B200 (sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
// Prints one line per experiment: name, warps/SM, FMA lanes per clock per SM, TFLOP/s.
// The measured FFMA peak is the FP32 roofline denominator used by bench.py.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

// ---------------------------------------------------------------- register-only FMA peak
template <int RT, int RC, bool PACKED>
__global__ void __launch_bounds__(256) k_fma_reg(float* out, int iters, float seed) {
    float a[RT], w[RC];
    for (int i = 0; i < RT; ++i) a[i] = seed + threadIdx.x * 1e-6f + i;
    for (int j = 0; j < RC; ++j) w[j] = seed * 0.5f + j * 1e-3f;
    float2 acc[RT][RC / 2];
    for (int i = 0; i < RT; ++i) for (int j = 0; j < RC / 2; ++j) acc[i][j] = make_float2(0.f, 0.f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < RT; ++i) {
#pragma unroll
            for (int j = 0; j < RC / 2; ++j) {
                if (PACKED) {
                    acc[i][j] = __ffma2_rn(make_float2(a[i], a[i]), make_float2(w[2 * j], w[2 * j + 1]), acc[i][j]);
                } else {
                    acc[i][j].x = fmaf(a[i], w[2 * j], acc[i][j].x);
                    acc[i][j].y = fmaf(a[i], w[2 * j + 1], acc[i][j].y);
                }
            }
        }
        // keep operands changing so nothing is hoisted/folded
#pragma unroll
        for (int i = 0; i < RT; ++i) a[i] = a[i] * 1.0000001f;
    }
    float s = 0.f;
    for (int i = 0; i < RT; ++i) for (int j = 0; j < RC / 2; ++j) s += acc[i][j].x + acc[i][j].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// ---------------------------------------------------------------- LDS.128 broadcast patterns
// pattern: byte offset of lane l = (l / div) * stride  (16-byte aligned)
__global__ void __launch_bounds__(256) k_lds(float* out, int iters, int div, int stride, long long* cyc) {
    extern __shared__ float4 sm4[];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm4[i] = make_float4(i, 1.f, 2.f, 3.f);
    __syncthreads();
    int lane = threadIdx.x & 31;
    int base = ((lane / div) * stride) / 16 + (threadIdx.x >> 5) * 64;
    float4 s = make_float4(0, 0, 0, 0);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            float4 v = sm4[(base + u * 128 + it) & 4095];
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s.x + s.y + s.z + s.w;
}

// ---------------------------------------------------------------- smem-fed outer-product tile
// warp = TG trajectory groups x CG column groups (TG*CG == 32); thread tile RT x RC.
// A[k][TG*RT] activations, W[k][CG*RC] weights, K = 64 rows, looped `iters` times.
template <int RT, int RC, int TG, bool PACKED>
__global__ void __launch_bounds__(256) k_tile(float* out, int iters, long long* cyc) {
    constexpr int CG = 32 / TG, K = 64, LDA = TG * RT, LDW = CG * RC;
    extern __shared__ float sm[];
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int nw = blockDim.x >> 5;
    float* A = sm + warp * K * LDA;          // warp-private activations
    float* W = sm + nw * K * LDA;            // shared weights
    for (int i = threadIdx.x; i < nw * K * LDA + K * LDW; i += blockDim.x) sm[i] = (i % 97) * 1e-3f;
    __syncthreads();
    int tg = lane / CG, cg = lane % CG;
    float2 acc[RT][RC / 2];
    for (int i = 0; i < RT; ++i) for (int j = 0; j < RC / 2; ++j) acc[i][j] = make_float2(0.f, 0.f);
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll 8
        for (int k = 0; k < K; ++k) {
            float a[RT], w[RC];
#pragma unroll
            for (int i = 0; i < RT; i += 4) *(float4*)&a[i] = *(const float4*)&A[k * LDA + tg * RT + i];
#pragma unroll
            for (int j = 0; j < RC; j += 4) *(float4*)&w[j] = *(const float4*)&W[k * LDW + cg * RC + j];
#pragma unroll
            for (int i = 0; i < RT; ++i)
#pragma unroll
                for (int j = 0; j < RC / 2; ++j) {
                    if (PACKED) acc[i][j] = __ffma2_rn(make_float2(a[i], a[i]), make_float2(w[2 * j], w[2 * j + 1]), acc[i][j]);
                    else { acc[i][j].x = fmaf(a[i], w[2 * j], acc[i][j].x); acc[i][j].y = fmaf(a[i], w[2 * j + 1], acc[i][j].y); }
                }
        }
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    float s = 0.f;
    for (int i = 0; i < RT; ++i) for (int j = 0; j < RC / 2; ++j) s += acc[i][j].x + acc[i][j].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

static float* d_out; static long long* d_cyc; static int n_sm;

template <typename F> static float time_ms(F launch) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(); CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) {
        CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}
static double max_cyc(int n) {
    static long long h[4096]; CK(cudaMemcpy(h, d_cyc, n * sizeof(long long), cudaMemcpyDeviceToHost));
    long long m = 0; for (int i = 0; i < n; ++i) if (h[i] > m) m = h[i]; return (double)m;
}

template <int RT, int RC, bool P> static void run_reg(int threads, int bps) {
    int iters = 20000; int grid = n_sm * bps;
    float ms = time_ms([&] { k_fma_reg<RT, RC, P><<<grid, threads>>>(d_out, iters, 1.0f); });
    double fma = (double)grid * threads * iters * RT * RC;
    printf("reg   %s tile %dx%-2d thr %3d blk/SM %d : %7.2f TFLOP/s  (%.3f ms)\n", P ? "FFMA2" : "FFMA ", RT, RC, threads, bps, 2 * fma / ms * 1e-9, ms);
}
template <int RT, int RC, int TG, bool P> static void run_tile(int threads, int bps) {
    constexpr int CG = 32 / TG, K = 64;
    int iters = 400; int grid = n_sm * bps; int nw = threads / 32;
    size_t smem = (size_t)(nw * K * TG * RT + K * CG * RC) * 4;
    CK(cudaFuncSetAttribute(k_tile<RT, RC, TG, P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    float ms = time_ms([&] { k_tile<RT, RC, TG, P><<<grid, threads, smem>>>(d_out, iters, d_cyc); });
    double fma = (double)grid * threads * iters * K * RT * RC;
    double cyc = max_cyc(grid);
    printf("tile  %s tile %dx%-2d warp %dx%-2d thr %3d blk/SM %d smem %3zuKB : %7.2f TFLOP/s  %6.1f FMA/clk/SM (%.3f ms)\n",
           P ? "FFMA2" : "FFMA ", RT, RC, TG, CG, threads, bps, smem >> 10, 2 * fma / ms * 1e-9,
           (double)bps * threads * iters * K * RT * RC / cyc, ms);
}
static void run_lds(int div, int stride, int threads, const char* what) {
    int iters = 2000; int grid = n_sm;
    CK(cudaFuncSetAttribute(k_lds, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    float ms = time_ms([&] { k_lds<<<grid, threads, 65536>>>(d_out, iters, div, stride, d_cyc); });
    double cyc = max_cyc(grid);
    double n_lds = (double)(threads / 32) * iters * 16;
    printf("lds128 %-28s thr %3d : %5.2f clk per warp-LDS.128 per SM (%.3f ms)\n", what, threads, cyc / n_lds, ms);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0)); n_sm = p.multiProcessorCount;
    printf("device %s  SMs %d  clock %.0f MHz\n", p.name, n_sm, p.clockRate * 1e-3);
    CK(cudaMalloc(&d_out, 64 << 20)); CK(cudaMalloc(&d_cyc, 4096 * sizeof(long long)));
    run_reg<8, 8, false>(256, 1); run_reg<8, 8, false>(256, 2); run_reg<8, 8, false>(128, 1);
    run_reg<8, 8, true>(256, 1);  run_reg<8, 8, true>(256, 2);  run_reg<8, 8, true>(128, 1);
    run_reg<4, 8, false>(256, 2); run_reg<4, 8, true>(256, 2);
    run_reg<4, 4, false>(256, 4); run_reg<4, 4, true>(256, 4);
    run_lds(32, 0, 256, "all lanes same 16B");
    run_lds(8, 32, 256, "4 groups x16B (stride 32B)");
    run_lds(4, 48, 256, "8 groups x16B (stride 48B)");
    run_lds(2, 48, 256, "16 groups x16B (stride 48B)");
    run_lds(1, 16, 256, "32 distinct x16B (512B)");
    run_tile<8, 8, 4, false>(256, 1); run_tile<8, 8, 4, true>(256, 1);
    run_tile<8, 8, 4, true>(128, 1);  run_tile<8, 8, 4, true>(128, 2);
    run_tile<8, 8, 2, true>(256, 1);  run_tile<8, 8, 8, true>(256, 1);
    run_tile<8, 12, 4, true>(256, 1); run_tile<8, 12, 2, true>(256, 1);
    run_tile<8, 16, 4, true>(256, 1); run_tile<8, 16, 2, true>(256, 1);
    run_tile<4, 8, 8, true>(256, 1);  run_tile<4, 8, 8, true>(256, 2);
    run_tile<4, 12, 8, true>(256, 1); run_tile<4, 16, 8, true>(256, 1);
    run_tile<4, 4, 8, true>(256, 2);  run_tile<4, 4, 8, false>(256, 2);
    run_tile<4, 4, 8, true>(128, 1);  run_tile<4, 4, 8, false>(128, 1); run_tile<4, 4, 8, true>(128, 4);
    run_tile<4, 8, 8, true>(128, 1); run_tile<4, 8, 8, false>(128, 1); run_tile<8, 8, 4, false>(128, 1);
    run_tile<8, 4, 4, true>(256, 2);
    printf("done\n");
    return 0;
}
