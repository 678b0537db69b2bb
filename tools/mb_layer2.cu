// Micro-benchmark (development aid, round 2): one dense 64 x 64 layer step of the LARGE-tile MLP rollout with the packed
// FFMA2 instruction (fma.rn.f32x2), for register tiles 8x8 / 16x8 / 8x16 and tiles of 256 / 512 trajectories per CTA
// (512 = one activation buffer updated in place: barrier, write, barrier).  Prints cycles per layer and FMA/clk/SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mb_layer2 tools/mb_layer2.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)
constexpr int HID = 64;
__constant__ float Wc[HID * HID + HID];      // CONSTW variants: the layer's weights as constant-bank operands

__device__ __forceinline__ uint64_t pack2(float lo, float hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ void ffma2(uint64_t& acc, uint64_t a, uint64_t b) { asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b)); }

template <int N> __device__ __forceinline__ void ldv(float (&d)[N], const float* p) {
#pragma unroll
    for (int c = 0; c < N / 4; ++c) { float4 v = *reinterpret_cast<const float4*>(p + 4 * c); d[4 * c] = v.x; d[4 * c + 1] = v.y; d[4 * c + 2] = v.z; d[4 * c + 3] = v.w; }
}
template <int RT> __device__ __forceinline__ void lda(float (&d)[RT], const float* p, int hs) {
#pragma unroll
    for (int c = 0; c < RT / 4; ++c) { float4 v = *reinterpret_cast<const float4*>(p + c * hs); d[4 * c] = v.x; d[4 * c + 1] = v.y; d[4 * c + 2] = v.z; d[4 * c + 3] = v.w; }
}

// PACKED = 1: FFMA2 over column pairs, 0: scalar FFMA, 2: scalar FFMA with the weights read from the constant bank at a
// warp-uniform index (no shared-memory -> register fill for the weight operand).  INPLACE: one activation buffer.
template <int RT, int RJ, int UNR, int PACKED, int INPLACE>
__global__ void __launch_bounds__(256) k_layer(float* out, int iters, int TN, long long* cyc) {
    extern __shared__ __align__(16) float sm[];
    float* W = sm;
    float* act0 = W + HID * HID + HID;
    float* act1 = INPLACE ? act0 : act0 + HID * TN;
    for (int i = threadIdx.x; i < HID * HID + HID; i += blockDim.x) W[i] = ((i * 37) % 101 - 50) * 2e-3f;
    for (int i = threadIdx.x; i < (INPLACE ? 1 : 2) * HID * TN; i += blockDim.x) act0[i] = (i % 13) * 0.01f;
    __syncthreads();
    const int n_tg = TN / RT;
    const int tg = threadIdx.x % n_tg, jg = threadIdx.x / n_tg;
    const int hs = 4 * n_tg;
    constexpr int RP = RJ / 2;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const float* in = (it & 1) ? act1 : act0;
        float* o = (it & 1) ? act0 : act1;
        float acc[RT][RJ];
        {
            float b[RJ];
            ldv<RJ>(b, W + HID * HID + jg * RJ);
#pragma unroll
            for (int i = 0; i < RT; ++i)
#pragma unroll
                for (int r = 0; r < RJ; ++r) acc[i][r] = b[r];
        }
        const float* ap = in + tg * 4;
        const float* wp = W + jg * RJ;
        if (PACKED == 2) {
            const int jgu = __shfl_sync(0xffffffffu, jg, 0) * RJ;         // warp-uniform (n_tg is a multiple of 32)
#pragma unroll UNR
            for (int k = 0; k < HID; ++k) {
                float a[RT];
                lda<RT>(a, ap + k * TN, hs);
#pragma unroll
                for (int i = 0; i < RT; ++i)
#pragma unroll
                    for (int r = 0; r < RJ; ++r) acc[i][r] = fmaf(a[i], Wc[k * HID + jgu + r], acc[i][r]);
            }
        } else if (PACKED) {
            uint64_t pa[RT][RP];
#pragma unroll
            for (int i = 0; i < RT; ++i)
#pragma unroll
                for (int r = 0; r < RP; ++r) pa[i][r] = pack2(acc[i][2 * r], acc[i][2 * r + 1]);
#pragma unroll UNR
            for (int k = 0; k < HID; ++k) {
                float a[RT], w[RJ];
                lda<RT>(a, ap + k * TN, hs);
                ldv<RJ>(w, wp + k * HID);
                uint64_t bp[RP];
#pragma unroll
                for (int r = 0; r < RP; ++r) bp[r] = pack2(w[2 * r], w[2 * r + 1]);
#pragma unroll
                for (int i = 0; i < RT; ++i) {
                    const uint64_t ad = pack2(a[i], a[i]);
#pragma unroll
                    for (int r = 0; r < RP; ++r) ffma2(pa[i][r], ad, bp[r]);
                }
            }
#pragma unroll
            for (int i = 0; i < RT; ++i)
#pragma unroll
                for (int r = 0; r < RP; ++r) unpack2(pa[i][r], acc[i][2 * r], acc[i][2 * r + 1]);
        } else {
#pragma unroll UNR
            for (int k = 0; k < HID; ++k) {
                float a[RT], w[RJ];
                lda<RT>(a, ap + k * TN, hs);
                ldv<RJ>(w, wp + k * HID);
#pragma unroll
                for (int i = 0; i < RT; ++i)
#pragma unroll
                    for (int r = 0; r < RJ; ++r) acc[i][r] = fmaf(a[i], w[r], acc[i][r]);
            }
        }
        if (INPLACE) __syncthreads();
#pragma unroll
        for (int r = 0; r < RJ; ++r) {
            float* d = o + (jg * RJ + r) * TN + tg * 4;
#pragma unroll
            for (int c = 0; c < RT / 4; ++c)
                *reinterpret_cast<float4*>(d + c * hs) = make_float4(fmaxf(acc[4 * c][r], 0.f), fmaxf(acc[4 * c + 1][r], 0.f),
                                                                    fmaxf(acc[4 * c + 2][r], 0.f), fmaxf(acc[4 * c + 3][r], 0.f));
        }
        __syncthreads();
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
    if (threadIdx.x < TN) out[blockIdx.x * TN + threadIdx.x] = act0[threadIdx.x] + act1[threadIdx.x];
}

static float* d_out; static long long* d_cyc; static int n_sm;

template <int RT, int RJ, int UNR, int PACKED, int INPLACE> static void run(int TN) {
    int iters = 1000;
    int threads = (TN / RT) * (HID / RJ);
    size_t smem = (size_t)(HID * HID + HID + (INPLACE ? 1 : 2) * HID * TN) * 4;
    auto k = k_layer<RT, RJ, UNR, PACKED, INPLACE>;
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, k));
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int rep = 0; rep < 2; ++rep) { k<<<n_sm, threads, smem>>>(d_out, iters, TN, d_cyc); CK(cudaDeviceSynchronize()); }
    static long long h[1024];
    CK(cudaMemcpy(h, d_cyc, n_sm * sizeof(long long), cudaMemcpyDeviceToHost));
    long long m = 0; for (int i = 0; i < n_sm; ++i) if (h[i] > m) m = h[i];
    double cpl = (double)m / iters;
    printf("TN %3d tile %2dx%-2d unr %2d %s %s thr %3d regs %3d : %8.1f cyc/layer  %6.1f FMA/clk/SM  (%.3f of 128)\n", TN, RT, RJ, UNR,
           PACKED == 2 ? "CONSTW" : PACKED ? "FFMA2" : "FFMA ", INPLACE ? "inplace" : "pingpong", threads, fa.numRegs, cpl, (double)TN * HID * HID / cpl,
           (double)TN * HID * HID / cpl / 128.0);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0)); n_sm = p.multiProcessorCount;
    CK(cudaMalloc(&d_out, 16 << 20)); CK(cudaMalloc(&d_cyc, 1024 * sizeof(long long)));
    { static float h[HID * HID + HID]; for (int i = 0; i < HID * HID + HID; ++i) h[i] = ((i * 37) % 101 - 50) * 2e-3f; CK(cudaMemcpyToSymbol(Wc, h, sizeof(h))); }
    run<8, 8, 8, 2, 0>(256);  run<8, 8, 16, 2, 0>(256); run<8, 8, 8, 2, 0>(128); run<16, 8, 8, 2, 0>(256);
    run<8, 8, 8, 0, 0>(256);  run<8, 8, 8, 1, 0>(256);  run<8, 8, 4, 1, 0>(256);  run<8, 8, 16, 1, 0>(256);
    run<16, 8, 4, 1, 0>(256); run<16, 8, 8, 1, 0>(256); run<8, 16, 4, 1, 0>(256); run<8, 16, 8, 1, 0>(256);
    run<16, 8, 4, 0, 0>(256); run<8, 16, 4, 0, 0>(256);
    run<16, 8, 4, 1, 1>(512); run<16, 8, 8, 1, 1>(512); run<8, 16, 4, 1, 1>(512); run<8, 16, 8, 1, 1>(512);
    run<8, 8, 8, 1, 1>(256);  run<16, 16, 4, 1, 1>(512); run<16, 16, 2, 1, 1>(512);
    run<8, 8, 8, 1, 0>(128);  run<8, 8, 8, 1, 0>(224);  run<16, 8, 4, 1, 0>(128);
    printf("done\n");
    return 0;
}
