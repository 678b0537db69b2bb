// tcgen05 probe (development aid): a chain of L dense 64->64 ReLU layers on one 128-row tile per CTA,
//   D[128x64] (TMEM, FP32) = A[128x64] . W^T,  kind::tf32, 1 pass (TF32) or 3 passes (3xTF32 ~ FP32 accuracy)
// A from TMEM (.ts) or shared memory (.ss); W resident in shared memory, canonical K-major no-swizzle layout.
// Checks the descriptors/layouts of csrc/umma.cuh against a CPU reference and measures cycles per layer.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/umma_probe tools/umma_probe.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../mpc-deep-learning-sp24_b200/csrc/umma.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

using namespace mpcdl;
constexpr int HID = 64, M = 128;
constexpr int W_FLOATS = HID * HID;                 // one 64x64 matrix in canonical layout
constexpr int LBO_B = (HID / 8) * 128, SBO = 128;   // B: N=64 rows -> 8 core matrices per k-chunk
constexpr int LBO_A = (M / 8) * 128;                // A in smem: 128 rows -> 16 core matrices per k-chunk

// canonical K-major offset (floats) of element (row r, col k) for a matrix with `rows` rows
__host__ __device__ inline int canon(int r, int k, int rows) { return (k / 4) * (rows / 8) * 32 + (r / 8) * 32 + (r % 8) * 4 + (k % 4); }

template <int PASSES, bool A_TMEM>
__global__ void __launch_bounds__(128, 1) probe_kernel(const float* __restrict__ Wc /* [L][2][4096] canonical hi,lo */,
                                                       const float* __restrict__ bias /* [L][64] */, const float* __restrict__ A0 /* [grid][128][64] */,
                                                       float* __restrict__ out, int L, int reps, long long* cyc) {
    extern __shared__ __align__(1024) float sm[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    float* Ws = sm;                                  // [L][2][4096]
    float* As = sm + L * 2 * W_FLOATS;               // [2][128*64] hi, lo (only .ss)
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < L * 2 * W_FLOATS; i += 128) Ws[i] = Wc[i];
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) { umma::tmem_alloc(&tmem_base_s, 256); umma::tmem_relinquish(); }
    umma::fence_proxy_async_smem();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tbase = tmem_base_s;
    const uint32_t lane_addr = tbase + (static_cast<uint32_t>(warp * 32) << 16);   // this warp's 32 TMEM lanes
    const uint32_t D_COL = 0, AH_COL = 64, AL_COL = 128;
    constexpr uint32_t IDESC = umma::idesc_tf32(M, HID);

    float a[HID];
    uint32_t parity = 0;
    long long t0 = clock64();
    for (int rep = 0; rep < reps; ++rep) {
#pragma unroll
        for (int k = 0; k < HID; ++k) a[k] = A0[(static_cast<size_t>(blockIdx.x) * M + tid) * HID + k];
        for (int l = 0; l < L; ++l) {
            // ---- publish the A operand (this thread's row) as hi / lo TF32 parts
#pragma unroll
            for (int c = 0; c < HID / 8; ++c) {
                uint32_t hi[8], lo[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) umma::split_tf32(a[c * 8 + i], hi[i], lo[i]);
                if (A_TMEM) {
                    umma::tmem_st8(lane_addr + AH_COL + c * 8, hi);
                    if (PASSES >= 3) umma::tmem_st8(lane_addr + AL_COL + c * 8, lo);
                } else {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        float* d = As + canon(tid, c * 8 + h * 4, M);
                        *reinterpret_cast<uint4*>(d) = make_uint4(hi[h * 4], hi[h * 4 + 1], hi[h * 4 + 2], hi[h * 4 + 3]);
                        if (PASSES >= 3) *reinterpret_cast<uint4*>(d + M * HID) = make_uint4(lo[h * 4], lo[h * 4 + 1], lo[h * 4 + 2], lo[h * 4 + 3]);
                    }
                }
            }
            if (A_TMEM) umma::wait_st(); else umma::fence_proxy_async_smem();
            umma::fence_before_sync();
            __syncthreads();
            // ---- one thread issues the layer's MMAs: D = Ahi.Bhi (+ Alo.Bhi + Ahi.Blo)
            if (tid == 0) {
                umma::fence_after_sync();
                const uint32_t bh = smem_u32(Ws + (l * 2 + 0) * W_FLOATS), bl = smem_u32(Ws + (l * 2 + 1) * W_FLOATS);
                const uint32_t ah = smem_u32(As), al = smem_u32(As + M * HID);
                uint32_t accum = 0;
#pragma unroll
                for (int p = 0; p < PASSES; ++p) {
                    const bool a_lo = p == 1 || p == 3, b_lo = p == 2 || p == 3;
#pragma unroll
                    for (int kk = 0; kk < HID / 8; ++kk) {
                        const uint64_t bd = umma::smem_desc_kmajor((b_lo ? bl : bh) + kk * 2 * LBO_B, LBO_B, SBO);
                        if (A_TMEM) {
                            umma::mma_tf32_ts(tbase + D_COL, tbase + (a_lo ? AL_COL : AH_COL) + kk * 8, bd, IDESC, accum);
                        } else {
                            const uint64_t ad = umma::smem_desc_kmajor((a_lo ? al : ah) + kk * 2 * LBO_A, LBO_A, SBO);
                            umma::mma_tf32_ss(tbase + D_COL, ad, bd, IDESC, accum);
                        }
                        accum = 1;
                    }
                }
                umma::mma_commit(&bar);
            }
            mbar_wait(&bar, parity);
            parity ^= 1u;
            umma::fence_after_sync();
            // ---- epilogue: D row -> registers, + bias, relu
#pragma unroll
            for (int c = 0; c < HID / 8; ++c) {
                float v[8];
                umma::tmem_ld8(lane_addr + D_COL + c * 8, v);
                umma::wait_ld();
#pragma unroll
                for (int i = 0; i < 8; ++i) a[c * 8 + i] = fmaxf(v[i] + __ldg(bias + l * HID + c * 8 + i), 0.f);
            }
            umma::fence_before_sync();      // order these TMEM reads before the next layer's MMA writes
        }
    }
    long long t1 = clock64();
    if (tid == 0) cyc[blockIdx.x] = t1 - t0;
#pragma unroll
    for (int k = 0; k < HID; ++k) out[(static_cast<size_t>(blockIdx.x) * M + tid) * HID + k] = a[k];
    __syncthreads();
    if (warp == 0) umma::tmem_dealloc(tbase, 256);
}

// ------------------------------------------------------------------------------------------ host
static float tf32_rna_host(float x) {
    uint32_t u; memcpy(&u, &x, 4);
    u = (u + 0x1000u) & 0xFFFFE000u;
    float r; memcpy(&r, &u, 4); return r;
}
#include <cstring>

template <int PASSES, bool A_TMEM>
static void run(int L, int grid, int reps, const std::vector<float>& W, const std::vector<float>& B, const std::vector<float>& A0,
                const std::vector<double>& ref64, const std::vector<float>& ref32) {
    std::vector<float> Wc((size_t)L * 2 * W_FLOATS);
    for (int l = 0; l < L; ++l)
        for (int n = 0; n < HID; ++n)
            for (int k = 0; k < HID; ++k) {
                float w = W[((size_t)l * HID + n) * HID + k];
                float hi = tf32_rna_host(w), lo = tf32_rna_host(w - hi);
                Wc[((size_t)l * 2 + 0) * W_FLOATS + canon(n, k, HID)] = hi;
                Wc[((size_t)l * 2 + 1) * W_FLOATS + canon(n, k, HID)] = lo;
            }
    float *dW, *dB, *dA, *dO; long long* dC;
    CK(cudaMalloc(&dW, Wc.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dA, A0.size() * 4));
    CK(cudaMalloc(&dO, A0.size() * 4)); CK(cudaMalloc(&dC, grid * sizeof(long long)));
    CK(cudaMemcpy(dW, Wc.data(), Wc.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dA, A0.data(), A0.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(dO, 0, A0.size() * 4));
    size_t smem = ((size_t)L * 2 * W_FLOATS + 2 * M * HID) * 4;
    auto k = probe_kernel<PASSES, A_TMEM>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, 128, smem>>>(dW, dB, dA, dO, L, reps, dC);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("passes %d a_tmem %d : KERNEL FAILED %s\n", PASSES, (int)A_TMEM, cudaGetErrorString(e)); exit(2); }
    std::vector<float> out(A0.size()); std::vector<long long> cyc(grid);
    CK(cudaMemcpy(out.data(), dO, out.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cyc.data(), dC, grid * sizeof(long long), cudaMemcpyDeviceToHost));
    double mx = 0, ek = 0, e32 = 0, sk = 0, s32 = 0;
    for (size_t i = 0; i < out.size(); ++i) mx = fmax(mx, fabs(ref64[i]));
    for (size_t i = 0; i < out.size(); ++i) {
        double dk = fabs(out[i] - ref64[i]), d3 = fabs(ref32[i] - ref64[i]);
        ek = fmax(ek, dk); e32 = fmax(e32, d3); sk += dk; s32 += d3;
    }
    long long cm = 0; for (auto c : cyc) cm = c > cm ? c : cm;
    printf("passes %d A_%s L %d grid %3d : max rel err %.3e (mean %.3e) | fp32-fma ref %.3e (mean %.3e) | %.0f cyc/layer\n", PASSES,
           A_TMEM ? "tmem" : "smem", L, grid, ek / mx, sk / out.size() / mx, e32 / mx, s32 / out.size() / mx, (double)cm / reps / L);
    cudaFree(dW); cudaFree(dB); cudaFree(dA); cudaFree(dO); cudaFree(dC);
}

int main(int argc, char** argv) {
    const int L = 4, grid = argc > 1 ? atoi(argv[1]) : 148, reps = 50;
    srand(1);
    auto rnd = [] { return (rand() / (float)RAND_MAX) * 2.f - 1.f; };
    std::vector<float> W((size_t)L * HID * HID), B((size_t)L * HID), A0((size_t)grid * M * HID);
    for (auto& w : W) w = rnd() * 0.35f;
    for (auto& b : B) b = rnd() * 0.3f;
    for (auto& a : A0) a = rnd() * 2.f;
    std::vector<double> ref64(A0.size()); std::vector<float> ref32(A0.size());
    for (size_t r = 0; r < (size_t)grid * M; ++r) {
        double x[HID], y[HID]; float xf[HID], yf[HID];
        for (int k = 0; k < HID; ++k) { x[k] = A0[r * HID + k]; xf[k] = A0[r * HID + k]; }
        for (int l = 0; l < L; ++l) {
            for (int n = 0; n < HID; ++n) {
                double s = 0; float sf = 0.f;
                for (int k = 0; k < HID; ++k) { s += (double)W[((size_t)l * HID + n) * HID + k] * x[k]; sf = fmaf(W[((size_t)l * HID + n) * HID + k], xf[k], sf); }
                y[n] = fmax(s + B[l * HID + n], 0.0); yf[n] = fmaxf(sf + B[l * HID + n], 0.f);
            }
            for (int k = 0; k < HID; ++k) { x[k] = y[k]; xf[k] = yf[k]; }
        }
        for (int k = 0; k < HID; ++k) { ref64[r * HID + k] = x[k]; ref32[r * HID + k] = xf[k]; }
    }
    run<1, false>(L, grid, reps, W, B, A0, ref64, ref32);
    run<3, false>(L, grid, reps, W, B, A0, ref64, ref32);
    run<1, true>(L, grid, reps, W, B, A0, ref64, ref32);
    run<3, true>(L, grid, reps, W, B, A0, ref64, ref32);
    run<4, true>(L, grid, reps, W, B, A0, ref64, ref32);
    printf("done\n");
    return 0;
}
