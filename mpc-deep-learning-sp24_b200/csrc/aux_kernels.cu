// K6..K8: the small entry points around the rollout kernels that complete the reference call surface.
//   K6 noise_forward   FeedforwardNoiseModel.forward / get_prediction   model_scripts/noise_model.py:34-45
//                      (build_mlp topology utils/pytorch_utils.py:19-57): out = mu + softplus(p) * eps
//   K7 dynnn_rollout   DynamicsNN.get_prediction fed back on itself (the run_dynamics_model(open_loop=True)
//                      loop, scripts/test_model.py:67-87, with model_scripts/noise_model.py:113-125 as the model):
//                      q_{t+1} = noise_net(q history, u history), history deques left-padded by repetition
//                      (noise_model.py:117-121), batched over N trajectories and fused over H steps
//   K8 error_metrics   the per-sample MAE / RMSE of eval_models.calc_error  model_scripts/eval_models.py:91-92
// Same building blocks as the fused rollout kernels (dense_resident tiles, Philox keyed by the global
// trajectory index); all FP32, bias first, k ascending.
#include <algorithm>
#include "owner.cuh"
#include "launch.h"

namespace mpcdl {

struct AuxSmem { int blob, nin, a0, a1, zb, total; };
__host__ __device__ inline AuxSmem aux_layout(int blob_floats, int in_dim, int size, int TN) {
    AuxSmem L;
    int o = 0;
    L.blob = o; o += align_up(blob_floats, 32);
    L.nin = o;  o += in_dim * TN;
    L.a0 = o;   o += size * TN;
    L.a1 = o;   o += size * TN;
    L.zb = o;   o += 12 * TN;
    L.total = o;
    return L;
}

// all layers of the noise net on the CTA's tile: nin [in_dim][TN] -> zb [12][TN] (mu rows 0..5, pre-scale rows 6..11)
__device__ __forceinline__ void noise_net(const NoiseParams& nz, const float* blob, const float* nin, float* a0, float* a1,
                                          float* zb, int TN, int tid, int nthreads) {
    const float* W = blob;
    int K = nz.in_dim;
    const float* in = nin;
    for (int l = 0; l <= nz.nlayers; ++l) {
        const bool last = l == nz.nlayers;
        const int outw = last ? 12 : nz.size;
        float* out = last ? zb : ((l & 1) ? a1 : a0);
        dense_resident<4, 4>(in, K, W, outw, outw, last ? ACT_IDENTITY : nz.act, out, TN, tid, nthreads);
        __syncthreads();
        W += (K + 1) * outw;
        K = outw;
        in = out;
    }
}

__device__ __forceinline__ void draw6(const AuxParams& p, long long n, int t, int H, float (&e)[6]) {
    if (p.eps) {
        const float* ep = p.eps + (n * H + t) * 6;
#pragma unroll
        for (int i = 0; i < 6; ++i) e[i] = __ldg(ep + i);
    } else {
        philox_normal6(p.seed, static_cast<uint64_t>(p.traj_offset + n), static_cast<uint32_t>(p.step0 + t), e);
    }
}

// ---------------------------------------------------------------------------------------------- K6
__global__ void __launch_bounds__(AUX_THREADS, 1) noise_forward_kernel(const __grid_constant__ AuxParams p) {
    extern __shared__ __align__(128) float sm[];
    constexpr int TN = AUX_TN;
    const AuxSmem L = aux_layout(p.nz.blob_floats, p.nz.in_dim, p.nz.size, TN);
    const int tid = threadIdx.x, hist = p.nz.hist;
    for (int i = tid; i < p.nz.blob_floats; i += blockDim.x) sm[L.blob + i] = p.nz.blob[i];
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        __syncthreads();
        // input = concat([q (6*hist), u (2*hist)])  (noise_model.py:35), one column per sample
        for (int i = tid; i < p.nz.in_dim * TN; i += blockDim.x) {
            const int pos = i % TN, k = i / TN;
            long long n = (long long)tile * TN + pos;
            if (n >= p.N) n = p.N - 1;
            sm[L.nin + k * TN + pos] = k < 6 * hist ? __ldg(p.q + n * 6 * hist + k) : __ldg(p.u + n * 2 * hist + (k - 6 * hist));
        }
        __syncthreads();
        noise_net(p.nz, sm + L.blob, sm + L.nin, sm + L.a0, sm + L.a1, sm + L.zb, TN, tid, blockDim.x);
        const long long n = (long long)tile * TN + tid;
        if (tid < TN && n < p.N) {
            float e[6];
            draw6(p, n, 0, 1, e);
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const float mu = sm[L.zb + i * TN + tid], sc = softplus_f(sm[L.zb + (6 + i) * TN + tid]);
                p.out[n * 6 + i] = __fadd_rn(mu, __fmul_rn(e[i], sc));       // Normal(mu, softplus(p)).rsample()
                if (p.mu) p.mu[n * 6 + i] = mu;
                if (p.scale) p.scale[n * 6 + i] = sc;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------- K7
__global__ void __launch_bounds__(AUX_THREADS, 1) dynnn_rollout_kernel(const __grid_constant__ AuxParams p) {
    extern __shared__ __align__(128) float sm[];
    constexpr int TN = AUX_TN;
    const AuxSmem L = aux_layout(p.nz.blob_floats, p.nz.in_dim, p.nz.size, TN);
    const int tid = threadIdx.x, hist = p.nz.hist, H = p.H;
    float* const nin = sm + L.nin;
    for (int i = tid; i < p.nz.blob_floats; i += blockDim.x) sm[L.blob + i] = p.nz.blob[i];
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const bool owner = tid < TN;
        long long n = (long long)tile * TN + tid;
        const bool valid = owner && n < p.N;
        if (n >= p.N) n = p.N - 1;
        float q[6];
        if (owner) {
#pragma unroll
            for (int i = 0; i < 6; ++i) q[i] = __ldg(p.q + n * 6 + i);
        }
        __syncthreads();
        for (int t = 0; t < H; ++t) {
            if (owner) {
                const float2 uu = __ldg(reinterpret_cast<const float2*>(p.u) + n * H + t);
                // deque append; the first call left-pads by repetition (noise_model.py:117-121).  q is already in
                // the noise model's order, so the rows are written directly (noise_push permutes a frame-order state)
                const int e0 = t == 0 ? 0 : hist - 1;
                if (t > 0)
                    for (int e = 0; e + 1 < hist; ++e) {
#pragma unroll
                        for (int i = 0; i < 6; ++i) nin[(e * 6 + i) * TN + tid] = nin[((e + 1) * 6 + i) * TN + tid];
                        nin[(6 * hist + e * 2) * TN + tid] = nin[(6 * hist + (e + 1) * 2) * TN + tid];
                        nin[(6 * hist + e * 2 + 1) * TN + tid] = nin[(6 * hist + (e + 1) * 2 + 1) * TN + tid];
                    }
                for (int e = e0; e < hist; ++e) {
#pragma unroll
                    for (int i = 0; i < 6; ++i) nin[(e * 6 + i) * TN + tid] = q[i];
                    nin[(6 * hist + e * 2) * TN + tid] = uu.x;
                    nin[(6 * hist + e * 2 + 1) * TN + tid] = uu.y;
                }
            }
            __syncthreads();
            noise_net(p.nz, sm + L.blob, nin, sm + L.a0, sm + L.a1, sm + L.zb, TN, tid, blockDim.x);
            if (owner) {
                float e[6];
                draw6(p, n, t, H, e);
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const float mu = sm[L.zb + i * TN + tid], sc = softplus_f(sm[L.zb + (6 + i) * TN + tid]);
                    q[i] = __fadd_rn(mu, __fmul_rn(e[i], sc));
                }
                if (valid) {
                    float2* o = reinterpret_cast<float2*>(p.out + (n * H + t) * 6);
                    o[0] = make_float2(q[0], q[1]);
                    o[1] = make_float2(q[2], q[3]);
                    o[2] = make_float2(q[4], q[5]);
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------- K8
// per sample: mae = mean |o - l|, rmse = sqrt(mean (o - l)^2) over the 6 outputs  (eval_models.py:91-92)
__global__ void error_metrics_kernel(const float* __restrict__ output, const float* __restrict__ labels, long long B,
                                     float* __restrict__ mae, float* __restrict__ rmse) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= B) return;
    float sa = 0.f, ss = 0.f;
#pragma unroll
    for (int d = 0; d < 6; ++d) {
        const float df = __fsub_rn(__ldg(output + n * 6 + d), __ldg(labels + n * 6 + d));
        sa = __fadd_rn(sa, fabsf(df));
        ss = __fadd_rn(ss, __fmul_rn(df, df));
    }
    mae[n] = __fdiv_rn(sa, 6.0f);
    rmse[n] = __fsqrt_rn(__fdiv_rn(ss, 6.0f));
}

// ---------------------------------------------------------------------------------------------- K9 MPPI update
// (new: SURVEY.md 8f-1; the closest reference caller is the feedback loop of run_dynamics_model, test_model.py:67-87)
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__global__ void __launch_bounds__(MPPI_THREADS) mppi_min_kernel(const float* __restrict__ costs, long long N, float* __restrict__ part) {
    __shared__ float wm[MPPI_THREADS / 32];
    float m = INFINITY;
    for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (long long)gridDim.x * blockDim.x)
        m = fminf(m, __ldg(costs + n));
    m = warp_min(m);
    if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < MPPI_THREADS / 32; ++w) m = fminf(m, wm[w]);
        part[blockIdx.x] = m;
    }
}
__global__ void mppi_min_final_kernel(const float* __restrict__ part, int nparts, float* __restrict__ out) {
    float m = INFINITY;
    for (int i = threadIdx.x; i < nparts; i += 32) m = fminf(m, part[i]);
    m = warp_min(m);
    if (threadIdx.x == 0) *out = m;
}

// stage 1: per CTA, sum_n w_n * (u_{n,t} - U[t]) for every (t, channel) and sum_n w_n, in a fixed order:
// warp-shuffle tree over the 32 trajectories of a warp pass, per-warp running sums in shared memory, warps summed 0..7
__global__ void __launch_bounds__(MPPI_THREADS) mppi_partial_kernel(const __grid_constant__ MppiParams p) {
    extern __shared__ float acc[];                       // [warps][2H+1]
    const int H = p.H, W = 2 * H + 1, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* mine = acc + warp * W;
    for (int i = lane; i < W; i += 32) mine[i] = 0.f;
    __syncwarp();
    const float cmin = __ldg(p.cmin);
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long rounds = (p.N + stride - 1) / stride;
    for (long long r = 0; r < rounds; ++r) {
        const long long n = r * stride + (long long)blockIdx.x * blockDim.x + threadIdx.x;
        const float w = n < p.N ? __expf(-(__ldg(p.costs + n) - cmin) * p.inv_lambda) : 0.f;
        const float ws = warp_sum(w);
        if (lane == 0) mine[2 * H] += ws;
        for (int t = 0; t < H; ++t) {
            float d0 = 0.f, d1 = 0.f;
            if (w > 0.f) {                               // zero-weight trajectories contribute nothing: skip their draws
                const float2 ub = __ldg(reinterpret_cast<const float2*>(p.U) + t);
                float z0, z1;
                philox_normal2(p.ctrl_seed, static_cast<uint64_t>(p.traj_offset + n), static_cast<uint32_t>(t), z0, z1);
                d0 = w * (fminf(fmaxf(fmaf(p.sig[0], z0, ub.x), p.ulo[0]), p.uhi[0]) - ub.x);
                d1 = w * (fminf(fmaxf(fmaf(p.sig[1], z1, ub.y), p.ulo[1]), p.uhi[1]) - ub.y);
            }
            d0 = warp_sum(d0);
            d1 = warp_sum(d1);
            if (lane == 0) { mine[2 * t] += d0; mine[2 * t + 1] += d1; }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < W; i += blockDim.x) {
        float s = 0.f;
        for (int w = 0; w < MPPI_THREADS / 32; ++w) s += acc[w * W + i];
        p.workspace[(size_t)blockIdx.x * (W + 1) + i] = s;
    }
}
// stage 2: CTA partials summed in CTA order (double accumulation), one thread per entry
__global__ void mppi_final_kernel(const __grid_constant__ MppiParams p, int nparts) {
    const int W = 2 * p.H + 1;
    for (int i = threadIdx.x; i < W; i += blockDim.x) {
        double s = 0.0;
        for (int b = 0; b < nparts; ++b) s += (double)p.workspace[(size_t)b * (W + 1) + i];
        p.sums[i] = (float)s;
    }
    if (threadIdx.x == 0) p.sums[W] = __ldg(p.cmin);
}

// ---------------------------------------------------------------------------------------------- launchers
cudaError_t launch_mppi_min(const float* costs, long long N, float* workspace, float* cmin_out, cudaStream_t st) {
    const int grid = (int)std::min<long long>(MPPI_GRID, (N + MPPI_THREADS - 1) / MPPI_THREADS);
    mppi_min_kernel<<<grid, MPPI_THREADS, 0, st>>>(costs, N, workspace);
    mppi_min_final_kernel<<<1, 32, 0, st>>>(workspace, grid, cmin_out);
    return cudaGetLastError();
}

cudaError_t launch_mppi_reduce(const MppiParams& p, cudaStream_t st) {
    const int grid = (int)std::min<long long>(MPPI_GRID, (p.N + MPPI_THREADS - 1) / MPPI_THREADS);
    const int smem = (MPPI_THREADS / 32) * (2 * p.H + 1) * 4;
    mppi_partial_kernel<<<grid, MPPI_THREADS, smem, st>>>(p);
    mppi_final_kernel<<<1, 256, 0, st>>>(p, grid);
    return cudaGetLastError();
}

int aux_smem_bytes(int blob_floats, int in_dim, int size) { return aux_layout(blob_floats, in_dim, size, AUX_TN).total * 4; }

cudaError_t launch_aux(const AuxParams& p, bool closed_loop, int grid, int smem_bytes, cudaStream_t st) {
    auto k = closed_loop ? dynnn_rollout_kernel : noise_forward_kernel;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    k<<<grid, AUX_THREADS, smem_bytes, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_error_metrics(const float* output, const float* labels, long long B, float* mae, float* rmse, cudaStream_t st) {
    const int threads = 256;
    error_metrics_kernel<<<(unsigned)((B + threads - 1) / threads), threads, 0, st>>>(output, labels, B, mae, rmse);
    return cudaGetLastError();
}

}  // namespace mpcdl
