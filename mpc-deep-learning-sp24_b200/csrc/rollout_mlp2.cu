// K1b: weight-resident FP32 rollout of the ReLU-MLP forward-dynamics model (the C2 / C5 hot kernel).
//   reference: Basic_Model.forward  model_scripts/basic_model.py:58-61 (ctor 45-56)
//              z-score               model_scripts/basic_model.py:28-35
//              closed loop           oracle/rollout_ref.py rollout_mlp
// Same blob, same arithmetic order (bias first, k ascending, FP32 FMA) as rollout_mlp.cu, but every
// loop bound is a template constant and the mapping follows the measurements in profiles/r01a:
//   * the whole packed blob (<= 70 KB for 64x4) is TMA-bulk-copied into shared memory ONCE per CTA;
//   * a hidden layer is an RT x RJ register tile per thread over a fully unrolled K slice; for small
//     tiles (32 trajectories per SM, the 4096 x 50 planner call) K is split over KS thread sets whose
//     partial sums meet in shared memory -- 2 warps per scheduler hide the LDS latency that a single
//     warp cannot (2009 vs 3149 cycles per dependent 64x64 layer);
//   * large tiles use 8x8 register tiles (64 FFMA per 4 LDS.128: shared-memory fill rate stops binding);
//   * the output layer (width 8) is split over K quarters; the trajectory's owner thread sums the
//     quarters, de-normalises, stores the state, accumulates the cost and re-normalises the next input,
//     so state and cost never leave its registers.
// Noise injection and non-resident (hidden >= 128) models stay on the generic kernel (rollout_mlp.cu).
#include "owner.cuh"
#include "launch.h"

namespace mpcdl {

struct Mlp2Smem {     // float offsets; xin .. ybuf are group 0's, group g's follow at + g * gstride
    int bars, blob, norm, xin, act0, act1, red, ybuf, gstride, total;
};
__host__ __device__ inline Mlp2Smem mlp2_layout(int blob_floats, int hid, int TN, int ks, int kq, int ng) {
    Mlp2Smem L;
    int o = 0;
    L.bars = o;  o += 32;
    L.blob = o;  o += align_up(blob_floats, 32);
    L.norm = o;  o += 32;
    const int g0 = o;
    L.xin = o;   o += 8 * TN;
    L.act0 = o;  o += hid * TN;
    L.act1 = o;  o += hid * TN;
    L.red = o;   o += (ks - 1) * hid * TN;
    L.ybuf = o;  o += kq * 8 * TN;
    L.gstride = align_up(o - g0, 32);
    L.total = g0 + ng * L.gstride;
    return L;
}

template <int RT>
__device__ __forceinline__ void ld_traj(float (&a)[RT], const float* p, int hs) {
#pragma unroll
    for (int c = 0; c < RT / 4; ++c) {
        float4 v = *reinterpret_cast<const float4*>(p + c * hs);
        a[4 * c] = v.x; a[4 * c + 1] = v.y; a[4 * c + 2] = v.z; a[4 * c + 3] = v.w;
    }
}

// acc[i][r] += sum_{k<KLEN} in[k][pos(i)] * w[k][r]     (k ascending, one FMA per term; issued as FFMA2 on column pairs)
template <int RT, int RJ, int KLEN, int LDW>
__device__ __forceinline__ void tile_gemm(float (&acc)[RT][RJ], const float* __restrict__ in, int TN, int hs,
                                          const float* __restrict__ w) {
    static_assert(RJ % 2 == 0, "column pairs");
    constexpr int RP = RJ / 2;
    // straight-line code while it stays I-cache friendly (<= 1024 FMA), else 16 k-steps per trip (512 FFMA2 + 64 LDS for
    // the 8 x 8 tile; tools/mb_layer2: 10 967 cycles per layer against 11 494 at 8 k-steps per trip)
    constexpr int UNR = KLEN * RT * RJ <= 1024 ? KLEN : 16;
    uint64_t pa[RT][RP];
#pragma unroll
    for (int i = 0; i < RT; ++i)
#pragma unroll
        for (int r = 0; r < RP; ++r) pa[i][r] = pack2(acc[i][2 * r], acc[i][2 * r + 1]);
#pragma unroll UNR
    for (int k = 0; k < KLEN; ++k) {
        float a[RT], b[RJ];
        ld_traj<RT>(a, in + k * TN, hs);
        load_cols<RJ>(b, w + k * LDW);
        uint64_t bp[RP];
#pragma unroll
        for (int r = 0; r < RP; ++r) bp[r] = pack2(b[2 * r], b[2 * r + 1]);
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
}

template <int RT, int RJ>
__device__ __forceinline__ void store_relu(float* out, int TN, int hs, const float (&acc)[RT][RJ]) {
#pragma unroll
    for (int r = 0; r < RJ; ++r)
#pragma unroll
        for (int c = 0; c < RT / 4; ++c)
            *reinterpret_cast<float4*>(out + r * TN + c * hs) =
                make_float4(fmaxf(acc[4 * c][r], 0.f), fmaxf(acc[4 * c + 1][r], 0.f), fmaxf(acc[4 * c + 2][r], 0.f),
                            fmaxf(acc[4 * c + 3][r], 0.f));
}

template <int HID, int RT, int RJ, int KS>
struct Mlp2Cfg {
    static constexpr int N_JG = HID / RJ;
    static constexpr int KLEN = HID / KS;
    static constexpr int KQ = HID * KS / (RT * RJ);          // threads per trajectory = K slices of the output layer
    static constexpr int QLEN = HID / KQ;
    static constexpr int MAX_THREADS = (HID == 16 && RT == 8) ? 512 : 256;   // <= 256 threads: 255 registers for 8x8 tiles
    static_assert(KQ >= 1 && HID % KQ == 0 && HID % KS == 0 && RT % 4 == 0, "unsupported MLP tile configuration");
};

// NG = 2: the CTA is two independent groups of threads, each rolling out its own tile of TN trajectories with its own
// activation buffers and a named barrier of its own, sharing the weights.  The two groups drift apart, so the per-layer
// epilogue of one (bias, ReLU, transposed stores, barrier: ~940 of ~5700 clocks, tools/mb_layer2) runs under the k-loop of
// the other on every scheduler instead of idling the FMA pipe.
template <int HID, int RT, int RJ, int KS, int NG>
__global__ void __launch_bounds__((Mlp2Cfg<HID, RT, RJ, KS>::MAX_THREADS), 1) rollout_mlp2_kernel(const __grid_constant__ RolloutParams p) {
    using C = Mlp2Cfg<HID, RT, RJ, KS>;
    extern __shared__ __align__(128) float sm[];
    const ModelParams& m = p.m;
    const int TN = m.TN, n_tg = m.n_tg, NL = m.layers, H = p.io.H;
    const Mlp2Smem L = mlp2_layout((int)m.data_floats, HID, TN, KS, C::KQ, NG);
    const int GT = (int)blockDim.x / NG;                       // threads per group
    const int grp = NG == 1 ? 0 : (int)threadIdx.x / GT;
    const int tid = (int)threadIdx.x - grp * GT;               // thread index INSIDE the group: every mapping below uses it
    uint64_t* full = reinterpret_cast<uint64_t*>(sm + L.bars);
    float* const blob = sm + L.blob;
    float* const normS = sm + L.norm;
    float* const xin = sm + L.xin + grp * L.gstride;
    float* const act0 = sm + L.act0 + grp * L.gstride;
    const int act_flip = L.act1 - L.act0;
    float* const red = sm + L.red + grp * L.gstride;
    float* const ybuf = sm + L.ybuf + grp * L.gstride;
    auto group_sync = [&]() {
        if (NG == 1) __syncthreads();
        else named_bar_sync(1 + grp, GT);
    };

    // ---- one-time: weights -> shared memory (TMA bulk copies on one mbarrier), statistics
    if (threadIdx.x == 0) {
        mbar_init(full, 1);
        mbar_fence_init();
        mbar_expect_tx(full, m.data_floats * 4u);
        for (uint32_t o = 0; o < m.data_floats; o += 8192u)
            bulk_g2s(blob + o, m.blob + o, min(8192u, m.data_floats - o) * 4u, full);
    }
    for (int i = threadIdx.x; i < 28; i += blockDim.x) normS[i] = p.io.norm[i];       // f_mean8 f_std8 l_mean6 l_std6
    __syncthreads();
    mbar_wait(full, 0);

    const int per = n_tg * C::N_JG;
    const int ks = tid / per, tt = tid - ks * per;
    const int tg = tt % n_tg, jg = tt / n_tg;
    const int hs = 4 * n_tg;
    const int otraj = tid % TN, okq = tid / TN;                // output-layer mapping
    const bool owner = tid < TN;
    const PartDesc& pin = m.parts[0];
    const PartDesc& pout = m.parts[NL + 1];

    // pass k hands tiles [k NG G, (k+1) NG G) out GROUP-major (G = gridDim.x): a last, partial pass first fills group 0 of every
    // CTA, and a CTA running one group finishes its tile in 0.6 of the two-group time (131 072 x 50: 4.70 -> 4.26 ms)
    for (int tile = blockIdx.x + grp * gridDim.x; tile < p.io.ntiles; tile += gridDim.x * NG) {
        long long n = (long long)tile * TN + tid;
        const bool valid = owner && n < p.io.N;
        if (n >= p.io.N) n = p.io.N - 1;
        float s[6], cost = 0.f;
        float2 unext = make_float2(0.f, 0.f);
        if (owner) {
#pragma unroll
            for (int d = 0; d < 6; ++d) s[d] = __ldg(p.io.init + n * p.io.init_mul * 6 + d);
            unext = load_control(p.io, n, 0);
        }
        for (int t = 0; t < H; ++t) {
            float u[2] = {unext.x, unext.y};
            if (owner) {
                if (t + 1 < H) unext = load_control(p.io, n, t + 1);      // prefetch: hidden behind this step's layers
#pragma unroll
                for (int k = 0; k < 6; ++k) xin[k * TN + tid] = __fdiv_rn(__fsub_rn(s[k], normS[k]), normS[8 + k]);
#pragma unroll
                for (int k = 0; k < 2; ++k) xin[(6 + k) * TN + tid] = __fdiv_rn(__fsub_rn(u[k], normS[6 + k]), normS[14 + k]);
            }
            group_sync();
            // ---- input layer 8 -> HID (K = 8: not worth splitting)
            if (ks == 0) {
                float acc[RT][RJ], b[RJ];
                load_cols<RJ>(b, blob + pin.off + jg * RJ);
#pragma unroll
                for (int i = 0; i < RT; ++i)
#pragma unroll
                    for (int r = 0; r < RJ; ++r) acc[i][r] = b[r];
                tile_gemm<RT, RJ, 8, HID>(acc, xin + tg * 4, TN, hs, blob + pin.off + pin.bias_floats + jg * RJ);
                store_relu<RT, RJ>(act0 + (jg * RJ) * TN + tg * 4, TN, hs, acc);
            }
            group_sync();
            // ---- hidden layers HID -> HID
            for (int l = 1; l <= NL; ++l) {
                const float* in = act0 + ((l - 1) & 1) * act_flip;
                float* out = act0 + (l & 1) * act_flip;
                const float* wl = blob + m.parts[l].off;
                float acc[RT][RJ];
                if (ks == 0) {
                    float b[RJ];
                    load_cols<RJ>(b, wl + jg * RJ);
#pragma unroll
                    for (int i = 0; i < RT; ++i)
#pragma unroll
                        for (int r = 0; r < RJ; ++r) acc[i][r] = b[r];
                } else {
#pragma unroll
                    for (int i = 0; i < RT; ++i)
#pragma unroll
                        for (int r = 0; r < RJ; ++r) acc[i][r] = 0.f;
                }
                tile_gemm<RT, RJ, C::KLEN, HID>(acc, in + (ks * C::KLEN) * TN + tg * 4, TN, hs,
                                                wl + HID + (ks * C::KLEN) * HID + jg * RJ);
                if (KS > 1) {
                    if (ks > 0) {
                        float* d = red + ((ks - 1) * HID + jg * RJ) * TN + tg * 4;
#pragma unroll
                        for (int r = 0; r < RJ; ++r)
#pragma unroll
                            for (int c = 0; c < RT / 4; ++c)
                                *reinterpret_cast<float4*>(d + r * TN + c * hs) =
                                    make_float4(acc[4 * c][r], acc[4 * c + 1][r], acc[4 * c + 2][r], acc[4 * c + 3][r]);
                    }
                    group_sync();
                    if (ks == 0) {
#pragma unroll
                        for (int q = 0; q < KS - 1; ++q)
#pragma unroll
                            for (int r = 0; r < RJ; ++r) {
                                float part[RT];
                                ld_traj<RT>(part, red + (q * HID + jg * RJ + r) * TN + tg * 4, hs);
#pragma unroll
                                for (int i = 0; i < RT; ++i) acc[i][r] = __fadd_rn(acc[i][r], part[i]);
                            }
                    }
                }
                if (ks == 0) store_relu<RT, RJ>(out + (jg * RJ) * TN + tg * 4, TN, hs, acc);
                group_sync();
            }
            // ---- output layer HID -> 6 (blob width 8): K quarters per trajectory, owner sums them
            {
                const float* in = act0 + (NL & 1) * act_flip + (okq * C::QLEN) * TN + otraj;
                const float* w = blob + pout.off + pout.bias_floats + (okq * C::QLEN) * 8;
                float y[8];
                if (okq == 0) load_cols<8>(y, blob + pout.off);
                else {
#pragma unroll
                    for (int d = 0; d < 8; ++d) y[d] = 0.f;
                }
#pragma unroll
                for (int k = 0; k < C::QLEN; ++k) {
                    float a = in[k * TN], b[8];
                    load_cols<8>(b, w + k * 8);
#pragma unroll
                    for (int d = 0; d < 6; ++d) y[d] = fmaf(a, b[d], y[d]);
                }
#pragma unroll
                for (int d = 0; d < 6; ++d) ybuf[(okq * 8 + d) * TN + otraj] = y[d];
            }
            group_sync();
            if (owner) {
#pragma unroll
                for (int d = 0; d < 6; ++d) {
                    float y = ybuf[d * TN + tid];
#pragma unroll
                    for (int q = 1; q < C::KQ; ++q) y = __fadd_rn(y, ybuf[(q * 8 + d) * TN + tid]);
                    ybuf[d * TN + tid] = y;
                }
                float nxt[6];
                owner_finish(p, normS + 16, ybuf, nullptr, TN, tid, n, valid, t, u, nxt, cost);
#pragma unroll
                for (int d = 0; d < 6; ++d) s[d] = nxt[d];
            }
        }
        if (valid && p.io.costs) p.io.costs[n] = cost;
        group_sync();     // ybuf / xin of this tile are dead before the next tile's owners write
    }
}

// ------------------------------------------------------------------ launcher
template <int HID, int RT, int RJ, int KS, int NG>
static cudaError_t launch_mlp2_t(const RolloutParams& p, int grid, int smem_bytes, cudaStream_t st) {
    auto k = rollout_mlp2_kernel<HID, RT, RJ, KS, NG>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    k<<<grid, p.m.n_cthreads, smem_bytes, st>>>(p);      // n_cthreads = threads of ALL groups
    return cudaGetLastError();
}

int mlp2_smem_bytes(int blob_floats, int hid, int TN, int ks, int rt, int rj, int ng) {
    return mlp2_layout(blob_floats, hid, TN, ks, hid * ks / (rt * rj), ng).total * 4;
}

cudaError_t launch_mlp2(const RolloutParams& p, int rt, int rj, int ks, int ng, int grid, int smem_bytes, cudaStream_t st) {
    if (ng == 2 && p.m.hid == 64 && rt == 8 && rj == 8 && ks == 1) return launch_mlp2_t<64, 8, 8, 1, 2>(p, grid, smem_bytes, st);
    if (ng != 1) return cudaErrorInvalidConfiguration;
#define MPCDL_CASE(H_, RT_, RJ_, KS_) \
    if (p.m.hid == H_ && rt == RT_ && rj == RJ_ && ks == KS_) return launch_mlp2_t<H_, RT_, RJ_, KS_, 1>(p, grid, smem_bytes, st);
    MPCDL_MLP2_CONFIGS(MPCDL_CASE)
#undef MPCDL_CASE
    return cudaErrorInvalidConfiguration;
}

}  // namespace mpcdl
