// K5: tensor-core (tcgen05 / TMEM) fused whole-horizon rollout of the ReLU-MLP forward-dynamics model.
//   reference: Basic_Model.forward  model_scripts/basic_model.py:58-61 (ctor 45-56)
//              z-score               model_scripts/basic_model.py:28-35
//              closed loop           oracle/rollout_ref.py rollout_mlp
// The large-N planner call (1 M trajectories x 50 steps) is a chain of dense [128 x 64].[64 x 64] layers per
// tile, FMA-pipe bound at 0.71 of the FFMA peak in rollout_mlp2.cu; here the layers run on tcgen05:
//   * all weights (tf32 hi/lo, canonical K-major UMMA order) are TMA-copied into shared memory ONCE per CTA;
//   * a CTA owns TWO tiles of 128 trajectories; each tile has its own TMEM slice (D 64 + A_hi 64 + A_lo 64
//     columns), its own 8 epilogue warps (TMEM lane = trajectory, two threads per trajectory with 32 hidden
//     units each) and its own MMA-issuing warp, so
//     the MMA phase of one tile overlaps the epilogue phase of the other (both fit: 2 x 192 <= 512 columns);
//   * the epilogue reads D, adds the bias, applies ReLU and writes the activations back as the next layer's
//     A operand (tcgen05.st, tf32 hi/lo split) -- activations never touch shared memory; the 64 -> 6 output
//     layer, de-normalisation, cost and state feedback stay FP32 on the CUDA cores in the trajectory's thread;
//   * the issuing warp runs converged and elects one lane only for the tcgen05 instructions, so descriptor
//     arithmetic stays on the uniform datapath (N = 64 MMAs are issue-bound otherwise).
#include "owner.cuh"
#include "umma.cuh"
#include "tc_params.h"
#include "launch.h"

namespace mpcdl {

constexpr uint32_t MC_D = 0, MC_A_HI = 64, MC_A_LO = 128;

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}

// one layer: D[128 x 64] = A[128 x K] . W^T  (W: N = 64 rows, K-major; hi block at sb, lo block at sb + BLK)
template <bool X3, int K>
__device__ __forceinline__ void mc_issue_layer(uint32_t tcol, uint32_t sb) {
    constexpr uint32_t LBO = 64 / 8 * 128, BLK = K / 4 * LBO;
    constexpr uint32_t IDESC = umma::idesc_tf32(128, 64);
    const uint64_t d0 = umma::smem_desc_kmajor(sb, LBO, 128u);
    uint32_t acc = X3 ? 0u : 1u;       // one-pass TF32: the epilogue has pre-loaded D with the layer's bias, every MMA accumulates
#pragma unroll
    for (int pass = X3 ? 0 : 2; pass < 3; ++pass) {
#pragma unroll
        for (int kk = 0; kk < K / 8; ++kk) {
            const uint64_t bd = d0 + (((pass == 1 ? BLK : 0u) + kk * 2u * LBO) >> 4);
            umma::mma_tf32_ts(tcol + MC_D, tcol + (pass == 0 ? MC_A_LO : MC_A_HI) + kk * 8, bd, IDESC, acc);
            acc = 1;
        }
    }
}

template <bool X3>
__device__ __forceinline__ void mc_stage16(uint32_t t_hi, uint32_t t_lo, const float (&v)[16]) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        // one-pass TF32: the tensor core reads the top 19 bits of the operand, so adding half a tf32 ulp is the whole rounding
        // (same value as tf32_round_fast, one ALU instruction less per unit); tf32x3 needs the masked hi for the exact residual
        hi[i] = X3 ? umma::tf32_round_fast(v[i]) : __float_as_uint(v[i]) + 0x1000u;
        lo[i] = X3 ? __float_as_uint(v[i] - __uint_as_float(hi[i])) : 0u;
    }
    umma::tmem_st16(t_hi, hi);
    if (X3) umma::tmem_st16(t_lo, lo);
}

// TILES x EW: 2 tiles with 8 epilogue warps each (two threads per trajectory, 256 TMEM columns per tile: both precisions, noise
// net), or -- TF32 without the noise net, enough tiles -- 4 tiles with 4 epilogue warps each (one thread per trajectory and all
// 64 units, 128 columns per tile: D + A_hi): a layer of a tile is a latency chain (MMAs -> commit -> tcgen05.ld -> bias / ReLU
// -> tcgen05.st -> arrive, ~1900 clocks) and four chains in flight keep the tensor pipe busier than two (1 M x 50: 8.15 ->
// 5.50 ms).  Two tcgen05.ld blocks in flight per wait were measured on top: slower (5.75 ms, register pressure at 640 threads).
template <int PREC, int TILES, int EW>
__global__ void __launch_bounds__(mc_threads(TILES), 1) rollout_mlp_tc_kernel(const __grid_constant__ McRolloutParams p) {
    static_assert((TILES == 2 && EW == 8) || (TILES == 4 && EW == 4 && PREC == PREC_TF32), "tile configurations of tc_params.h");
    constexpr int TCOLS = 512 / TILES;               // TMEM columns per tile
    constexpr int UPT = TC_HID / (EW / 4);           // hidden units per epilogue thread
    constexpr bool X3 = PREC == PREC_TF32X3;
    constexpr int HID = TC_HID, TN = TC_TN;
    extern __shared__ __align__(1024) float sm[];
    const McModel& m = p.m;
    const McSmem L = mc_layout((int)m.w_floats, m.layers, p.nz, m.tiles_per_cta);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = p.io.H, NL = m.layers;
    const long long N = p.io.N;

    uint64_t* wfull = reinterpret_cast<uint64_t*>(sm + L.bars);
    uint64_t* opnd = wfull + 1;                      // [TILES] operands of the next layer are in TMEM
    uint64_t* accb = opnd + TILES;                   // [TILES] the layer's MMAs have completed
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accb + TILES);

    if (tid == 0) {
        mbar_init(wfull, 1);
        for (int t = 0; t < TILES; ++t) { mbar_init(&opnd[t], EW * 32); mbar_init(&accb[t], 1); }
        mbar_fence_init();
        // weights -> shared memory, once (hi blocks, and lo blocks for tf32x3)
        uint32_t bytes = 0;
        for (int l = 0; l <= NL; ++l) bytes += (l == 0 ? 8 : HID) * HID * 4u * (X3 ? 2u : 1u);
        mbar_expect_tx(wfull, bytes);
        for (int l = 0; l <= NL; ++l) {
            const uint32_t blk = (l == 0 ? 8 : HID) * HID, n = blk * (X3 ? 2u : 1u);
            for (uint32_t o = 0; o < n; o += 4096u)
                bulk_g2s(sm + L.w + m.w_smem[l] + o, m.blob + m.w_off[l] + o, min(4096u, n - o) * 4u, wfull);
        }
    }
    if (warp == TILES * EW) { umma::tmem_alloc(tmem_slot, 512); umma::tmem_relinquish(); }
    for (int i = tid; i < (NL + 1) * HID; i += blockDim.x) sm[L.bias + i] = m.blob[m.bias_off + i];
    for (int i = tid; i < HID * 8 + 8; i += blockDim.x) sm[L.out + i] = m.blob[m.out_off + i];
    for (int i = tid; i < 28; i += blockDim.x) sm[L.norm + i] = p.io.norm[i];       // f_mean8 f_std8 l_mean6 l_std6
    if (p.nz.enabled)
        for (int i = tid; i < p.nz.blob_floats; i += blockDim.x) sm[L.nblob + i] = p.nz.blob[i];
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tbase = *tmem_slot;
    const int tpc = m.tiles_per_cta;
    const int ngroups = (p.io.ntiles + tpc - 1) / tpc;                 // tile groups, one per CTA iteration

    constexpr int NEPI = TILES * EW;
    if (warp >= NEPI) {
        // ============================================================== MMA issuer of tile slot (warp - NEPI)
        const int ts = warp - NEPI;
        const uint32_t tcol = tbase + ts * TCOLS;
        const uint32_t wbase = smem_u32(sm + L.w);
        uint32_t par = 0;
        mbar_wait(wfull, 0);
        if (ts < tpc)
            for (int g = blockIdx.x; g < ngroups; g += gridDim.x) {
                if (g * tpc + ts >= p.io.ntiles) break;
                for (int t = 0; t < H; ++t)
                    for (int l = 0; l <= NL; ++l) {
                        mbar_wait(&opnd[ts], par);
                        par ^= 1u;
                        umma::fence_after_sync();
                        if (elect_one()) {
                            const uint32_t sb = wbase + m.w_smem[l] * 4u;
                            if (l == 0) mc_issue_layer<X3, 8>(tcol, sb);
                            else mc_issue_layer<X3, HID>(tcol, sb);
                            umma::mma_commit(&accb[ts]);
                        }
                        __syncwarp();
                    }
            }
    } else {
        // ============================================================== epilogue warps: two threads per trajectory
        const int ts = warp / EW, we = warp % EW;
        const int q = we & 3, uh = we >> 2;                 // TMEM lane quarter, hidden-unit half
        const int row = q * 32 + lane;
        const int u0 = uh * UPT;
        const bool owner = uh == 0;
        const uint32_t tlane = tbase + ts * TCOLS + (static_cast<uint32_t>(q * 32) << 16);
        const float* const normS = sm + L.norm;
        const float* const outS = sm + L.out;
        float* const ypart = sm + L.ypart + ts * 6 * TN;
        // residual-noise net (noise_model.py:34-40, 86-103): per-tile buffers, evaluated by the tile's 256 epilogue threads
        float* const nin = sm + L.ntile + ts * L.ntile_floats;
        float* const nact0 = nin + p.nz.in_dim * TN;
        const int nact_flip = p.nz.size * TN;
        float* const nzb = nact0 + 2 * nact_flip;
        const int etid = tid - ts * EW * 32;
        uint32_t par = 0;
        // One-pass TF32: the bias is not added in the epilogue -- D is pre-loaded with it (4 tcgen05.st of 16 columns instead of 64
        // FADDs per thread and layer) and the layer's MMAs accumulate on top; the epilogue's per-unit work is then one VIADDMNMX.
        auto put_bias = [&](int l, int b_lo, int b_hi) {
            const float4* bl = reinterpret_cast<const float4*>(sm + L.bias + l * HID + u0);
            for (int b = b_lo; b < b_hi; ++b) {
                uint32_t r[16];
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    const float4 bb = bl[b * 4 + i / 4];
                    r[i] = __float_as_uint(bb.x); r[i + 1] = __float_as_uint(bb.y); r[i + 2] = __float_as_uint(bb.z); r[i + 3] = __float_as_uint(bb.w);
                }
                umma::tmem_st16(tlane + MC_D + u0 + b * 16, r);
            }
        };
        if (ts < tpc)
            for (int g = blockIdx.x; g < ngroups; g += gridDim.x) {
                const int tile = g * tpc + ts;
                if (tile >= p.io.ntiles) break;
                long long n = (long long)tile * TN + row;
                const bool valid = n < N;
                if (n >= N) n = N - 1;
                float s[6], cost = 0.f;
#pragma unroll
                for (int d = 0; d < 6; ++d) s[d] = owner ? __ldg(p.io.init + n * p.io.init_mul * 6 + d) : 0.f;
                float2 unext = owner ? load_control(p.io, n, 0) : make_float2(0.f, 0.f);
                for (int t = 0; t < H; ++t) {
                    const float u[2] = {unext.x, unext.y};
                    if (owner) {
                        if (t + 1 < H) unext = load_control(p.io, n, t + 1);        // prefetch: hidden behind this step's layers
                        // z-score ([s,u] - f_mean) / f_std  (basic_model.py:34) -> A operand columns 0..7
                        float xn[8];
#pragma unroll
                        for (int k = 0; k < 6; ++k) xn[k] = __fdiv_rn(__fsub_rn(s[k], normS[k]), normS[8 + k]);
#pragma unroll
                        for (int k = 0; k < 2; ++k) xn[6 + k] = __fdiv_rn(__fsub_rn(u[k], normS[6 + k]), normS[14 + k]);
                        uint32_t hi[8], lo[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            hi[i] = X3 ? umma::tf32_round_fast(xn[i]) : __float_as_uint(xn[i]) + 0x1000u;
                            lo[i] = X3 ? __float_as_uint(xn[i] - __uint_as_float(hi[i])) : 0u;
                        }
                        umma::tmem_st8(tlane + MC_A_HI, hi);
                        if (X3) umma::tmem_st8(tlane + MC_A_LO, lo);
                    }
                    if (!X3) put_bias(0, 0, UPT / 16);                   // D <- bias of the input layer (this thread's columns)
                    umma::wait_st();
                    umma::fence_before_sync();
                    mbar_arrive(&opnd[ts]);
                    if (p.nz.enabled) {
                        // The net sees the history up to (s_t, u_t) only, so it runs here, under the MMAs of the input layer.
                        if (owner) noise_push(nin, TN, row, p.nz.hist, t == 0, s, u);
                        named_bar_sync(1 + ts, EW * 32);
                        const float* W = sm + L.nblob;
                        int K = p.nz.in_dim;
                        for (int ph = 0; ph <= p.nz.nlayers; ++ph) {
                            const bool lastp = ph == p.nz.nlayers;
                            const int outw = lastp ? 12 : p.nz.size;
                            dense_resident<4, 4>(ph == 0 ? nin : nact0 + ((ph - 1) & 1) * nact_flip, K, W, outw, outw,
                                                 lastp ? ACT_IDENTITY : p.nz.act, lastp ? nzb : nact0 + (ph & 1) * nact_flip, TN, etid, EW * 32);
                            W += (K + 1) * outw; K = p.nz.size;
                            named_bar_sync(1 + ts, EW * 32);
                        }
                    }
                    float y[6];
#pragma unroll
                    for (int d = 0; d < 6; ++d) y[d] = owner ? outS[HID * 8 + d] : 0.f;
                    for (int l = 0; l <= NL; ++l) {
                        const bool last = l == NL;
                        const float4* bl = reinterpret_cast<const float4*>(sm + L.bias + l * HID + u0);
                        mbar_wait(&accb[ts], par);
                        par ^= 1u;
                        umma::fence_after_sync();
#pragma unroll
                        for (int b = 0; b < UPT / 16; ++b) {
                            float v[16];
                            umma::tmem_ld16(tlane + MC_D + u0 + b * 16, v);
                            umma::wait_ld();
                            if (!X3 && !last) put_bias(l + 1, b, b + 1);     // these D columns are read: the next layer's bias goes in
                            if (!X3 && !last) {
                                // one-pass TF32, hidden layer: ReLU and the operand's rounding in ONE ALU instruction (the epilogue is
                                // bound by the half-rate ALU, profiles/r02/mlp_tc4): max(bits(t) + 0x1000, 0x1000) as signed integers
                                // is bits(relu(t)) + 0x1000 for every finite t (negative t and -0 give 0x1000, which the tensor core
                                // truncates to +0 like the bits of relu(t) = 0 plus the rounding constant): the same operand bits
                                uint32_t hi[16];
#pragma unroll
                                for (int i = 0; i < 16; ++i) hi[i] = (uint32_t)__viaddmax_s32(__float_as_int(v[i]), 0x1000, 0x1000);
                                umma::tmem_st16(tlane + MC_A_HI + u0 + b * 16, hi);
                                continue;
                            }
#pragma unroll
                            for (int i = 0; i < 16; i += 4) {
                                const float4 bb = X3 ? bl[b * 4 + i / 4] : make_float4(0.f, 0.f, 0.f, 0.f);   // TF32: D already holds the bias
                                v[i] = fmaxf(v[i] + bb.x, 0.f); v[i + 1] = fmaxf(v[i + 1] + bb.y, 0.f);
                                v[i + 2] = fmaxf(v[i + 2] + bb.z, 0.f); v[i + 3] = fmaxf(v[i + 3] + bb.w, 0.f);
                            }
                            if (!last) {
                                mc_stage16<X3>(tlane + MC_A_HI + u0 + b * 16, tlane + MC_A_LO + u0 + b * 16, v);
                            } else {                                     // output layer 64 -> 6 in FP32, this thread's 32 units
#pragma unroll
                                for (int i = 0; i < 16; ++i) {
                                    const float4 w0 = *reinterpret_cast<const float4*>(outS + (u0 + b * 16 + i) * 8);
                                    const float2 w1 = *reinterpret_cast<const float2*>(outS + (u0 + b * 16 + i) * 8 + 4);
                                    y[0] = fmaf(v[i], w0.x, y[0]); y[1] = fmaf(v[i], w0.y, y[1]); y[2] = fmaf(v[i], w0.z, y[2]);
                                    y[3] = fmaf(v[i], w0.w, y[3]); y[4] = fmaf(v[i], w1.x, y[4]); y[5] = fmaf(v[i], w1.y, y[5]);
                                }
                            }
                        }
                        if (!last) {
                            umma::wait_st();
                            umma::fence_before_sync();
                            mbar_arrive(&opnd[ts]);
                        } else {
                            umma::fence_before_sync();               // D reads are ordered before the next step's MMAs
                        }
                    }
                    // the unit-half-1 thread hands its partial output sums to the trajectory's owner thread
                    if (EW == 8) {
                        if (!owner) {
#pragma unroll
                            for (int d = 0; d < 6; ++d) ypart[d * TN + row] = y[d];
                        }
                        named_bar_sync(1 + ts, EW * 32);
                    }
                    if (owner) {
                        if (EW == 8) {
#pragma unroll
                            for (int d = 0; d < 6; ++d) y[d] += ypart[d * TN + row];
                        }
                        float nxt[6], nzr[12];
                        if (p.nz.enabled) {
#pragma unroll
                            for (int i = 0; i < 12; ++i) nzr[i] = nzb[i * TN + row];
                        }
                        owner_finish(p, normS + 16, y, nzr, 1, 0, n, valid, t, u, nxt, cost);
#pragma unroll
                        for (int d = 0; d < 6; ++d) s[d] = nxt[d];
                    }
                }
                if (owner && valid && p.io.costs) p.io.costs[n] = cost;
            }
    }
    // ------------------------------------------------------------------ teardown
    umma::fence_before_sync();
    __syncthreads();
    if (warp == TILES * EW) umma::tmem_dealloc(tbase, 512);
}

cudaError_t launch_mlp_tc(const McRolloutParams& p, int grid, int smem_bytes, cudaStream_t st) {
    auto launch = [&](auto k, int threads) {
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
        if (e != cudaSuccess) return e;
        k<<<grid, threads, smem_bytes, st>>>(p);
        return cudaGetLastError();
    };
    if (p.m.tiles_per_cta == 4) {
        if (p.m.prec != PREC_TF32 || p.nz.enabled) return cudaErrorInvalidConfiguration;
        return launch(rollout_mlp_tc_kernel<PREC_TF32, 4, 4>, mc_threads(4));
    }
    if (p.m.prec == PREC_TF32X3) return launch(rollout_mlp_tc_kernel<PREC_TF32X3, 2, 8>, mc_threads(2));
    if (p.m.prec == PREC_TF32) return launch(rollout_mlp_tc_kernel<PREC_TF32, 2, 8>, mc_threads(2));
    return cudaErrorInvalidConfiguration;
}

}  // namespace mpcdl
