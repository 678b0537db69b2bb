// K10: BF16 tensor-core (tcgen05 / TMEM) fused whole-horizon rollout of the LSTM window model with a WIDE hidden
// state (hidden 256: BASELINE.json config 4, "LSTM hidden 256, BF16 tensor-core path vs FP32 tolerance study").
//   reference: LSTM_Model.forward  model_scripts/deep_models.py:81-92 (ctor 70-79), per-position z-score :29-36,
//              closed loop         oracle/rollout_ref.py rollout_rnn
// The FP32 FFMA kernel needs 3.36 s for 65536 x 100 (14.7 MFLOP per trajectory-step); here the gate GEMMs
// [128 x 256|16] . [256|16 x 1024] run on tcgen05 with BF16 operands and FP32 accumulation:
//   * tile = 128 trajectories = MMA M = one TMEM lane per trajectory; a cell's 1024 gate columns are produced in
//     four chunks of 64 hidden units x [i|f|g|o] = 256 TMEM columns, DOUBLE-BUFFERED (2 x 256 = all 512 columns):
//     the 16 epilogue warps turn chunk c into (c, h) while the tensor core already accumulates chunk c + 1;
//   * both operands come from shared memory (.ss): the A tiles x_f and h_{f-1} ([128 x 256] bf16, 64 KB each, canonical
//     K-major no-swizzle UMMA layout) and the weights, streamed from L2 through a 4-slot TMA ring of [256 x 32] slices;
//   * the cell state c stays FP32 in the epilogue threads' registers (64 per thread); h is rounded to BF16 once and
//     written in operand layout to a per-CTA scratch (L2-resident, 2 x T tiles): it is the next frame's recurrent
//     operand and the next layer's input, re-loaded by TMA -- layer-major order keeps on-chip state to one layer;
//   * layer 0 (K = 8 features) uses K = 16: [x_hi | x_lo] against [W | W], i.e. the inputs keep 16 mantissa bits;
//   * activations on MUFU.TANH (5 per unit); full_c head, de-normalisation, cost and feedback stay FP32.
// Stateful models: (h, c) of every layer come from / go to the carry planes [2L][N][256] (FP32); the carried h is rounded to
// BF16 when it is staged as frame 0's recurrent operand, like every other h.  Residual-noise net: evaluated by the epilogue
// warps on the CUDA cores at the end of a step, its buffers in the A_h region (idle then), its history in the scratch.
#include "owner.cuh"
#include "umma.cuh"
#include "lstm_tc_params.h"
#include "launch.h"

namespace mpcdl {
namespace {

__device__ __forceinline__ float lt_tanh(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float lt_sig(float x) { return fmaf(lt_tanh(0.5f * x), 0.5f, 0.5f); }
__device__ __forceinline__ bool lt_elect() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred != 0;
}
// two floats -> packed bf16x2 (e0 in the low half = lower address)
__device__ __forceinline__ uint32_t lt_pack(float e0, float e1) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(e1), "f"(e0));
    return r;
}
__device__ __forceinline__ float lt_bf16_round(float x) { return __uint_as_float(lt_pack(0.f, x) & 0xFFFF0000u); }
// kind::f16, A = B = BF16 (format 1), D = F32, both K-major, dense
__host__ __device__ constexpr uint32_t lt_idesc(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
__device__ __forceinline__ void lt_mma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void lt_fence_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

__device__ __forceinline__ bool lt_try_wait_cluster(uint64_t* bar, uint32_t parity);
#ifdef MPCDL_LT_DEBUG
// development aid: p.dbg is HOST-mapped pinned memory; a wait that times out leaves (code, count) there before the trap
#define LT_WAIT(bar, par, code, cnt) lt_wait_dbg<false>(bar, par, p.dbg, code, cnt)
#define LT_WAITC(bar, par, code, cnt) lt_wait_dbg<false>(bar, par, p.dbg, code, cnt)
#define LT_MARK(i, v) do { if (p.dbg) { ((volatile unsigned int*)p.dbg)[i] = (v); } } while (0)
template <bool CL>
__device__ __forceinline__ void lt_wait_dbg(uint64_t* bar, uint32_t parity, unsigned int* dbg, unsigned int code, unsigned int cnt) {
    uint32_t spins = 0;
    while (!(CL ? lt_try_wait_cluster(bar, parity) : mbar_try_wait(bar, parity))) {
        if (++spins > (1u << 20)) {
            if (dbg && (threadIdx.x & 31) == 0) { volatile unsigned int* d = dbg; code += 8 * (blockIdx.x & 1); d[16 + 2 * code] = 0xDEAD0000u | blockIdx.x; d[17 + 2 * code] = cnt; __threadfence_system(); }
            __nanosleep(1000000);
            __trap();
        }
    }
}
#else
#define LT_WAIT(bar, par, code, cnt) mbar_wait(bar, par)
#define LT_WAITC(bar, par, code, cnt) mbar_wait(bar, par)      // cta-scope acquire: the data is read by the MMA (async proxy), not by this thread
__device__ __forceinline__ void lt_wait_cluster(uint64_t* bar, uint32_t parity) {      // bounded spin like mbar_wait
    uint32_t spins = 0;
    while (!lt_try_wait_cluster(bar, parity)) {
        if (++spins > (1u << 26)) __trap();
    }
}
#define LT_MARK(i, v) do { } while (0)
#endif

struct LtBars {
    uint64_t full[LT_NSLOTS], empty[LT_NSLOTS], ax_full, ah_full[2], d_full[2], d_empty[2];   // ah_full: units 0..191 | 192..255
    uint32_t tmem_slot;
};

// ---- CTA-pair (cta_group::2) plumbing
__device__ __forceinline__ uint32_t lt_cta_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void lt_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive (default semantics, as CUTLASS ClusterBarrier::arrive: a release.cluster arrive costs ~1200 cycles here) on the mbarrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void lt_remote_arrive(uint64_t* bar, uint32_t cta) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)), "r"(cta)
        : "memory");
}
// wait with cluster-scope acquire (the leader's barriers also collect arrivals of the peer CTA)
__device__ __forceinline__ bool lt_try_wait_cluster(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void lt_mma2(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// all MMAs issued so far by this thread arrive (once) on the mbarrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void lt_commit2(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
                 "h"(static_cast<uint16_t>(3))
                 : "memory");
}
__device__ __forceinline__ void lt_tmem_alloc2(uint32_t* smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void lt_tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

}  // namespace

// One CTA PAIR (cluster of 2, cta_group::2) per two SMs: each CTA owns a tile of 128 trajectories (its A operands, its
// 128 TMEM lanes, its epilogue) and streams HALF of every weight slot; the leader CTA issues M = 256 MMAs that read both
// CTAs' shared memory, so every weight byte fetched from L2 feeds 256 trajectories.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(LT_THREADS, 1) rollout_lstm_tc_kernel(const __grid_constant__ LtRolloutParams p) {
    extern __shared__ __align__(1024) unsigned char smb[];
    const LtModel& m = p.m;
    const int T = m.window, NL = m.layers, H = p.io.H;
    const LtSmem L = lt_layout(T);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = lt_cta_rank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int npt = (p.io.ntiles + 1) >> 1;                            // pair tiles (2 x 128 trajectories)
    LtBars* bars = reinterpret_cast<LtBars*>(smb + L.bars);
    unsigned char* const ax = smb + L.ax;
    unsigned char* const ah = smb + L.ah;
    unsigned char* const ring = smb + L.ring;
    // raw window frames [T][8][128] FP32: row-private (only the owner thread of a trajectory touches its column), kept in
    // the per-CTA scratch (L1/L2) so that shared memory holds a fifth ring slot
    const int nin_dim = p.nz.enabled ? p.nz.in_dim : 0;
    const bool stateful = m.stateful != 0;
    float* const xwin = reinterpret_cast<float*>(p.scratch + (size_t)blockIdx.x * lt_scratch_bytes_per_cta(T, nin_dim) + (size_t)T * LT_A_BYTES);
    float* const biasS = reinterpret_cast<float*>(smb + L.bias);
    const float* const fcG = m.blob + m.fc_off;                         // full_c head [256][8] + bias[8]: read through L1 (once per step)
    float* const normS = reinterpret_cast<float*>(smb + L.norm);
    float* const ypart = reinterpret_cast<float*>(ax + 8192);          // head partials: [3][6][128], beside the layer-0 tile

    if (tid == 0) {
        // the leader's barriers that gate MMA issue also take ONE relayed arrival of the peer CTA per phase
        const uint32_t extra = rank == 0 ? 1u : 0u;
        for (int i = 0; i < LT_NSLOTS; ++i) { mbar_init(&bars->full[i], 1 + extra); mbar_init(&bars->empty[i], 1); }
        mbar_init(&bars->ax_full, 1 + extra);
        mbar_init(&bars->ah_full[0], 1 + extra);
        mbar_init(&bars->ah_full[1], 1 + extra);
        for (int i = 0; i < 2; ++i) { mbar_init(&bars->d_full[i], 1); mbar_init(&bars->d_empty[i], LT_EPI_WARPS * 32 + extra); }
        mbar_fence_init();
    }
    if (warp == LT_EPI_WARPS + 1) lt_tmem_alloc2(&bars->tmem_slot, 512);
    for (int i = tid; i < 2 * T * 8 + 12; i += blockDim.x) normS[i] = p.io.norm[i];
    umma::fence_before_sync();
    __syncthreads();
    lt_cluster_sync();                       // both CTAs' barriers and TMEM exist before any cross-CTA arrive / MMA
    umma::fence_after_sync();
    const uint32_t tbase = bars->tmem_slot;
    const int my_tiles = (npt - pair + npairs - 1) / npairs;
    const unsigned char* const wbase = reinterpret_cast<const unsigned char*>(m.blob + m.w_off);

    if (warp >= LT_EPI_WARPS) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
      if (warp == LT_EPI_WARPS) {
        // ================================================================ weight producer: one thread feeds this CTA's half of the ring
        if (lane == 0) {
            uint32_t cnt = 0;
            auto produce = [&](const unsigned char* src, uint32_t bytes) {      // src = slot base in the blob ([rank][...])
                const uint32_t slot = cnt % LT_NSLOTS;
                if (cnt >= LT_NSLOTS) LT_WAIT(&bars->empty[slot], ((cnt / LT_NSLOTS) - 1u) & 1u, 0, cnt);
                mbar_expect_tx(&bars->full[slot], bytes);
                bulk_g2s(ring + slot * LT_SLOT_BYTES, src + rank * bytes, bytes, &bars->full[slot]);
                ++cnt;
                if (blockIdx.x == 0) LT_MARK(2, cnt);
            };
            for (int it = 0; it < my_tiles; ++it)
                for (int t = 0; t < H; ++t)
                    for (int l = 0; l < NL; ++l)
                        for (int f = 0; f < T; ++f)
                            for (int c = 0; c < LT_NCH; ++c) {
                                const unsigned char* src = wbase + lt_layer_off_bytes(l) + (uint32_t)c * lt_chunk_bytes(l);
                                if (l == 0) { produce(src, lt_x0_slot_bytes()); src += 2 * lt_x0_slot_bytes(); }
                                else { for (int s = 0; s < LT_SPC; ++s) produce(src + s * 2 * LT_SLOT_BYTES, LT_SLOT_BYTES); src += LT_SPC * 2 * LT_SLOT_BYTES; }
                                if (f > 0 || stateful) for (int s = 0; s < LT_SPC; ++s) produce(src + s * 2 * LT_SLOT_BYTES, LT_SLOT_BYTES);
                            }
        }
      } else if (warp == LT_EPI_WARPS + 1) {
        // ================================================================ MMA warp.  Both CTAs walk the same sequence of
        // sync points; the peer (rank 1) only RELAYS each of its local events to the leader's barrier, the leader
        // (rank 0) waits for both and issues the M = 256 MMAs (one elected lane) and the multicast commits.
        constexpr uint32_t IDESC = lt_idesc(256, 256);
        const uint32_t ax_u = smem_u32(ax), ah_u = smem_u32(ah), ring_u = smem_u32(ring);
        uint32_t cnt = 0, dq = 0, axp = 0, ahp = 0;
#ifdef MPCDL_LT_DEBUG
        long long wcyc[6] = {0, 0, 0, 0, 0, 0};
        const long long tstart = clock64();
#endif
        auto await = [&](uint64_t* bar, uint32_t parity, unsigned code, unsigned n) {
#ifdef MPCDL_LT_DEBUG
            const long long t0 = clock64();
            LT_WAITC(bar, parity, code, n);
            wcyc[code] += clock64() - t0;
#else
            LT_WAITC(bar, parity, code, n);
#endif
            if (rank != 0) {
                if (lane == 0) lt_remote_arrive(bar, 0);
                __syncwarp();
            }
        };
        // one ring slot = `nk` K=16 MMAs of A k-chunks [kc0, kc0 + 2 nk) against the slot
        auto slot_mmas = [&](uint32_t a_u, uint32_t kc0, int nk, uint32_t dcol, uint32_t& acc) {
            const uint32_t slot = cnt % LT_NSLOTS;
            await(&bars->full[slot], (cnt / LT_NSLOTS) & 1u, 1, cnt);
            umma::fence_after_sync();
            if (rank == 0 && lt_elect()) {
                for (int kk = 0; kk < nk; ++kk) {
                    const uint64_t ad = umma::smem_desc_kmajor(a_u + (kc0 + 2u * kk) * LT_A_LBO, LT_A_LBO, 128u);
                    const uint64_t bd = umma::smem_desc_kmajor(ring_u + slot * LT_SLOT_BYTES + 2u * kk * LT_B_LBO, LT_B_LBO, 128u);
                    lt_mma2(dcol, ad, bd, IDESC, acc);
                    acc = 1u;
                }
                lt_commit2(&bars->empty[slot]);
            }
            __syncwarp();
            ++cnt;
        };
        for (int it = 0; it < my_tiles; ++it)
            for (int t = 0; t < H; ++t)
                for (int l = 0; l < NL; ++l)
                    for (int f = 0; f < T; ++f)
                        for (int c = 0; c < LT_NCH; ++c) {
                            const uint32_t buf = dq & 1u;
                            if (dq >= 2u) await(&bars->d_empty[buf], ((dq >> 1) - 1u) & 1u, 2, dq);
                            ++dq;
                            umma::fence_after_sync();
                            const uint32_t dcol = tbase + buf * 256u;
                            uint32_t acc = 0u;
                            if (c == 0) { await(&bars->ax_full, axp, 3, dq); axp ^= 1u; umma::fence_after_sync(); }
                            if (l == 0) slot_mmas(ax_u, 0u, 1, dcol, acc);
                            else for (uint32_t s = 0; s < (uint32_t)LT_SPC; ++s) slot_mmas(ax_u, 8u * s, 4, dcol, acc);
                            if (f > 0 || stateful) {
                                for (uint32_t s = 0; s < (uint32_t)LT_SPC; ++s) {
                                    // h_{f-1} (frame 0 of a stateful model: the carried h) arrives in two pieces: units 0..191 while the previous cell's last chunk is still in
                                    // its epilogue, units 192..255 (that chunk's output) right after it
                                    if (c == 0 && s == 0) { await(&bars->ah_full[0], ahp, 4, dq); umma::fence_after_sync(); }
                                    if (c == 0 && s == LT_SPC - 1) { await(&bars->ah_full[1], ahp, 4, dq); ahp ^= 1u; umma::fence_after_sync(); }
                                    slot_mmas(ah_u, 8u * s, 4, dcol, acc);
                                }
                            }
                            if (rank == 0 && lt_elect()) lt_commit2(&bars->d_full[buf]);
                            __syncwarp();
                            if (blockIdx.x == 0 && lane == 0) LT_MARK(1, dq);
                        }
#ifdef MPCDL_LT_DEBUG
        if (blockIdx.x < 2 && lane == 0 && p.dbg) {     // cycles the MMA warp of CTA 0 / 1 spent in each kind of wait
            for (int i = 0; i < 6; ++i) p.dbg[48 + 8 * blockIdx.x + i] = (unsigned int)(wcyc[i] >> 10);
            p.dbg[48 + 8 * blockIdx.x + 6] = (unsigned int)((clock64() - tstart) >> 10);
        }
#endif
      }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");
#include "rollout_lstm_tc_epi.inl"
    }
    // ------------------------------------------------------------------ teardown
    umma::fence_before_sync();
    __syncthreads();
    lt_cluster_sync();                       // the peer may still read this CTA's shared memory / signal its barriers until here
    if (warp == LT_EPI_WARPS + 1) lt_tmem_dealloc2(tbase, 512);
}

cudaError_t launch_lstm_tc(const LtRolloutParams& p, int grid, int smem_bytes, cudaStream_t st) {
    auto k = rollout_lstm_tc_kernel;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    k<<<grid, LT_THREADS, smem_bytes, st>>>(p);      // grid is even: __cluster_dims__(2, 1, 1)
    return cudaGetLastError();
}

}  // namespace mpcdl
