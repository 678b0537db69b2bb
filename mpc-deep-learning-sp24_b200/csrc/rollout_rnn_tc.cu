// K4: tensor-core (tcgen05 / TMEM) fused whole-horizon rollout of the GRU / LSTM window models.
//   reference: GRU_Model.forward   model_scripts/deep_models.py:57-66
//              LSTM_Model.forward  model_scripts/deep_models.py:81-92
//              per-position z-score model_scripts/deep_models.py:29-36
//              closed loop         oracle/rollout_ref.py rollout_rnn
// ncu on the FP32 kernel (profiles/r01b/ncu_c3.txt) shows the step is a dense contraction
// ([128 x 64] . [64 x 192|256] per cell, fma pipe 43 % active), so the gate GEMMs move to tcgen05:
//   * one CTA owns a tile of 128 trajectories = the MMA M dimension = one TMEM lane per trajectory;
//   * D[128 x 256] (gate pre-activations, FP32) lives in TMEM columns 0..255; the A operands (layer input
//     x_f and recurrent h_{f-1}, 64 tf32 columns each, hi and lo parts) live in TMEM columns 256..511 and
//     are written by the epilogue threads straight from registers (tcgen05.st) -- activations never
//     touch shared memory;
//   * B = the layer's weight rows, pre-split on the host into tf32 hi/lo and pre-arranged in the canonical
//     K-major UMMA layout, streamed from L2 through a TMA ring (or kept resident per layer when it fits);
//   * precision TF32X3: D = A_lo.B_hi + A_hi.B_lo + A_hi.B_hi (FP32-class accuracy, stated bound in
//     DESIGN.md); precision TF32: one pass + MUFU.TANH activations;
//   * 8 epilogue warps (lane quarter = warp % 4, hidden-unit half = warp / 4) turn D into the new hidden
//     state with the layer's whole output sequence seq[T][32] held in REGISTERS, and re-publish it as
//     the next A operands; one control thread issues TMA and MMA.
#include "owner.cuh"
#include "umma.cuh"
#include "tc_params.h"
#include "launch.h"
#include <mutex>

namespace mpcdl {

enum { CELL_GRU = 0, CELL_LSTM = 1 };

constexpr uint32_t TC_D = 0, TC_AX_HI = 256, TC_AX_LO = 320, TC_AH_HI = 384, TC_AH_LO = 448;

__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// activations on PRE-SCALED arguments (the scale is folded into the bias FMA of the epilogue):
//   accurate: t = -log2e x -> sigmoid = 1/(1 + 2^t);  t = 2 log2e x -> tanh = 1 - 2/(1 + 2^t)   (abs err ~1e-7)
//   FAST    : t = x/2 -> sigmoid = 0.5 tanh(t) + 0.5; t = x -> MUFU.TANH                        (abs err ~5e-4)
template <bool FAST> __device__ __forceinline__ float act_sig_scaled(float t) {
    return FAST ? fmaf(tanh_fast(t), 0.5f, 0.5f) : rcp_approx(1.0f + ex2_approx(t));
}
template <bool FAST> __device__ __forceinline__ float act_tanh_scaled(float t) {
    return FAST ? tanh_fast(t) : fmaf(-2.0f, rcp_approx(1.0f + ex2_approx(t)), 1.0f);
}

// registers -> A operand columns [col, col+16) of this thread's TMEM lane, as tf32 hi (and lo)
template <bool X3>
__device__ __forceinline__ void stage16(uint32_t t_hi, uint32_t t_lo, const float* v) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        // one-pass TF32: the tensor core reads the top 19 bits of the operand, so adding half a tf32 ulp is the whole rounding
        // (the value tf32_round_fast gives, without its mask); tf32x3 needs the masked hi for the exact residual
        hi[i] = X3 ? umma::tf32_round_fast(v[i]) : __float_as_uint(v[i]) + 0x1000u;
        lo[i] = X3 ? __float_as_uint(v[i] - __uint_as_float(hi[i])) : 0u;
    }
    umma::tmem_st16(t_hi, hi);
    if (X3) umma::tmem_st16(t_lo, lo);
}
template <bool X3>
__device__ __forceinline__ void stage8(uint32_t t_hi, uint32_t t_lo, const float (&v)[8]) {
    uint32_t hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        hi[i] = X3 ? umma::tf32_round_fast(v[i]) : __float_as_uint(v[i]) + 0x1000u;
        lo[i] = X3 ? __float_as_uint(v[i] - __uint_as_float(hi[i])) : 0u;
    }
    umma::tmem_st8(t_hi, hi);
    if (X3) umma::tmem_st8(t_lo, lo);
}

// ------------------------------------------------------------------ control thread: ring + MMA issue
// All ring bookkeeping is incremental (slot / parity counters that wrap): the control thread is ONE scalar
// thread, and a 32-bit division per chunk costs more than the chunk's MMAs.
struct TcCtl {
    const TcModel& m;
    float* ring;
    uint64_t *full, *empty;
    uint32_t ld_count = 0, outstanding = 0;             // loads issued; loads issued and not yet released
    uint32_t ld_slot = 0, ld_par = 0, ld_wrapped = 0;   // next load's slot, parity of its empty barrier
    int li_it = 0, li_t = 0, li_l = 0, li_f = 0, li_part = 0, li_c = 0;   // load iterator
    bool li_done;
    int my_tiles, H;
    // alias mode: loads of step s > 0 (tile-major step index) wait until the owner threads released the ring after step s - 1
    uint64_t* ring_free = nullptr;
    uint32_t gate_done = 0;

    __device__ TcCtl(const TcModel& m_, float* ring_, uint64_t* full_, uint64_t* empty_, int my_tiles_, int H_)
        : m(m_), ring(ring_), full(full_), empty(empty_), my_tiles(my_tiles_), H(H_) {
        li_done = my_tiles_ == 0;
    }
    __device__ bool part_used(int f, int part) const { return part == 0 || f > 0 || m.stateful; }
    __device__ int ngroups(const TcPart& p) const { return p.nchunks >= m.group ? p.nchunks / m.group : 1; }
    __device__ void li_advance() {
        const TcPart& p = m.parts[2 * li_l + li_part];
        if (++li_c < ngroups(p)) return;
        li_c = 0;
        if (m.resident) {                       // per layer visit: x chunks, then h chunks (if ever used)
            if (li_part == 0 && (m.window > 1 || m.stateful)) { li_part = 1; return; }
            li_part = 0;
        } else {                                // per cell: x chunks, then h chunks (if used by this frame)
            if (li_part == 0 && part_used(li_f, 1)) { li_part = 1; return; }
            li_part = 0;
            if (++li_f < m.window) return;
            li_f = 0;
        }
        if (++li_l < m.layers) return;
        li_l = 0;
        if (++li_t < H) return;
        li_t = 0;
        if (++li_it < my_tiles) return;
        li_done = true;
    }
    // Issue ring loads while slots are free.  Loads up to index `need` are waited for; loads beyond it are only
    // issued when their slot is already free, so MMA issue never stalls behind MMAs that are still in flight.
    __device__ void top_up(uint32_t need) {
        while (!li_done && outstanding < (uint32_t)m.nslots) {
            if (ring_free) {
                const uint32_t step = (uint32_t)(li_it * H + li_t);
                if (step > gate_done) {
                    const uint32_t par = (step - 1u) & 1u;
                    if (need != 0xFFFFFFFFu && ld_count <= need) mbar_wait(ring_free, par);
                    else if (!mbar_try_wait(ring_free, par)) break;
                    gate_done = step;
                }
            }
            if (ld_wrapped) {
                if (need != 0xFFFFFFFFu && ld_count <= need) mbar_wait(&empty[ld_slot], ld_par);
                else if (!mbar_try_wait(&empty[ld_slot], ld_par)) break;
            }
            const TcPart& p = m.parts[2 * li_l + li_part];
            const uint32_t g = p.nchunks >= m.group ? m.group : 1;          // 16-row chunks in this group
            const float* src = m.blob + p.off + li_c * g * p.chunk_floats;
            float* dst = ring + ld_slot * m.slot_floats;
            if (m.prec == PREC_TF32X3) {                // [hi|lo] x g, contiguous in the blob
                mbar_expect_tx(&full[ld_slot], g * p.chunk_floats * 4u);
                bulk_g2s(dst, src, g * p.chunk_floats * 4u, &full[ld_slot]);
            } else {                                    // hi blocks only, compacted
                const uint32_t hb = p.chunk_floats / 2;
                mbar_expect_tx(&full[ld_slot], g * hb * 4u);
                for (uint32_t i = 0; i < g; ++i) bulk_g2s(dst + i * hb, src + i * p.chunk_floats, hb * 4u, &full[ld_slot]);
            }
            ++ld_count;
            ++outstanding;
            if (++ld_slot == (uint32_t)m.nslots) { ld_slot = 0; ld_par ^= ld_wrapped; ld_wrapped = 1; }
            li_advance();
        }
    }
};

// All MMAs of one ring slot = G weight chunks of KC rows of K (both operands' hi/lo parts): D[:, DCOL..DCOL+N) (+)= A . W^T.
//   GRU D columns: [nx | r | z | nh]; weight_ih rows are packed [n; r; z] -> DCOL 0, weight_hh rows [r; z; n]
//   -> DCOL 64, so each part is ONE N = 192 instruction per k-step (nh is zeroed by the epilogue after it
//   is read, which lets the recurrent part always accumulate).  LSTM: [i | f | g | o], N = 256, DCOL 0.
// Pass order lo.hi, hi.lo, hi.hi: the small terms are summed first.
template <int CELL, bool X3, int PART, int KC, int G>
__device__ __forceinline__ void issue_group(uint32_t tbase, uint32_t sb, uint32_t a_hi, uint32_t a_lo, uint32_t& acc) {
    constexpr int N = (CELL == CELL_GRU ? 3 : 4) * 64;
    constexpr uint32_t LBO = N / 8 * 128, BLK = KC / 4 * LBO;          // one hi (or lo) block of KC rows
    constexpr uint32_t SUB = X3 ? 2 * BLK : BLK;                       // bytes per 16-row chunk inside the slot
    constexpr uint32_t DCOL = (CELL == CELL_GRU && PART == 1) ? 64 : 0;
    constexpr uint32_t IDESC = umma::idesc_tf32(128, N);
    const uint64_t d0 = umma::smem_desc_kmajor(sb, LBO, 128u);
#pragma unroll
    for (int s = 0; s < G; ++s) {
#pragma unroll
        for (int pass = X3 ? 0 : 2; pass < 3; ++pass) {
#pragma unroll
            for (int kk = 0; kk < KC / 8; ++kk) {
                const uint64_t bd = d0 + ((s * SUB + (pass == 1 ? BLK : 0u) + kk * 2u * LBO) >> 4);
                umma::mma_tf32_ts(tbase + TC_D + DCOL, (pass == 0 ? a_lo : a_hi) + s * KC + kk * 8, bd, IDESC, acc);
                acc = 1;
            }
        }
    }
}
template <int CELL, bool X3, int PART>
__device__ __forceinline__ void issue_groups(int G, uint32_t tbase, uint32_t sb, uint32_t a_hi, uint32_t a_lo, uint32_t& acc) {
    if (G == 4) issue_group<CELL, X3, PART, TC_KS, 4>(tbase, sb, a_hi, a_lo, acc);
    else if (G == 2) issue_group<CELL, X3, PART, TC_KS, 2>(tbase, sb, a_hi, a_lo, acc);
    else issue_group<CELL, X3, PART, TC_KS, 1>(tbase, sb, a_hi, a_lo, acc);
}

template <int CELL, int PREC, int T>
__global__ void __launch_bounds__(TC_THREADS, 1) rollout_rnn_tc_kernel(const __grid_constant__ TcRolloutParams p) {
    constexpr bool X3 = PREC == PREC_TF32X3;
    constexpr bool FAST = PREC == PREC_TF32;
    constexpr int HID = TC_HID, TN = TC_TN, CW = TC_EPI_THREADS / 32;      // CW = first warp of the control warpgroup
    extern __shared__ __align__(1024) float sm[];
    const TcModel& m = p.m;
    const bool alias = m.noise_alias != 0 && p.nz.enabled;     // noise-net buffers inside the weight ring (tc_params.h)
    const TcSmem L = tc_layout(m.nslots, m.slot_floats, m.layers, T, p.nz, alias);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int H = p.io.H, NL = m.layers;
    const long long N = p.io.N;
    const bool stateful = m.stateful != 0;

    uint64_t* full = reinterpret_cast<uint64_t*>(sm + L.bars);
    uint64_t* empty = full + TC_MAX_SLOTS;
    uint64_t* opnd_ready = empty + TC_MAX_SLOTS;
    uint64_t* acc_ready = opnd_ready + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_ready + 1);
    uint64_t* ring_free = acc_ready + 2;         // alias mode: the owner threads are done with the noise buffers of a step

    if (tid == 0) {
        for (int i = 0; i < m.nslots; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_init(opnd_ready, TC_EPI_THREADS);
        mbar_init(acc_ready, 1);
        mbar_init(ring_free, TC_TN);
        mbar_fence_init();
    }
    if (warp == CW) { umma::tmem_alloc(tmem_slot, 512); umma::tmem_relinquish(); }
    // tile-invariant constants -> shared memory.  Gate biases are pre-scaled for the exp2-based activations:
    //   sigmoid(g + b) = 1 / (1 + 2^(-log2e g - log2e b)),  tanh(g + b) = 1 - 2 / (1 + 2^(2 log2e g + 2 log2e b));
    //   the MUFU.TANH mode (TF32) uses sigmoid(x) = 0.5 tanh(0.5 x) + 0.5 -> 0.5 b for sigmoid gates, b for tanh gates.
    for (int i = tid; i < NL * HID * 4; i += blockDim.x) {
        const int g = i & 3;
        const bool tanh_gate = CELL == CELL_GRU ? g >= 2 : g == 2;
        const float sc = FAST ? (tanh_gate ? 1.0f : 0.5f) : (tanh_gate ? 2.8853900817779268f : -1.4426950408889634f);
        sm[L.bias + i] = m.blob[m.bias_off + i] * sc;
    }
    for (int i = tid; i < HID * 8 + 8; i += blockDim.x) sm[L.fc + i] = m.blob[m.fc_off + i];
    for (int i = tid; i < 2 * T * 8 + 12; i += blockDim.x) sm[L.norm + i] = p.io.norm[i];
    if (p.nz.enabled && !alias)
        for (int i = tid; i < p.nz.blob_floats; i += blockDim.x) sm[L.nblob + i] = p.nz.blob[i];
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tbase = *tmem_slot;
    const int my_tiles = (p.io.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    if (warp >= CW) {
        // ============================================================== control warpgroup (one active thread)
        // the register pool of a CTA is what it was launched with (threads x 96 here; 168 for 8 epilogue warps): the two
        // setmaxnreg values must add up to at most that, or the .inc never returns
        if (TC_UG == 2) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 32;");
        if (warp == CW && lane == 0) {
            TcCtl c(m, sm + L.ring, full, empty, my_tiles, H);
            if (alias) c.ring_free = ring_free;
            uint32_t use_count = 0, cellpar = 0;
            uint32_t us = 0, up = 0;                         // slot / full-barrier parity of the next chunk to use
            const uint32_t ring_s = smem_u32(sm + L.ring);
            const uint32_t slot_bytes = m.slot_floats * 4u;
            const uint32_t nslots = m.nslots;
            const bool resident = m.resident != 0;
            for (int it = 0; it < my_tiles; ++it)
                for (int t = 0; t < H; ++t)
                    for (int l = 0; l < NL; ++l) {
                        const int G = m.group, kg = G * TC_KS;
                        const int nkx = c.ngroups(m.parts[2 * l]), nkh = c.ngroups(m.parts[2 * l + 1]);
                        const uint32_t layer_s = us, layer_p = up, layer_count = use_count;
                        for (int f = 0; f < T; ++f) {
                            if (resident) { us = layer_s; up = layer_p; use_count = layer_count; }
                            c.top_up(0xFFFFFFFFu);
                            mbar_wait(opnd_ready, cellpar);
                            cellpar ^= 1u;
                            umma::fence_after_sync();
                            uint32_t acc = 0;
                            const bool release = !resident || f == T - 1;
                            const int nch = nkx + (c.part_used(f, 1) ? nkh : 0);
                            for (int ch = 0; ch < nch; ++ch) {
                                c.top_up(use_count);
                                mbar_wait(&full[us], up);
                                umma::fence_after_sync();
                                const uint32_t sb = ring_s + us * slot_bytes;
                                if (ch < nkx) {              // input part: weight_ih_l{l} . x_f (one-pass TF32, window 4: x double-buffered by f & 1)
                                    const uint32_t ax = tbase + TC_AX_HI + ((FAST && T == TC_T) ? (uint32_t)(f & 1) * 64u : 0u);
                                    if (l == 0) issue_group<CELL, X3, 0, 8, 1>(tbase, sb, ax, tbase + TC_AX_LO, acc);
                                    else issue_groups<CELL, X3, 0>(G, tbase, sb, ax + ch * kg, tbase + TC_AX_LO + ch * kg, acc);
                                } else {                     // recurrent part: weight_hh_l{l} . h_{f-1} (always accumulates)
                                    uint32_t acc_h = 1;
                                    const int k0 = (ch - nkx) * kg;
                                    issue_groups<CELL, X3, 1>(G, tbase, sb, tbase + TC_AH_HI + k0, tbase + TC_AH_LO + k0, acc_h);
                                }
                                if (release) { umma::mma_commit(&empty[us]); --c.outstanding; }
                                ++use_count;
                                if (++us == nslots) { us = 0; up ^= 1u; }
                            }
                            umma::mma_commit(acc_ready);
                        }
                    }
        }
        __syncwarp();
    } else {
        if (TC_UG == 2) asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        else asm volatile("setmaxnreg.inc.sync.aligned.u32 112;");        // 512 x 112 + 128 x 32 = 640 x 96
#include "rollout_rnn_tc_epi.inl"
    }
    // ------------------------------------------------------------------ teardown
    umma::fence_before_sync();
    __syncthreads();
    if (warp == CW) umma::tmem_dealloc(tbase, 512);
}

// ------------------------------------------------------------------ launcher
template <int CELL, int PREC, int T>
static cudaError_t launch_rnn_tc_t(const TcRolloutParams& p, int grid, int smem_bytes, cudaStream_t st) {
    auto k = rollout_rnn_tc_kernel<CELL, PREC, T>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    k<<<grid, TC_THREADS, smem_bytes, st>>>(p);
    return cudaGetLastError();
}
template <int T>
static cudaError_t launch_rnn_tc_w(const TcRolloutParams& p, int grid, int smem_bytes, cudaStream_t st) {
    const bool gru = p.m.cell == CELL_GRU;
    if (p.m.prec == PREC_TF32X3)
        return gru ? launch_rnn_tc_t<CELL_GRU, PREC_TF32X3, T>(p, grid, smem_bytes, st) : launch_rnn_tc_t<CELL_LSTM, PREC_TF32X3, T>(p, grid, smem_bytes, st);
    if (p.m.prec == PREC_TF32)
        return gru ? launch_rnn_tc_t<CELL_GRU, PREC_TF32, T>(p, grid, smem_bytes, st) : launch_rnn_tc_t<CELL_LSTM, PREC_TF32, T>(p, grid, smem_bytes, st);
    return cudaErrorInvalidConfiguration;
}

cudaError_t launch_rnn_tc(const TcRolloutParams& p, int grid, int smem_bytes, cudaStream_t st) {
    if (p.m.window == TC_T) return launch_rnn_tc_w<TC_T>(p, grid, smem_bytes, st);
    if (p.m.window == TC_T_LONG && p.seqbuf) return launch_rnn_tc_w<TC_T_LONG>(p, grid, smem_bytes, st);
    return cudaErrorInvalidConfiguration;
}

// ------------------------------------------------------------------ static scratch (mpcdl_rollout_static_scratch)
// Callers that may not allocate (CUDA graph capture, serving loops) get the window-8 sequences from this array instead of a
// caller buffer.  Every GPU that loads the library holds its 16 MB until exit; every call on a device shares it, so calls
// that overlap on the GPU race.  Nothing in it has to start at zero: a thread only ever loads words it stored earlier in
// the same rollout step (rollout_rnn_tc_epi.inl), so stale contents of the previous call are never read.
static __device__ __align__(256) float g_tc_static_seq[(size_t)TC_STATIC_CTAS * TC_T_LONG * TC_HID * TC_TN];

// In host code `g_tc_static_seq` is only a handle: the device address comes from cudaGetSymbolAddress, once per device, with
// that device current.
static float* g_tc_static_ptr[64];
static std::mutex g_tc_static_mu;

cudaError_t launch_rnn_tc_static(TcRolloutParams p, int grid, int smem_bytes, cudaStream_t st) {
    if (p.m.window != TC_T_LONG || grid < 1 || grid > TC_STATIC_CTAS) return cudaErrorInvalidConfiguration;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    float* ptr;
    {
        std::lock_guard<std::mutex> lk(g_tc_static_mu);
        if (!g_tc_static_ptr[dev]) {
            void* q = nullptr;
            e = cudaGetSymbolAddress(&q, g_tc_static_seq);
            if (e != cudaSuccess) return e;
            g_tc_static_ptr[dev] = static_cast<float*>(q);
        }
        ptr = g_tc_static_ptr[dev];
    }
    p.seqbuf = ptr;
    return launch_rnn_tc_w<TC_T_LONG>(p, grid, smem_bytes, st);
}

}  // namespace mpcdl
