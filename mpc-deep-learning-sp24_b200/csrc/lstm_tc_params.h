// Parameter block, blob layout and shared-memory carve-up of the BF16 tcgen05 LSTM rollout kernel for wide
// hidden states (rollout_lstm_tc.cu; BASELINE.json config 4: LSTM hidden 256).  Host + device.
#pragma once
#include <cstdint>
#include "params.h"

namespace mpcdl {

constexpr int LT_TN = 128;             // trajectories per CTA tile (one TMEM lane per trajectory); a CTA PAIR runs M = 256 MMAs
constexpr int LT_HID = 256;            // hidden units
constexpr int LT_CH = 64;              // hidden units per gate chunk: D chunk = [i|f|g|o] x 64 = 256 TMEM columns
constexpr int LT_NCH = LT_HID / LT_CH; // 4 chunks per cell
constexpr int LT_KS = 64;              // K rows per weight slot: per CTA [128 gate rows][64 k] bf16 = 16 KB = four K=16 MMAs
constexpr int LT_SPC = LT_HID / LT_KS; // slots per 256-deep operand (4)
constexpr int LT_SLOT_BYTES = 128 * LT_KS * 2;        // per CTA of the pair (its half of the 256 gate rows)
constexpr int LT_NSLOTS = 5;           // 80 KB ring: the weight stream is latency-bound (L2 -> SM ~ 1.9 k cycles under load)
constexpr int LT_EPI_WARPS = 16;       // lane quarter = warp % 4, unit group (16 of a chunk's 64) = warp / 4
constexpr int LT_THREADS = (LT_EPI_WARPS + 4) * 32;   // + one control warpgroup: weight producer warp, MMA issuer warp, two idle
                                                      // (setmaxnreg moves registers between whole warpgroups)
constexpr int LT_A_BYTES = LT_TN * LT_HID * 2;        // one A operand tile [128 x 256] bf16 = 64 KB
constexpr int LT_A_LBO = (LT_TN / 8) * 128;           // bytes between 16-B k-chunks of an A tile (2048)
constexpr int LT_B_LBO = (128 / 8) * 128;             // same for a CTA's half of a weight slot (2048)

// blob (float units; the weight region holds bf16 pairs):
//   [bias  L x 1024 : [l][chunk][gate][64]  (b_ih + b_hh)]
//   [fc    256 x 8 (W^T, cols 6..7 zero) + 8 (bias)]        [pad to 32 floats]
//   [layer 0: per chunk {x slot K = 16 ([x_hi | x_lo] against [W | W]) 2 x 4 KB, 4 h slots of 2 x 16 KB}]
//   [layer l > 0: per chunk {4 x slots, 4 h slots}]
// a slot is [CTA rank 2][k/8][n/8][8 rows][8 k] bf16: each CTA of the pair streams ITS half (128 of the chunk's 256 gate
// rows: rank 0 = i, f; rank 1 = g, o) as a canonical K-major no-swizzle UMMA block; row n of chunk c is
// gate n / 64, unit 64 c + n % 64  (torch row gate * 256 + unit, gate order i, f, g, o).
struct LtModel {
    const float* blob;
    uint32_t bias_off, fc_off, w_off;      // float offsets
    int layers, window;
    int stateful;                          // (h, c) of every layer carried from step to step / call to call (deep_models.py:88-91)
};
__host__ __device__ inline uint32_t lt_x0_slot_bytes() { return 128 * 16 * 2; }      // per CTA
__host__ __device__ inline uint32_t lt_layer_bytes(int l) {
    return l == 0 ? LT_NCH * 2 * (lt_x0_slot_bytes() + LT_SPC * LT_SLOT_BYTES) : LT_NCH * 2 * 2 * LT_SPC * LT_SLOT_BYTES;
}
__host__ __device__ inline uint32_t lt_layer_off_bytes(int l) {       // from w_off
    return l == 0 ? 0u : lt_layer_bytes(0) + (uint32_t)(l - 1) * lt_layer_bytes(1);
}
__host__ __device__ inline uint32_t lt_chunk_bytes(int l) { return lt_layer_bytes(l) / LT_NCH; }

struct LtRolloutParams {
    LtModel m;
    NoiseParams nz;             // residual-noise net (noise_model.py:34-40), evaluated on the CUDA cores at the end of every step
    IoParams io;
    unsigned int* dbg;          // development aid (MPCDL_LT_DEBUG builds): host-mapped words, else unused
    unsigned char* scratch;     // per CTA: [T frames] A tiles of 64 KB: the current layer's output sequence, written IN PLACE
                                // over the previous layer's (cell (l, f) overwrites slot f after loading it as its x operand)
};
// + the raw window frames [T][8][128] FP32 and the noise net's history [nin_dim][128] FP32 (both row-private)
__host__ __device__ inline size_t lt_scratch_bytes_per_cta(int window, int nin_dim) {
    return (size_t)window * LT_A_BYTES + (size_t)window * 8 * LT_TN * 4 + (size_t)nin_dim * LT_TN * 4;
}
// noise-net buffers inside the (idle at the end of a step) A_h region: [blob][nin][nact0][nact1][nzb][ybuf], float offsets
struct LtNoiseSmem { int blob, nin, nact0, nact1, nzb, ybuf, total; };
__host__ __device__ inline LtNoiseSmem lt_noise_layout(const NoiseParams& nz) {
    LtNoiseSmem L;
    int o = 0;
    L.blob = o;  o += align_up(nz.blob_floats, 32);
    L.nin = o;   o += nz.in_dim * LT_TN;
    L.nact0 = o; o += nz.size * LT_TN;
    L.nact1 = o; o += nz.size * LT_TN;
    L.nzb = o;   o += 12 * LT_TN;
    L.ybuf = o;  o += 8 * LT_TN;
    L.total = o;
    return L;
}

struct LtSmem {        // byte offsets
    int ax, ah, ring, bias, norm, bars, total;
};
__host__ __device__ inline LtSmem lt_layout(int window) {
    LtSmem L;
    int o = 0;
    L.ax = o;    o += LT_A_BYTES;                       // also: layer-0 tile [0, 4 KB), head partials [8 KB, 17 KB)
    L.ah = o;    o += LT_A_BYTES;
    L.ring = o;  o += LT_NSLOTS * LT_SLOT_BYTES;
    L.bias = o;  o += 1024 * 4;                         // current layer
    L.norm = o;  o += ((2 * window * 8 + 12 + 3) / 4 * 4) * 4;
    L.bars = o;  o += 256;
    L.total = o;
    return L;
}

}  // namespace mpcdl
