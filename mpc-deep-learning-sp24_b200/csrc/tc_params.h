// Parameter block and shared-memory carve-up of the tcgen05 (tensor-core) GRU/LSTM rollout kernel
// (rollout_rnn_tc.cu).  Host + device.
#pragma once
#include <cstdint>
#include "params.h"

namespace mpcdl {

constexpr int TC_TN = 128;            // trajectories per CTA tile = MMA M (one TMEM lane per trajectory)
constexpr int TC_HID = 64;            // hidden units (the shipped GRU/LSTM checkpoints)
constexpr int TC_T = 4;               // window frames (deep_models.py:99 FT_SIZE): a layer's output sequence stays in registers
constexpr int TC_T_LONG = 8;          // the other ft_size of the reference's training grid (deep_models.py:184): the sequence
                                      //   goes through a per-CTA scratch in global memory (L2), each thread re-reading its own writes
#ifndef MPCDL_TC_UG
#define MPCDL_TC_UG 2
#endif
constexpr int TC_UG = MPCDL_TC_UG;    // hidden-unit groups (2 or 4): an epilogue thread owns one trajectory row x TC_UPT units.  4 groups
constexpr int TC_UPT = TC_HID / TC_UG;//   (16 warps x 16 units, 112 registers) measured no faster than 2 (8 warps x 32 units): profiles/r02/epi16
constexpr int TC_EPI_THREADS = 128 * TC_UG;   // epilogue warps: lane quarter = warp % 4, unit group = warp / 4
constexpr int TC_THREADS = TC_EPI_THREADS + 128;   // + 1 control warpgroup (TMA producer + MMA issuer = one elected lane of its first warp);
                                      //   a whole warpgroup so that setmaxnreg can move its registers to the epilogue warps
constexpr int TC_KS = 16;             // K rows per weight chunk (2 tf32 MMA k-steps)

enum { PREC_FP32 = 0, PREC_TF32X3 = 1, PREC_TF32 = 2 };

// One weight matrix of the packed tensor-core blob: W[N rows = gate units][K] stored as `nchunks`
// K-slices; a chunk is [hi block][lo block], a block is the canonical K-major no-swizzle UMMA layout
// [k/4][n/8][8 rows][4 k] (core matrix = 8 rows x 16 B).  hi = tf32(w), lo = tf32(w - hi).
struct TcPart {
    uint32_t off;            // float offset of chunk 0 in the blob
    uint32_t chunk_floats;   // hi + lo
    uint16_t K, kc;          // total K, K per chunk
    uint16_t nchunks, N;
};

struct TcModel {
    const float* blob;
    uint32_t bias_off;       // [L][HID][4] per-unit gate biases (GRU: b_r, b_z, b_in, b_hn; LSTM: i, f, g, o)
    uint32_t fc_off;         // full_c: [HID][8] (W^T, cols 6..7 zero) then bias[8]
    TcPart parts[2 * MAX_LAYERS];      // [2l] = weight_ih_l{l}, [2l+1] = weight_hh_l{l}
    int cell, prec, layers, window, stateful;
    int nslots, slot_floats, resident; // ring geometry; resident = a layer's chunks stay for all T frames
    int group;                         // 16-row weight chunks per ring slot (1, 2 or 4): one TMA/issue unit
    int noise_alias;                   // the residual-noise net's buffers live INSIDE the weight ring (resident mode): at the end of a
                                       // step every slot is dead until the next step's layer 0 is loaded, see tc_layout
};

struct TcRolloutParams {
    TcModel m;
    IoParams io;
    NoiseParams nz;
    float* seqbuf;           // window > TC_T: [grid][window][TC_HID][TC_TN] layer output sequences, else unused
};

// mpcdl_rollout_static_scratch: the window-8 sequences live in a file-scope __device__ array of rollout_rnn_tc.cu sized for
// this many CTAs (64 x 256 KB = 16 MB per GPU that loads the library) -> at most 64 x 128 = 8192 trajectories per call.
constexpr int TC_STATIC_CTAS = 64;
constexpr long long TC_STATIC_MAX_TRAJ = (long long)TC_STATIC_CTAS * TC_TN;
__host__ __device__ inline bool tc_window_ok(int window) { return window == TC_T || window == TC_T_LONG; }
__host__ __device__ inline size_t tc_seq_bytes_per_cta(int window) { return window > TC_T ? (size_t)window * TC_HID * TC_TN * 4 : 0; }

struct TcSmem {       // float offsets
    int bars, ring, bias, fc, norm, xwin, ypart, nblob, nin, nact0, nact1, nzb, total;
};

// alias: the noise net's weights, inputs, activations and outputs (59 KB for the default net) are carved out of the ring
// instead of their own region, which is what lets a layer's tf32x3 weights (196 KB) stay resident with the noise net on.
__host__ __device__ inline int tc_noise_floats(const NoiseParams& nz) {
    return nz.enabled ? align_up(nz.blob_floats, 4) + (nz.in_dim + 2 * nz.size + 12) * TC_TN : 0;
}
__host__ __device__ inline TcSmem tc_layout(int nslots, int slot_floats, int layers, int window, const NoiseParams& nz, bool alias = false) {
    TcSmem L;
    int o = 0;
    L.bars = o;   o += 64;                                   // full[16] empty[16]... as uint64 + tmem ptr: 256 B
    o = align_up(o, 32);
    // barrier block is 64 floats = 32 uint64: full[<=12], empty[<=12], opnd, acc, tmem slot
    L.ring = o;   o += nslots * slot_floats;
    L.bias = o;   o += layers * TC_HID * 4;
    L.fc = o;     o += TC_HID * 8 + 8;
    L.norm = o;   o += align_up(2 * window * 8 + 12, 4);
    L.xwin = o;   o += window * 8 * TC_TN;
    L.ypart = o;  o += 4 * 6 * TC_TN;                        // head partial sums [unit group][6][TN] (same 12 KB for 2 or 4 groups)
    const int end = o;
    if (alias) o = L.ring;
    L.nblob = o;  o += nz.enabled ? align_up(nz.blob_floats, 4) : 0;
    L.nin = o;    o += nz.enabled ? nz.in_dim * TC_TN : 0;
    L.nact0 = o;  o += nz.enabled ? nz.size * TC_TN : 0;
    L.nact1 = o;  o += nz.enabled ? nz.size * TC_TN : 0;
    L.nzb = o;    o += nz.enabled ? 12 * TC_TN : 0;
    L.total = alias ? end : o;
    return L;
}
constexpr int TC_MAX_SLOTS = 12;

}  // namespace mpcdl

// ------------------------------------------------------------------ tensor-core MLP kernel (rollout_mlp_tc.cu)
namespace mpcdl {

constexpr int MC_TILES = 2;           // trajectory tiles per CTA: the MMA phase of one overlaps the epilogue of the other;
                                      //   8 epilogue warps per tile (lane quarter = warp % 4, unit half = warp / 4: 2 threads per trajectory),
                                      //   256 TMEM columns per tile: D [0,64), A_hi [64,128), A_lo [128,192)
constexpr int MC_TILES_MAX = 4;       // TF32 without the noise net: 4 tiles of 128 columns (D, A_hi), 4 epilogue warps per tile
// threads of a CTA with `tiles` tile slots: the epilogue warps + one MMA-issuing warp per tile
__host__ __device__ constexpr int mc_threads(int tiles) { return tiles == 4 ? (4 * 4 + 4) * 32 : (2 * 8 + 2) * 32; }

// blob: [bias (L+1) x 64][out: W^T [64][8], bias [8]][pad to 32] then per layer l = 0..L the matrix
// W_l [64 rows = units][K = 8 | 64] as [hi block][lo block] in the canonical K-major UMMA order (N = 64).
struct McModel {
    const float* blob;
    uint32_t bias_off, out_off;
    uint32_t w_off[MAX_LAYERS + 1];   // float offset of layer l's hi block (lo block follows)
    uint32_t w_smem[MAX_LAYERS + 1];  // float offset of layer l inside the shared-memory weight region
    uint32_t w_floats;                // shared-memory weight region size
    int layers, prec, tiles_per_cta;
};
struct McRolloutParams {
    McModel m;
    IoParams io;
    NoiseParams nz;                   // residual-noise net on the tile's epilogue warps (CUDA cores), per-tile buffers
};
struct McSmem { int bars, w, bias, out, norm, ypart, nblob, ntile, ntile_floats, total; };
// per-tile noise buffers (ntile + ts * ntile_floats): nin [in_dim][TN], nact0 / nact1 [size][TN], nzb [12][TN]
__host__ __device__ inline McSmem mc_layout(int w_floats, int layers, const NoiseParams& nz, int tiles) {
    McSmem L;
    int o = 0;
    L.bars = o;  o += 64;
    L.w = o;     o += align_up(w_floats, 32);
    L.bias = o;  o += (layers + 1) * TC_HID;
    L.out = o;   o += TC_HID * 8 + 8;
    L.norm = o;  o += 32;
    L.ypart = o; o += MC_TILES * 6 * TC_TN;     // output-layer partial sums of the unit-half-1 threads
    L.nblob = o; o += nz.enabled ? align_up(nz.blob_floats, 4) : 0;
    L.ntile = o;
    L.ntile_floats = nz.enabled ? (nz.in_dim + 2 * nz.size + 12) * TC_TN : 0;
    o += tiles * L.ntile_floats;
    L.total = o;
    return L;
}

}  // namespace mpcdl
