// Kernel parameter blocks and the shared-memory carve-up (host + device).
#pragma once
#include <cstdint>
#include "tile.cuh"

namespace mpcdl {

constexpr int MAX_LAYERS = 8;
constexpr int MAX_WINDOW = 8;
constexpr int MAX_PARTS = 2 * MAX_LAYERS + 2;
constexpr int MAX_NOISE_LAYERS = 4;
constexpr int SMEM_LIMIT = 232448;          // 227 KB opt-in maximum per CTA on sm_100

struct NoiseParams {
    const float* blob;       // gmem: per layer W^T[K][ldw] + bias[ldw]
    int enabled, hist, size, nlayers, act;
    int blob_floats;
    int in_dim;              // 8*hist
};

struct IoParams {
    const float* norm;       // f_mean[T*8] f_std[T*8] l_mean[6] l_std[6]   (T = max(window,1))
    const float* init;       // [N,6] or [N,T,8]
    const float* U;          // [N,H,2]
    float* states;           // [N,H,6] or null
    float* costs;            // [N] or null
    const float* cost_ref;   // [H,6] or null
    float qw[6], rw[2];
    const float* eps;        // [N,H,6] or null (-> Philox(seed))
    unsigned long long seed;
    long long traj_offset;   // global index of trajectory 0 (shard invariance of the Philox stream)
    float* carry;            // stateful RNN: [L][N][HID] h (+ [L][N][HID] c for LSTM), in/out
    long long N;
    int H;
    int ntiles;
    // MPPI control sampling (sample != 0): U is the NOMINAL sequence [H,2]; trajectory n uses
    // u_t = clamp(U[t] + sig * z(ctrl_seed, traj_offset + n, t), ulo, uhi), z ~ N(0,1) from Philox (counter word 3 = 2)
    int sample;
    int init_mul;            // 1: init is per trajectory; 0: one initial state / window shared by all trajectories
    float sig[2], ulo[2], uhi[2];
    unsigned long long ctrl_seed;
    // stacked models (validation sweep): N = n_models * npm trajectories, model mi owns [mi*npm, (mi+1)*npm)
    // and tiles [mi*tpm, (mi+1)*tpm); its weights are blob + mi * blob_stride.  share_inputs: every model reads the SAME
    // npm initial states / control sequences (init [npm,..], U [npm,H,2]).  Single model: npm = N, tpm = ntiles.
    long long npm;
    int tpm, share_inputs;
};

struct ModelParams {
    const float* blob;
    PartDesc parts[MAX_PARTS];
    int nparts;
    int hid, layers, window, stateful;
    uint32_t data_off, data_floats;          // ring-streamed region of the blob
    uint32_t blob_stride;                    // floats between the blobs of stacked models (0 = one model)
    int n_tg, TN;                            // trajectory groups / trajectories per CTA tile
    int n_cthreads;                          // consumer threads (multiple of 32); producer warp follows
    int nslots, slot_floats, resident;
};

struct RolloutParams {
    ModelParams m;
    IoParams io;
    NoiseParams nz;
};

// fused training epoch of the forward-dynamics MLP (train_kernels.cu; basic_model.py:100-110): one CTA per model
struct TrainParams {
    float* theta;            // [n_models][n_params] parameters in torch order (per Linear: weight out x in, bias), in/out
    float* adam_m;           // [n_models][n_params] exp_avg, in/out
    float* adam_v;           // [n_models][n_params] exp_avg_sq, in/out
    const float* featrs;     // [n_rows, 8]  (shared by all models)
    const float* labels;     // [n_rows, 6]
    const int* perm;         // [n_models][n_samples] row index of every sample of this epoch, in training order
    float* losses;           // [n_models][ceil(n_samples / batch)] per-batch L1 loss, or null
    int n_samples, batch, hidden, layers;
    long long step0;         // optimizer steps taken before this launch
    double lr, beta1, beta2, eps;
};

// fused training epoch of the recurrent models (train_rnn_kernels.cu; deep_models.py:106-146): one CTA per model
constexpr int TRAIN_RNN_MAXB = 16;     // batch rows per optimizer step (the reference's BATCH_S grid is 8 / 16)
struct TrainRnnParams {
    float* theta;            // [n_models][param_stride] parameters in torch order (per layer w_ih, w_hh, b_ih, b_hh; full_c.weight, .bias), in/out
    float* adam_m;           // [n_models][param_stride] exp_avg, in/out
    float* adam_v;           // [n_models][param_stride] exp_avg_sq, in/out
    const float* featrs;     // [n_rows, window, 8]  (shared by all models)
    const float* labels;     // [n_rows, 6]
    const int* perm;         // [n_models][n_samples]
    float* losses;           // [n_models][ceil(n_samples / batch)] or null
    float* work;             // [n_models][work_floats] scratch (k-major weight copies, saved activations, gate gradients)
    long long work_floats, param_stride;
    int n_samples, batch, kind, hidden, layers, window;
    long long step0;
    double lr, beta1, beta2, eps;
};

// standalone noise-net entry points (aux_kernels.cu): forward (H = 1, explicit history) and DynamicsNN closed loop
struct AuxParams {
    NoiseParams nz;
    const float* q;          // forward: [N, 6*hist] history (oldest first); closed loop: q0 [N,6]
    const float* u;          // forward: [N, 2*hist];                          closed loop: U [N,H,2]
    const float* eps;        // [N,H,6] or null (-> Philox(seed, traj_offset + n, step0 + t))
    unsigned long long seed;
    long long traj_offset;
    int step0;
    float* out;              // [N,H,6]
    float* mu;               // forward only, optional [N,6]
    float* scale;            // forward only, optional [N,6]
    long long N;
    int H;
    int ntiles;
};

// MPPI update (aux_kernels.cu): softmin-weighted sum of the sampled control perturbations
constexpr int MPPI_GRID = 296, MPPI_THREADS = 256, MPPI_MAX_H = 512;
struct MppiParams {
    const float* costs;
    long long N, traj_offset;
    int H;
    const float* U;          // nominal [H,2]
    float sig[2], ulo[2], uhi[2];
    unsigned long long ctrl_seed;
    float inv_lambda;
    const float* cmin;       // device [1]
    float* workspace;        // [MPPI_GRID][2H+2]
    float* sums;             // [2H+2]
};

// dataset build / statistics / evaluation inputs (dataset_kernels.cu): all bags concatenated into one row array
struct DatasetParams {
    const float* rows;              // [R,8]
    const long long* bag_row_off;   // [n_bags]   first row of every bag
    const long long* samp_off;      // [n_bags]   first sample of every bag (windows never cross a bag boundary)
    int n_bags;
    int window;                     // 0 = basic_model.RobotData, >= 1 = deep_models.RobotData(size)
    long long n;                    // samples in total
};
// DynamicsDataset resampling (utils/data_utils.py:21-37, 52-72), float64
struct ResampleParams {
    const double* t;                // [M] seconds, ascending
    const double* q;                // [M,6]
    const double* u;                // [M,2]
    long long M;
    double dt;                      // DATASET_DT
    int step;                       // int(dt // DATASET_DT)
    long long n_q, n_u;             // outputs: q_out [n_q,6] = q_interp[::step], u_out [n_u,2] = pooled u_interp
    double* q_out;
    double* u_out;
};

// ------------------------------------------------------------------ shared-memory layout (float offsets)
struct SmemLayout {
    int bars;        // 2*nslots uint64
    int ring;
    int norm;        // 2*T*8 + 12
    int xin;         // MLP: [8][TN]      RNN: xn [T][8][TN]
    int xwin;        // RNN: raw window [T][8][TN]
    int act0, act1;  // MLP: [HID][TN] x2  RNN: act0 = seq [T][HID][TN], act1 = hc [HID][TN] (stateful)
    int ybuf;        // [8][TN]
    int nblob, nin, nact0, nact1, nzb;   // noise net
    int total;       // floats
};

__host__ __device__ inline int align_up(int x, int a) { return (x + a - 1) / a * a; }

__host__ __device__ inline SmemLayout make_layout(const ModelParams& m, const NoiseParams& nz, bool rnn) {
    SmemLayout L;
    int o = 0;
    const int T = m.window > 0 ? m.window : 1, TN = m.TN;
    L.bars = o;   o += align_up(4 * m.nslots, 32);                       // 2*nslots u64 = 4*nslots floats
    L.ring = o;   o += align_up(m.resident ? (int)m.data_floats : m.nslots * m.slot_floats, 32);
    L.norm = o;   o += align_up(2 * T * 8 + 12, 32);
    L.xin = o;    o += (rnn ? T : 1) * 8 * TN;
    L.xwin = o;   o += rnn ? T * 8 * TN : 0;
    L.act0 = o;   o += (rnn ? T : 1) * m.hid * TN;
    L.act1 = o;   o += (rnn ? (m.stateful ? 1 : 0) : 1) * m.hid * TN;
    L.ybuf = o;   o += 8 * TN;
    L.nblob = o;  o += nz.enabled ? align_up(nz.blob_floats, 32) : 0;
    L.nin = o;    o += nz.enabled ? nz.in_dim * TN : 0;
    L.nact0 = o;  o += nz.enabled ? nz.size * TN : 0;
    L.nact1 = o;  o += nz.enabled ? nz.size * TN : 0;
    L.nzb = o;    o += nz.enabled ? 12 * TN : 0;
    L.total = o;
    return L;
}

}  // namespace mpcdl
