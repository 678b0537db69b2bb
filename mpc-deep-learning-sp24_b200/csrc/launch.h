// Kernel instantiation lists and launcher prototypes (host side of csrc/).
#pragma once
#include <cuda_runtime.h>
#include "params.h"
#include "tc_params.h"
#include "lstm_tc_params.h"

namespace mpcdl {

// (hidden, RT, RJ): RT trajectories x RJ hidden units per thread.  RJ is fixed per (family,
// hidden) because the packed blob interleaves gates in groups of RJ units.
#define MPCDL_MLP_CONFIGS(X) \
    X(16, 2, 2) X(16, 8, 2) X(32, 2, 4) X(32, 8, 4) X(64, 2, 4) X(64, 4, 4) X(64, 8, 4) \
    X(128, 2, 4) X(128, 8, 4) X(256, 2, 8) X(256, 8, 8)
#define MPCDL_RNN_CONFIGS(X) \
    X(16, 2, 4) X(16, 4, 4) X(32, 2, 4) X(32, 4, 4) X(64, 2, 4) X(64, 4, 4) X(64, 8, 4) \
    X(128, 2, 4) X(128, 4, 4) X(256, 2, 4) X(256, 4, 4)

// weight-resident MLP kernel (rollout_mlp2.cu): (hidden, RT, RJ, KSPLIT)
#define MPCDL_MLP2_CONFIGS(X) \
    X(64, 4, 4, 2) X(64, 8, 8, 2) X(64, 8, 8, 1) X(32, 4, 4, 2) X(32, 8, 4, 1) X(16, 4, 4, 2) X(16, 8, 2, 1)

// CTA sizes are kept at 8 or 16 warps INCLUDING the producer warp: a 17th (or 9th) warp would put a
// 5th (3rd) warp on one SM sub-partition and cut the per-thread register budget from 128 to 96.
constexpr int MLP_MAX_THREADS = 512;                 // 480 consumers + producer warp, <= 128 regs
__host__ __device__ constexpr int mlp_max_consumers() { return MLP_MAX_THREADS - 32; }
__host__ __device__ constexpr int rnn_max_consumers(int rt, int rj) { return rt * rj * 4 > 64 ? 224 : 480; }

inline int mlp_rj(int hid) { return hid <= 16 ? 2 : hid <= 128 ? 4 : 8; }
inline int rnn_rj(int) { return 4; }

cudaError_t launch_mlp(const RolloutParams& p, int rt, int rj, int grid, int smem_bytes, cudaStream_t st);
cudaError_t launch_mlp2(const RolloutParams& p, int rt, int rj, int ks, int ng, int grid, int smem_bytes, cudaStream_t st);   // ng = thread groups per CTA
int mlp2_smem_bytes(int blob_floats, int hid, int TN, int ks, int rt, int rj, int ng);
cudaError_t launch_rnn(const RolloutParams& p, int cell, int rt, int rj, int grid, int smem_bytes, cudaStream_t st);
cudaError_t launch_rnn_tc(const TcRolloutParams& p, int grid, int smem_bytes, cudaStream_t st);
// window 8 with the sequences in the library's static device array (grid <= TC_STATIC_CTAS); allocates nothing
cudaError_t launch_rnn_tc_static(TcRolloutParams p, int grid, int smem_bytes, cudaStream_t st);
cudaError_t launch_lstm_tc(const LtRolloutParams& p, int grid, int smem_bytes, cudaStream_t st);
cudaError_t launch_mlp_tc(const McRolloutParams& p, int grid, int smem_bytes, cudaStream_t st);
constexpr int AUX_TN = 128, AUX_THREADS = 256;       // noise-net entry points: 128 samples per CTA tile
int aux_smem_bytes(int blob_floats, int in_dim, int size);
cudaError_t launch_aux(const AuxParams& p, bool closed_loop, int grid, int smem_bytes, cudaStream_t st);
cudaError_t launch_mppi_min(const float* costs, long long N, float* workspace, float* cmin_out, cudaStream_t st);
cudaError_t launch_mppi_reduce(const MppiParams& p, cudaStream_t st);
cudaError_t launch_error_metrics(const float* output, const float* labels, long long B, float* mae, float* rmse, cudaStream_t st);
cudaError_t launch_dataset_stats(const DatasetParams& d, float* stats, cudaStream_t st);
cudaError_t launch_dataset_build(const DatasetParams& d, const float* stats, float* featrs, float* labels, cudaStream_t st);
cudaError_t launch_rollout_inputs(const DatasetParams& d, int H, float* init, float* U, float* target, cudaStream_t st);
cudaError_t launch_curve_metrics(const float* pred, const float* target, const float* lstat, int n_models, long long n, int H,
                                 float* mae, float* rmse, cudaStream_t st);
cudaError_t launch_resample(const ResampleParams& p, cudaStream_t st);
cudaError_t launch_cdr_decode(const unsigned char* blob, const long long* offsets, long long n, double* out, int* status, cudaStream_t st);
int train_smem_bytes(int hidden, int layers);
cudaError_t launch_train_mlp(const TrainParams& p, int n_models, int smem_bytes, cudaStream_t st);
long long train_rnn_param_count(int kind, int hidden, int layers);
long long train_rnn_work_floats(int kind, int hidden, int layers, int window);
int train_rnn_smem_bytes(int kind, int hidden);
cudaError_t launch_train_rnn(const TrainRnnParams& p, int n_models, int num_sms, int smem_bytes, cudaStream_t st);
cudaError_t launch_l1_batches(const float* output, const float* labels, long long n, int batch, float* losses, cudaStream_t st);
cudaError_t launch_poison_smem(int num_sms, cudaStream_t st);
cudaError_t launch_ffma_peak(float* out, int iters, int grid, bool packed, cudaStream_t st);

}  // namespace mpcdl
