/* mpcdl.h -- C ABI of the fused B200 (sm_100a) rollout kernels for the learned BARC vehicle-dynamics
 * models of itsjoshzhang/MPC-deep-learning-sp24.
 *
 * The reference exposes this path only as Python nn.Modules (it has no C ABI and no batched or
 * multi-step rollout).  Each entry point below names the reference interface
 * it replaces (file:line relative to the reference repo).  All pointers marked "device" are CUDA
 * device pointers owned by the caller; the library borrows them for the duration of the call,
 * launches on the given stream and returns without synchronising.  Status: 0 = ok, < 0 = error
 * (text via mpcdl_last_error(), thread-local).  There is no CPU fallback.
 *
 * Row / state conventions (scripts/setup_data.py:70-80):
 *   frame   = [x, y, psi, w_psi, v_long, v_tran, u_a, u_steer]   state = frame[0:6], control = frame[6:8]
 *   noise q = [v_long, v_tran, w_psi, x, y, psi]                  (model_scripts/noise_model.py:108)
 */
#ifndef MPCDL_H
#define MPCDL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPCDL_VERSION 105

enum { MPCDL_MLP = 0, MPCDL_GRU = 1, MPCDL_LSTM = 2 };
/* Arithmetic of the gate GEMMs.  FP32 = FFMA kernels (the parity path, <= 1e-5 one-step).  The tensor-core
 * modes run the [128 x 64].[64 x 64|192|256] layer / gate contractions on tcgen05 (hidden 64; GRU/LSTM window 4 or 8,
 * window 8 needs the device scratch):
 * TF32X3 = three TF32 passes on hi/lo split operands (FP32-class accuracy, bound stated in DESIGN.md 5),
 * TF32 = one TF32 pass + MUFU.TANH activations.  State, accumulation, statistics and cost stay FP32.
 * BF16 = the wide LSTM (hidden 256; stateless or stateful, with or without the noise net): gate GEMMs [128 x 256].[256 x 1024] on
 * tcgen05 with BF16 operands, FP32 accumulation, FP32 cell state; needs a device scratch (mpcdl_rollout_scratch_bytes). */
enum { MPCDL_PREC_FP32 = 0, MPCDL_PREC_TF32X3 = 1, MPCDL_PREC_TF32 = 2, MPCDL_PREC_BF16 = 3 };
enum { MPCDL_ACT_IDENTITY = 0, MPCDL_ACT_RELU = 1, MPCDL_ACT_TANH = 2, MPCDL_ACT_SIGMOID = 3 };
enum {
    MPCDL_OK = 0,
    MPCDL_ERR_ARG = -1,        /* bad shape / null pointer / unsupported size           */
    MPCDL_ERR_UNSUPPORTED = -2,/* hidden size / layer count without a compiled kernel    */
    MPCDL_ERR_CUDA = -3,       /* CUDA runtime error (text has cudaGetErrorString)       */
    MPCDL_ERR_DEVICE = -4      /* not an sm_100 device                                   */
};

/* Architecture of a dynamics model.
 *   MLP : Basic_Model(inputs=8, hidden, output=6, layers, do_drop)    model_scripts/basic_model.py:45-56
 *   GRU : GRU_Model(8, hidden, 6, layers, state_f)                    model_scripts/deep_models.py:46-55
 *   LSTM: LSTM_Model(8, hidden, 6, layers, state_f)                   model_scripts/deep_models.py:70-79
 * window = FT_SIZE frames per sample for GRU/LSTM (deep_models.py:99), 0 for the MLP. */
typedef struct {
    int32_t kind, hidden, layers, window;
} mpcdl_model_desc;

/* Size in floats of the packed weight blob for `desc`, or 0 if unsupported. */
size_t mpcdl_blob_floats(const mpcdl_model_desc* desc);

/* Pack a Basic_Model state_dict (host pointers, torch layout W[out][in], FP32) into `blob` (host).
 * W/b hold layers+2 entries: Linear(8,h), layers x Linear(h,h), Linear(h,6) -- the nn.Linear
 * entries of `self.layers` in order, Dropout entries skipped (basic_model.py:48-56). */
int mpcdl_pack_mlp(const mpcdl_model_desc* desc, const float* const* W, const float* const* b, float* blob);

/* Pack a GRU_Model / LSTM_Model state_dict: w_ih[l] = {gru,lstm}.weight_ih_l{l} [G*h][in_l],
 * w_hh[l] [G*h][h], b_ih[l], b_hh[l] [G*h] (G = 3: r,z,n / 4: i,f,g,o), fc_w = full_c.weight [6][h],
 * fc_b [6]  (deep_models.py:54-55, 78-79; key names in SURVEY.md App. B). */
int mpcdl_pack_rnn(const mpcdl_model_desc* desc, const float* const* w_ih, const float* const* w_hh,
                   const float* const* b_ih, const float* const* b_hh, const float* fc_w, const float* fc_b,
                   float* blob);

/* Tensor-core blob of a GRU_Model / LSTM_Model (same arguments as mpcdl_pack_rnn): every weight matrix is
 * split into tf32 hi/lo parts and laid out in the canonical K-major UMMA core-matrix order, in K-slices
 * the kernel streams with TMA.  Only hidden = 64, window = 4 is supported (0 / MPCDL_ERR_UNSUPPORTED else).
 * mpcdl_blob_floats_tc also sizes the MLP blob of mpcdl_pack_mlp_tc. */
size_t mpcdl_blob_floats_tc(const mpcdl_model_desc* desc);
int mpcdl_pack_rnn_tc(const mpcdl_model_desc* desc, const float* const* w_ih, const float* const* w_hh,
                      const float* const* b_ih, const float* const* b_hh, const float* fc_w, const float* fc_b,
                      float* blob);

/* Tensor-core blob of a Basic_Model (same arguments as mpcdl_pack_mlp; hidden = 64): per layer the matrix as
 * tf32 hi/lo blocks in UMMA operand order; the 64 -> 6 output layer stays FP32. */
int mpcdl_pack_mlp_tc(const mpcdl_model_desc* desc, const float* const* W, const float* const* b, float* blob);

/* Residual-noise net = utils/pytorch_utils.py:19-57 build_mlp((6+2)*history -> size x n_layers -> 12)
 * as used by FeedforwardNoiseModel (model_scripts/noise_model.py:19-40).  W[i] [out][in], b[i] [out],
 * n_layers+1 entries.  Returns floats written (mpcdl_noise_blob_floats gives the size). */
size_t mpcdl_noise_blob_floats(int history, int size, int n_layers);
int mpcdl_pack_noise(int history, int size, int n_layers, const float* const* W, const float* const* b, float* blob);

typedef struct {
    mpcdl_model_desc model;
    int32_t stateful;          /* GRU/LSTM state_f: carry hidden state step to step (deep_models.py:61-65) */
    const float* wblob;        /* device: mpcdl_pack_* output                                              */
    const float* norm;         /* device: f_mean[T*8] f_std[T*8] l_mean[6] l_std[6], T = max(window,1)
                                  (RobotData.norm statistics, basic_model.py:28-35 / deep_models.py:29-36) */
    const float* init;         /* device: s0 [N,6] (MLP) or window [N,T,8] (GRU/LSTM)                      */
    const float* U;            /* device: controls [N,H,2]                                                 */
    float* states;             /* device: out [N,H,6] predicted states (states[:,t] = s_{t+1}), or NULL    */
    float* costs;              /* device: out [N] accumulated cost, or NULL                                */
    const float* cost_ref;     /* device: [H,6] reference path for the quadratic tracking cost, or NULL    */
    float cost_qw[6], cost_rw[2];
    /* residual-noise injection (noise_model.py:86-103): next = pred + mu + softplus(p) * eps */
    const float* noise_wblob;  /* device: mpcdl_pack_noise output, or NULL = no noise                      */
    int32_t noise_history, noise_size, noise_layers, noise_act;
    const float* eps;          /* device: [N,H,6] standard-normal draws, or NULL = in-kernel Philox(seed)   */
    uint64_t seed;
    int64_t traj_offset;       /* global index of trajectory 0 of this shard (Philox key)                  */
    float* carry;              /* device, stateful only: [L][N][hidden] h (+ [L][N][hidden] c for LSTM),
                                  initial state in, final state out (deep_models.py h_state)               */
    int64_t N;
    int32_t H;
    /* tuning overrides, 0 = automatic; tune_kc < 0 forces the generic ring-fed kernel */
    int32_t tune_rt, tune_ntg, tune_nslots, tune_kc;
    int32_t precision;         /* MPCDL_PREC_*; the tensor-core modes need wblob_tc                         */
    const float* wblob_tc;     /* device: mpcdl_pack_rnn_tc output (precision != FP32), else NULL           */
    /* MPPI planner call (generalises run_dynamics_model's feedback loop, scripts/test_model.py:67-87):
     * sample_controls != 0 -> U is the NOMINAL control sequence [H,2] and trajectory n applies
     *   u_t = clamp(U[t] + ctrl_sigma * z, ctrl_lo, ctrl_hi),  z ~ N(0,1) from Philox(ctrl_seed, traj_offset + n, t),
     * generated in registers (no [N,H,2] tensor exists); shared_init != 0 -> init holds ONE state [6] / window [T,8]
     * that every trajectory starts from. */
    int32_t sample_controls, shared_init;
    float ctrl_sigma[2], ctrl_lo[2], ctrl_hi[2];
    uint64_t ctrl_seed;
    /* MPCDL_PREC_BF16, and the TF32 modes with window 8: device workspace of >= mpcdl_rollout_scratch_bytes(args) bytes (the
     * per-CTA layer output sequences: BF16 operand layout / FP32); borrowed for the duration of the launch, contents undefined afterwards. */
    void* scratch;
    uint64_t scratch_bytes;
    /* Stacked models (validation sweep over several checkpoints of one architecture, model_scripts/eval_models.py:14-65
     * run as ONE launch): n_models > 1 -> wblob holds n_models packed blobs back to back (mpcdl_blob_floats each), the
     * N = n_models * (N / n_models) trajectories are model-major (model m owns [m*N/n_models, (m+1)*N/n_models)) and so
     * are states / costs / eps.  share_inputs != 0 -> init and U hold ONE set of N / n_models trajectories that every
     * model rolls out.  FP32 precision only, stateless, no MPPI sampling; 0 and 1 both mean a single model. */
    int32_t n_models, share_inputs;
} mpcdl_rollout_args;

/* The fused closed loop.  Replaces, batched over N trajectories and H steps in ONE launch:
 *   Basic_Model.forward   basic_model.py:58-61      GRU_Model.forward  deep_models.py:60-66
 *   LSTM_Model.forward    deep_models.py:85-92      z-score            basic_model.py:34 / deep_models.py:35
 *   FeedforwardNoiseModel.forward noise_model.py:34-40, history buffering noise_model.py:95-103
 *   the feedback loop of run_dynamics_model(open_loop=True)  scripts/test_model.py:67-87
 * H = 1 with identity statistics is the Module.forward used by eval_models.calc_error
 * (model_scripts/eval_models.py:86). `stream` is a cudaStream_t (NULL = default stream). */
int mpcdl_rollout(const mpcdl_rollout_args* args, void* stream);

/* mpcdl_rollout for callers that may not allocate device memory (CUDA graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same arguments, results and
 * status codes as mpcdl_rollout, for the one kernel family whose scratch it replaces: GRU / LSTM, hidden 64, window 8,
 * precision MPCDL_PREC_TF32X3 or MPCDL_PREC_TF32.  args->scratch / scratch_bytes are ignored: the per-CTA layer output
 * sequences live in a static device array inside the library (16 MB on every GPU that loads it, held until exit), which
 * bounds a call to N <= MPCDL_STATIC_SCRATCH_MAX_TRAJ trajectories; larger N and other shapes return MPCDL_ERR_UNSUPPORTED.
 * The call allocates and frees nothing.  All calls on a device share the array (contents of the previous call remain, none
 * of it needs clearing): two calls that overlap on the GPU (different streams) race; serialise them. */
#define MPCDL_STATIC_SCRATCH_MAX_TRAJ 8192
int mpcdl_rollout_static_scratch(const mpcdl_rollout_args* args, void* stream);

/* Bytes of args->scratch that mpcdl_rollout needs for these args (0 unless precision = MPCDL_PREC_BF16,
 * or a TF32 mode with model.window = 8). */
size_t mpcdl_rollout_scratch_bytes(const mpcdl_rollout_args* args);

/* What mpcdl_rollout would launch for these args (no launch): tile size, grid, smem, config. */
typedef struct {
    int32_t rt, rj, n_tg, tile_traj, threads, grid, smem_bytes, resident, nslots, slot_floats, ntiles;
    int32_t variant;           /* 0 = generic ring-fed kernel, 2 = weight-resident MLP kernel,
                                  4 = tcgen05 GRU/LSTM kernel, 5 = tcgen05 MLP kernel, 6 = BF16 tcgen05 LSTM-256                     */
    int32_t ksplit;            /* K slices of a hidden layer (weight-resident MLP kernel)                   */
} mpcdl_plan;
int mpcdl_rollout_plan(const mpcdl_rollout_args* args, mpcdl_plan* plan);

/* Stand-alone residual-noise net (the reference's FeedforwardNoiseModel call surface).
 *   closed_loop = 0: FeedforwardNoiseModel.forward / get_prediction  model_scripts/noise_model.py:34-45
 *       q [N, 6*history] (history oldest first, q order), u [N, 2*history]  ->  out [N,6] = mu + softplus(p) * eps,
 *       optional mu / scale [N,6] (the Normal's parameters);  H must be 1, eps is [N,6].
 *   closed_loop = 1: DynamicsNN.get_prediction (noise_model.py:113-125) fed back on itself for H steps, i.e. the
 *       open-loop branch of run_dynamics_model (scripts/test_model.py:67-87) batched over N trajectories:
 *       q [N,6] initial state (q order), u [N,H,2] controls  ->  out [N,H,6] (out[:,t] = q_{t+1});
 *       the history deques start left-padded with (q0, u0) (noise_model.py:119-121); eps is [N,H,6].
 * eps = NULL draws Philox(seed, traj_offset + n, step0 + t), the same stream as mpcdl_rollout. */
typedef struct {
    const float* noise_wblob;  /* device: mpcdl_pack_noise output */
    int32_t history, size, n_layers, act;
    int32_t closed_loop;
    const float* q;            /* device */
    const float* u;            /* device */
    const float* eps;          /* device or NULL */
    uint64_t seed;
    int64_t traj_offset;
    int32_t step0;
    float* out;                /* device */
    float* mu;                 /* device or NULL (closed_loop = 0 only) */
    float* scale;              /* device or NULL (closed_loop = 0 only) */
    int64_t N;
    int32_t H;
} mpcdl_noise_args;
int mpcdl_noise_forward(const mpcdl_noise_args* args, void* stream);

/* MPPI update from the costs of sampled rollouts (sample_controls above): with w_n = exp(-(cost_n - cmin) / lambda),
 *   sums[0 .. 2H) = sum_n w_n * (u_{n,t} - U[t])   (the sampled controls are re-generated from the same Philox keys),
 *   sums[2H] = sum_n w_n,  sums[2H+1] = cmin.
 * The new nominal sequence is clamp(U + sums[0..2H) / sums[2H]).  Sharded use: all-reduce MIN the local minima
 * (mpcdl_mppi_min), pass the global one as cmin, all-reduce SUM the first 2H+1 entries.  Deterministic (fixed-order
 * two-stage reduction, no atomics).  workspace: device, >= mpcdl_mppi_workspace_floats(H) floats. */
typedef struct {
    const float* costs;        /* device [N] */
    int64_t N, traj_offset;
    int32_t H;
    const float* U;            /* device [H,2] nominal sequence */
    float ctrl_sigma[2], ctrl_lo[2], ctrl_hi[2];
    uint64_t ctrl_seed;
    float lambda;
    const float* cmin;         /* device [1] global minimum cost, or NULL = minimum of these N costs */
    float* workspace;          /* device */
    float* sums;               /* device out [2H+2] */
} mpcdl_mppi_args;
size_t mpcdl_mppi_workspace_floats(int H);
int mpcdl_mppi_min(const float* costs, int64_t N, float* workspace, float* cmin_out, void* stream);
int mpcdl_mppi_reduce(const mpcdl_mppi_args* args, void* stream);

/* Per-sample error metrics of eval_models.calc_error (model_scripts/eval_models.py:91-92):
 * mae[n] = mean_d |output[n,d] - labels[n,d]|, rmse[n] = sqrt(mean_d (output - labels)^2), d < 6; all device, FP32. */
int mpcdl_error_metrics(const float* output, const float* labels, int64_t B, float* mae, float* rmse, void* stream);

/* ---- Data side of the path on the device --------------------------------------------------
 * A dataset is described by the recorded rows of all bags concatenated (FP32: the pickle's doubles rounded to nearest,
 * as torch.tensor(list) does) and, per bag, its first row and its first sample; windows never cross a bag boundary.
 *   window = 0 : basic_model.RobotData  (model_scripts/basic_model.py:13-26)   sample i of a bag = (row i, label row i+1);
 *                                        n_b = len_b - 1
 *   window = T : deep_models.RobotData  (model_scripts/deep_models.py:13-27)   sample i = (rows i..i+T-1, label row i+T);
 *                                        n_b = len_b - T - 1 */
typedef struct {
    const float* rows;            /* device [R,8]                                   */
    const int64_t* bag_row_off;   /* device [n_bags] first row of every bag         */
    const int64_t* samp_off;      /* device [n_bags] first sample of every bag      */
    int32_t n_bags, window;
    int64_t n;                    /* samples in total                               */
} mpcdl_dataset_desc;

/* RobotData.norm statistics (basic_model.py:28-32, deep_models.py:29-33): mean and UNBIASED std over the samples, per
 * window position, written in the layout of mpcdl_rollout_args.norm: f_mean[T*8] f_std[T*8] l_mean[6] l_std[6]
 * (T = max(window,1)); device, FP32.  Two passes in double precision; within one FP32 rounding of the FP64 statistics (std <= 1e-6 relative to torch.std, mean <= 1e-6 of max(|mean|, std) against torch.mean). */
int mpcdl_dataset_stats(const mpcdl_dataset_desc* d, float* stats, void* stream);
/* featrs [n,T,8] (window = 0: [n,8]) and labels [n,6], device.  stats = NULL -> raw values, else the z-score
 * (x - mean) / std of RobotData.norm (basic_model.py:34-35, deep_models.py:35-36: subtract THEN divide). */
int mpcdl_dataset_build(const mpcdl_dataset_desc* d, const float* stats, float* featrs, float* labels, void* stream);
/* Open-loop evaluation inputs cut from the recorded rows, for samples that leave H recorded frames ahead (the caller
 * sizes samp_off for that: n_b = len_b - H (window 0) / len_b - T - H + 1): init [n,6] or [n,T,8], the recorded controls
 * U [n,H,2] and (optional) the recorded states target [n,H,6] the predictions are compared with. */
int mpcdl_rollout_inputs(const mpcdl_dataset_desc* d, int32_t H, float* init, float* U, float* target, void* stream);
/* Error curves of an open-loop validation sweep: pred [n_models,n,H,6] raw predicted states, target [n,H,6],
 * lstat = l_mean[6] l_std[6]; mae / rmse [n_models,H] = mean over the n trajectories of eval_models.calc_error's
 * per-sample metrics (model_scripts/eval_models.py:91-92) on label-normalised values, one value per horizon step. */
int mpcdl_curve_metrics(const float* pred, const float* target, const float* lstat, int32_t n_models, int64_t n, int32_t H,
                        float* mae, float* rmse, void* stream);

/* DynamicsDataset resampling (utils/data_utils.py:21-37 and :52-72), float64 like numpy, bit-exact:
 *   t_i = np.arange(t[0], t[-1], dataset_dt)[i];  x_i = np.interp(t_i, t, x) per channel;
 *   q_out [n_q,6] = q_interp[::step]  (n_q = ceil(n_interp / step));  u_out [n_u,2] = mean of `step` consecutive
 *   u_interp rows (n_u = n_interp / step).  t [M] ascending seconds, q [M,6], u [M,2]; all device pointers. */
typedef struct {
    const double* t;
    const double* q;
    const double* u;
    int64_t M;
    double dataset_dt;
    int32_t step;
    int64_t n_interp;
    double* q_out;
    double* u_out;
} mpcdl_resample_args;
int mpcdl_resample(const mpcdl_resample_args* args, void* stream);

/* ---- Training step of the forward-dynamics MLP ---------------------------------------------------
 * One launch = one EPOCH of train_model (model_scripts/basic_model.py:100-110) for n_models models at once (one CTA per
 * model): for every batch of `batch` samples in the order `perm` gives (DataLoader(shuffle=True), the last batch may be
 * short): forward, nn.L1Loss, backward, torch.optim.Adam step (single-tensor formulas, FP32).  Parameters and both Adam
 * moments are updated in place; `losses` receives the per-batch loss values.  hidden in {16, 32, 64}, layers 1..8. */
typedef struct {
    float* theta;              /* device [n_models, n_params]: Basic_Model.parameters() flattened in torch order          */
    float* adam_m;             /* device [n_models, n_params] exp_avg                                                    */
    float* adam_v;             /* device [n_models, n_params] exp_avg_sq                                                 */
    const float* featrs;       /* device [n_rows, 8] z-scored features (RobotData.featrs), shared by the models          */
    const float* labels;       /* device [n_rows, 6]                                                                     */
    const int32_t* perm;       /* device [n_models, n_samples] row of every training sample of this epoch, in order      */
    float* losses;             /* device [n_models, ceil(n_samples / batch)] or NULL                                     */
    int64_t n_rows;
    int32_t n_models, n_samples, batch, hidden, layers;
    int64_t step0;             /* optimizer steps already taken (Adam's bias correction uses step0 + batch index + 1)    */
    double lr, beta1, beta2, eps;
} mpcdl_train_args;
int64_t mpcdl_train_param_count(int32_t hidden, int32_t layers);
int mpcdl_train_mlp_epoch(const mpcdl_train_args* args, void* stream);
/* ---- Training step of the recurrent models ("fused fwd+bwd for MLP/GRU") ----------------------------
 * One launch = one EPOCH of train_model (model_scripts/deep_models.py:106-146) for n_models models at once (one CTA per
 * model): for every batch of `batch` windows (above 16: micro-batches of 16 whose gradients are summed) in the order `perm` gives (DataLoader(shuffle=False) over the
 * random_split subset; the last batch may be short): forward of GRU_Model / LSTM_Model (deep_models.py:57-66, 81-92) from
 * h0 = 0 (init_hidden runs before every batch, :118-119), nn.L1Loss, back-propagation through the window, torch.optim.Adam
 * step (single-tensor formulas, FP32).  kind 1 = GRU, 2 = LSTM; hidden in {16, 32, 64, 128}; layers 1..8; window 1..8.
 * theta / adam_m / adam_v hold the model's parameters()  flattened in torch order (per layer weight_ih, weight_hh, bias_ih,
 * bias_hh; then full_c.weight, full_c.bias), one model every `param_stride` floats (>= mpcdl_train_rnn_param_count, a
 * multiple of 4; all three 16-byte aligned).  work = caller-allocated scratch of n_models * mpcdl_train_rnn_work_floats
 * floats (16-byte aligned; contents need no initialisation). */
typedef struct {
    float* theta;              /* device [n_models, param_stride]                                                        */
    float* adam_m;             /* device [n_models, param_stride] exp_avg                                                */
    float* adam_v;             /* device [n_models, param_stride] exp_avg_sq                                             */
    const float* featrs;       /* device [n_rows, window, 8] z-scored windows (deep_models.RobotData.featrs)             */
    const float* labels;       /* device [n_rows, 6]                                                                     */
    const int32_t* perm;       /* device [n_models, n_samples] row of every training sample of this epoch, in order      */
    float* losses;             /* device [n_models, ceil(n_samples / batch)] or NULL                                     */
    float* work;               /* device [n_models, mpcdl_train_rnn_work_floats(...)]                                    */
    int64_t n_rows, param_stride;
    int32_t n_models, n_samples, batch, kind, hidden, layers, window;
    int64_t step0;             /* optimizer steps already taken                                                          */
    double lr, beta1, beta2, eps;
} mpcdl_train_rnn_args;
int64_t mpcdl_train_rnn_param_count(int32_t kind, int32_t hidden, int32_t layers);
int64_t mpcdl_train_rnn_work_floats(int32_t kind, int32_t hidden, int32_t layers, int32_t window);
int mpcdl_train_rnn_epoch(const mpcdl_train_rnn_args* args, void* stream);
/* Validation loss of basic_model.py:112-122: losses[i] = L1 loss of batch i of `batch` consecutive rows of output / labels [n,6]. */
int mpcdl_l1_batches(const float* output, const float* labels, int64_t n, int32_t batch, float* losses, void* stream);

/* ---- rosbag2 CDR decode ----------------------------------------------------------------------------
 * The `messages.data` blobs of one topic of a rosbag2 sqlite3 file (data/rosbags.py:100-124 fetches them; the reference
 * hands them to rclpy's deserialize_message with mpclab_msgs/msg/VehicleStateMsg) decoded on the device: blob = the
 * messages back to back (each with its 4-byte CDR encapsulation header), offsets [n + 1] their byte offsets.
 * out [n, 9] float64 = [t, x.x, x.y, e.psi, w.w_psi, v.v_long, v.v_tran, u.u_a, u.u_steer] (the fields
 * scripts/setup_data.py:70-80 keeps, + the sort key t); status [n] int32: 0 ok, 1 truncated / not little-endian CDR. */
int mpcdl_cdr_decode(const uint8_t* blob, const int64_t* offsets, int64_t n, double* out, int32_t* status, void* stream);

/* Test aid: fills the shared memory of every SM with NaN patterns (one CTA per SM, maximum dynamic shared memory), so a
 * kernel that reads shared memory it has not written shows up as NaN / changed bits in the determinism tests
 * (compute-sanitizer is not available on the GPU pool). */
int mpcdl_debug_poison_smem(void* stream);

/* FP32 FFMA peak micro-kernel (register-only FMA chains on every SM); used by bench.py to measure
 * the FP32 roofline denominator on the box it runs on.  out: device buffer of >= grid*256 floats.
 * FLOPs per launch = 2 * grid * 256 * iters * 64. */
int mpcdl_ffma_peak(float* out_device, int iters, int grid, void* stream);
/* The same tile issued as packed FFMA2 (fma.rn.f32x2, the form the rollout kernels use); same FLOP count. */
int mpcdl_ffma2_peak(float* out_device, int iters, int grid, void* stream);

const char* mpcdl_last_error(void);
int mpcdl_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MPCDL_H */
