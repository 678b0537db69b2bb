// torch.ops.mpcdl.* -- the compiled operator boundary of the fused rollout path (TORCH_LIBRARY, CUDA key only).
//
// Every operator checks its tensors (CUDA, float32, contiguous, same device, shapes), fills the plain-C argument
// struct of include/mpcdl.h and calls the C ABI on at::cuda::getCurrentCUDAStream().  No torch kernel runs here and
// there is no CPU implementation: CPU tensors fail in the dispatcher.  The ctypes binding (mpcdl/_lib.py) stays for
// hosts without torch and for the packers / planner.
//
// Reference surface replaced: Basic_Model.forward model_scripts/basic_model.py:58-61, GRU_Model.forward /
// LSTM_Model.forward model_scripts/deep_models.py:60-66 / 85-92, FeedforwardNoiseModel.forward model_scripts/noise_model.py:34-40,
// DynamicsNN.get_prediction noise_model.py:113-125, the feedback loop of run_dynamics_model scripts/test_model.py:67-87 and
// the per-sample metrics of eval_models.calc_error model_scripts/eval_models.py:91-92.
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>

#include <cstring>
#include <tuple>

#include "../../include/mpcdl.h"

namespace {

using at::Tensor;
using OptTensor = std::optional<Tensor>;

void check_status(int rc) { TORCH_CHECK(rc == MPCDL_OK, "mpcdl error ", rc, ": ", mpcdl_last_error()); }

// float32 + CUDA + contiguous + on `dev`; returns the raw device pointer
float* f32c(const Tensor& t, const char* name, const c10::Device& dev) {
    TORCH_CHECK(t.scalar_type() == at::kFloat, "mpcdl: ", name, " must be float32, got ", t.scalar_type());
    TORCH_CHECK(t.is_cuda(), "mpcdl: ", name, " must be a CUDA tensor (no CPU fallback), got ", t.device());
    TORCH_CHECK(t.device() == dev, "mpcdl: ", name, " is on ", t.device(), ", expected ", dev);
    TORCH_CHECK(t.is_contiguous(), "mpcdl: ", name, " must be contiguous");
    return static_cast<float*>(t.data_ptr());
}
float* f32c(const OptTensor& t, const char* name, const c10::Device& dev) {
    return t.has_value() && t->defined() ? f32c(*t, name, dev) : nullptr;
}
bool has(const OptTensor& t) { return t.has_value() && t->defined(); }

std::tuple<Tensor, Tensor> rollout(int kind, const OptTensor& wblob, const Tensor& norm, const Tensor& init, const Tensor& U,
                                   int64_t hidden, int64_t layers, int64_t window, bool stateful, const OptTensor& carry,
                                   const OptTensor& noise_wblob, at::IntArrayRef noise_cfg, const OptTensor& eps, int64_t seed,
                                   int64_t traj_offset, const OptTensor& cost_ref, at::ArrayRef<double> cost_w, bool return_states,
                                   bool return_costs, at::IntArrayRef tune, int64_t precision, const OptTensor& wblob_tc,
                                   at::ArrayRef<double> ctrl, int64_t ctrl_seed, int64_t n_samples, bool shared_init,
                                   const OptTensor& out_states, const OptTensor& out_costs, int64_t n_models, bool share_inputs) {
    TORCH_CHECK(U.is_cuda(), "mpcdl: U must be a CUDA tensor (no CPU fallback), got ", U.device());
    const c10::Device dev = U.device();
    const int T = kind == MPCDL_MLP ? 0 : (int)window;
    int64_t N, H;
    if (!ctrl.empty()) {
        // MPPI call: U is the nominal sequence [H,2]; the N sampled sequences exist only in registers
        TORCH_CHECK(ctrl.size() == 6 && U.dim() == 2 && U.size(1) == 2 && n_samples >= 1,
                    "mpcdl: sampled controls need U [H,2], ctrl = sigma[2]+lo[2]+hi[2] and n_samples >= 1; got U ", U.sizes());
        N = n_samples; H = U.size(0);
    } else {
        TORCH_CHECK(U.dim() == 3 && U.size(2) == 2, "mpcdl: U must be [N,H,2], got ", U.sizes());
        N = U.size(0); H = U.size(1);
    }
    const int64_t n_in = N;                               // trajectories the inputs hold
    if (n_models > 1 && share_inputs) N *= n_models;      // stacked models all rolling out the same n_in trajectories
    {
        const int64_t lead = shared_init ? 0 : 1;
        const bool ok = kind == MPCDL_MLP ? (init.dim() == lead + 1 && init.size(lead) == 6)
                                          : (init.dim() == lead + 2 && init.size(lead) == T && init.size(lead + 1) == 8);
        TORCH_CHECK(ok && (shared_init || init.size(0) == n_in), "mpcdl: init must be ", shared_init ? "" : "[N] + ",
                    kind == MPCDL_MLP ? "[6]" : "[window,8]", " with N = ", n_in, ", got ", init.sizes());
    }
    const int nt = T > 0 ? T : 1;
    TORCH_CHECK(norm.numel() == 2 * nt * 8 + 12, "mpcdl: norm must hold ", 2 * nt * 8 + 12, " floats (f_mean,f_std [", nt,
                ",8]; l_mean,l_std [6])");
    mpcdl_rollout_args a;
    std::memset(&a, 0, sizeof(a));
    a.model.kind = kind; a.model.hidden = (int32_t)hidden; a.model.layers = (int32_t)layers; a.model.window = T;
    a.precision = (int32_t)precision;
    if (a.precision != MPCDL_PREC_FP32) {
        const size_t need = mpcdl_blob_floats_tc(&a.model);
        if (need == 0) check_status(MPCDL_ERR_UNSUPPORTED);
        TORCH_CHECK(has(wblob_tc) && (size_t)wblob_tc->numel() == need, "mpcdl: tensor-core precision needs wblob_tc with ", need,
                    " floats (pack_rnn(..., tensor_core=True))");
        a.wblob_tc = f32c(wblob_tc, "wblob_tc", dev);
    } else {
        const size_t need = mpcdl_blob_floats(&a.model) * (size_t)(n_models > 1 ? n_models : 1);
        if (need == 0) check_status(MPCDL_ERR_UNSUPPORTED);
        TORCH_CHECK(has(wblob) && (size_t)wblob->numel() == need, "mpcdl: wblob has ", has(wblob) ? wblob->numel() : 0,
                    " floats, model needs ", need);
        a.wblob = f32c(wblob, "wblob", dev);
    }
    a.stateful = stateful && kind != MPCDL_MLP;
    a.norm = f32c(norm, "norm", dev);
    a.init = f32c(init, "init", dev);
    a.U = f32c(U, "U", dev);
    // caller-owned result buffers (a pipeline that overlaps the copy-out keeps one pair per slot) or fresh ones
    Tensor states, costs;
    if (return_states && has(out_states)) {
        TORCH_CHECK(out_states->dim() == 3 && out_states->size(0) == N && out_states->size(1) == H && out_states->size(2) == 6,
                    "mpcdl: out_states must be [", N, ",", H, ",6], got ", out_states->sizes());
        f32c(*out_states, "out_states", dev);
        states = *out_states;
    } else {
        states = return_states ? at::empty({N, H, 6}, U.options()) : at::empty({0}, U.options());
    }
    if (return_costs && has(out_costs)) {
        TORCH_CHECK(out_costs->dim() == 1 && out_costs->size(0) == N, "mpcdl: out_costs must be [", N, "], got ", out_costs->sizes());
        f32c(*out_costs, "out_costs", dev);
        costs = *out_costs;
    } else {
        costs = return_costs ? at::empty({N}, U.options()) : at::empty({0}, U.options());
    }
    a.states = return_states ? static_cast<float*>(states.data_ptr()) : nullptr;
    a.costs = return_costs ? static_cast<float*>(costs.data_ptr()) : nullptr;
    if (has(cost_ref)) {
        TORCH_CHECK(cost_ref->dim() == 2 && cost_ref->size(0) == H && cost_ref->size(1) == 6 && cost_w.size() == 8,
                    "mpcdl: cost_ref must be [H,6] and cost_w = qw[6] + rw[2]");
        a.cost_ref = f32c(cost_ref, "cost_ref", dev);
    }
    if (cost_w.size() == 8) {
        for (int i = 0; i < 6; ++i) a.cost_qw[i] = (float)cost_w[i];
        for (int i = 0; i < 2; ++i) a.cost_rw[i] = (float)cost_w[6 + i];
    }
    if (has(noise_wblob)) {
        TORCH_CHECK(noise_cfg.size() == 4, "mpcdl: noise_cfg = [history, size, n_layers, act]");
        TORCH_CHECK((size_t)noise_wblob->numel() == mpcdl_noise_blob_floats((int)noise_cfg[0], (int)noise_cfg[1], (int)noise_cfg[2]),
                    "mpcdl: noise_wblob size does not match noise_cfg");
        a.noise_wblob = f32c(noise_wblob, "noise_wblob", dev);
        a.noise_history = (int32_t)noise_cfg[0]; a.noise_size = (int32_t)noise_cfg[1];
        a.noise_layers = (int32_t)noise_cfg[2]; a.noise_act = (int32_t)noise_cfg[3];
        if (has(eps)) {
            TORCH_CHECK(eps->dim() == 3 && eps->size(0) == N && eps->size(1) == H && eps->size(2) == 6,
                        "mpcdl: eps must be [N,H,6], got ", eps->sizes());
            a.eps = f32c(eps, "eps", dev);
        }
    }
    a.seed = (uint64_t)seed;
    a.traj_offset = traj_offset;
    if (a.stateful) {
        const int64_t planes = layers * (kind == MPCDL_LSTM ? 2 : 1);
        TORCH_CHECK(has(carry) && carry->dim() == 3 && carry->size(0) == planes && carry->size(1) == N && carry->size(2) == hidden,
                    "mpcdl: stateful rollout needs carry [", planes, ",", N, ",", hidden, "]");
        a.carry = f32c(carry, "carry", dev);
    }
    a.N = N; a.H = (int32_t)H;
    if (tune.size() == 4) { a.tune_rt = (int32_t)tune[0]; a.tune_ntg = (int32_t)tune[1]; a.tune_nslots = (int32_t)tune[2]; a.tune_kc = (int32_t)tune[3]; }
    if (!ctrl.empty()) {
        a.sample_controls = 1;
        for (int i = 0; i < 2; ++i) { a.ctrl_sigma[i] = (float)ctrl[i]; a.ctrl_lo[i] = (float)ctrl[2 + i]; a.ctrl_hi[i] = (float)ctrl[4 + i]; }
        a.ctrl_seed = (uint64_t)ctrl_seed;
    }
    a.shared_init = shared_init;
    a.n_models = (int32_t)n_models; a.share_inputs = share_inputs;
    c10::cuda::CUDAGuard guard(dev);                     // the launch goes to the tensors' device
    Tensor scratch;
    if (a.precision != MPCDL_PREC_FP32) {
        // per-CTA layer output sequences of the BF16 LSTM-256 kernel and of the window-8 tcgen05 GRU / LSTM kernel; from the
        // caching allocator, so the buffer is tied to the current stream like every other tensor of the call
        const size_t nscratch = mpcdl_rollout_scratch_bytes(&a);
        if (nscratch) {
            scratch = at::empty({(int64_t)nscratch}, U.options().dtype(at::kByte));
            a.scratch = scratch.data_ptr(); a.scratch_bytes = nscratch;
        }
    }
    check_status(mpcdl_rollout(&a, at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return {states, costs};
}

#define MPCDL_ROLLOUT_ARGS                                                                                                          \
    const OptTensor &wblob, const Tensor &norm, const Tensor &init, const Tensor &U, int64_t hidden, int64_t layers, int64_t window, \
        bool stateful, const OptTensor &carry, const OptTensor &noise_wblob, at::IntArrayRef noise_cfg, const OptTensor &eps,        \
        int64_t seed, int64_t traj_offset, const OptTensor &cost_ref, at::ArrayRef<double> cost_w, bool return_states,              \
        bool return_costs, at::IntArrayRef tune, int64_t precision, const OptTensor &wblob_tc, at::ArrayRef<double> ctrl,           \
        int64_t ctrl_seed, int64_t n_samples, bool shared_init, const OptTensor &out_states, const OptTensor &out_costs,            \
        int64_t n_models, bool share_inputs
#define MPCDL_ROLLOUT_PASS                                                                                                          \
    wblob, norm, init, U, hidden, layers, window, stateful, carry, noise_wblob, noise_cfg, eps, seed, traj_offset, cost_ref, cost_w, \
        return_states, return_costs, tune, precision, wblob_tc, ctrl, ctrl_seed, n_samples, shared_init, out_states, out_costs,      \
        n_models, share_inputs

std::tuple<Tensor, Tensor> rollout_mlp(MPCDL_ROLLOUT_ARGS) { return rollout(MPCDL_MLP, MPCDL_ROLLOUT_PASS); }
std::tuple<Tensor, Tensor> rollout_gru(MPCDL_ROLLOUT_ARGS) { return rollout(MPCDL_GRU, MPCDL_ROLLOUT_PASS); }
std::tuple<Tensor, Tensor> rollout_lstm(MPCDL_ROLLOUT_ARGS) { return rollout(MPCDL_LSTM, MPCDL_ROLLOUT_PASS); }

// ---------------------------------------------------------------------------- noise net / error metrics / MPPI update
void noise_args(mpcdl_noise_args& a, const Tensor& noise_wblob, at::IntArrayRef cfg, const c10::Device& dev) {
    std::memset(&a, 0, sizeof(a));
    TORCH_CHECK(cfg.size() == 4, "mpcdl: noise_cfg = [history, size, n_layers, act]");
    TORCH_CHECK((size_t)noise_wblob.numel() == mpcdl_noise_blob_floats((int)cfg[0], (int)cfg[1], (int)cfg[2]),
                "mpcdl: noise_wblob size does not match noise_cfg");
    a.noise_wblob = f32c(noise_wblob, "noise_wblob", dev);
    a.history = (int32_t)cfg[0]; a.size = (int32_t)cfg[1]; a.n_layers = (int32_t)cfg[2]; a.act = (int32_t)cfg[3];
}

// FeedforwardNoiseModel.forward (model_scripts/noise_model.py:34-40): q [B,6*hist], u [B,2*hist] -> [B,6]
std::tuple<Tensor, Tensor, Tensor> noise_forward(const Tensor& noise_wblob, at::IntArrayRef cfg, const Tensor& q, const Tensor& u,
                                                 const OptTensor& eps, int64_t seed, int64_t traj_offset, int64_t step0,
                                                 bool return_params) {
    TORCH_CHECK(q.is_cuda(), "mpcdl: q must be a CUDA tensor (no CPU fallback), got ", q.device());
    const c10::Device dev = q.device();
    mpcdl_noise_args a;
    noise_args(a, noise_wblob, cfg, dev);
    const int64_t hist = a.history;
    TORCH_CHECK(q.dim() == 2 && q.size(1) == 6 * hist && u.dim() == 2 && u.size(0) == q.size(0) && u.size(1) == 2 * hist,
                "mpcdl: noise forward needs q [B,", 6 * hist, "] and u [B,", 2 * hist, "], got ", q.sizes(), " ", u.sizes());
    const int64_t B = q.size(0);
    a.q = f32c(q, "q", dev); a.u = f32c(u, "u", dev);
    if (has(eps)) {
        TORCH_CHECK(eps->dim() == 2 && eps->size(0) == B && eps->size(1) == 6, "mpcdl: eps must be [B,6], got ", eps->sizes());
        a.eps = f32c(eps, "eps", dev);
    }
    a.seed = (uint64_t)seed; a.traj_offset = traj_offset; a.step0 = (int32_t)step0;
    Tensor out = at::empty({B, 6}, q.options());
    Tensor mu = return_params ? at::empty({B, 6}, q.options()) : at::empty({0}, q.options());
    Tensor scale = return_params ? at::empty({B, 6}, q.options()) : at::empty({0}, q.options());
    a.out = static_cast<float*>(out.data_ptr());
    if (return_params) { a.mu = static_cast<float*>(mu.data_ptr()); a.scale = static_cast<float*>(scale.data_ptr()); }
    a.N = B; a.H = 1;
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_noise_forward(&a, at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return {out, mu, scale};
}

// DynamicsNN.get_prediction fed back on itself (noise_model.py:113-125, test_model.py:67-87), fused over H
Tensor dynamics_nn_rollout(const Tensor& noise_wblob, at::IntArrayRef cfg, const Tensor& q0, const Tensor& U, const OptTensor& eps,
                           int64_t seed, int64_t traj_offset) {
    TORCH_CHECK(U.is_cuda(), "mpcdl: U must be a CUDA tensor (no CPU fallback), got ", U.device());
    const c10::Device dev = U.device();
    mpcdl_noise_args a;
    noise_args(a, noise_wblob, cfg, dev);
    TORCH_CHECK(U.dim() == 3 && U.size(2) == 2 && q0.dim() == 2 && q0.size(0) == U.size(0) && q0.size(1) == 6,
                "mpcdl: DynamicsNN rollout needs q0 [N,6] and U [N,H,2], got ", q0.sizes(), " ", U.sizes());
    const int64_t N = U.size(0), H = U.size(1);
    a.closed_loop = 1;
    a.q = f32c(q0, "q0", dev); a.u = f32c(U, "U", dev);
    if (has(eps)) {
        TORCH_CHECK(eps->dim() == 3 && eps->size(0) == N && eps->size(1) == H && eps->size(2) == 6, "mpcdl: eps must be [N,H,6], got ",
                    eps->sizes());
        a.eps = f32c(eps, "eps", dev);
    }
    a.seed = (uint64_t)seed; a.traj_offset = traj_offset;
    Tensor out = at::empty({N, H, 6}, U.options());
    a.out = static_cast<float*>(out.data_ptr());
    a.N = N; a.H = (int32_t)H;
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_noise_forward(&a, at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return out;
}

// per-sample MAE / RMSE of eval_models.calc_error (model_scripts/eval_models.py:91-92)
std::tuple<Tensor, Tensor> error_metrics(const Tensor& output, const Tensor& labels) {
    TORCH_CHECK(output.is_cuda(), "mpcdl: output must be a CUDA tensor (no CPU fallback), got ", output.device());
    const c10::Device dev = output.device();
    TORCH_CHECK(output.dim() == 2 && output.size(1) == 6 && output.sizes() == labels.sizes(),
                "mpcdl: error_metrics needs output and labels [B,6], got ", output.sizes(), " ", labels.sizes());
    const int64_t B = output.size(0);
    Tensor mae = at::empty({B}, output.options()), rmse = at::empty({B}, output.options());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_error_metrics(f32c(output, "output", dev), f32c(labels, "labels", dev), B, static_cast<float*>(mae.data_ptr()),
                                     static_cast<float*>(rmse.data_ptr()), at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return {mae, rmse};
}

Tensor mppi_min(const Tensor& costs) {
    TORCH_CHECK(costs.is_cuda(), "mpcdl: costs must be a CUDA tensor (no CPU fallback), got ", costs.device());
    const c10::Device dev = costs.device();
    Tensor ws = at::empty({(int64_t)mpcdl_mppi_workspace_floats(1)}, costs.options());
    Tensor out = at::empty({1}, costs.options());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_mppi_min(f32c(costs, "costs", dev), costs.numel(), static_cast<float*>(ws.data_ptr()),
                                static_cast<float*>(out.data_ptr()), at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return out;
}

// sums [2H+2] = [sum_n w_n (u_n - U) per (t, channel), sum_n w_n, cmin]; include/mpcdl.h mpcdl_mppi_reduce
Tensor mppi_reduce(const Tensor& costs, const Tensor& U, at::ArrayRef<double> ctrl, int64_t ctrl_seed, int64_t traj_offset, double lam,
                   const OptTensor& cmin) {
    TORCH_CHECK(costs.is_cuda(), "mpcdl: costs must be a CUDA tensor (no CPU fallback), got ", costs.device());
    const c10::Device dev = costs.device();
    TORCH_CHECK(U.dim() == 2 && U.size(1) == 2 && ctrl.size() == 6 && costs.dim() == 1,
                "mpcdl: mppi_reduce needs costs [N], U [H,2], ctrl = sigma[2]+lo[2]+hi[2]");
    const int64_t H = U.size(0);
    mpcdl_mppi_args a;
    std::memset(&a, 0, sizeof(a));
    a.costs = f32c(costs, "costs", dev); a.N = costs.numel(); a.traj_offset = traj_offset; a.H = (int32_t)H;
    a.U = f32c(U, "U", dev);
    for (int i = 0; i < 2; ++i) { a.ctrl_sigma[i] = (float)ctrl[i]; a.ctrl_lo[i] = (float)ctrl[2 + i]; a.ctrl_hi[i] = (float)ctrl[4 + i]; }
    a.ctrl_seed = (uint64_t)ctrl_seed; a.lambda = (float)lam;
    if (has(cmin)) a.cmin = f32c(cmin, "cmin", dev);
    Tensor ws = at::empty({(int64_t)mpcdl_mppi_workspace_floats((int)H)}, costs.options());
    Tensor sums = at::empty({2 * H + 2}, costs.options());
    a.workspace = static_cast<float*>(ws.data_ptr()); a.sums = static_cast<float*>(sums.data_ptr());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_mppi_reduce(&a, at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return sums;
}

// ---------------------------------------------------------------------------- dataset build / statistics / evaluation inputs
const int64_t* i64c(const Tensor& t, const char* name, const c10::Device& dev) {
    TORCH_CHECK(t.scalar_type() == at::kLong && t.is_cuda() && t.device() == dev && t.is_contiguous(), "mpcdl: ", name,
                " must be a contiguous int64 CUDA tensor on ", dev);
    return static_cast<const int64_t*>(t.data_ptr());
}
mpcdl_dataset_desc dataset_desc(const Tensor& rows, const Tensor& bag_row_off, const Tensor& samp_off, int64_t window, int64_t n) {
    TORCH_CHECK(rows.is_cuda(), "mpcdl: rows must be a CUDA tensor (no CPU fallback), got ", rows.device());
    const c10::Device dev = rows.device();
    TORCH_CHECK(rows.dim() == 2 && rows.size(1) == 8, "mpcdl: rows must be [R,8], got ", rows.sizes());
    TORCH_CHECK(bag_row_off.dim() == 1 && bag_row_off.sizes() == samp_off.sizes() && bag_row_off.numel() >= 1,
                "mpcdl: bag_row_off and samp_off must be [n_bags]");
    mpcdl_dataset_desc d;
    d.rows = f32c(rows, "rows", dev);
    d.bag_row_off = i64c(bag_row_off, "bag_row_off", dev);
    d.samp_off = i64c(samp_off, "samp_off", dev);
    d.n_bags = (int32_t)bag_row_off.numel(); d.window = (int32_t)window; d.n = n;
    return d;
}

// RobotData.norm statistics (basic_model.py:28-32, deep_models.py:29-33) in the rollout's `norm` layout
Tensor dataset_stats(const Tensor& rows, const Tensor& bag_row_off, const Tensor& samp_off, int64_t window, int64_t n) {
    const mpcdl_dataset_desc d = dataset_desc(rows, bag_row_off, samp_off, window, n);
    const int64_t T = window > 0 ? window : 1;
    Tensor stats = at::empty({2 * T * 8 + 12}, rows.options());
    c10::cuda::CUDAGuard guard(rows.device());
    check_status(mpcdl_dataset_stats(&d, static_cast<float*>(stats.data_ptr()), at::cuda::getCurrentCUDAStream(rows.device().index()).stream()));
    return stats;
}
// RobotData.__init__ (+ norm when stats is given): featrs [n,T,8] / [n,8], labels [n,6]
std::tuple<Tensor, Tensor> dataset_build(const Tensor& rows, const Tensor& bag_row_off, const Tensor& samp_off, int64_t window, int64_t n,
                                         const OptTensor& stats) {
    const mpcdl_dataset_desc d = dataset_desc(rows, bag_row_off, samp_off, window, n);
    const c10::Device dev = rows.device();
    const int64_t T = window > 0 ? window : 1;
    if (has(stats)) TORCH_CHECK(stats->numel() == 2 * T * 8 + 12, "mpcdl: stats must hold ", 2 * T * 8 + 12, " floats");
    Tensor featrs = window > 0 ? at::empty({n, T, 8}, rows.options()) : at::empty({n, 8}, rows.options());
    Tensor labels = at::empty({n, 6}, rows.options());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_dataset_build(&d, f32c(stats, "stats", dev), static_cast<float*>(featrs.data_ptr()), static_cast<float*>(labels.data_ptr()),
                                     at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return {featrs, labels};
}
// open-loop evaluation inputs: init, recorded controls U [n,H,2], recorded states target [n,H,6]
std::tuple<Tensor, Tensor, Tensor> rollout_inputs(const Tensor& rows, const Tensor& bag_row_off, const Tensor& samp_off, int64_t window,
                                                  int64_t n, int64_t H) {
    const mpcdl_dataset_desc d = dataset_desc(rows, bag_row_off, samp_off, window, n);
    const c10::Device dev = rows.device();
    TORCH_CHECK(H >= 1, "mpcdl: H must be >= 1");
    Tensor init = window > 0 ? at::empty({n, window, 8}, rows.options()) : at::empty({n, 6}, rows.options());
    Tensor U = at::empty({n, H, 2}, rows.options()), target = at::empty({n, H, 6}, rows.options());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_rollout_inputs(&d, (int32_t)H, static_cast<float*>(init.data_ptr()), static_cast<float*>(U.data_ptr()),
                                      static_cast<float*>(target.data_ptr()), at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return {init, U, target};
}
// per-step MAE / RMSE curves of stacked models: pred [M,n,H,6] (or [n,H,6]), target [n,H,6], lstat [12] -> ([M,H], [M,H])
std::tuple<Tensor, Tensor> curve_metrics(const Tensor& pred, const Tensor& target, const Tensor& lstat) {
    TORCH_CHECK(pred.is_cuda(), "mpcdl: pred must be a CUDA tensor (no CPU fallback), got ", pred.device());
    const c10::Device dev = pred.device();
    TORCH_CHECK(target.dim() == 3 && target.size(2) == 6 && (pred.dim() == 3 || pred.dim() == 4) && lstat.numel() == 12,
                "mpcdl: curve_metrics needs pred [M,n,H,6] | [n,H,6], target [n,H,6], lstat [12]");
    const int64_t M = pred.dim() == 4 ? pred.size(0) : 1, n = target.size(0), H = target.size(1);
    TORCH_CHECK(pred.numel() == M * n * H * 6, "mpcdl: pred ", pred.sizes(), " does not match target ", target.sizes());
    Tensor mae = at::empty({M, H}, pred.options()), rmse = at::empty({M, H}, pred.options());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_curve_metrics(f32c(pred, "pred", dev), f32c(target, "target", dev), f32c(lstat, "lstat", dev), (int32_t)M, n, (int32_t)H,
                                     static_cast<float*>(mae.data_ptr()), static_cast<float*>(rmse.data_ptr()),
                                     at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return {mae, rmse};
}
// DynamicsDataset resampling (utils/data_utils.py:21-37, 52-72): float64 t [M], q [M,6], u [M,2] -> (q_ds, u_ds)
std::tuple<Tensor, Tensor> resample(const Tensor& t, const Tensor& q, const Tensor& u, double dataset_dt, int64_t step, int64_t n_interp) {
    TORCH_CHECK(t.is_cuda(), "mpcdl: t must be a CUDA tensor (no CPU fallback), got ", t.device());
    const c10::Device dev = t.device();
    for (const Tensor* x : {&t, &q, &u})
        TORCH_CHECK(x->scalar_type() == at::kDouble && x->is_cuda() && x->device() == dev && x->is_contiguous(),
                    "mpcdl: resample needs contiguous float64 CUDA tensors on one device");
    const int64_t M = t.numel();
    TORCH_CHECK(t.dim() == 1 && q.dim() == 2 && q.size(0) == M && q.size(1) == 6 && u.dim() == 2 && u.size(0) == M && u.size(1) == 2,
                "mpcdl: resample needs t [M], q [M,6], u [M,2]");
    TORCH_CHECK(step >= 1 && n_interp >= 0, "mpcdl: resample needs step >= 1 and n_interp >= 0");
    mpcdl_resample_args a;
    a.t = static_cast<const double*>(t.data_ptr()); a.q = static_cast<const double*>(q.data_ptr()); a.u = static_cast<const double*>(u.data_ptr());
    a.M = M; a.dataset_dt = dataset_dt; a.step = (int32_t)step; a.n_interp = n_interp;
    Tensor qo = at::empty({(n_interp + step - 1) / step, 6}, t.options()), uo = at::empty({n_interp / step, 2}, t.options());
    a.q_out = static_cast<double*>(qo.data_ptr()); a.u_out = static_cast<double*>(uo.data_ptr());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_resample(&a, at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return {qo, uo};
}
// one epoch of train_model (model_scripts/basic_model.py:100-110) for n_models models in one launch: theta / adam_m / adam_v
// [n_models, n_params] are updated in place; perm [n_models, n_samples] int32 row order; -> per-batch losses [n_models, n_batches]
Tensor train_mlp_epoch(Tensor theta, Tensor adam_m, Tensor adam_v, const Tensor& featrs, const Tensor& labels, const Tensor& perm,
                       int64_t hidden, int64_t layers, int64_t batch, int64_t step0, double lr, double beta1, double beta2, double eps) {
    TORCH_CHECK(theta.is_cuda(), "mpcdl: theta must be a CUDA tensor (no CPU fallback), got ", theta.device());
    const c10::Device dev = theta.device();
    const int64_t np = mpcdl_train_param_count((int32_t)hidden, (int32_t)layers);
    TORCH_CHECK(theta.dim() == 2 && theta.size(1) == np, "mpcdl: theta must be [n_models, ", np, "] for hidden=", hidden, " layers=", layers, ", got ", theta.sizes());
    TORCH_CHECK(adam_m.sizes() == theta.sizes() && adam_v.sizes() == theta.sizes(), "mpcdl: adam_m / adam_v must have theta's shape");
    TORCH_CHECK(featrs.dim() == 2 && featrs.size(1) == 8 && labels.dim() == 2 && labels.size(1) == 6 && labels.size(0) == featrs.size(0),
                "mpcdl: featrs [n,8] and labels [n,6] expected, got ", featrs.sizes(), " ", labels.sizes());
    TORCH_CHECK(perm.scalar_type() == at::kInt && perm.is_cuda() && perm.device() == dev && perm.is_contiguous() && perm.dim() == 2 &&
                perm.size(0) == theta.size(0) && perm.size(1) >= 1, "mpcdl: perm must be a contiguous int32 CUDA tensor [n_models, n_samples]");
    TORCH_CHECK(batch >= 1 && step0 >= 0, "mpcdl: batch >= 1 and step0 >= 0 expected");
    mpcdl_train_args a;
    memset(&a, 0, sizeof(a));
    a.theta = f32c(theta, "theta", dev); a.adam_m = f32c(adam_m, "adam_m", dev); a.adam_v = f32c(adam_v, "adam_v", dev);
    a.featrs = f32c(featrs, "featrs", dev); a.labels = f32c(labels, "labels", dev);
    a.perm = static_cast<const int32_t*>(perm.data_ptr());
    a.n_rows = featrs.size(0); a.n_models = (int32_t)theta.size(0); a.n_samples = (int32_t)perm.size(1); a.batch = (int32_t)batch;
    a.hidden = (int32_t)hidden; a.layers = (int32_t)layers; a.step0 = step0; a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
    Tensor losses = at::empty({theta.size(0), (perm.size(1) + batch - 1) / batch}, theta.options());
    a.losses = static_cast<float*>(losses.data_ptr());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_train_mlp_epoch(&a, at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return losses;
}
// one epoch of train_model (model_scripts/deep_models.py:106-146) for n_models recurrent models in one launch: theta / adam_m /
// adam_v [n_models, param_stride] updated in place; featrs [n, window, 8]; perm [n_models, n_samples] int32; work = scratch
// [n_models, mpcdl_train_rnn_work_floats]; -> per-batch losses [n_models, n_batches]
Tensor train_rnn_epoch(Tensor theta, Tensor adam_m, Tensor adam_v, const Tensor& featrs, const Tensor& labels, const Tensor& perm, Tensor work,
                       int64_t kind, int64_t hidden, int64_t layers, int64_t batch, int64_t step0, double lr, double beta1, double beta2, double eps) {
    TORCH_CHECK(theta.is_cuda(), "mpcdl: theta must be a CUDA tensor (no CPU fallback), got ", theta.device());
    const c10::Device dev = theta.device();
    TORCH_CHECK(featrs.dim() == 3 && featrs.size(2) == 8 && labels.dim() == 2 && labels.size(1) == 6 && labels.size(0) == featrs.size(0),
                "mpcdl: featrs [n,window,8] and labels [n,6] expected, got ", featrs.sizes(), " ", labels.sizes());
    const int64_t window = featrs.size(1);
    const int64_t np = mpcdl_train_rnn_param_count((int32_t)kind, (int32_t)hidden, (int32_t)layers);
    const int64_t wf = mpcdl_train_rnn_work_floats((int32_t)kind, (int32_t)hidden, (int32_t)layers, (int32_t)window);
    TORCH_CHECK(np > 0 && wf > 0, "mpcdl: kind must be 1 (GRU) or 2 (LSTM), hidden / layers / window positive");
    TORCH_CHECK(theta.dim() == 2 && theta.size(1) >= np && theta.size(1) % 4 == 0, "mpcdl: theta must be [n_models, stride] with stride a multiple of 4 and >= ",
                np, " for hidden=", hidden, " layers=", layers, ", got ", theta.sizes());
    TORCH_CHECK(adam_m.sizes() == theta.sizes() && adam_v.sizes() == theta.sizes(), "mpcdl: adam_m / adam_v must have theta's shape");
    TORCH_CHECK(perm.scalar_type() == at::kInt && perm.is_cuda() && perm.device() == dev && perm.is_contiguous() && perm.dim() == 2 &&
                perm.size(0) == theta.size(0) && perm.size(1) >= 1, "mpcdl: perm must be a contiguous int32 CUDA tensor [n_models, n_samples]");
    TORCH_CHECK(work.numel() >= theta.size(0) * wf, "mpcdl: work must hold n_models x ", wf, " floats, got ", work.numel());
    TORCH_CHECK(batch >= 1 && step0 >= 0, "mpcdl: batch >= 1 and step0 >= 0 expected");
    mpcdl_train_rnn_args a;
    memset(&a, 0, sizeof(a));
    a.theta = f32c(theta, "theta", dev); a.adam_m = f32c(adam_m, "adam_m", dev); a.adam_v = f32c(adam_v, "adam_v", dev);
    a.featrs = f32c(featrs, "featrs", dev); a.labels = f32c(labels, "labels", dev); a.work = f32c(work, "work", dev);
    a.perm = static_cast<const int32_t*>(perm.data_ptr());
    a.n_rows = featrs.size(0); a.param_stride = theta.size(1); a.n_models = (int32_t)theta.size(0); a.n_samples = (int32_t)perm.size(1);
    a.batch = (int32_t)batch; a.kind = (int32_t)kind; a.hidden = (int32_t)hidden; a.layers = (int32_t)layers; a.window = (int32_t)window;
    a.step0 = step0; a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
    Tensor losses = at::empty({theta.size(0), (perm.size(1) + batch - 1) / batch}, theta.options());
    a.losses = static_cast<float*>(losses.data_ptr());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_train_rnn_epoch(&a, at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return losses;
}
// per-batch L1 loss of consecutive batches (the validation loop of basic_model.py:112-122): output / labels [n,6] -> [ceil(n / batch)]
Tensor l1_batches(const Tensor& output, const Tensor& labels, int64_t batch) {
    TORCH_CHECK(output.is_cuda(), "mpcdl: output must be a CUDA tensor (no CPU fallback), got ", output.device());
    const c10::Device dev = output.device();
    TORCH_CHECK(output.dim() == 2 && output.size(1) == 6 && labels.sizes() == output.sizes() && batch >= 1 && output.size(0) >= 1,
                "mpcdl: l1_batches needs output and labels [n,6] and batch >= 1");
    Tensor losses = at::empty({(output.size(0) + batch - 1) / batch}, output.options());
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_l1_batches(f32c(output, "output", dev), f32c(labels, "labels", dev), output.size(0), (int32_t)batch,
                                  static_cast<float*>(losses.data_ptr()), at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return losses;
}
// rosbag2 CDR decode: blob uint8 [bytes], offsets int64 [n + 1] -> (rows float64 [n, 9] = [t, x, y, psi, w_psi, v_long, v_tran, u_a, u_steer],
// status int32 [n])
std::tuple<Tensor, Tensor> cdr_decode(const Tensor& blob, const Tensor& offsets) {
    TORCH_CHECK(blob.is_cuda(), "mpcdl: blob must be a CUDA tensor (no CPU fallback), got ", blob.device());
    const c10::Device dev = blob.device();
    TORCH_CHECK(blob.scalar_type() == at::kByte && blob.is_contiguous() && blob.dim() == 1, "mpcdl: blob must be a contiguous uint8 vector");
    TORCH_CHECK(offsets.scalar_type() == at::kLong && offsets.is_cuda() && offsets.device() == dev && offsets.is_contiguous() && offsets.dim() == 1 &&
                offsets.numel() >= 2, "mpcdl: offsets must be a contiguous int64 CUDA vector [n + 1] on the blob's device");
    const int64_t n = offsets.numel() - 1;
    Tensor rows = at::empty({n, 9}, blob.options().dtype(at::kDouble)), status = at::empty({n}, blob.options().dtype(at::kInt));
    c10::cuda::CUDAGuard guard(dev);
    check_status(mpcdl_cdr_decode(static_cast<const uint8_t*>(blob.data_ptr()), static_cast<const int64_t*>(offsets.data_ptr()), n,
                                  static_cast<double*>(rows.data_ptr()), static_cast<int32_t*>(status.data_ptr()),
                                  at::cuda::getCurrentCUDAStream(dev.index()).stream()));
    return {rows, status};
}
// test aid: NaN-fill the shared memory of every SM of `like`'s device (determinism tests)
void debug_poison_smem(const Tensor& like) {
    TORCH_CHECK(like.is_cuda(), "mpcdl: debug_poison_smem needs a CUDA tensor to name the device");
    c10::cuda::CUDAGuard guard(like.device());
    check_status(mpcdl_debug_poison_smem(at::cuda::getCurrentCUDAStream(like.device().index()).stream()));
}

}  // namespace

#define MPCDL_ROLLOUT_SCHEMA                                                                                                          \
    "(Tensor? wblob, Tensor norm, Tensor init, Tensor U, int hidden, int layers, int window, bool stateful, Tensor(a!)? carry, "       \
    "Tensor? noise_wblob, int[] noise_cfg, Tensor? eps, int seed, int traj_offset, Tensor? cost_ref, float[] cost_w, "                \
    "bool return_states, bool return_costs, int[] tune, int precision, Tensor? wblob_tc, float[] ctrl, int ctrl_seed, "               \
    "int n_samples, bool shared_init, Tensor(b!)? out_states, Tensor(c!)? out_costs, int n_models=1, bool share_inputs=False) "       \
    "-> (Tensor, Tensor)"

TORCH_LIBRARY(mpcdl, m) {
    m.def("rollout_mlp" MPCDL_ROLLOUT_SCHEMA);
    m.def("rollout_gru" MPCDL_ROLLOUT_SCHEMA);
    m.def("rollout_lstm" MPCDL_ROLLOUT_SCHEMA);
    m.def("noise_forward(Tensor noise_wblob, int[] noise_cfg, Tensor q, Tensor u, Tensor? eps, int seed, int traj_offset, int step0, "
          "bool return_params) -> (Tensor, Tensor, Tensor)");
    m.def("dynamics_nn_rollout(Tensor noise_wblob, int[] noise_cfg, Tensor q0, Tensor U, Tensor? eps, int seed, int traj_offset) -> Tensor");
    m.def("error_metrics(Tensor output, Tensor labels) -> (Tensor, Tensor)");
    m.def("mppi_reduce(Tensor costs, Tensor U, float[] ctrl, int ctrl_seed, int traj_offset, float lam, Tensor? cmin) -> Tensor");
    m.def("mppi_min(Tensor costs) -> Tensor");
    m.def("dataset_stats(Tensor rows, Tensor bag_row_off, Tensor samp_off, int window, int n) -> Tensor");
    m.def("dataset_build(Tensor rows, Tensor bag_row_off, Tensor samp_off, int window, int n, Tensor? stats) -> (Tensor, Tensor)");
    m.def("rollout_inputs(Tensor rows, Tensor bag_row_off, Tensor samp_off, int window, int n, int H) -> (Tensor, Tensor, Tensor)");
    m.def("curve_metrics(Tensor pred, Tensor target, Tensor lstat) -> (Tensor, Tensor)");
    m.def("resample(Tensor t, Tensor q, Tensor u, float dataset_dt, int step, int n_interp) -> (Tensor, Tensor)");
    m.def("train_mlp_epoch(Tensor(a!) theta, Tensor(b!) adam_m, Tensor(c!) adam_v, Tensor featrs, Tensor labels, Tensor perm, int hidden, int layers, "
          "int batch, int step0, float lr, float beta1, float beta2, float eps) -> Tensor");
    m.def("train_rnn_epoch(Tensor(a!) theta, Tensor(b!) adam_m, Tensor(c!) adam_v, Tensor featrs, Tensor labels, Tensor perm, Tensor(d!) work, "
          "int kind, int hidden, int layers, int batch, int step0, float lr, float beta1, float beta2, float eps) -> Tensor");
    m.def("l1_batches(Tensor output, Tensor labels, int batch) -> Tensor");
    m.def("cdr_decode(Tensor blob, Tensor offsets) -> (Tensor, Tensor)");
    m.def("debug_poison_smem(Tensor like) -> ()");
}

TORCH_LIBRARY_IMPL(mpcdl, CUDA, m) {
    m.impl("rollout_mlp", &rollout_mlp);
    m.impl("rollout_gru", &rollout_gru);
    m.impl("rollout_lstm", &rollout_lstm);
    m.impl("noise_forward", &noise_forward);
    m.impl("dynamics_nn_rollout", &dynamics_nn_rollout);
    m.impl("error_metrics", &error_metrics);
    m.impl("mppi_reduce", &mppi_reduce);
    m.impl("mppi_min", &mppi_min);
    m.impl("dataset_stats", &dataset_stats);
    m.impl("dataset_build", &dataset_build);
    m.impl("rollout_inputs", &rollout_inputs);
    m.impl("curve_metrics", &curve_metrics);
    m.impl("resample", &resample);
    m.impl("train_mlp_epoch", &train_mlp_epoch);
    m.impl("train_rnn_epoch", &train_rnn_epoch);
    m.impl("l1_batches", &l1_batches);
    m.impl("cdr_decode", &cdr_decode);
    m.impl("debug_poison_smem", &debug_poison_smem);
}
