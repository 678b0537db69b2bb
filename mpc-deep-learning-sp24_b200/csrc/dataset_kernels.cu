// Data side of the rollout path on the device:
//   K11 dataset_stats_kernel    per-position mean / UNBIASED std of the features and labels that RobotData builds
//                               model_scripts/basic_model.py:13-35, model_scripts/deep_models.py:13-36
//   K12 dataset_build_kernel    the feature windows / labels themselves (optionally z-scored: subtract THEN divide)
//   K13 rollout_inputs_kernel   closed-loop evaluation inputs cut from a recorded bag: start state / window, the recorded
//                               controls of the next H frames (what an open-loop validation sweep feeds the rollout)
//   K14 curve_metrics_kernel    per-step MAE / RMSE of predicted states against the recorded ones in label-normalised
//                               units -- eval_models.calc_error's metric (model_scripts/eval_models.py:91-92) per horizon step
//   K15 resample_kernel         DynamicsDataset: np.interp to the 20 ms grid, q[::step], mean-pooled u
//                               utils/data_utils.py:21-37 (subsample / pool), :52-72 (interpolation)
// All reductions run in a fixed order (per-thread serial sums, then a shared-memory tree): results are reproducible.
#include <algorithm>

#include "launch.h"

namespace mpcdl {

namespace {

constexpr int DS_THREADS = 256;

// sample s -> first row of its window; bags are concatenated, windows never cross a bag boundary
__device__ __forceinline__ long long sample_row(const DatasetParams& d, long long s) {
    int lo = 0, hi = d.n_bags - 1;                       // last bag with samp_off[b] <= s
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(d.samp_off + mid) <= s) lo = mid; else hi = mid - 1;
    }
    return __ldg(d.bag_row_off + lo) + (s - __ldg(d.samp_off + lo));
}

// fixed-order CTA reduction of one double per thread (DS_THREADS threads): tree over shared memory
__device__ __forceinline__ double cta_sum(double v, double* red) {
    red[threadIdx.x] = v;
    __syncthreads();
    for (int o = DS_THREADS / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    const double r = red[0];
    __syncthreads();
    return r;
}

}  // namespace

// ---------------------------------------------------------------------------------------------- K11
// one CTA per statistic column: columns [0, T*8) = feature (position f, channel k), [T*8, T*8+6) = label channel.
// Two passes in double (mean, then sum of squared deviations): at least as accurate as torch's float32 mean and
// double Welford std; results rounded to float once.  stats layout = the kernels' `norm` vector.
__global__ void __launch_bounds__(DS_THREADS) dataset_stats_kernel(const __grid_constant__ DatasetParams d, float* __restrict__ stats) {
    __shared__ double red[DS_THREADS];
    const int T = d.window > 0 ? d.window : 1;
    const int col = blockIdx.x;
    int f, k;
    if (col < T * 8) { f = col / 8; k = col % 8; }
    else { f = d.window > 0 ? d.window : 1; k = col - T * 8; }      // label = the row after the window (MLP: row i+1)
    double s = 0.0;
    for (long long i = threadIdx.x; i < d.n; i += DS_THREADS) s += (double)__ldg(d.rows + (sample_row(d, i) + f) * 8 + k);
    const double mean = cta_sum(s, red) / (double)d.n;
    double q = 0.0;
    for (long long i = threadIdx.x; i < d.n; i += DS_THREADS) {
        const double v = (double)__ldg(d.rows + (sample_row(d, i) + f) * 8 + k) - mean;
        q += v * v;
    }
    const double var = cta_sum(q, red) / (double)(d.n - 1);          // unbiased (torch.std default)
    if (threadIdx.x == 0) {
        if (col < T * 8) { stats[col] = (float)mean; stats[T * 8 + col] = (float)sqrt(var); }
        else { stats[2 * T * 8 + k] = (float)mean; stats[2 * T * 8 + 6 + k] = (float)sqrt(var); }
    }
}

// ---------------------------------------------------------------------------------------------- K12
// featrs [n, T, 8] (MLP: [n, 8]), labels [n, 6]; stats != null -> (x - mean) / std with torch's two roundings
__global__ void __launch_bounds__(DS_THREADS) dataset_build_kernel(const __grid_constant__ DatasetParams d, const float* __restrict__ stats,
                                                                    float* __restrict__ featrs, float* __restrict__ labels) {
    const int T = d.window > 0 ? d.window : 1, per = T * 8 + 6;
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= d.n * per) return;
    const long long s = e / per;
    const int c = (int)(e - s * per);
    const long long r0 = sample_row(d, s);
    if (c < T * 8) {
        float v = __ldg(d.rows + r0 * 8 + c);                      // window rows are contiguous: (r0 + f) * 8 + k = r0 * 8 + c
        if (stats) v = __fdiv_rn(__fsub_rn(v, __ldg(stats + c)), __ldg(stats + T * 8 + c));
        featrs[s * (T * 8) + c] = v;
    } else {
        const int k = c - T * 8;
        float v = __ldg(d.rows + (r0 + T) * 8 + k);
        if (stats) v = __fdiv_rn(__fsub_rn(v, __ldg(stats + 2 * T * 8 + k)), __ldg(stats + 2 * T * 8 + 6 + k));
        labels[s * 6 + k] = v;
    }
}

// ---------------------------------------------------------------------------------------------- K13
// Open-loop evaluation inputs for start samples s = 0 .. n-1 (n = samples that leave H recorded frames ahead):
//   MLP  init[s] = row r [:6]            U[s,t] = row (r + t)     [6:8]   target[s,t] = row (r + t + 1) [:6]
//   RNN  init[s] = rows r .. r+T-1       U[s,t] = row (r + T + t) [6:8]   target[s,t] = row (r + T + t) [:6]
// (App. A.4 convention of the rollout: the frame appended at step t pairs the predicted state with the control recorded
//  in the same message.)
__global__ void __launch_bounds__(DS_THREADS) rollout_inputs_kernel(const __grid_constant__ DatasetParams d, int H, float* __restrict__ init,
                                                                     float* __restrict__ U, float* __restrict__ target) {
    const int T = d.window, ni = T > 0 ? T * 8 : 6, per = ni + H * 8;
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= d.n * per) return;
    const long long s = e / per;
    const int c = (int)(e - s * per);
    const long long r0 = sample_row(d, s);
    if (c < ni) {
        init[s * ni + c] = __ldg(d.rows + r0 * 8 + c);     // window rows are contiguous; the MLP start state is row[:6]
        return;
    }
    const int t = (c - ni) / 8, k = (c - ni) % 8;
    if (k < 6) {
        if (target) target[(s * H + t) * 6 + k] = __ldg(d.rows + (r0 + (T > 0 ? T + t : t + 1)) * 8 + k);
    } else {
        U[(s * H + t) * 2 + (k - 6)] = __ldg(d.rows + (r0 + (T > 0 ? T + t : t)) * 8 + k);
    }
}

// ---------------------------------------------------------------------------------------------- K14
// pred [M, n, H, 6] raw predicted states of M stacked models, target [n, H, 6] raw recorded states, lstat = l_mean[6] l_std[6]
// out_mae / out_rmse [M, H]: mean over the n trajectories of the per-sample metric of calc_error,
//   d = (pred - l_mean) / l_std - (target - l_mean) / l_std ;  mae = mean_6 |d| ;  rmse = sqrt(mean_6 d^2)
// One CTA per (model, step): per-thread serial sums over its trajectories, then the fixed tree.
__global__ void __launch_bounds__(DS_THREADS) curve_metrics_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                                                    const float* __restrict__ lstat, long long n, int H,
                                                                    float* __restrict__ out_mae, float* __restrict__ out_rmse) {
    __shared__ double red[DS_THREADS];
    const int m = blockIdx.x / H, t = blockIdx.x % H;
    float lm[6], ls[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) { lm[k] = __ldg(lstat + k); ls[k] = __ldg(lstat + 6 + k); }
    double sa = 0.0, sr = 0.0;
    for (long long i = threadIdx.x; i < n; i += DS_THREADS) {
        const float* p = pred + (((long long)m * n + i) * H + t) * 6;
        const float* g = target + (i * H + t) * 6;
        float a = 0.f, q = 0.f;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const float pn = __fdiv_rn(__fsub_rn(__ldg(p + k), lm[k]), ls[k]);
            const float gn = __fdiv_rn(__fsub_rn(__ldg(g + k), lm[k]), ls[k]);
            const float df = __fsub_rn(pn, gn);
            a = __fadd_rn(a, fabsf(df));
            q = __fadd_rn(q, __fmul_rn(df, df));
        }
        sa += (double)__fdiv_rn(a, 6.0f);
        sr += (double)__fsqrt_rn(__fdiv_rn(q, 6.0f));
    }
    const double ta = cta_sum(sa, red), tr = cta_sum(sr, red);
    if (threadIdx.x == 0) {
        out_mae[blockIdx.x] = (float)(ta / (double)n);
        out_rmse[blockIdx.x] = (float)(tr / (double)n);
    }
}

// ---------------------------------------------------------------------------------------------- K15
// DynamicsDataset.from_dict + __init__ (utils/data_utils.py:21-37, 52-72), float64 throughout like numpy:
//   t_i  = np.arange(t[0], t[-1], dt)[i]        = t0 (i = 0), t0 + dt (i = 1), t0 + i * ((t0 + dt) - t0) (i >= 2)
//   v_i  = np.interp(t_i, t, v)                 = fp[j] if xp[j] == x else ((fp[j+1]-fp[j])/(xp[j+1]-xp[j])) * (x - xp[j]) + fp[j]
//   q_out[k] = q_interp[k * step]               (q[::step])
//   u_out[k] = (((u_i0 + u_i1) + ...) + u_i(step-1)) / step over i = k*step .. k*step+step-1     (np.mean(axis=1), sequential)
// No fused multiply-add anywhere: numpy's C loops are compiled without contraction.
__device__ __forceinline__ double grid_time(double t0, double dt, long long i) {
    if (i == 0) return t0;
    const double t1 = __dadd_rn(t0, dt);
    if (i == 1) return t1;
    return __dadd_rn(t0, __dmul_rn((double)i, __dsub_rn(t1, t0)));
}
__device__ __forceinline__ double interp1(const double* __restrict__ xp, const double* __restrict__ fp, int stride, long long M, double x) {
    if (x < xp[0]) return fp[0];
    if (x > xp[M - 1]) return fp[(M - 1) * stride];
    long long lo = 0, hi = M - 1;                        // largest j with xp[j] <= x
    while (lo < hi) {
        const long long mid = (lo + hi + 1) >> 1;
        if (xp[mid] <= x) lo = mid; else hi = mid - 1;
    }
    const long long j = lo;
    if (j == M - 1 || xp[j] == x) return fp[j * stride];
    const double slope = __ddiv_rn(__dsub_rn(fp[(j + 1) * stride], fp[j * stride]), __dsub_rn(xp[j + 1], xp[j]));
    double r = __dadd_rn(__dmul_rn(slope, __dsub_rn(x, xp[j])), fp[j * stride]);
    if (isnan(r)) {
        r = __dadd_rn(__dmul_rn(slope, __dsub_rn(x, xp[j + 1])), fp[(j + 1) * stride]);
        if (isnan(r) && fp[j * stride] == fp[(j + 1) * stride]) r = fp[j * stride];
    }
    return r;
}
__global__ void __launch_bounds__(DS_THREADS) resample_kernel(const __grid_constant__ ResampleParams p) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long nq = p.n_q * 6;
    if (e < nq) {
        const long long k = e / 6;
        const int c = (int)(e % 6);
        p.q_out[e] = interp1(p.t, p.q + c, 6, p.M, grid_time(p.t[0], p.dt, k * p.step));
        return;
    }
    const long long e2 = e - nq;
    if (e2 >= p.n_u * 2) return;
    const long long k = e2 / 2;
    const int c = (int)(e2 % 2);
    double s = interp1(p.t, p.u + c, 2, p.M, grid_time(p.t[0], p.dt, k * p.step));
    for (int i = 1; i < p.step; ++i) s = __dadd_rn(s, interp1(p.t, p.u + c, 2, p.M, grid_time(p.t[0], p.dt, k * p.step + i)));
    p.u_out[e2] = __ddiv_rn(s, (double)p.step);
}

// ---------------------------------------------------------------------------------------------- launchers
cudaError_t launch_dataset_stats(const DatasetParams& d, float* stats, cudaStream_t st) {
    const int T = d.window > 0 ? d.window : 1;
    dataset_stats_kernel<<<T * 8 + 6, DS_THREADS, 0, st>>>(d, stats);
    return cudaGetLastError();
}
cudaError_t launch_dataset_build(const DatasetParams& d, const float* stats, float* featrs, float* labels, cudaStream_t st) {
    const int T = d.window > 0 ? d.window : 1;
    const long long total = d.n * (T * 8 + 6);
    dataset_build_kernel<<<(unsigned)((total + DS_THREADS - 1) / DS_THREADS), DS_THREADS, 0, st>>>(d, stats, featrs, labels);
    return cudaGetLastError();
}
cudaError_t launch_rollout_inputs(const DatasetParams& d, int H, float* init, float* U, float* target, cudaStream_t st) {
    const long long total = d.n * ((d.window > 0 ? d.window * 8 : 6) + (long long)H * 8);
    rollout_inputs_kernel<<<(unsigned)((total + DS_THREADS - 1) / DS_THREADS), DS_THREADS, 0, st>>>(d, H, init, U, target);
    return cudaGetLastError();
}
cudaError_t launch_curve_metrics(const float* pred, const float* target, const float* lstat, int n_models, long long n, int H,
                                 float* mae, float* rmse, cudaStream_t st) {
    curve_metrics_kernel<<<n_models * H, DS_THREADS, 0, st>>>(pred, target, lstat, n, H, mae, rmse);
    return cudaGetLastError();
}
cudaError_t launch_resample(const ResampleParams& p, cudaStream_t st) {
    const long long total = p.n_q * 6 + p.n_u * 2;
    if (total == 0) return cudaSuccess;
    resample_kernel<<<(unsigned)((total + DS_THREADS - 1) / DS_THREADS), DS_THREADS, 0, st>>>(p);
    return cudaGetLastError();
}

}  // namespace mpcdl
