// K16/K17: the training step of the small forward-dynamics MLP, fused and persistent.
//   reference: train_model  model_scripts/basic_model.py:76-110
//       for featrs, labels in t_load:            # DataLoader(batch_size=16, shuffle=True), last batch may be short
//           optimizer.zero_grad(); output = model(featrs); loss = nn.L1Loss()(output, labels)
//           loss.backward(); optimizer.step()    # optim.Adam(model.parameters(), lr=LEARN_R)
//   and the validation loss of :112-122 (mean over batches of the per-batch L1 loss).
// The reference pays ~0.3 ms of Python / ATen dispatch per batch of 16 for 17.6 k parameters.  Here ONE launch runs a whole
// epoch: a CTA owns one model (an ensemble member / one point of the hidden x layers grid of basic_model.py:141-147 per
// CTA), keeps its parameters and gradients in shared memory for all batches, and walks the batches in the given sample
// order: forward, sign(output - label) / (B * 6), backward, Adam -- nothing but the Adam moments (L2-resident, one
// read-modify-write per parameter and step) and the batch rows touch global memory.
// Parameter vector layout = torch order of Basic_Model.parameters(): per Linear (out x in weight, row-major) then bias.
// Adam arithmetic follows torch.optim.Adam's single-tensor path op for op in FP32:
//   m = m + (g - m) * (1 - b1);  v = v * b2 + (1 - b2) * g * g;  denom = sqrt(v) / sqrt(1 - b2^t) + eps;
//   p = p + (-lr / (1 - b1^t)) * m / denom
// The sums of the matrix products run in a fixed order (k ascending; batch ascending), so a launch is reproducible; they
// are not MKL's order, so parity with torch is to FP32 rounding per step (tests/test_gpu_train.py states the bounds).
#include "launch.h"

namespace mpcdl {

namespace {

constexpr int TR_THREADS = 256;
constexpr int TR_MAXB = 16;          // batch rows held on chip (the reference's BATCH_S grid is 16 / 32 / 64: larger batches are
                                     // processed as micro-batches of 16 whose gradients accumulate before the Adam step)

struct TrLayer { int in, out, w_g, b_g, w_s, b_s; };      // offsets: _g in the torch-ordered vector, _s in shared memory (row stride in + 1)

__device__ __forceinline__ void tr_sync() { __syncthreads(); }

}  // namespace

// theta / adam_m / adam_v: [n_models][n_params]; perm: [n_models][n_samples] sample order of this epoch (int32);
// losses: [n_models][n_batches] per-batch L1 loss (the value loss.item() would print)
__global__ void __launch_bounds__(TR_THREADS, 1) train_mlp_kernel(const TrainParams p) {
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x;
    const int hid = p.hidden, NL = p.layers + 2;          // Linear layers: in, `layers` hidden, out
    TrLayer L[MAX_LAYERS + 2];
    int og = 0, os = 0;
    for (int l = 0; l < NL; ++l) {
        L[l].in = l == 0 ? 8 : hid;
        L[l].out = l == NL - 1 ? 6 : hid;
        L[l].w_g = og; og += L[l].in * L[l].out;
        L[l].b_g = og; og += L[l].out;
        L[l].w_s = os; os += L[l].out * (L[l].in + 1);
        L[l].b_s = os; os += L[l].out;
    }
    const int n_params = og, n_smem_params = (os + 3) & ~3;
    float* const W = sm;                                   // parameters, padded rows
    float* const G = W + n_smem_params;                    // gradients, same layout
    float* const act = G + n_smem_params;                  // act[0] = batch inputs [16][8]; act[l + 1] = output of layer l [16][out]
    float* const dlt = act + TR_MAXB * (8 + (NL - 1) * hid + 8);      // two delta buffers [16][hid]
    float* const red = dlt + 2 * TR_MAXB * hid;            // [TR_THREADS] loss reduction
    const long long mi = blockIdx.x;
    float* const theta = p.theta + mi * n_params;
    float* const am = p.adam_m + mi * n_params;
    float* const av = p.adam_v + mi * n_params;
    const int* const perm = p.perm + mi * p.n_samples;

    for (int l = 0; l < NL; ++l) {                         // parameters -> shared memory (padded rows)
        for (int i = tid; i < L[l].in * L[l].out; i += TR_THREADS) W[L[l].w_s + (i / L[l].in) * (L[l].in + 1) + i % L[l].in] = theta[L[l].w_g + i];
        for (int i = tid; i < L[l].out; i += TR_THREADS) W[L[l].b_s + i] = theta[L[l].b_g + i];
    }
    tr_sync();
    auto act_of = [&](int l) { return act + (l == 0 ? 0 : TR_MAXB * (8 + (l - 1) * hid)); };      // input of layer l

    const int n_batches = (p.n_samples + p.batch - 1) / p.batch;
    for (int bi = 0; bi < n_batches; ++bi) {
        const int b0 = bi * p.batch, B = min(p.batch, p.n_samples - b0);          // DataLoader(drop_last=False): the last batch may be short
        for (int i = tid; i < n_smem_params; i += TR_THREADS) G[i] = 0.f;        // optimizer.zero_grad()
        float loss_acc = 0.f;
        for (int mb0 = 0; mb0 < B; mb0 += TR_MAXB) {                              // micro-batches of <= 16 rows
            const int MB = min(TR_MAXB, B - mb0);
            // ---- gather the rows of this micro-batch
            for (int i = tid; i < MB * 8; i += TR_THREADS) {
                const int b = i >> 3, k = i & 7;
                act[b * 8 + k] = __ldg(p.featrs + (long long)perm[b0 + mb0 + b] * 8 + k);
            }
            tr_sync();
            // ---- forward: out[b][j] = relu(bias[j] + sum_k in[b][k] W[j][k])   (bias first, k ascending)
            for (int l = 0; l < NL; ++l) {
                const int in = L[l].in, out = L[l].out;
                const float* a = act_of(l);
                float* o = act_of(l + 1);
                // thread (j, r): unit j, batch rows r, r + R, r + 2R, r + 3R (R = threads / out): four independent FMA chains share
                // every weight load; each output still sums bias first, k ascending
                const int R = TR_THREADS / out > 0 ? TR_THREADS / out : 1;
                for (int idx = tid; idx < R * out; idx += TR_THREADS) {
                    const int r = idx / out, j = idx - r * out;
                    const float* w = W + L[l].w_s + j * (in + 1);
                    const float bias = W[L[l].b_s + j];
                    for (int b = r; b < MB; b += 4 * R) {
                        const int b1 = min(b + R, MB - 1), b2 = min(b + 2 * R, MB - 1), b3 = min(b + 3 * R, MB - 1);
                        float s0 = bias, s1 = bias, s2 = bias, s3 = bias;
                        for (int k = 0; k < in; ++k) {
                            const float wk = w[k];
                            s0 = fmaf(a[b * in + k], wk, s0); s1 = fmaf(a[b1 * in + k], wk, s1);
                            s2 = fmaf(a[b2 * in + k], wk, s2); s3 = fmaf(a[b3 * in + k], wk, s3);
                        }
                        const bool last = l == NL - 1;
                        o[b * out + j] = last ? s0 : fmaxf(s0, 0.f);
                        if (b + R < MB) o[(b + R) * out + j] = last ? s1 : fmaxf(s1, 0.f);
                        if (b + 2 * R < MB) o[(b + 2 * R) * out + j] = last ? s2 : fmaxf(s2, 0.f);
                        if (b + 3 * R < MB) o[(b + 3 * R) * out + j] = last ? s3 : fmaxf(s3, 0.f);
                    }
                }
                tr_sync();
            }
            // ---- L1 loss: mean over the B x 6 elements of the WHOLE batch; d loss / d out = sign(out - label) / (B * 6)
            float* d_out = dlt;
            {
                const float* o = act_of(NL);
                const float inv = 1.0f / (float)(B * 6);
                for (int idx = tid; idx < MB * 6; idx += TR_THREADS) {
                    const int b = idx / 6, d = idx - b * 6;
                    const float diff = o[b * 6 + d] - __ldg(p.labels + (long long)perm[b0 + mb0 + b] * 6 + d);
                    loss_acc += fabsf(diff);
                    d_out[b * 6 + d] = diff > 0.f ? inv : diff < 0.f ? -inv : 0.f;
                }
            }
            tr_sync();
            // ---- backward, layer by layer: dW[j][k] += sum_b d[b][j] a[b][k], db[j] += sum_b d[b][j], d_in = (d W) * relu'(a)
            for (int l = NL - 1; l >= 0; --l) {
                const int in = L[l].in, out = L[l].out;
                const float* a = act_of(l);
                const float* d = (NL - 1 - l) & 1 ? dlt + TR_MAXB * hid : dlt;
                float* d_in = (NL - 1 - l) & 1 ? dlt : dlt + TR_MAXB * hid;
                // dW: four entries (j, k), (j + J, k), ... per trip (J = threads / in rows of the weight matrix apart): independent chains
                {
                    const int J = TR_THREADS / in > 0 ? TR_THREADS / in : 1;
                    for (int idx = tid; idx < J * in; idx += TR_THREADS) {
                        const int r = idx / in, k = idx - r * in;
                        for (int j = r; j < out; j += 4 * J) {
                            const int j1 = min(j + J, out - 1), j2 = min(j + 2 * J, out - 1), j3 = min(j + 3 * J, out - 1);
                            float* g0 = G + L[l].w_s + j * (in + 1) + k;
                            float* g1 = G + L[l].w_s + j1 * (in + 1) + k;
                            float* g2 = G + L[l].w_s + j2 * (in + 1) + k;
                            float* g3 = G + L[l].w_s + j3 * (in + 1) + k;
                            float s0 = *g0, s1 = *g1, s2 = *g2, s3 = *g3;
                            for (int b = 0; b < MB; ++b) {
                                const float ab = a[b * in + k];
                                s0 = fmaf(d[b * out + j], ab, s0); s1 = fmaf(d[b * out + j1], ab, s1);
                                s2 = fmaf(d[b * out + j2], ab, s2); s3 = fmaf(d[b * out + j3], ab, s3);
                            }
                            *g0 = s0;
                            if (j + J < out) *g1 = s1;
                            if (j + 2 * J < out) *g2 = s2;
                            if (j + 3 * J < out) *g3 = s3;
                        }
                    }
                }
                for (int j = tid; j < out; j += TR_THREADS) {
                    float s = G[L[l].b_s + j];
                    for (int b = 0; b < MB; ++b) s += d[b * out + j];
                    G[L[l].b_s + j] = s;
                }
                if (l > 0) {
                    const int R = TR_THREADS / in > 0 ? TR_THREADS / in : 1;
                    for (int idx = tid; idx < R * in; idx += TR_THREADS) {
                        const int r = idx / in, k = idx - r * in;
                        const float* wc = W + L[l].w_s + k;
                        for (int b = r; b < MB; b += 4 * R) {
                            const int b1 = min(b + R, MB - 1), b2 = min(b + 2 * R, MB - 1), b3 = min(b + 3 * R, MB - 1);
                            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
                            for (int j = 0; j < out; ++j) {
                                const float wj = wc[j * (in + 1)];
                                s0 = fmaf(d[b * out + j], wj, s0); s1 = fmaf(d[b1 * out + j], wj, s1);
                                s2 = fmaf(d[b2 * out + j], wj, s2); s3 = fmaf(d[b3 * out + j], wj, s3);
                            }
                            d_in[b * in + k] = a[b * in + k] > 0.f ? s0 : 0.f;             // relu'(pre-activation) = (output > 0)
                            if (b + R < MB) d_in[(b + R) * in + k] = a[(b + R) * in + k] > 0.f ? s1 : 0.f;
                            if (b + 2 * R < MB) d_in[(b + 2 * R) * in + k] = a[(b + 2 * R) * in + k] > 0.f ? s2 : 0.f;
                            if (b + 3 * R < MB) d_in[(b + 3 * R) * in + k] = a[(b + 3 * R) * in + k] > 0.f ? s3 : 0.f;
                        }
                    }
                }
                tr_sync();
            }
        }
        // ---- loss of the batch (fixed-order tree over the threads' partial sums)
        red[tid] = loss_acc;
        tr_sync();
        for (int o = TR_THREADS / 2; o > 0; o >>= 1) {
            if (tid < o) red[tid] += red[tid + o];
            tr_sync();
        }
        if (tid == 0 && p.losses) p.losses[mi * n_batches + bi] = red[0] / (float)(B * 6);
        // ---- Adam (torch.optim.Adam, single-tensor path): step t = step0 + bi + 1
        {
            const double t = (double)(p.step0 + bi + 1);
            // python-double scalars of torch.optim.adam._single_tensor_adam, cast to FP32 where torch hands them to a kernel
            const double bc1 = 1.0 - pow(p.beta1, t), bc2 = 1.0 - pow(p.beta2, t);
            const float step_size = (float)(-(p.lr / bc1)), bc2_sqrt = (float)sqrt(bc2);
            const float w1 = (float)(1.0 - p.beta1), w2 = (float)(1.0 - p.beta2), b2f = (float)p.beta2, epsf = (float)p.eps;
            for (int l = 0; l < NL; ++l) {
                const int in = L[l].in, out = L[l].out;
                for (int i = tid; i < in * out + out; i += TR_THREADS) {
                    const bool is_w = i < in * out;
                    const int si = is_w ? L[l].w_s + (i / in) * (in + 1) + i % in : L[l].b_s + (i - in * out);
                    const int gi = is_w ? L[l].w_g + i : L[l].b_g + (i - in * out);
                    const float g = G[si];
                    float m = am[gi], v = av[gi];
                    m = fmaf(w1, __fsub_rn(g, m), m);                                               // exp_avg.lerp_(grad, 1 - beta1): fmadd(weight, end - self, self)
                    v = __fadd_rn(__fmul_rn(v, b2f), __fmul_rn(__fmul_rn(w2, g), g));               // mul_(beta2).addcmul_(grad, grad, value = 1 - beta2)
                    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), epsf);
                    W[si] = __fadd_rn(W[si], __fdiv_rn(__fmul_rn(step_size, m), denom));            // addcdiv_(exp_avg, denom, value = -step_size)
                    am[gi] = m; av[gi] = v;
                }
            }
        }
        tr_sync();
    }
    for (int l = 0; l < NL; ++l) {                         // parameters back to the torch-ordered vector
        for (int i = tid; i < L[l].in * L[l].out; i += TR_THREADS) theta[L[l].w_g + i] = W[L[l].w_s + (i / L[l].in) * (L[l].in + 1) + i % L[l].in];
        for (int i = tid; i < L[l].out; i += TR_THREADS) theta[L[l].b_g + i] = W[L[l].b_s + i];
    }
}

// validation loss of basic_model.py:112-122: per batch of `batch` consecutive samples the L1 loss (mean over B x 6), one
// CTA per batch; the caller averages the batches (v_loss /= len(v_load)).  output / labels [n,6].
__global__ void __launch_bounds__(128) l1_batches_kernel(const float* __restrict__ output, const float* __restrict__ labels, long long n,
                                                         int batch, float* __restrict__ losses) {
    __shared__ float red[128];
    const long long b0 = (long long)blockIdx.x * batch;
    const int B = (int)min((long long)batch, n - b0);
    float s = 0.f;
    for (int i = threadIdx.x; i < B * 6; i += 128) s += fabsf(output[b0 * 6 + i] - labels[b0 * 6 + i]);
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) losses[blockIdx.x] = red[0] / (float)(B * 6);
}

int train_smem_bytes(int hidden, int layers) {
    const int NL = layers + 2;
    int os = 0;
    for (int l = 0; l < NL; ++l) {
        const int in = l == 0 ? 8 : hidden, out = l == NL - 1 ? 6 : hidden;
        os += out * (in + 1) + out;
    }
    const int n_smem_params = (os + 3) & ~3;
    return (2 * n_smem_params + TR_MAXB * (8 + (NL - 1) * hidden + 8) + 2 * TR_MAXB * hidden + TR_THREADS) * 4;
}

cudaError_t launch_train_mlp(const TrainParams& p, int n_models, int smem_bytes, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(train_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    train_mlp_kernel<<<n_models, TR_THREADS, smem_bytes, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_l1_batches(const float* output, const float* labels, long long n, int batch, float* losses, cudaStream_t st) {
    const long long nb = (n + batch - 1) / batch;
    l1_batches_kernel<<<(unsigned)nb, 128, 0, st>>>(output, labels, n, batch, losses);
    return cudaGetLastError();
}

}  // namespace mpcdl
