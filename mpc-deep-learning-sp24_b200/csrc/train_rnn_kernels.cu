// K19: the training step of the recurrent dynamics models ("fused fwd+bwd for MLP/GRU"), fused and persistent.
//   reference: train_model  model_scripts/deep_models.py:106-146
//       for featrs, labels in t_load:            # DataLoader(batch_size=BATCH_S, shuffle=False) over a random_split subset
//           if STATE_F: model.init_hidden(batch_s)        # -> zeros every batch: stateful or not, a batch starts from h = 0
//           optimizer.zero_grad(); output = model(featrs); loss = nn.L1Loss()(output, labels)
//           loss.backward(); optimizer.step()             # optim.Adam(model.parameters(), lr=LEARN_R)
//   model(featrs) = full_c(nn.GRU | nn.LSTM(featrs [B,T,8], h0 = 0)[:, -1, :])   deep_models.py:57-66, 81-92
// ONE launch runs a whole epoch of one model per CTA (the reference's hidden x layers x ft_size x batch grid, deep_models.py:
// 182-190, or an ensemble, trains side by side): per batch the forward pass through L layers x T window frames, the L1 loss,
// back-propagation through time over the window, and the Adam step -- the gradient of a weight matrix is summed over the
// window and the batch in registers and goes straight into Adam (a gradient buffer exists only for batches of more than 16 rows,
// which run as micro-batches).
// Parameters (358 KB for GRU 64 x 4) do not fit shared memory: they stay in global memory (L2 / L1) in torch order
// ([gate rows][in]: lanes over `in` read it coalesced in the backward-data products) and as a k-major copy in the work buffer
// ([in][gate rows]: lanes over hidden units read it coalesced in the forward pass); Adam writes both.
// Cell formulas = torch's:  GRU rows [r; z; n], n = tanh(W_in x + b_in + r (W_hn h + b_hn)),
// h' = (1 - z) n + z h;  LSTM rows [i; f; g; o], c' = f c + i g, h' = o tanh(c').  Sums run in a fixed order (k ascending, then
// frame, then batch row), so a launch is reproducible; parity with torch's autograd is to FP32 rounding per step
// (tests/test_gpu_train_rnn.py states the bounds).
#include "launch.h"

namespace mpcdl {

namespace {

constexpr int TRN_THREADS = 256;
constexpr int RB = TRAIN_RNN_MAXB;     // batch rows held on chip (the reference's grid trains with BATCH_S 8 / 16; larger batches: micro-batches)

struct RnnOff {                        // float offsets inside the torch-ordered parameter vector
    int wih[MAX_LAYERS], whh[MAX_LAYERS], bih[MAX_LAYERS], bhh[MAX_LAYERS], fcw, fcb, total;
};

__device__ __forceinline__ float sigm(float x) { return 1.0f / (1.0f + expf(-x)); }

struct AdamC { float step_size, bc2_sqrt, w1, w2, b2, eps; };

// torch.optim.Adam, single-tensor path (same op order as train_kernels.cu)
__device__ __forceinline__ float adam_update(float w, float g, float& m, float& v, const AdamC& c) {
    m = fmaf(c.w1, __fsub_rn(g, m), m);
    v = __fadd_rn(__fmul_rn(v, c.b2), __fmul_rn(__fmul_rn(c.w2, g), g));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), c.bc2_sqrt), c.eps);
    return __fadd_rn(w, __fdiv_rn(__fmul_rn(c.step_size, m), denom));
}

}  // namespace

// KIND: 1 = GRU, 2 = LSTM.  theta / adam_m / adam_v [n_models][n_params]; featrs [n_rows][T][8]; labels [n_rows][6];
// perm [n_models][n_samples]; work [n_models][work_floats]; losses [n_models][n_batches]
// MINB = CTAs per SM the register allocation allows: 1 (up to 255 registers: fastest for a single model, 367 us per step of
// GRU 64 x 4) when there are no more models than SMs, 2 (128 registers) beyond that -- two models per SM hide each other's
// load latency: 2.8e5 -> 3.8e5 model-steps/s for 296 models (profiles/r02/train_rnn/timing_two_per_sm.log)
template <int KIND, int MINB>
__global__ void __launch_bounds__(TRN_THREADS, MINB) train_rnn_kernel(const TrainRnnParams p) {
    constexpr int G = KIND == 1 ? 3 : 4;          // gate blocks of a weight matrix
    constexpr int NS = KIND == 1 ? 4 : 5;         // saved planes per cell: GRU r z n nh / LSTM i f g o c
    extern __shared__ __align__(16) float sm[];
    const int tid = threadIdx.x;
    const int h = p.hidden, L = p.layers, T = p.window, GH = G * h;
    RnnOff O;
    {
        int o = 0;
        for (int l = 0; l < L; ++l) {
            const int in = l == 0 ? 8 : h;
            O.wih[l] = o; o += GH * in;
            O.whh[l] = o; o += GH * h;
            O.bih[l] = o; o += GH;
            O.bhh[l] = o; o += GH;
        }
        O.fcw = o; o += 6 * h;
        O.fcb = o; o += 6;
        O.total = o;
    }
    const long long mi = blockIdx.x;
    float* const theta = p.theta + mi * p.param_stride;      // stride = a multiple of 4 floats: float4 access in the Adam pass
    float* const am = p.adam_m + mi * p.param_stride;
    float* const av = p.adam_v + mi * p.param_stride;
    const int* const perm = p.perm + mi * p.n_samples;
    // work buffer
    float* const WT = p.work + mi * p.work_floats;                 // k-major copies of the weight matrices, at the offsets of the originals
    float* const Y = WT + ((O.total + 3) & ~3);                    // [L][T][RB][h]   layer outputs
    float* const S = Y + (size_t)L * T * RB * h;                   // [L][T][RB][NS h] saved gate values
    float* const GI = S + (size_t)L * T * RB * NS * h;             // [T][RB][GH]     gate gradients of the layer in flight (input side)
    float* const GR = GI + (size_t)T * RB * GH;                    // [T][RB][GH]     (recurrent side; differs from GI for the GRU's n gate only)
    float* const DY = GR + (size_t)T * RB * GH;                    // [T][RB][h]      gradient of the layer's output sequence
    float* const X0 = DY + (size_t)T * RB * h;                     // [T][RB][8]      the batch's windows
    float* const GA = X0 + (size_t)T * RB * 8;                     // [n_params]      gradient sums of the earlier micro-batches of a batch > RB rows
    // shared memory
    float* const xs = sm;                                          // [RB][h]  cell input
    float* const hs = xs + RB * h;                                 // [2][RB][h] h_{f-1} / h_f
    float* const gsi = hs + 2 * RB * h;                            // [RB][GH] gate gradients (input side)
    float* const gsr = gsi + RB * GH;                              // [RB][GH] (recurrent side)
    float* const dhs = gsr + RB * GH;                              // [RB][h]  gradient arriving from frame f + 1 through h
    float* const dcs = dhs + RB * h;                               // [RB][h]  ... through c (LSTM)
    float* const dhd = dcs + RB * h;                               // [RB][h]  direct part of dh_{f-1} (GRU: dh z)
    float* const outs = dhd + RB * h;                              // [RB][8]  head outputs, then d loss / d out
    float* const red = outs + RB * 8;                              // [TRN_THREADS]

    // k-major copies of the weight matrices
    for (int l = 0; l < L; ++l) {
        const int in = l == 0 ? 8 : h;
        for (int i = tid; i < GH * in; i += TRN_THREADS) WT[O.wih[l] + (i % in) * GH + i / in] = theta[O.wih[l] + i];
        for (int i = tid; i < GH * h; i += TRN_THREADS) WT[O.whh[l] + (i % h) * GH + i / h] = theta[O.whh[l] + i];
    }
    __syncthreads();

    const int n_batches = (p.n_samples + p.batch - 1) / p.batch;
    for (int bi = 0; bi < n_batches; ++bi) {
        const int b0 = bi * p.batch, B = min(p.batch, p.n_samples - b0);       // DataLoader(drop_last=False): the last batch may be short
        // ---- Adam's scalars for this step: python doubles in torch.optim.adam._single_tensor_adam, cast where torch casts them
        AdamC ac;
        {
            const double t = (double)(p.step0 + bi + 1);
            const double bc1 = 1.0 - pow(p.beta1, t), bc2 = 1.0 - pow(p.beta2, t);
            ac.step_size = (float)(-(p.lr / bc1)); ac.bc2_sqrt = (float)sqrt(bc2);
            ac.w1 = (float)(1.0 - p.beta1); ac.w2 = (float)(1.0 - p.beta2); ac.b2 = (float)p.beta2; ac.eps = (float)p.eps;
        }
        // A batch of more than RB rows (the shipped checkpoints were trained with BATCH_S 64) runs as micro-batches of RB rows:
        // forward + backward per micro-batch, parameter gradients summed in GA, ONE Adam step with the last micro-batch.
        float loss_sum = 0.f;                              // thread 0: sum of |out - label| over the batch
        for (int mb0 = 0; mb0 < B; mb0 += RB) {
        const int MB = min(RB, B - mb0);
        const bool first = mb0 == 0, lastmb = mb0 + RB >= B;
        // ---- gather the windows of this micro-batch: X0[f][b][:] (rows >= MB are zero and get zero gradients)
        for (int i = tid; i < T * RB * 8; i += TRN_THREADS) {
            const int f = i / (RB * 8), b = (i / 8) % RB, k = i & 7;
            X0[i] = b < MB ? __ldg(p.featrs + ((long long)perm[b0 + mb0 + b] * T + f) * 8 + k) : 0.f;
        }
        __syncthreads();
        // ================= forward: layer-major, frames ascending
        int cur = 0;
        for (int l = 0; l < L; ++l) {
            const int in = l == 0 ? 8 : h;
            const float* const wihT = WT + O.wih[l];
            const float* const whhT = WT + O.whh[l];
            const float* const bih = theta + O.bih[l];
            const float* const bhh = theta + O.bhh[l];
            for (int f = 0; f < T; ++f) {
                const float* xin = l == 0 ? X0 + f * RB * 8 : Y + ((size_t)(l - 1) * T + f) * RB * h;
                for (int i = tid; i < RB * in; i += TRN_THREADS) xs[i] = xin[i];
                __syncthreads();
                const float* hp = hs + cur * RB * h;
                float* hn = hs + (cur ^ 1) * RB * h;
                float* const Yo = Y + ((size_t)l * T + f) * RB * h;
                float* const So = S + ((size_t)l * T + f) * RB * NS * h;
                const float* const Sp = S + ((size_t)l * T + f - 1) * RB * NS * h;       // previous frame (LSTM c), f > 0 only
                // task = (4 batch rows, hidden unit jj): all gate pre-activations of the unit for the four rows in registers
                for (int task = tid; task < (RB / 4) * h; task += TRN_THREADS) {
                    const int bq = task / h, jj = task - bq * h, r0 = bq * 4;
                    float a[4][4];          // [gate][row]; GRU: r, z, n_x, n_h
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        if (KIND == 1) {
                            a[0][r] = bih[jj] + bhh[jj]; a[1][r] = bih[h + jj] + bhh[h + jj]; a[2][r] = bih[2 * h + jj]; a[3][r] = bhh[2 * h + jj];
                        } else {
#pragma unroll
                            for (int g = 0; g < 4; ++g) a[g][r] = bih[g * h + jj] + bhh[g * h + jj];
                        }
                    }
                    const float* wp = wihT + jj;
#pragma unroll 2
                    for (int k = 0; k < in; k += 4) {
                        float4 xv[4];
#pragma unroll
                        for (int r = 0; r < 4; ++r) xv[r] = *reinterpret_cast<const float4*>(xs + (r0 + r) * in + k);
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) {
                            float wg[G];
#pragma unroll
                            for (int g = 0; g < G; ++g) wg[g] = wp[g * h];
                            wp += GH;
#pragma unroll
                            for (int r = 0; r < 4; ++r) {
                                const float x = kk == 0 ? xv[r].x : kk == 1 ? xv[r].y : kk == 2 ? xv[r].z : xv[r].w;
#pragma unroll
                                for (int g = 0; g < G; ++g) a[g][r] = fmaf(x, wg[g], a[g][r]);
                            }
                        }
                    }
                    if (f > 0) {            // h_{-1} = 0: the recurrent product of frame 0 vanishes
                        const float* wq = whhT + jj;
#pragma unroll 2
                        for (int k = 0; k < h; k += 4) {
                            float4 xv[4];
#pragma unroll
                            for (int r = 0; r < 4; ++r) xv[r] = *reinterpret_cast<const float4*>(hp + (r0 + r) * h + k);
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk) {
                                float wg[G];
#pragma unroll
                                for (int g = 0; g < G; ++g) wg[g] = wq[g * h];
                                wq += GH;
#pragma unroll
                                for (int r = 0; r < 4; ++r) {
                                    const float x = kk == 0 ? xv[r].x : kk == 1 ? xv[r].y : kk == 2 ? xv[r].z : xv[r].w;
                                    if (KIND == 1) {
                                        a[0][r] = fmaf(x, wg[0], a[0][r]); a[1][r] = fmaf(x, wg[1], a[1][r]); a[3][r] = fmaf(x, wg[2], a[3][r]);
                                    } else {
#pragma unroll
                                        for (int g = 0; g < G; ++g) a[g][r] = fmaf(x, wg[g], a[g][r]);
                                    }
                                }
                            }
                        }
                    }
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const int b = r0 + r;
                        float hv;
                        if (KIND == 1) {
                            const float rg = sigm(a[0][r]), zg = sigm(a[1][r]), ng = tanhf(fmaf(rg, a[3][r], a[2][r]));
                            const float hprev = f > 0 ? hp[b * h + jj] : 0.f;
                            hv = fmaf(zg, hprev - ng, ng);                              // (1 - z) n + z h
                            float* s = So + (size_t)b * NS * h + jj;
                            s[0] = rg; s[h] = zg; s[2 * h] = ng; s[3 * h] = a[3][r];
                        } else {
                            const float ig = sigm(a[0][r]), fg = sigm(a[1][r]), gg = tanhf(a[2][r]), og = sigm(a[3][r]);
                            const float cp = f > 0 ? Sp[(size_t)b * NS * h + 4 * h + jj] : 0.f;
                            const float c = fmaf(fg, cp, ig * gg);
                            hv = og * tanhf(c);
                            float* s = So + (size_t)b * NS * h + jj;
                            s[0] = ig; s[h] = fg; s[2 * h] = gg; s[3 * h] = og; s[4 * h] = c;
                        }
                        hn[b * h + jj] = hv;
                        Yo[b * h + jj] = hv;
                    }
                }
                __syncthreads();
                cur ^= 1;
            }
        }
        // ================= head, L1 loss and its gradient
        const float* const htop = hs + cur * RB * h;
        float loss_part = 0.f;
        if (tid < RB * 6) {
            const int b = tid / 6, d = tid - b * 6;
            float s = theta[O.fcb + d];
            const float* w = theta + O.fcw + d * h;
            for (int k = 0; k < h; ++k) s = fmaf(htop[b * h + k], w[k], s);
            float dl = 0.f;
            if (b < MB) {
                const float diff = s - __ldg(p.labels + (long long)perm[b0 + mb0 + b] * 6 + d);
                const float inv = 1.0f / (float)(B * 6);
                loss_part = fabsf(diff);
                dl = diff > 0.f ? inv : diff < 0.f ? -inv : 0.f;
            }
            outs[b * 8 + d] = dl;
        }
        red[tid] = loss_part;
        __syncthreads();
        for (int o = TRN_THREADS / 2; o > 0; o >>= 1) {
            if (tid < o) red[tid] += red[tid + o];
            __syncthreads();
        }
        if (tid == 0) {
            loss_sum += red[0];
            if (lastmb && p.losses) p.losses[mi * n_batches + bi] = loss_sum / (float)(B * 6);
        }
        // ---- d loss / d h_top into DY[T-1] (the other frames of the top layer receive nothing from above)
        for (int i = tid; i < T * RB * h; i += TRN_THREADS) {
            const int f = i / (RB * h), b = (i / h) % RB, k = i % h;
            float s = 0.f;
            if (f == T - 1) {
#pragma unroll
                for (int d = 0; d < 6; ++d) s = fmaf(outs[b * 8 + d], theta[O.fcw + d * h + k], s);
            }
            DY[i] = s;
        }
        __syncthreads();
        // ---- head parameters: gradient in registers -> Adam
        for (int i = tid; i < 6 * h + 6; i += TRN_THREADS) {
            float g = 0.f;
            if (i < 6 * h) {
                const int d = i / h, k = i - d * h;
                for (int b = 0; b < RB; ++b) g = fmaf(outs[b * 8 + d], htop[b * h + k], g);
            } else {
                for (int b = 0; b < RB; ++b) g += outs[b * 8 + (i - 6 * h)];
            }
            const int gi = O.fcw + i;
            if (!first) g += GA[gi];
            if (lastmb) {
                float m = am[gi], v = av[gi];
                theta[gi] = adam_update(theta[gi], g, m, v, ac);
                am[gi] = m; av[gi] = v;
            } else {
                GA[gi] = g;
            }
        }
        // ================= backward through time, top layer first
        for (int l = L - 1; l >= 0; --l) {
            const int in = l == 0 ? 8 : h;
            const float* const wih = theta + O.wih[l];
            const float* const whh = theta + O.whh[l];
            for (int i = tid; i < RB * h; i += TRN_THREADS) { dhs[i] = 0.f; dcs[i] = 0.f; }
            __syncthreads();
            for (int f = T - 1; f >= 0; --f) {
                const float* const So = S + ((size_t)l * T + f) * RB * NS * h;
                const float* const Sp = S + ((size_t)l * T + f - 1) * RB * NS * h;
                const float* const Yp = Y + ((size_t)l * T + f - 1) * RB * h;          // h_{f-1}, f > 0 only
                float* const GIo = GI + (size_t)f * RB * GH;
                float* const GRo = GR + (size_t)f * RB * GH;
                // -- phase A: gate gradients of the cell, one (row, unit) per thread
                for (int i = tid; i < RB * h; i += TRN_THREADS) {
                    const int b = i / h, jj = i - b * h;
                    const float dh = DY[(size_t)f * RB * h + i] + dhs[i];
                    const float* s = So + (size_t)b * NS * h + jj;
                    if (KIND == 1) {
                        const float rg = s[0], zg = s[h], ng = s[2 * h], nh = s[3 * h];
                        const float hprev = f > 0 ? Yp[i] : 0.f;
                        const float dn = dh * (1.f - zg), dz = dh * (hprev - ng);
                        const float dan = dn * (1.f - ng * ng);
                        const float dar = dan * nh * rg * (1.f - rg), daz = dz * zg * (1.f - zg);
                        float* gi = gsi + b * GH + jj;
                        float* gr = gsr + b * GH + jj;
                        gi[0] = dar; gi[h] = daz; gi[2 * h] = dan;
                        gr[0] = dar; gr[h] = daz; gr[2 * h] = dan * rg;
                        float* go = GIo + b * GH + jj;
                        go[0] = dar; go[h] = daz; go[2 * h] = dan;
                        go = GRo + b * GH + jj;
                        go[0] = dar; go[h] = daz; go[2 * h] = dan * rg;
                        dhd[i] = dh * zg;
                    } else {
                        const float ig = s[0], fg = s[h], gg = s[2 * h], og = s[3 * h], c = s[4 * h];
                        const float cp = f > 0 ? Sp[(size_t)b * NS * h + 4 * h + jj] : 0.f;
                        const float tc = tanhf(c);
                        const float dc = dcs[i] + dh * og * (1.f - tc * tc);
                        const float dai = dc * gg * ig * (1.f - ig), daf = dc * cp * fg * (1.f - fg);
                        const float dag = dc * ig * (1.f - gg * gg), dao = dh * tc * og * (1.f - og);
                        float* gi = gsi + b * GH + jj;
                        gi[0] = dai; gi[h] = daf; gi[2 * h] = dag; gi[3 * h] = dao;
                        float* go = GIo + b * GH + jj;
                        go[0] = dai; go[h] = daf; go[2 * h] = dag; go[3 * h] = dao;
                        dcs[i] = dc * fg;
                        dhd[i] = 0.f;
                    }
                }
                __syncthreads();
                // -- phase B: dh_{f-1} = direct + g_r W_hh  (f > 0);  d x_f = g_i W_ih -> DY[f] of the layer below (l > 0).
                //    task = (2 rows, 4 columns k0..k0+3): lanes over column groups read W[j][k0..] as coalesced float4; running
                //    pointers, two j-steps of four rows in flight (the loop waits on L2 / L1 latency, not on bandwidth)
                const float* const gr_s = KIND == 1 ? gsr : gsi;
                const int nt_h = f > 0 ? (RB / 2) * (h / 4) : 0, nt_x = l > 0 ? (RB / 2) * (in / 4) : 0;
                for (int task = tid; task < nt_h + nt_x; task += TRN_THREADS) {
                    const bool rec = task < nt_h;
                    const int t2 = rec ? task : task - nt_h;
                    const int ncol = rec ? h : in, ncq = ncol >> 2;
                    const int bq = t2 / ncq, k0 = (t2 - bq * ncq) * 4, r0 = bq * 2;
                    const float* wp = (rec ? whh : wih) + k0;
                    const float* g0 = (rec ? gr_s : gsi) + r0 * GH;
                    const float* g1 = g0 + GH;
                    float a0[4] = {0.f, 0.f, 0.f, 0.f}, a1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 2
                    for (int j = 0; j < GH; j += 4) {
                        const float4 ga = *reinterpret_cast<const float4*>(g0 + j), gb = *reinterpret_cast<const float4*>(g1 + j);
                        float4 w[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) { w[q] = *reinterpret_cast<const float4*>(wp); wp += ncol; }
                        const float gav[4] = {ga.x, ga.y, ga.z, ga.w}, gbv[4] = {gb.x, gb.y, gb.z, gb.w};
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            a0[0] = fmaf(gav[q], w[q].x, a0[0]); a0[1] = fmaf(gav[q], w[q].y, a0[1]);
                            a0[2] = fmaf(gav[q], w[q].z, a0[2]); a0[3] = fmaf(gav[q], w[q].w, a0[3]);
                            a1[0] = fmaf(gbv[q], w[q].x, a1[0]); a1[1] = fmaf(gbv[q], w[q].y, a1[1]);
                            a1[2] = fmaf(gbv[q], w[q].z, a1[2]); a1[3] = fmaf(gbv[q], w[q].w, a1[3]);
                        }
                    }
                    if (rec) {
                        float* d0 = dhs + r0 * h + k0;
                        const float* e0 = dhd + r0 * h + k0;
#pragma unroll
                        for (int c = 0; c < 4; ++c) { d0[c] = e0[c] + a0[c]; d0[h + c] = e0[h + c] + a1[c]; }
                    } else {
                        float* d0 = DY + (size_t)f * RB * h + r0 * h + k0;
                        *reinterpret_cast<float4*>(d0) = make_float4(a0[0], a0[1], a0[2], a0[3]);
                        *reinterpret_cast<float4*>(d0 + h) = make_float4(a1[0], a1[1], a1[2], a1[3]);
                    }
                }
                __syncthreads();
            }
            // -- phase C: parameter gradients of the layer, summed over frames and rows in registers, straight into Adam.
            //    W_ih [GH][in]: x_f = X0 / Y[l-1];  W_hh [GH][h]: h_{f-1} = Y[l][f-1] (frame 0 contributes nothing).
            //    task = a 4 x 4 tile (rows j0.., columns k0..)
            {
                const float* const Xl = l == 0 ? X0 : Y + (size_t)(l - 1) * T * RB * h;
                const float* const Yl = Y + (size_t)l * T * RB * h;
                const int tiles_ih = (GH / 4) * (in / 4), tiles_hh = (GH / 4) * (h / 4);
                for (int task = tid; task < tiles_ih + tiles_hh; task += TRN_THREADS) {
                    const bool ih = task < tiles_ih;
                    const int t2 = ih ? task : task - tiles_ih;
                    const int ncol = ih ? in : h;
                    const int jt = t2 / (ncol / 4), kt = t2 - jt * (ncol / 4), j0 = jt * 4, k0 = kt * 4;
                    float acc[4][4];
#pragma unroll
                    for (int a = 0; a < 4; ++a)
#pragma unroll
                        for (int c = 0; c < 4; ++c) acc[a][c] = 0.f;
                    const float* gsrc = ih ? GI : GR;
                    if (KIND != 1) gsrc = GI;
                    // frames and rows are contiguous: one flat loop over (f, b) with running pointers; W_hh pairs g_f with h_{f-1}
                    const float* gp = gsrc + (ih ? 0 : (size_t)RB * GH) + j0;
                    const float* xp = (ih ? Xl : Yl) + k0;
                    const int nfb = (ih ? T : T - 1) * RB;
#pragma unroll 4
                    for (int n = 0; n < nfb; ++n) {
                        const float4 gv = *reinterpret_cast<const float4*>(gp);
                        const float4 xv = *reinterpret_cast<const float4*>(xp);
                        gp += GH; xp += ncol;
                        const float ga[4] = {gv.x, gv.y, gv.z, gv.w}, xa[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
                        for (int a = 0; a < 4; ++a)
#pragma unroll
                            for (int c = 0; c < 4; ++c) acc[a][c] = fmaf(ga[a], xa[c], acc[a][c]);
                    }
                    const int base = ih ? O.wih[l] : O.whh[l];
#pragma unroll
                    for (int a = 0; a < 4; ++a) {
                        const int gi = base + (j0 + a) * ncol + k0;
                        if (!first) {
                            const float4 gp4 = *reinterpret_cast<const float4*>(GA + gi);
                            acc[a][0] += gp4.x; acc[a][1] += gp4.y; acc[a][2] += gp4.z; acc[a][3] += gp4.w;
                        }
                        if (!lastmb) {
                            *reinterpret_cast<float4*>(GA + gi) = make_float4(acc[a][0], acc[a][1], acc[a][2], acc[a][3]);
                            continue;
                        }
                        float4 w = *reinterpret_cast<const float4*>(theta + gi);
                        float4 m = *reinterpret_cast<const float4*>(am + gi);
                        float4 v = *reinterpret_cast<const float4*>(av + gi);
                        w.x = adam_update(w.x, acc[a][0], m.x, v.x, ac); w.y = adam_update(w.y, acc[a][1], m.y, v.y, ac);
                        w.z = adam_update(w.z, acc[a][2], m.z, v.z, ac); w.w = adam_update(w.w, acc[a][3], m.w, v.w, ac);
                        *reinterpret_cast<float4*>(theta + gi) = w;
                        *reinterpret_cast<float4*>(am + gi) = m;
                        *reinterpret_cast<float4*>(av + gi) = v;
                        WT[base + (size_t)(k0 + 0) * GH + j0 + a] = w.x; WT[base + (size_t)(k0 + 1) * GH + j0 + a] = w.y;
                        WT[base + (size_t)(k0 + 2) * GH + j0 + a] = w.z; WT[base + (size_t)(k0 + 3) * GH + j0 + a] = w.w;
                    }
                }
                // biases: b_ih[j] <- sum g_i, b_hh[j] <- sum g_r
                for (int i = tid; i < 2 * GH; i += TRN_THREADS) {
                    const bool ih = i < GH;
                    const int j = ih ? i : i - GH;
                    const float* gsrc = (ih || KIND != 1) ? GI : GR;
                    float g = 0.f;
                    for (int f = 0; f < T; ++f)
                        for (int b = 0; b < RB; ++b) g += gsrc[((size_t)f * RB + b) * GH + j];
                    const int gi = (ih ? O.bih[l] : O.bhh[l]) + j;
                    if (!first) g += GA[gi];
                    if (lastmb) {
                        float m = am[gi], v = av[gi];
                        theta[gi] = adam_update(theta[gi], g, m, v, ac);
                        am[gi] = m; av[gi] = v;
                    } else {
                        GA[gi] = g;
                    }
                }
            }
            __syncthreads();
        }
        }   // micro-batches
    }
}

long long train_rnn_param_count(int kind, int hidden, int layers) {
    const long long G = kind == 1 ? 3 : 4, h = hidden, GH = G * h;
    long long o = 0;
    for (int l = 0; l < layers; ++l) o += GH * (l == 0 ? 8 : h) + GH * h + 2 * GH;
    return o + 6 * h + 6;
}

long long train_rnn_work_floats(int kind, int hidden, int layers, int window) {
    const long long G = kind == 1 ? 3 : 4, NS = kind == 1 ? 4 : 5, h = hidden, GH = G * h, T = window, L = layers;
    const long long np = (train_rnn_param_count(kind, hidden, layers) + 3) & ~3LL;
    return np + L * T * RB * h + L * T * RB * NS * h + 2 * T * RB * GH + T * RB * h + T * RB * 8 + np;      // ... + the gradient sums GA
}

int train_rnn_smem_bytes(int kind, int hidden) {
    const int G = kind == 1 ? 3 : 4, h = hidden, GH = G * h;
    return (RB * h + 2 * RB * h + 2 * RB * GH + 3 * RB * h + RB * 8 + TRN_THREADS) * 4;
}

template <int KIND, int MINB>
static cudaError_t launch_train_rnn_t(const TrainRnnParams& p, int n_models, int smem_bytes, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(train_rnn_kernel<KIND, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    train_rnn_kernel<KIND, MINB><<<n_models, TRN_THREADS, smem_bytes, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_train_rnn(const TrainRnnParams& p, int n_models, int num_sms, int smem_bytes, cudaStream_t st) {
    const bool two = n_models > num_sms && 2 * (smem_bytes + 1024) <= SMEM_LIMIT;       // two CTAs fit one SM's shared memory
    if (p.kind == 1) return two ? launch_train_rnn_t<1, 2>(p, n_models, smem_bytes, st) : launch_train_rnn_t<1, 1>(p, n_models, smem_bytes, st);
    return two ? launch_train_rnn_t<2, 2>(p, n_models, smem_bytes, st) : launch_train_rnn_t<2, 1>(p, n_models, smem_bytes, st);
}

}  // namespace mpcdl
