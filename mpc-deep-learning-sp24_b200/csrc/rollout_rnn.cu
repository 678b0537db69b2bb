// K2/K3: fused whole-horizon rollout of the GRU / LSTM window models.
//   reference: GRU_Model.forward   model_scripts/deep_models.py:57-66
//              LSTM_Model.forward  model_scripts/deep_models.py:81-92
//              per-position z-score model_scripts/deep_models.py:29-36
//              closed loop         oracle/rollout_ref.py rollout_rnn
// Per rollout step the CTA re-encodes its tile's T-frame window layer by layer (layer-outer,
// frame-inner): the layer's output sequence overwrites its input sequence in shared memory
// (seq[T][HID][TN]), the LSTM cell state lives in registers across the frames of a layer, and the
// x- and h-projections of every cell stream through the TMA weight ring from L2.
#include "owner.cuh"
#include "launch.h"

namespace mpcdl {

enum { CELL_GRU = 0, CELL_LSTM = 1 };

template <int HID, int RT, int RJ>
struct RnnCfg {
    static constexpr int N_JG = HID / RJ;
    static constexpr int RTO = RT >= 4 ? 4 : 1;
    static constexpr int NEED = (RT * 8 * RJ + RTO * HID - 1) / (RTO * HID);
    static constexpr int RJO = NEED <= 1 ? 1 : NEED <= 2 ? 2 : NEED <= 4 ? 4 : 8;
    static constexpr int RTN = RT >= 4 ? 4 : RT;
};

template <int CELL, int HID, int RT, int RJ>
__global__ void __launch_bounds__(rnn_max_consumers(RT, RJ) + 32, 1) rollout_rnn_kernel(const __grid_constant__ RolloutParams p) {
    using C = RnnCfg<HID, RT, RJ>;
    constexpr int NGW = CELL == CELL_GRU ? 3 : 4;     // gates per weight row
    constexpr int NGA = 4;                            // accumulator blocks: GRU r,z,nx,nh / LSTM i,f,g,o
    extern __shared__ __align__(128) float sm[];
    const ModelParams& m = p.m;
    const SmemLayout L = make_layout(m, p.nz, true);
    const int TN = m.TN, n_tg = m.n_tg, nct = m.n_cthreads, T = m.window, NL = m.layers;
    const int tid = threadIdx.x;
    const int H = p.io.H;
    const long long N = p.io.N;
    const uint32_t kc_floats = m.slot_floats;
    const bool stateful = m.stateful != 0;

    uint64_t* full = reinterpret_cast<uint64_t*>(sm + L.bars);
    uint64_t* empty = full + m.nslots;
    Ring ring{sm + L.ring, full, empty, m.blob, 0u, (uint32_t)m.nslots, (uint32_t)m.slot_floats, m.data_off, m.resident != 0};

    if (tid == 0) {
        for (int i = 0; i < m.nslots; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], nct / 32);
        }
        mbar_fence_init();
    }
    float* normS = sm + L.norm;                       // f_mean[T*8] f_std[T*8] l_mean[6] l_std[6]
    for (int i = tid; i < 2 * T * 8 + 12; i += blockDim.x) normS[i] = p.io.norm[i];
    if (p.nz.enabled)
        for (int i = tid; i < p.nz.blob_floats; i += blockDim.x) sm[L.nblob + i] = p.nz.blob[i];
    __syncthreads();

    const int my_tiles = (p.io.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const PartDesc& fc_part = m.parts[2 * NL];

    // ------------------------------------------------------------------ producer warp
    if (tid >= nct) {
        if (tid == nct) {
            if (m.resident) {
                ring_produce_resident(ring, m.data_floats);
            } else {
                for (int it = 0; it < my_tiles; ++it) {
                    ring.blob = m.blob + (size_t)(((int)blockIdx.x + it * (int)gridDim.x) / p.io.tpm) * m.blob_stride;   // the tile's model
                    for (int t = 0; t < H; ++t) {
                        for (int l = 0; l < NL; ++l)
                            for (int f = 0; f < T; ++f) {
                                ring_produce_part(ring, m.parts[2 * l], kc_floats);
                                if (f > 0 || stateful) ring_produce_part(ring, m.parts[2 * l + 1], kc_floats);
                            }
                        ring_produce_part(ring, fc_part, kc_floats);
                    }
                }
            }
        }
        return;
    }
    if (m.resident) mbar_wait(&full[0], 0);

    // ------------------------------------------------------------------ consumers
    constexpr int VEC = TrajVec<RT>::VEC, NH = TrajVec<RT>::NH;
    const int hstride = VEC * n_tg;
    const bool active = tid < n_tg * C::N_JG;
    const int tg = tid % n_tg, jg = tid / n_tg;
    constexpr int VECO = TrajVec<C::RTO>::VEC;
    const int n_tgo = TN / C::RTO, hstride_o = VECO * n_tgo;
    const bool active_o = tid < n_tgo * (8 / C::RJO);
    const int tgo = tid % n_tgo, jgo = tid / n_tgo;

    float* xn = sm + L.xin;
    float* xwin = sm + L.xwin;
    float* seq = sm + L.act0;
    float* hc = sm + L.act1;
    float* ybuf = sm + L.ybuf;
    float* nin = sm + L.nin;
    float* const nact0 = sm + L.nact0;
    const int nact_flip = L.nact1 - L.nact0;
    float* nzb = sm + L.nzb;
    const int seq_stride = HID * TN;
    const int CARRY_C = NL;                            // LSTM: c planes follow the h planes

    for (int tile = blockIdx.x; tile < p.io.ntiles; tile += gridDim.x) {
        const int mi = tile / p.io.tpm;                // stacked models: the tile's model (0 for one model)
        const long long lim = (long long)(mi + 1) * p.io.npm;
        const long long tile0 = (long long)mi * p.io.npm + (long long)(tile - mi * p.io.tpm) * TN;
        const bool owner = tid < TN;
        long long n = tile0 + tid;
        const bool valid = owner && n < lim;
        if (n >= lim) n = lim - 1;
        const long long n_in = p.io.share_inputs ? n - (long long)mi * p.io.npm : n;
        float cost = 0.f;
        if (owner) {                                   // initial window -> xwin[f][k][pos]
            const float4* w4 = reinterpret_cast<const float4*>(p.io.init + n_in * p.io.init_mul * T * 8);
            for (int f = 0; f < T; ++f) {
                float4 a = __ldg(w4 + 2 * f), b = __ldg(w4 + 2 * f + 1);
                float* d = xwin + f * 8 * TN + tid;
                d[0] = a.x; d[TN] = a.y; d[2 * TN] = a.z; d[3 * TN] = a.w;
                d[4 * TN] = b.x; d[5 * TN] = b.y; d[6 * TN] = b.z; d[7 * TN] = b.w;
            }
        }
        named_bar_sync(1, nct);
        int base = 0;                                  // physical slot of logical frame 0
        for (int t = 0; t < H; ++t) {
            // ---- P0: per-position z-score of the window; noise history push
            for (int idx = tid; idx < T * 8 * TN; idx += nct) {
                int f = idx / (8 * TN), rem = idx - f * 8 * TN, k = rem / TN;
                int pf = base + f; if (pf >= T) pf -= T;
                xn[idx] = __fdiv_rn(__fsub_rn(xwin[pf * 8 * TN + rem], normS[f * 8 + k]), normS[T * 8 + f * 8 + k]);
            }
            float u[2] = {0.f, 0.f};
            if (owner) {
                float2 uu = load_control(p.io, n_in, t);
                u[0] = uu.x; u[1] = uu.y;
                if (p.nz.enabled) {                    // newest frame feeds the noise history
                    int pf = base + T - 1; if (pf >= T) pf -= T;
                    float s[6], uf[2];
#pragma unroll
                    for (int d = 0; d < 6; ++d) s[d] = xwin[(pf * 8 + d) * TN + tid];
                    uf[0] = xwin[(pf * 8 + 6) * TN + tid]; uf[1] = xwin[(pf * 8 + 7) * TN + tid];
                    noise_push(nin, TN, tid, p.nz.hist, t == 0, s, uf);
                }
            }
            named_bar_sync(1, nct);
            // ---- recurrent stack
            for (int l = 0; l < NL; ++l) {
                float cst[RT][RJ];                     // LSTM cell state, register resident
#pragma unroll
                for (int i = 0; i < RT; ++i)
#pragma unroll
                    for (int r = 0; r < RJ; ++r) cst[i][r] = 0.f;
                if (stateful) {                        // carry -> hc (smem) / cst (registers)
                    if (active) {
#pragma unroll
                        for (int i = 0; i < RT; ++i) {
                            long long q = tile0 + (i / VEC) * hstride + tg * VEC + (i % VEC);
                            if (q >= lim) q = lim - 1;
                            const float* ch = p.io.carry + ((long long)l * N + q) * HID + jg * RJ;
                            int pos = (i / VEC) * hstride + tg * VEC + (i % VEC);
#pragma unroll
                            for (int r = 0; r < RJ; ++r) hc[(jg * RJ + r) * TN + pos] = ch[r];
                            if (CELL == CELL_LSTM) {
                                const float* cc = p.io.carry + ((long long)(CARRY_C + l) * N + q) * HID + jg * RJ;
#pragma unroll
                                for (int r = 0; r < RJ; ++r) cst[i][r] = cc[r];
                            }
                        }
                    }
                    named_bar_sync(1, nct);
                }
                for (int f = 0; f < T; ++f) {
                    float acc[RT][NGA * RJ];
                    if (CELL == CELL_GRU) {            // nh block has no x-part contribution
#pragma unroll
                        for (int i = 0; i < RT; ++i)
#pragma unroll
                            for (int r = 0; r < RJ; ++r) acc[i][3 * RJ + r] = 0.f;
                    }
                    const float* xin_l = (l == 0 ? xn + f * 8 * TN : seq + f * seq_stride) + tg * VEC;
                    consume_part<RT, RJ, NGW, NGA, GateMap<0, 1, 2, 3>, 4>(acc, ring, m.parts[2 * l], kc_floats, xin_l, TN,
                                                                           hstride, jg * NGW * RJ, jg * 4 * RJ, active);
                    const bool has_h = f > 0 || stateful;
                    const float* hprev = (f > 0 ? seq + (f - 1) * seq_stride : hc);
                    if (has_h) {
                        if (CELL == CELL_GRU)
                            consume_part<RT, RJ, NGW, NGA, GateMap<0, 1, 3, 3>, 4>(acc, ring, m.parts[2 * l + 1], kc_floats,
                                                                                   hprev + tg * VEC, TN, hstride,
                                                                                   jg * NGW * RJ, 0, active);
                        else
                            consume_part<RT, RJ, NGW, NGA, GateMap<0, 1, 2, 3>, 4>(acc, ring, m.parts[2 * l + 1], kc_floats,
                                                                                   hprev + tg * VEC, TN, hstride,
                                                                                   jg * NGW * RJ, 0, active);
                    }
                    named_bar_sync(1, nct);            // every read of seq[f] (layer l-1 output) is done
                    if (active) {
                        float* out = seq + f * seq_stride;
#pragma unroll
                        for (int r = 0; r < RJ; ++r) {
                            const int j = jg * RJ + r;
                            float hold[RT], hnew[RT];
                            if (has_h) load_traj<RT>(hold, hprev + j * TN + tg * VEC, hstride);
                            else {
#pragma unroll
                                for (int i = 0; i < RT; ++i) hold[i] = 0.f;
                            }
#pragma unroll
                            for (int i = 0; i < RT; ++i) {
                                if (CELL == CELL_GRU) {
                                    // torch GRU cell: n = tanh(i_n + r*h_n); h' = (h - n)*z + n
                                    float rg = sigmoid_f(acc[i][r]);
                                    float zg = sigmoid_f(acc[i][RJ + r]);
                                    float ng = tanh_f(__fadd_rn(acc[i][2 * RJ + r], __fmul_rn(rg, acc[i][3 * RJ + r])));
                                    hnew[i] = __fadd_rn(__fmul_rn(__fsub_rn(hold[i], ng), zg), ng);
                                } else {
                                    // torch LSTM cell: c' = f*c + i*g; h' = o*tanh(c')
                                    float ig = sigmoid_f(acc[i][r]);
                                    float fg = sigmoid_f(acc[i][RJ + r]);
                                    float gg = tanh_f(acc[i][2 * RJ + r]);
                                    float og = sigmoid_f(acc[i][3 * RJ + r]);
                                    float cn = __fadd_rn(__fmul_rn(fg, cst[i][r]), __fmul_rn(ig, gg));
                                    cst[i][r] = cn;
                                    hnew[i] = __fmul_rn(og, tanh_f(cn));
                                }
                            }
                            store_traj<RT>(out + j * TN + tg * VEC, hstride, hnew);
                            if (stateful && f == T - 1) {
#pragma unroll
                                for (int i = 0; i < RT; ++i) {
                                    long long q = tile0 + (i / VEC) * hstride + tg * VEC + (i % VEC);
                                    if (q < lim) {
                                        p.io.carry[((long long)l * N + q) * HID + j] = hnew[i];
                                        if (CELL == CELL_LSTM)
                                            p.io.carry[((long long)(CARRY_C + l) * N + q) * HID + j] = cst[i][r];
                                    }
                                }
                            }
                        }
                    }
                    named_bar_sync(1, nct);
                }
            }
            // ---- head: full_c(out[:, -1, :]) -> ybuf ; residual-noise net
            {
                float acc[C::RTO][C::RJO];
                consume_part<C::RTO, C::RJO, 1, 1, GateMap<0>, 1>(acc, ring, fc_part, kc_floats,
                                                                   seq + (T - 1) * seq_stride + tgo * VECO, TN, hstride_o,
                                                                   jgo * C::RJO, jgo * C::RJO, active_o);
                if (active_o) {
#pragma unroll
                    for (int r = 0; r < C::RJO; ++r) {
                        float o[C::RTO];
#pragma unroll
                        for (int i = 0; i < C::RTO; ++i) o[i] = acc[i][r];
                        store_traj<C::RTO>(ybuf + (jgo * C::RJO + r) * TN + tgo * VECO, hstride_o, o);
                    }
                }
            }
            if (p.nz.enabled) {
                const float* W = sm + L.nblob;
                int K = p.nz.in_dim;
                for (int ph = 0; ph <= p.nz.nlayers; ++ph) {
                    const bool last = ph == p.nz.nlayers;
                    const int outw = last ? 12 : p.nz.size;
                    dense_resident<C::RTN, 4>(ph == 0 ? nin : nact0 + ((ph - 1) & 1) * nact_flip, K, W, outw, outw,
                                              last ? ACT_IDENTITY : p.nz.act, last ? nzb : nact0 + (ph & 1) * nact_flip, TN, tid, nct);
                    W += (K + 1) * outw; K = p.nz.size;
                    if (!last) named_bar_sync(1, nct);
                }
            }
            named_bar_sync(1, nct);
            // ---- owner: next state, shift the window (overwrite the oldest physical frame)
            if (owner) {
                float nxt[6];
                owner_finish(p, normS + 2 * T * 8, ybuf, nzb, TN, tid, n, valid, t, u, nxt, cost);
                float* d = xwin + base * 8 * TN + tid;
#pragma unroll
                for (int k = 0; k < 6; ++k) d[k * TN] = nxt[k];
                d[6 * TN] = u[0]; d[7 * TN] = u[1];
            }
            base = base + 1 == T ? 0 : base + 1;
            named_bar_sync(1, nct);
        }
        if (valid && p.io.costs) p.io.costs[n] = cost;
    }
}

// ------------------------------------------------------------------ launcher
template <int CELL, int HID, int RT, int RJ>
static cudaError_t launch_rnn_t(const RolloutParams& p, int grid, int smem_bytes, cudaStream_t st) {
    auto k = rollout_rnn_kernel<CELL, HID, RT, RJ>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    k<<<grid, p.m.n_cthreads + 32, smem_bytes, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_rnn(const RolloutParams& p, int cell, int rt, int rj, int grid, int smem_bytes, cudaStream_t st) {
#define MPCDL_CASE(H_, RT_, RJ_)                                                                         \
    if (p.m.hid == H_ && rt == RT_ && rj == RJ_)                                                         \
        return cell == CELL_GRU ? launch_rnn_t<CELL_GRU, H_, RT_, RJ_>(p, grid, smem_bytes, st)          \
                                : launch_rnn_t<CELL_LSTM, H_, RT_, RJ_>(p, grid, smem_bytes, st);
    MPCDL_RNN_CONFIGS(MPCDL_CASE)
#undef MPCDL_CASE
    return cudaErrorInvalidConfiguration;
}

}  // namespace mpcdl
