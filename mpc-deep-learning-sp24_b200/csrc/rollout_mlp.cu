// K1: fused whole-horizon rollout of the ReLU-MLP forward-dynamics model.
//   reference: Basic_Model.forward  model_scripts/basic_model.py:58-61 (ctor 45-56)
//              z-score               model_scripts/basic_model.py:28-35
//              closed loop           oracle/rollout_ref.py rollout_mlp
// One persistent CTA owns a tile of TN trajectories for all H steps; layer matrices arrive through
// the TMA weight ring (resident when they fit), activations ping-pong in shared memory as
// [k][TN], every consumer thread keeps an RT x RJ accumulator tile in registers, and the raw
// vehicle state + running cost stay in the registers of the trajectory's owner thread.
#include "owner.cuh"
#include "launch.h"

namespace mpcdl {

template <int HID, int RT, int RJ>
struct MlpCfg {
    static constexpr int N_JG = HID / RJ;
    static constexpr int RTO = RT >= 4 ? 4 : 1;                      // output-layer tile (width 8)
    static constexpr int NEED = (RT * 8 * RJ + RTO * HID - 1) / (RTO * HID);
    static constexpr int RJO = NEED <= 1 ? 1 : NEED <= 2 ? 2 : NEED <= 4 ? 4 : 8;
    static constexpr int RTN = RT >= 4 ? 4 : RT;                     // noise-net tile
};

template <int HID, int RT, int RJ>
__global__ void __launch_bounds__(MLP_MAX_THREADS, 1) rollout_mlp_kernel(const __grid_constant__ RolloutParams p) {
    using C = MlpCfg<HID, RT, RJ>;
    extern __shared__ __align__(128) float sm[];
    const ModelParams& m = p.m;
    const SmemLayout L = make_layout(m, p.nz, false);
    const int TN = m.TN, n_tg = m.n_tg, nct = m.n_cthreads;
    const int tid = threadIdx.x;
    const int H = p.io.H;
    const uint32_t kc_floats = m.slot_floats;

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
    // tile-invariant constants
    float* normS = sm + L.norm;
    for (int i = tid; i < 28; i += blockDim.x) normS[i] = p.io.norm[i];          // f_mean8 f_std8 l_mean6 l_std6
    if (p.nz.enabled)
        for (int i = tid; i < p.nz.blob_floats; i += blockDim.x) sm[L.nblob + i] = p.nz.blob[i];
    __syncthreads();

    const int my_tiles = (p.io.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    // ------------------------------------------------------------------ producer warp
    if (tid >= nct) {
        if (tid == nct) {
            if (m.resident) {
                ring_produce_resident(ring, m.data_floats);
            } else {
                for (int it = 0; it < my_tiles; ++it) {
                    ring.blob = m.blob + (size_t)(((int)blockIdx.x + it * (int)gridDim.x) / p.io.tpm) * m.blob_stride;   // the tile's model
                    for (int t = 0; t < H; ++t)
                        for (int i = 0; i < m.nparts; ++i) ring_produce_part(ring, m.parts[i], kc_floats);
                }
            }
        }
        return;
    }
    if (m.resident) mbar_wait(&full[0], 0);

    // ------------------------------------------------------------------ consumers
    constexpr int VEC = TrajVec<RT>::VEC;
    const int hstride = VEC * n_tg;
    const bool active = tid < n_tg * C::N_JG;
    const int tg = tid % n_tg, jg = tid / n_tg;
    // output layer mapping (width 8, RTO x RJO tiles)
    constexpr int VECO = TrajVec<C::RTO>::VEC;
    const int n_tgo = TN / C::RTO, hstride_o = VECO * n_tgo;
    const bool active_o = tid < n_tgo * (8 / C::RJO);
    const int tgo = tid % n_tgo, jgo = tid / n_tgo;

    float* xin = sm + L.xin;
    float* const act0 = sm + L.act0;
    const int act_flip = L.act1 - L.act0;           // act(i) = act0 + (i&1)*act_flip (no local arrays)
    float* ybuf = sm + L.ybuf;
    float* nin = sm + L.nin;
    float* const nact0 = sm + L.nact0;
    const int nact_flip = L.nact1 - L.nact0;
    float* nzb = sm + L.nzb;
    const int nphase = max(m.layers + 2, p.nz.enabled ? p.nz.nlayers + 1 : 0);

    for (int tile = blockIdx.x; tile < p.io.ntiles; tile += gridDim.x) {
        const bool owner = tid < TN;
        const int mi = tile / p.io.tpm;                                  // stacked models: the tile's model (0 for one model)
        const long long lim = (long long)(mi + 1) * p.io.npm;
        long long n = (long long)mi * p.io.npm + (long long)(tile - mi * p.io.tpm) * TN + tid;
        const bool valid = owner && n < lim;
        if (n >= lim) n = lim - 1;
        const long long n_in = p.io.share_inputs ? n - (long long)mi * p.io.npm : n;
        float s[6], cost = 0.f;
        if (owner) {
#pragma unroll
            for (int d = 0; d < 6; ++d) s[d] = __ldg(p.io.init + n_in * p.io.init_mul * 6 + d);
        }
        for (int t = 0; t < H; ++t) {
            float u[2] = {0.f, 0.f};
            if (owner) {
                float2 uu = load_control(p.io, n_in, t);
                u[0] = uu.x; u[1] = uu.y;
#pragma unroll
                for (int k = 0; k < 6; ++k) xin[k * TN + tid] = __fdiv_rn(__fsub_rn(s[k], normS[k]), normS[8 + k]);
#pragma unroll
                for (int k = 0; k < 2; ++k) xin[(6 + k) * TN + tid] = __fdiv_rn(__fsub_rn(u[k], normS[6 + k]), normS[14 + k]);
                if (p.nz.enabled) noise_push(nin, TN, tid, p.nz.hist, t == 0, s, u);
            }
            named_bar_sync(1, nct);
            for (int ph = 0; ph < nphase; ++ph) {
                if (ph <= m.layers) {                       // input layer (ph 0) and hidden layers: -> relu
                    float acc[RT][RJ];
                    const float* in = ph == 0 ? xin : act0 + ((ph - 1) & 1) * act_flip;
                    consume_part<RT, RJ, 1, 1, GateMap<0>, 1>(acc, ring, m.parts[ph], kc_floats, in + tg * VEC, TN, hstride,
                                                             jg * RJ, jg * RJ, active);
                    if (active) {
                        float* out = act0 + (ph & 1) * act_flip;
#pragma unroll
                        for (int r = 0; r < RJ; ++r) {
                            float o[RT];
#pragma unroll
                            for (int i = 0; i < RT; ++i) o[i] = fmaxf(acc[i][r], 0.f);
                            store_traj<RT>(out + (jg * RJ + r) * TN + tg * VEC, hstride, o);
                        }
                    }
                } else if (ph == m.layers + 1) {            // output layer, identity
                    float acc[C::RTO][C::RJO];
                    consume_part<C::RTO, C::RJO, 1, 1, GateMap<0>, 1>(acc, ring, m.parts[ph], kc_floats,
                                                                       act0 + ((ph - 1) & 1) * act_flip + tgo * VECO, TN, hstride_o,
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
                if (p.nz.enabled && ph <= p.nz.nlayers) {   // residual-noise net rides the same phases
                    const float* W = sm + L.nblob;
                    int K = p.nz.in_dim;
                    for (int j = 0; j < ph; ++j) { W += (K + 1) * p.nz.size; K = p.nz.size; }
                    const bool last = ph == p.nz.nlayers;
                    const int outw = last ? 12 : p.nz.size;
                    dense_resident<C::RTN, 4>(ph == 0 ? nin : nact0 + ((ph - 1) & 1) * nact_flip, K, W, outw, outw,
                                              last ? ACT_IDENTITY : p.nz.act, last ? nzb : nact0 + (ph & 1) * nact_flip, TN, tid, nct);
                }
                named_bar_sync(1, nct);
            }
            if (owner) {
                float nxt[6];
                owner_finish(p, normS + 16, ybuf, nzb, TN, tid, n, valid, t, u, nxt, cost);
#pragma unroll
                for (int d = 0; d < 6; ++d) s[d] = nxt[d];
            }
        }
        if (valid && p.io.costs) p.io.costs[n] = cost;
    }
}

// ------------------------------------------------------------------ launcher
template <int HID, int RT, int RJ>
static cudaError_t launch_mlp_t(const RolloutParams& p, int grid, int smem_bytes, cudaStream_t st) {
    auto k = rollout_mlp_kernel<HID, RT, RJ>;
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    k<<<grid, p.m.n_cthreads + 32, smem_bytes, st>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_mlp(const RolloutParams& p, int rt, int rj, int grid, int smem_bytes, cudaStream_t st) {
#define MPCDL_CASE(H_, RT_, RJ_) \
    if (p.m.hid == H_ && rt == RT_ && rj == RJ_) return launch_mlp_t<H_, RT_, RJ_>(p, grid, smem_bytes, st);
    MPCDL_MLP_CONFIGS(MPCDL_CASE)
#undef MPCDL_CASE
    return cudaErrorInvalidConfiguration;
}

}  // namespace mpcdl
