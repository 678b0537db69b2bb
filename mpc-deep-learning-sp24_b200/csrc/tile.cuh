// Register-tiled FP32 GEMM pieces shared by the MLP and GRU/LSTM rollout kernels, plus the
// weight ring (TMA bulk copies + full/empty mbarriers) that feeds them.
//
// Data layout in shared memory
//   activations  A[k][TN]      k-major, one column per trajectory of the CTA's tile
//   weights      W[k][width]   k-major rows of the (pre-transposed, gate-interleaved) layer matrix
// A thread owns an RT x (NG*RJ) register tile: RT trajectories x RJ hidden units x NG gates.
// Trajectory positions of thread-group `tg` (VEC = min(RT,4), NH = RT/VEC):
//   pos(h, i) = h*VEC*n_tg + tg*VEC + i      -> every LDS/STS is a contiguous, conflict-free vector
#pragma once
#include "common.cuh"

namespace mpcdl {

struct PartDesc {          // one layer matrix inside the packed blob (offsets in floats)
    uint32_t off;          // start: [bias block (bias_floats)] [nrows x width]
    uint16_t nrows;
    uint16_t width;
    uint32_t bias_floats;
};

// ------------------------------------------------------------------ weight ring
struct Ring {
    float* base;           // smem
    uint64_t* full;        // [nslots] (resident mode: only full[0])
    uint64_t* empty;       // [nslots]
    const float* blob;     // gmem
    uint32_t c;            // running chunk counter (identical in every thread)
    uint32_t nslots, slot_floats, data_off;
    bool resident;
};

// consumer side: returns smem pointer of the chunk that starts at blob float offset `off`
__device__ __forceinline__ const float* ring_acquire(Ring& r, uint32_t off) {
    if (r.resident) return r.base + (off - r.data_off);
    uint32_t slot = r.c % r.nslots;
    mbar_wait(&r.full[slot], (r.c / r.nslots) & 1u);
    return r.base + slot * r.slot_floats;
}
__device__ __forceinline__ void ring_release(Ring& r) {
    if (r.resident) return;
    uint32_t slot = r.c % r.nslots;
    __syncwarp();
    if ((threadIdx.x & 31) == 0) mbar_arrive(&r.empty[slot]);
    r.c++;
}
// producer side (one elected thread): stream `floats` starting at blob offset `off`
__device__ __forceinline__ void ring_produce(Ring& r, uint32_t off, uint32_t floats) {
    uint32_t slot = r.c % r.nslots;
    if (r.c >= r.nslots) mbar_wait(&r.empty[slot], ((r.c / r.nslots) - 1u) & 1u);
    mbar_expect_tx(&r.full[slot], floats * 4u);
    bulk_g2s(r.base + slot * r.slot_floats, r.blob + off, floats * 4u, &r.full[slot]);
    r.c++;
}
// producer: all chunks of one part, in the order the consumers acquire them
__device__ __forceinline__ void ring_produce_part(Ring& r, const PartDesc& p, uint32_t kc_floats) {
    if (p.bias_floats) ring_produce(r, p.off, p.bias_floats);
    uint32_t total = static_cast<uint32_t>(p.nrows) * p.width;
    uint32_t kc = (kc_floats / p.width) * p.width;           // whole rows per chunk
    for (uint32_t o = 0; o < total; o += kc) ring_produce(r, p.off + p.bias_floats + o, min(kc, total - o));
}
// producer, resident mode: the whole data region once, on full[0]
__device__ __forceinline__ void ring_produce_resident(Ring& r, uint32_t data_floats) {
    mbar_expect_tx(&r.full[0], data_floats * 4u);
    for (uint32_t o = 0; o < data_floats; o += 8192u)
        bulk_g2s(r.base + o, r.blob + r.data_off + o, min(8192u, data_floats - o) * 4u, &r.full[0]);
}

// ------------------------------------------------------------------ register tile
template <int RT>
struct TrajVec {
    static constexpr int VEC = RT >= 4 ? 4 : RT;
    static constexpr int NH = RT / VEC;
};

template <int RT>
__device__ __forceinline__ void load_traj(float (&a)[RT], const float* p, int hstride) {
    constexpr int VEC = TrajVec<RT>::VEC, NH = TrajVec<RT>::NH;
#pragma unroll
    for (int h = 0; h < NH; ++h) {
        if constexpr (VEC == 4) {
            float4 v = *reinterpret_cast<const float4*>(p + h * hstride);
            a[4 * h] = v.x; a[4 * h + 1] = v.y; a[4 * h + 2] = v.z; a[4 * h + 3] = v.w;
        } else if constexpr (VEC == 2) {
            float2 v = *reinterpret_cast<const float2*>(p + h * hstride);
            a[0] = v.x; a[1] = v.y;
        } else {
            a[0] = p[0];
        }
    }
}
template <int RT>
__device__ __forceinline__ void store_traj(float* p, int hstride, const float (&a)[RT]) {
    constexpr int VEC = TrajVec<RT>::VEC, NH = TrajVec<RT>::NH;
#pragma unroll
    for (int h = 0; h < NH; ++h) {
        if constexpr (VEC == 4) {
            *reinterpret_cast<float4*>(p + h * hstride) = make_float4(a[4 * h], a[4 * h + 1], a[4 * h + 2], a[4 * h + 3]);
        } else if constexpr (VEC == 2) {
            *reinterpret_cast<float2*>(p) = make_float2(a[0], a[1]);
        } else {
            p[0] = a[0];
        }
    }
}
template <int N>
__device__ __forceinline__ void load_cols(float (&b)[N], const float* p) {
    if constexpr (N % 4 == 0) {
#pragma unroll
        for (int c = 0; c < N / 4; ++c) {
            float4 v = *reinterpret_cast<const float4*>(p + 4 * c);
            b[4 * c] = v.x; b[4 * c + 1] = v.y; b[4 * c + 2] = v.z; b[4 * c + 3] = v.w;
        }
    } else if constexpr (N % 2 == 0) {
#pragma unroll
        for (int c = 0; c < N / 2; ++c) {
            float2 v = *reinterpret_cast<const float2*>(p + 2 * c);
            b[2 * c] = v.x; b[2 * c + 1] = v.y;
        }
    } else {
#pragma unroll
        for (int c = 0; c < N; ++c) b[c] = p[c];
    }
}

// gate g of the weight row accumulates into accumulator block MAP(g)
template <int M0, int M1 = 1, int M2 = 2, int M3 = 3>
struct GateMap {
    __device__ static constexpr int blk(int g) { return g == 0 ? M0 : g == 1 ? M1 : g == 2 ? M2 : M3; }
};

// acc[i][MAP(g)*RJ + r] += in[k][pos(i)] * w[k][g*RJ + r]   for k in [0, rows)
template <int RT, int RJ, int NGW, int NGA, class MAP>
__device__ __forceinline__ void fma_rows(float (&acc)[RT][NGA * RJ], const float* __restrict__ in, int ldin, int hstride,
                                         const float* __restrict__ w, int ldw, int rows) {
#pragma unroll 2
    for (int k = 0; k < rows; ++k) {
        float a[RT], b[NGW * RJ];
        load_traj<RT>(a, in + k * ldin, hstride);
        load_cols<NGW * RJ>(b, w + k * ldw);
#pragma unroll
        for (int g = 0; g < NGW; ++g)
#pragma unroll
            for (int i = 0; i < RT; ++i)
#pragma unroll
                for (int r = 0; r < RJ; ++r)
                    acc[i][MAP::blk(g) * RJ + r] = fmaf(a[i], b[g * RJ + r], acc[i][MAP::blk(g) * RJ + r]);
    }
}

// one layer matrix streamed through the ring: every consumer thread walks the chunks (so the
// empty barriers see all warps); only `active` threads do math.
//   w_col = thread's first column inside a row; the bias block (if any) is laid out like a row.
template <int RT, int RJ, int NGW, int NGA, class MAP, int NGB>
__device__ __forceinline__ void consume_part(float (&acc)[RT][NGA * RJ], Ring& ring, const PartDesc& p, uint32_t kc_floats,
                                             const float* in, int ldin, int hstride, int w_col, int b_col, bool active) {
    if (p.bias_floats) {
        const float* bb = ring_acquire(ring, p.off);
        if (active) {
            float b[NGB * RJ];
            load_cols<NGB * RJ>(b, bb + b_col);
#pragma unroll
            for (int i = 0; i < RT; ++i)
#pragma unroll
                for (int c = 0; c < NGB * RJ; ++c) acc[i][c] = b[c];
        }
        ring_release(ring);
    }
    const uint32_t width = p.width;
    const uint32_t kc = kc_floats / width;                    // rows per chunk
    const uint32_t base = p.off + p.bias_floats;
    for (uint32_t r0 = 0; r0 < p.nrows; r0 += kc) {
        uint32_t rows = min(kc, static_cast<uint32_t>(p.nrows) - r0);
        const float* w = ring_acquire(ring, base + r0 * width);
        if (active) fma_rows<RT, RJ, NGW, NGA, MAP>(acc, in + r0 * ldin, ldin, hstride, w + w_col, width, rows);
        ring_release(ring);
    }
}

// small dense layer with smem-resident weights (noise net): tiles looped over the CTA
//   W[k][ldw] followed by bias[ldw];  out[j][TN] = act(in^T W + b)
template <int RT, int RJ>
__device__ __forceinline__ void dense_resident(const float* in, int K, const float* W, int ldw, int outw, int act,
                                               float* out, int TN, int tid, int nthreads) {
    constexpr int VEC = TrajVec<RT>::VEC;
    const int n_tg = TN / RT, n_jg = outw / RJ, hstride = VEC * n_tg;
    const float* bias = W + K * ldw;
    for (int tile = tid; tile < n_tg * n_jg; tile += nthreads) {
        int tg = tile % n_tg, jg = tile / n_tg;
        float acc[RT][RJ];
        float b[RJ];
        load_cols<RJ>(b, bias + jg * RJ);
#pragma unroll
        for (int i = 0; i < RT; ++i)
#pragma unroll
            for (int r = 0; r < RJ; ++r) acc[i][r] = b[r];
        fma_rows<RT, RJ, 1, 1, GateMap<0>>(acc, in + tg * VEC, TN, hstride, W + jg * RJ, ldw, K);
#pragma unroll
        for (int r = 0; r < RJ; ++r) {
            float o[RT];
#pragma unroll
            for (int i = 0; i < RT; ++i) o[i] = apply_act(act, acc[i][r]);
            store_traj<RT>(out + (jg * RJ + r) * TN + tg * VEC, hstride, o);
        }
    }
}

}  // namespace mpcdl
