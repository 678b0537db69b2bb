// C ABI (include/mpcdl.h): weight packers, launch planning, argument checks, error strings.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/mpcdl.h"
#include "launch.h"

using namespace mpcdl;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

const int kMlpHid[] = {16, 32, 64, 128, 256};
bool hid_supported(int hid) { return std::find(std::begin(kMlpHid), std::end(kMlpHid), hid) != std::end(kMlpHid); }

int gates(int kind) { return kind == MPCDL_GRU ? 3 : kind == MPCDL_LSTM ? 4 : 1; }

int check_desc(const mpcdl_model_desc* d) {
    if (!d) return fail(MPCDL_ERR_ARG, "model desc is NULL");
    if (d->kind < MPCDL_MLP || d->kind > MPCDL_LSTM) return fail(MPCDL_ERR_ARG, "unknown model kind %d", d->kind);
    if (!hid_supported(d->hidden))
        return fail(MPCDL_ERR_UNSUPPORTED, "hidden=%d has no compiled kernel (supported: 16,32,64,128,256)", d->hidden);
    if (d->layers < 1 || d->layers > MAX_LAYERS) return fail(MPCDL_ERR_UNSUPPORTED, "layers=%d outside [1,%d]", d->layers, MAX_LAYERS);
    if (d->kind != MPCDL_MLP && (d->window < 1 || d->window > MAX_WINDOW))
        return fail(MPCDL_ERR_UNSUPPORTED, "window=%d outside [1,%d]", d->window, MAX_WINDOW);
    return MPCDL_OK;
}

// blob layout: a pure function of the model description
uint32_t build_parts(const mpcdl_model_desc& d, PartDesc* parts, int& nparts) {
    uint32_t off = 0;
    const int h = d.hidden;
    nparts = 0;
    auto add = [&](int nrows, int width, int bias) {
        PartDesc p;
        p.off = off; p.nrows = (uint16_t)nrows; p.width = (uint16_t)width; p.bias_floats = (uint32_t)bias;
        parts[nparts++] = p;
        off += bias + (uint32_t)nrows * width;
    };
    if (d.kind == MPCDL_MLP) {
        add(8, h, h);
        for (int l = 0; l < d.layers; ++l) add(h, h, h);
        add(h, 8, 8);
    } else {
        const int G = gates(d.kind);
        for (int l = 0; l < d.layers; ++l) {
            add(l == 0 ? 8 : h, G * h, 4 * h);
            add(h, G * h, 0);
        }
        add(h, 8, 8);
    }
    return off;
}

struct Plan {
    int rt, rj, n_tg, TN, nct, grid, smem_bytes, resident, nslots, slot_floats, ntiles;
    int variant, ks;      // variant 0 = generic ring kernel, 2 = weight-resident MLP kernel (rollout_mlp2.cu)
    int ng;               // weight-resident MLP kernel: independent thread groups per CTA (each owns a tile of TN trajectories)
};

// weight-resident MLP kernel: candidate tiles and their measured cost (cycles per dependent 64x64 layer,
// profiles/r01a/mb_layer.log); must match MPCDL_MLP2_CONFIGS in launch.h
struct Mlp2Cand { int hid, rt, rj, ks, tn_lo, tn_hi, tn_step, ng; };
const Mlp2Cand kMlp2[] = {
    {64, 4, 4, 2, 32, 32, 32, 1}, {64, 8, 8, 2, 64, 64, 32, 1}, {64, 8, 8, 1, 128, 256, 32, 1}, {64, 8, 8, 1, 128, 128, 32, 2},
    {32, 4, 4, 2, 32, 64, 32, 1}, {32, 8, 4, 1, 128, 256, 32, 1},
    {16, 4, 4, 2, 32, 64, 32, 1}, {16, 8, 2, 1, 128, 512, 32, 1},
};
constexpr double kMlp2TwoGroupCycles = 10300.0;  // per pass of 2 x 128 trajectories (1 M x 50: 32.8 ms against 35.7 ms for one group of 256,
                                                 // 65 536 x 50: 2.35 against 2.53 ms; profiles/r02/twogroup/ab.log)
double mlp2_cycles(const Mlp2Cand& c, int TN) {
    if (c.ng == 2) return kMlp2TwoGroupCycles;
    if (c.ks == 2 && c.rt == 4) return 2009.0 * TN / 32;     // TN 28 on 147 SMs (7 warps) measured SLOWER than TN 32 on 128 (r01c)
    if (c.ks == 2) return 4029.0 * TN / 64;
    return 6596.0 + (TN - 128) * 39.3;
}

// Attributes of the CURRENT CUDA device, queried once per device ordinal under a mutex (a process may drive several
// devices, and two threads may make their first call at once).  This table and the thread-local error string are the
// only process-wide state of the library.  A failed query is not cached.
struct DevInfo { int sms = 0, cc_major = 0; };
constexpr int kMaxDevices = 64;
DevInfo g_dev[kMaxDevices];
std::mutex g_dev_mu;

const DevInfo* device_info(bool quiet = false) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e == cudaSuccess && (dev < 0 || dev >= kMaxDevices)) e = cudaErrorInvalidDevice;
    if (e == cudaSuccess) {
        std::lock_guard<std::mutex> lk(g_dev_mu);
        DevInfo& d = g_dev[dev];
        if (d.sms) return &d;
        int sms = 0, cc = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e == cudaSuccess) e = cudaDeviceGetAttribute(&cc, cudaDevAttrComputeCapabilityMajor, dev);
        if (e == cudaSuccess && sms > 0) { d.cc_major = cc; d.sms = sms; return &d; }
    }
    if (quiet) (void)cudaGetLastError();
    else fail(MPCDL_ERR_CUDA, "CUDA device query failed: %s", cudaGetErrorString(e));
    return nullptr;
}
// SM count the planner sizes grids (and the BF16 scratch) for: the current device's, or a B200's 148 when there is no
// device (planning and the CPU-side tests work without a GPU)
int planning_sms() {
    const DevInfo* d = device_info(true);
    return d ? d->sms : 148;
}
// current device must be sm_100: status code + the device's attributes
int sm100_device(const DevInfo*& d) {
    d = device_info();
    if (!d) return MPCDL_ERR_CUDA;
    if (d->cc_major != 10) return fail(MPCDL_ERR_DEVICE, "mpcdl kernels are built for sm_100a only (device is sm_%d*)", d->cc_major);
    return MPCDL_OK;
}

void fill_model(ModelParams& m, const mpcdl_rollout_args& a, const Plan& pl, uint32_t total) {
    m.blob = a.wblob;
    m.hid = a.model.hidden; m.layers = a.model.layers;
    m.window = a.model.kind == MPCDL_MLP ? 0 : a.model.window;
    m.stateful = a.model.kind == MPCDL_MLP ? 0 : (a.stateful != 0);
    m.data_off = 0; m.data_floats = total;
    m.n_tg = pl.n_tg; m.TN = pl.TN; m.n_cthreads = pl.nct;
    m.nslots = pl.nslots; m.slot_floats = pl.slot_floats; m.resident = pl.resident;
}

void fill_io(IoParams& io, const mpcdl_rollout_args& a, int ntiles) {
    io.norm = a.norm; io.init = a.init; io.U = a.U; io.states = a.states; io.costs = a.costs;
    io.cost_ref = a.cost_ref;
    for (int i = 0; i < 6; ++i) io.qw[i] = a.cost_qw[i];
    for (int i = 0; i < 2; ++i) io.rw[i] = a.cost_rw[i];
    io.eps = a.eps; io.seed = a.seed; io.traj_offset = a.traj_offset; io.carry = a.carry;
    io.N = a.N; io.H = a.H; io.ntiles = ntiles;
    io.sample = a.sample_controls != 0; io.init_mul = a.shared_init ? 0 : 1; io.ctrl_seed = a.ctrl_seed;
    for (int i = 0; i < 2; ++i) { io.sig[i] = a.ctrl_sigma[i]; io.ulo[i] = a.ctrl_lo[i]; io.uhi[i] = a.ctrl_hi[i]; }
    io.npm = a.N; io.tpm = ntiles > 0 ? ntiles : 1; io.share_inputs = 0;      // single model; set_stacking() overrides
}

int n_models_of(const mpcdl_rollout_args& a) { return a.n_models > 1 ? a.n_models : 1; }
// stacked models: model-major trajectories, tiles never straddle two models
void set_stacking(RolloutParams& p, const mpcdl_rollout_args& a, int TN, uint32_t blob_floats) {
    const int nm = n_models_of(a);
    p.io.npm = a.N / nm;
    p.io.tpm = (int)((p.io.npm + TN - 1) / TN);
    p.io.ntiles = nm * p.io.tpm;
    p.io.share_inputs = nm > 1 && a.share_inputs;
    p.m.blob_stride = nm > 1 ? blob_floats : 0u;
}

void fill_noise(NoiseParams& nz, const mpcdl_rollout_args& a) {
    memset(&nz, 0, sizeof(nz));
    if (!a.noise_wblob) return;
    nz.blob = a.noise_wblob; nz.enabled = 1;
    nz.hist = a.noise_history; nz.size = a.noise_size; nz.nlayers = a.noise_layers; nz.act = a.noise_act;
    nz.in_dim = 8 * a.noise_history;
    nz.blob_floats = (int)mpcdl_noise_blob_floats(a.noise_history, a.noise_size, a.noise_layers);
}


// ---------------------------------------------------------------------------- tensor-core (tcgen05) path
bool tc_supported(const mpcdl_model_desc& d) { return d.kind != MPCDL_MLP && d.hidden == TC_HID && tc_window_ok(d.window); }

// blob: [bias L*64*4][fc 64*8+8][pad to 32 floats][per layer: x chunks, h chunks]
uint32_t build_tc_parts(const mpcdl_model_desc& d, TcPart* parts, uint32_t& bias_off, uint32_t& fc_off) {
    const int G = gates(d.kind);
    uint32_t off = 0;
    bias_off = off; off += (uint32_t)d.layers * TC_HID * 4;
    fc_off = off;   off += TC_HID * 8 + 8;
    off = (off + 31u) / 32u * 32u;
    for (int l = 0; l < d.layers; ++l)
        for (int part = 0; part < 2; ++part) {
            TcPart& p = parts[2 * l + part];
            p.K = (uint16_t)((part == 0 && l == 0) ? 8 : TC_HID);
            p.kc = (uint16_t)std::min<int>(p.K, TC_KS);
            p.nchunks = (uint16_t)(p.K / p.kc);
            p.N = (uint16_t)(G * TC_HID);
            p.chunk_floats = 2u * p.kc * p.N;
            p.off = off;
            off += p.nchunks * p.chunk_floats;
        }
    return off;
}

float tf32_rna_host(float x) {          // cvt.rna.tf32.f32: round to nearest, ties away, 10 explicit mantissa bits
    uint32_t u;
    memcpy(&u, &x, 4);
    u = (u + 0x1000u) & 0xFFFFE000u;
    float r;
    memcpy(&r, &u, 4);
    return r;
}

struct TcPlan { int nslots, slot_floats, resident, group, smem_bytes, ntiles, grid, alias; };

// Ring geometry: a slot holds `group` 16-row chunks of one weight matrix (one TMA + one MMA-issue unit).
// Big groups keep the single control thread off the critical path; residency (a layer's two matrices stay
// in shared memory for all T frames) divides the L2 -> SM weight traffic by T.
int make_tc_plan(const mpcdl_rollout_args& a, TcPlan& pl, int num_sms) {
    const mpcdl_model_desc& d = a.model;
    TcPart parts[2 * MAX_LAYERS];
    uint32_t bo, fo;
    build_tc_parts(d, parts, bo, fo);
    const bool x3 = a.precision == MPCDL_PREC_TF32X3;
    const uint32_t chunk = parts[1].chunk_floats / (x3 ? 1 : 2);        // floats per 16-row chunk in the ring (hi+lo | hi)
    const int per_part = TC_HID / TC_KS;
    NoiseParams nz;
    fill_noise(nz, a);
    int budget = SMEM_LIMIT - tc_layout(0, 0, d.layers, d.window, nz).total * 4;
    const int pref_x3[3] = {2, 4, 1}, pref_1[3] = {4, 2, 1};
    const int* pref = x3 ? pref_x3 : pref_1;
    int G = 0, nslots = 0, resident = 0, alias = 0;
    if (a.tune_kc >= 0)
        for (int i = 0; i < 3 && !G; ++i) {                             // resident, preferred group size first
            const int g = a.tune_ntg > 0 ? a.tune_ntg : pref[i], ns = std::min<int>(budget / (int)(g * chunk * 4), TC_MAX_SLOTS);
            if (ns >= 2 * per_part / g) { G = g; nslots = ns; resident = 1; }
        }
    // With the noise net on and no room for a resident layer beside it, keep the layer resident by lending the net the ring itself: at the end of a step
    // all weight slots are dead, so its buffers alias them (the history is rebuilt from the window buffer, which needs
    // history <= window).  C3 tf32x3: 16-row chunks streamed through 5 slots -> 2 matrices resident, 50 -> 33 ms.
    if (!G && a.tune_kc >= 0 && nz.enabled && nz.hist <= d.window && a.tune_nslots >= 0) {
        NoiseParams none;
        memset(&none, 0, sizeof(none));
        const int budget_alias = SMEM_LIMIT - tc_layout(0, 0, d.layers, d.window, none).total * 4;
        for (int i = 0; i < 3 && !G; ++i) {
            const int g = a.tune_ntg > 0 ? a.tune_ntg : pref[i], ns = std::min<int>(budget_alias / (int)(g * chunk * 4), TC_MAX_SLOTS);
            if (ns >= 2 * per_part / g && (long long)ns * g * chunk >= tc_noise_floats(nz)) { G = g; nslots = ns; resident = 1; alias = 1; }
        }
    }
    for (int g = 4; g >= 1 && !G; g >>= 1) {                            // streaming: largest group with >= 3 slots
        if (a.tune_ntg > 0 && g != a.tune_ntg) continue;
        const int ns = std::min<int>(budget / (int)(g * chunk * 4), TC_MAX_SLOTS);
        if (ns >= 3 || (g == 1 && ns >= 2)) { G = g; nslots = ns; }
    }
    if (!G) return fail(MPCDL_ERR_UNSUPPORTED, "tensor-core rollout: shared memory cannot hold the weight ring (noise net too large?)");
    if (a.tune_nslots > 0 && a.tune_nslots < nslots && !alias) { nslots = a.tune_nslots; if (nslots < 2 * per_part / G) resident = 0; }
    pl.group = G; pl.nslots = nslots; pl.resident = resident; pl.alias = alias;
    pl.slot_floats = (int)(G * chunk);
    pl.smem_bytes = std::max(tc_layout(nslots, pl.slot_floats, d.layers, d.window, nz, alias != 0).total * 4, 120 * 1024);   // > half an SM: one CTA (one TMEM owner) per SM
    pl.ntiles = (int)((a.N + TC_TN - 1) / TC_TN);
    pl.grid = std::min(pl.ntiles, num_sms);
    return MPCDL_OK;
}

// ---------------------------------------------------------------------------- tensor-core MLP (rollout_mlp_tc.cu)
bool mc_supported(const mpcdl_model_desc& d) { return d.kind == MPCDL_MLP && d.hidden == TC_HID; }

uint32_t build_mc_offsets(const mpcdl_model_desc& d, bool x3, McModel& m) {
    uint32_t off = 0, so = 0;
    m.bias_off = off; off += (uint32_t)(d.layers + 1) * TC_HID;
    m.out_off = off;  off += TC_HID * 8 + 8;
    off = (off + 31u) / 32u * 32u;
    for (int l = 0; l <= d.layers; ++l) {
        const uint32_t blk = (l == 0 ? 8u : (uint32_t)TC_HID) * TC_HID;
        m.w_off[l] = off;  off += 2 * blk;             // blob always holds hi and lo
        m.w_smem[l] = so;  so += blk * (x3 ? 2u : 1u); // shared memory holds what the precision needs
    }
    m.w_floats = so;
    m.layers = d.layers;
    return off;
}

struct McPlan { int smem_bytes, ntiles, grid, tiles_per_cta; };
int make_mc_plan(const mpcdl_rollout_args& a, McPlan& pl, int num_sms) {
    McModel m;
    build_mc_offsets(a.model, a.precision == MPCDL_PREC_TF32X3, m);
    NoiseParams nz;
    fill_noise(nz, a);
    pl.ntiles = (int)((a.N + TC_TN - 1) / TC_TN);
    pl.tiles_per_cta = (a.tune_ntg == 1 || a.tune_ntg == 2 || a.tune_ntg == 4) ? a.tune_ntg : (pl.ntiles > num_sms ? MC_TILES : 1);
    // four tile slots per CTA (128 TMEM columns each: D + A_hi): one-pass TF32 without the noise net, and enough tiles to fill them
    const bool four_ok = a.precision == MPCDL_PREC_TF32 && !nz.enabled;
    if (!a.tune_ntg && four_ok && pl.ntiles >= 8 * num_sms) pl.tiles_per_cta = MC_TILES_MAX;
    if (pl.tiles_per_cta == 4 && !four_ok) return fail(MPCDL_ERR_UNSUPPORTED, "tensor-core MLP: four tiles per CTA need MPCDL_PREC_TF32 without the noise net");
    // the noise net's buffers are per tile: when two sets do not fit beside the weights (tf32x3, 4 layers), one tile per CTA
    if (pl.tiles_per_cta == 2 && mc_layout((int)m.w_floats, a.model.layers, nz, 2).total * 4 > SMEM_LIMIT) pl.tiles_per_cta = 1;
    pl.smem_bytes = std::max(mc_layout((int)m.w_floats, a.model.layers, nz, pl.tiles_per_cta).total * 4, 120 * 1024);
    if (pl.smem_bytes > SMEM_LIMIT)
        return fail(MPCDL_ERR_UNSUPPORTED, "tensor-core MLP: %d layers of tf32 hi/lo weights%s (%d KB) do not fit shared memory; use MPCDL_PREC_TF32 or FP32",
                    a.model.layers, nz.enabled ? " + the noise net" : "", pl.smem_bytes / 1024);
    pl.grid = std::min((pl.ntiles + pl.tiles_per_cta - 1) / pl.tiles_per_cta, num_sms);
    return MPCDL_OK;
}

static inline bool resident_ring(const Plan& pl) { return pl.resident != 0; }

int make_plan(const mpcdl_rollout_args& a, Plan& best, int num_sms) {
    const mpcdl_model_desc& d = a.model;
    const bool rnn = d.kind != MPCDL_MLP;
    PartDesc parts[MAX_PARTS];
    int nparts;
    const uint32_t total = build_parts(d, parts, nparts);
    int width_max = 0, bias_max = 0;
    for (int i = 0; i < nparts; ++i) { width_max = std::max<int>(width_max, parts[i].width); bias_max = std::max<int>(bias_max, parts[i].bias_floats); }

    if (!rnn && !a.noise_wblob && !a.tune_nslots && a.tune_kc >= 0 && d.hidden <= 64 && n_models_of(a) == 1) {
        double best2 = 1e300;
        bool found2 = false;
        for (const Mlp2Cand& c : kMlp2) {
            if (c.hid != d.hidden || (a.tune_rt && a.tune_rt != c.rt)) continue;
            for (int TN = c.tn_lo; TN <= c.tn_hi; TN += c.tn_step) {
                if (a.tune_ntg && a.tune_ntg * c.rt != TN) continue;
                // tune_kc: 1 | 2 = K split of the one-group kernel, 4 = the two-group form (K split 1)
                if (a.tune_kc && a.tune_kc != (c.ng == 2 ? 4 : c.ks)) continue;
                Plan pl;
                memset(&pl, 0, sizeof(pl));
                pl.variant = 2; pl.ks = c.ks; pl.ng = c.ng; pl.rt = c.rt; pl.rj = c.rj; pl.TN = TN; pl.n_tg = TN / c.rt;
                const int per_group = pl.n_tg * (d.hidden / c.rj) * c.ks;
                pl.nct = per_group * c.ng;
                if (per_group % 32 || pl.nct > ((d.hidden == 16 && c.rt == 8) ? 512 : 256) || per_group < TN) continue;
                pl.resident = 1; pl.nslots = 1; pl.slot_floats = 0;
                pl.smem_bytes = mlp2_smem_bytes((int)total, d.hidden, TN, c.ks, c.rt, c.rj, c.ng);
                if (pl.smem_bytes > SMEM_LIMIT) continue;
                pl.ntiles = (int)((a.N + TN - 1) / TN);
                pl.grid = std::min(pl.ntiles, num_sms);                 // tiles go out group-major: a partial pass spreads over all SMs
                const double cost = (double)((pl.ntiles + num_sms * c.ng - 1) / (num_sms * c.ng)) * mlp2_cycles(c, TN);
                if (cost < best2) { best2 = cost; best = pl; found2 = true; }
            }
        }
        if (found2) return MPCDL_OK;
    }
    const int rts_mlp[] = {2, 4, 8}, rts_rnn[] = {2, 4, 8};
    const int rj = rnn ? rnn_rj(d.hidden) : mlp_rj(d.hidden);
    const int G = gates(d.kind), NGA = rnn ? 4 : 1;
    double best_cost = 1e300;
    bool found = false;
    for (int rt : (rnn ? rts_rnn : rts_mlp)) {
        // must match the instantiation lists in launch.h
        if (!rnn && rt == 4 && d.hidden != 64) continue;
        if (rnn && rt == 8 && d.hidden != 64) continue;
        if (a.tune_rt && rt != a.tune_rt) continue;
        const int n_jg = d.hidden / rj;
        if (rt > n_jg) continue;                                   // owner threads: TN <= consumer threads
        const int maxc = rnn ? rnn_max_consumers(rt, rj) : mlp_max_consumers();
        const int ntg_max = maxc / n_jg;
        for (int n_tg = ntg_max; n_tg >= 1; --n_tg)
        for (int nslots_try = 3; nslots_try >= 2; --nslots_try)          // a 2-slot ring where the third slot would cost tile rows
        for (int slot_kb = 32; slot_kb >= 16; slot_kb -= 8) {            // 32, 24 or 16 KB of weight rows per ring slot
            if (a.tune_ntg && n_tg != a.tune_ntg) continue;
            Plan pl;
            pl.variant = 0; pl.ks = 1; pl.ng = 1;
            pl.rt = rt; pl.rj = rj; pl.n_tg = n_tg; pl.TN = n_tg * rt;
            pl.nct = (n_tg * n_jg + 31) / 32 * 32;
            pl.resident = (!a.tune_nslots && total <= 24576u && n_models_of(a) == 1) ? 1 : 0;   // stacked models stream their own blob per tile
            if ((pl.resident || a.tune_nslots) && nslots_try != 3) continue;
            if ((pl.resident || a.tune_kc > 0 || !rnn) && slot_kb != 16) continue;
            // rows of K per ring slot, the hidden-deep matrices split into EQUAL chunks: 64 rows as 21 + 21 + 21 + 1 cost a whole
            // ring hand-off (~700 clocks: wait, release, refill latency) for one row (profiles/r02/c3_sweep.log)
            int kc_rows = std::max(1, slot_kb * 256 / width_max);
            if (d.hidden > kc_rows) {
                const int nch = (d.hidden + kc_rows - 1) / kc_rows;
                kc_rows = (d.hidden + nch - 1) / nch;
            }
            if (a.tune_kc > 0) kc_rows = a.tune_kc;
            pl.slot_floats = std::max(kc_rows * width_max, bias_max);
            pl.nslots = pl.resident ? 1 : (a.tune_nslots ? a.tune_nslots : nslots_try);
            ModelParams m;
            memset(&m, 0, sizeof(m));
            fill_model(m, a, pl, total);
            NoiseParams nz;
            fill_noise(nz, a);
            SmemLayout L = make_layout(m, nz, rnn);
            pl.smem_bytes = L.total * 4;
            if (pl.smem_bytes > SMEM_LIMIT) continue;
            pl.ntiles = n_models_of(a) * (int)((a.N / n_models_of(a) + pl.TN - 1) / pl.TN);
            pl.grid = std::min(pl.ntiles, num_sms);
            const int rounds = (pl.ntiles + num_sms - 1) / num_sms;
            double cost;
            if (rnn) {
                // time of one tile-step in microseconds, fitted to the C3 sweep (GRU 64 x 4, window 4, + noise net, 65 536 x 20 on a
                // B200): ring hand-offs 0.35 us each, 2.1 us per cell of barriers / activations / head share, and the FMA work at
                // 0.36 TFLOP/s per SM for 4-row tiles (8-row tiles: 7 consumer warps, 7 % slower per trajectory; 2-row: 1.6 x)
                const int T = d.window, Lc = d.layers, h = d.hidden;
                auto chunks = [&](int rows) { return (rows + kc_rows - 1) / kc_rows; };
                const int recur_frames = a.stateful ? T : T - 1;
                const double handoffs = Lc * T * 1.0 + T * chunks(8) + (Lc - 1) * T * chunks(h) + Lc * recur_frames * chunks(h) + 1 + chunks(h);
                const double flop = 2.0 * (T * G * h * 8.0 + (Lc - 1) * T * G * h * (double)h + Lc * recur_frames * G * h * (double)h + 6.0 * h);
                // 8-row tiles run 1..5 % faster per trajectory than 4-row tiles in the same configuration (LSTM 30.4 / 31.9 ms,
                // GRU 23.2 / 23.8 ms); an odd number of trajectory groups leaves half a warp idle and misaligns the warps with the
                // unit groups (stateful C3: 76-trajectory tiles 35.4 ms, 64-trajectory tiles 31.0 ms)
                // ... but only with all 7 consumer warps the 255-register tile allows: 6 warps 8 %, 4 warps 18 % slower (C3 sweep)
                const int cw = pl.nct / 32;
                const double f8 = cw >= 7 ? 0.99 : cw == 6 ? 1.08 : 1.18;       // 0.965 at 65 536, but 1 M x 100 prefers 120-trajectory 4-row tiles (1812 / 1860 ms)
                const double rate = 0.363e6 / (rt == 4 ? 1.0 : rt == 8 ? f8 : 1.6) / ((n_tg & 1) && n_tg > 1 ? 1.08 : 1.0);          // FLOP per microsecond per SM
                const double per_step = (resident_ring(pl) ? 0.0 : handoffs * 0.35) + Lc * T * 2.1 + pl.TN * flop / rate;
                // two slots expose the refill latency, more so with short chunks (sweep: 112 x 16 KB x 2 slots 26.0 ms, 96 x 24 KB x 2 24.9 ms)
                cost = rounds * per_step * (nslots_try == 3 ? 1.0 : slot_kb >= 24 ? 1.02 : 1.08);
            } else {
                const double instr = NGA * rt * rj + (rt + 3) / 4 + (G * rj + 3) / 4;
                cost = rounds * instr * std::max(pl.nct / 128.0, 2.0) * (1.0 + 0.01 * (8 - rt));
            }
            if (cost < best_cost) { best_cost = cost; best = pl; found = true; }
        }
    }
    if (!found) return fail(MPCDL_ERR_UNSUPPORTED, "no launch configuration fits (hidden=%d window=%d)", d.hidden, d.window);
    return MPCDL_OK;
}

// ---------------------------------------------------------------------------- BF16 tensor-core LSTM, hidden 256 (rollout_lstm_tc.cu)
bool lt_supported(const mpcdl_model_desc& d) {
    return d.kind == MPCDL_LSTM && d.hidden == LT_HID && d.window >= 1 && lt_layout(d.window).total <= SMEM_LIMIT;
}
// blob: [bias L x 1024][fc 256 x 8 + 8][pad to 32 floats][bf16 weight slots, lstm_tc_params.h]; returns total floats
uint32_t build_lt_offsets(const mpcdl_model_desc& d, LtModel& m) {
    uint32_t off = 0;
    m.bias_off = off; off += (uint32_t)d.layers * 4 * LT_HID;
    m.fc_off = off;   off += LT_HID * 8 + 8;
    off = (off + 31u) / 32u * 32u;
    m.w_off = off;
    off += (lt_layer_off_bytes(d.layers - 1) + lt_layer_bytes(d.layers - 1)) / 4;
    m.layers = d.layers; m.window = d.window;
    return off;
}
uint16_t bf16_rne_host(float x) {         // cvt.rn.bf16.f32 (finite inputs)
    uint32_t u;
    memcpy(&u, &x, 4);
    u += 0x7FFFu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}
struct LtPlan { int smem_bytes, ntiles, grid; };
void make_lt_plan(const mpcdl_rollout_args& a, LtPlan& pl, int num_sms) {
    pl.smem_bytes = lt_layout(a.model.window).total;
    pl.ntiles = (int)((a.N + LT_TN - 1) / LT_TN);
    pl.grid = 2 * std::min((pl.ntiles + 1) / 2, num_sms / 2);          // CTA pairs (cluster of 2): two tiles per pair
}
bool is_lt(const mpcdl_rollout_args& a) { return a.precision == MPCDL_PREC_BF16; }

bool tc_supported(const mpcdl_model_desc& d);
bool mc_supported(const mpcdl_model_desc& d);
int check_args(const mpcdl_rollout_args* a) {
    if (!a) return fail(MPCDL_ERR_ARG, "args is NULL");
    int rc = check_desc(&a->model);
    if (rc) return rc;
    if (a->N < 1 || a->H < 1) return fail(MPCDL_ERR_ARG, "N=%lld H=%d must be >= 1", (long long)a->N, a->H);
    if ((!a->wblob && a->precision == MPCDL_PREC_FP32) || !a->norm || !a->init || !a->U)
        return fail(MPCDL_ERR_ARG, "wblob/norm/init/U must be non-NULL device pointers");
    if (!a->states && !a->costs) return fail(MPCDL_ERR_ARG, "both outputs (states, costs) are NULL");
    if (a->model.kind != MPCDL_MLP && a->stateful && !a->carry) return fail(MPCDL_ERR_ARG, "stateful rollout needs a carry buffer");
    if (a->precision == MPCDL_PREC_BF16) {
        if (!lt_supported(a->model))
            return fail(MPCDL_ERR_UNSUPPORTED, "MPCDL_PREC_BF16 is the LSTM hidden=%d kernel; got kind=%d hidden=%d window=%d",
                        LT_HID, a->model.kind, a->model.hidden, a->model.window);
        if (a->noise_wblob) {
            NoiseParams nz;
            fill_noise(nz, *a);
            if (lt_noise_layout(nz).total * 4 > LT_A_BYTES)
                return fail(MPCDL_ERR_UNSUPPORTED, "MPCDL_PREC_BF16: the residual-noise net's buffers (%d B) exceed the %d B the kernel can lend them; use MPCDL_PREC_FP32",
                            lt_noise_layout(nz).total * 4, LT_A_BYTES);
        }
        if (!a->wblob_tc) return fail(MPCDL_ERR_ARG, "MPCDL_PREC_BF16 needs wblob_tc (mpcdl_pack_rnn_tc)");
        const size_t need = mpcdl_rollout_scratch_bytes(a);
        if (!a->scratch || a->scratch_bytes < need)
            return fail(MPCDL_ERR_ARG, "MPCDL_PREC_BF16 needs a device scratch of %zu bytes (mpcdl_rollout_scratch_bytes), got %llu", need, (unsigned long long)a->scratch_bytes);
    } else if (a->precision != MPCDL_PREC_FP32) {
        if (a->precision != MPCDL_PREC_TF32X3 && a->precision != MPCDL_PREC_TF32) return fail(MPCDL_ERR_ARG, "unknown precision %d", a->precision);
        if (!tc_supported(a->model) && !mc_supported(a->model))
            return fail(MPCDL_ERR_UNSUPPORTED, "tensor-core precision needs hidden=%d (GRU/LSTM: window=%d or %d); got kind=%d hidden=%d window=%d; use MPCDL_PREC_FP32",
                        TC_HID, TC_T, TC_T_LONG, a->model.kind, a->model.hidden, a->model.window);
        if (!a->wblob_tc) return fail(MPCDL_ERR_ARG, "tensor-core precision needs wblob_tc (mpcdl_pack_rnn_tc)");
        const size_t need = mpcdl_rollout_scratch_bytes(a);
        if (need && (!a->scratch || a->scratch_bytes < need))
            return fail(MPCDL_ERR_ARG, "tensor-core rollout with window=%d needs a device scratch of %zu bytes (mpcdl_rollout_scratch_bytes), got %llu",
                        a->model.window, need, (unsigned long long)a->scratch_bytes);
    }
    if (a->n_models > 1) {
        if (a->precision != MPCDL_PREC_FP32 || a->stateful || a->sample_controls || a->shared_init)
            return fail(MPCDL_ERR_UNSUPPORTED, "stacked models (n_models=%d): FP32 precision, stateless, explicit controls only", a->n_models);
        if (a->N % a->n_models) return fail(MPCDL_ERR_ARG, "N=%lld is not a multiple of n_models=%d", (long long)a->N, a->n_models);
    }
    if (a->sample_controls && !(a->ctrl_lo[0] <= a->ctrl_hi[0] && a->ctrl_lo[1] <= a->ctrl_hi[1] && a->ctrl_sigma[0] >= 0.f && a->ctrl_sigma[1] >= 0.f))
        return fail(MPCDL_ERR_ARG, "sample_controls needs ctrl_sigma >= 0 and ctrl_lo <= ctrl_hi");
    if (a->noise_wblob) {
        if (a->noise_history < 1 || a->noise_history > 8) return fail(MPCDL_ERR_UNSUPPORTED, "noise history=%d outside [1,8]", a->noise_history);
        if (a->noise_size < 4 || a->noise_size > 64 || a->noise_size % 4) return fail(MPCDL_ERR_UNSUPPORTED, "noise size=%d must be a multiple of 4 in [4,64]", a->noise_size);
        if (a->noise_layers < 0 || a->noise_layers > MAX_NOISE_LAYERS) return fail(MPCDL_ERR_UNSUPPORTED, "noise n_layers=%d outside [0,%d]", a->noise_layers, MAX_NOISE_LAYERS);
        if (a->noise_act < MPCDL_ACT_IDENTITY || a->noise_act > MPCDL_ACT_SIGMOID) return fail(MPCDL_ERR_UNSUPPORTED, "noise activation %d unsupported", a->noise_act);
    }
    return MPCDL_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------ C ABI
extern "C" {

const char* mpcdl_last_error(void) { return g_err; }
int mpcdl_version(void) { return MPCDL_VERSION; }

size_t mpcdl_blob_floats(const mpcdl_model_desc* desc) {
    if (check_desc(desc)) return 0;
    PartDesc parts[MAX_PARTS];
    int n;
    return build_parts(*desc, parts, n);
}

int mpcdl_pack_mlp(const mpcdl_model_desc* desc, const float* const* W, const float* const* b, float* blob) {
    int rc = check_desc(desc);
    if (rc) return rc;
    if (desc->kind != MPCDL_MLP || !W || !b || !blob) return fail(MPCDL_ERR_ARG, "mpcdl_pack_mlp: bad arguments");
    PartDesc parts[MAX_PARTS];
    int n;
    uint32_t total = build_parts(*desc, parts, n);
    std::fill(blob, blob + total, 0.f);
    const int h = desc->hidden;
    for (int p = 0; p < n; ++p) {
        const int in = p == 0 ? 8 : h, out = p == n - 1 ? 6 : h, width = parts[p].width;
        float* bias = blob + parts[p].off;
        float* rows = bias + parts[p].bias_floats;
        for (int j = 0; j < out; ++j) {
            bias[j] = b[p][j];
            for (int k = 0; k < in; ++k) rows[k * width + j] = W[p][j * in + k];
        }
    }
    return MPCDL_OK;
}

int mpcdl_pack_rnn(const mpcdl_model_desc* desc, const float* const* w_ih, const float* const* w_hh,
                   const float* const* b_ih, const float* const* b_hh, const float* fc_w, const float* fc_b, float* blob) {
    int rc = check_desc(desc);
    if (rc) return rc;
    if (desc->kind == MPCDL_MLP || !w_ih || !w_hh || !b_ih || !b_hh || !fc_w || !fc_b || !blob)
        return fail(MPCDL_ERR_ARG, "mpcdl_pack_rnn: bad arguments");
    PartDesc parts[MAX_PARTS];
    int n;
    uint32_t total = build_parts(*desc, parts, n);
    std::fill(blob, blob + total, 0.f);
    const int h = desc->hidden, G = gates(desc->kind), RJ = rnn_rj(h), n_jg = h / RJ;
    for (int l = 0; l < desc->layers; ++l) {
        const int in = l == 0 ? 8 : h;
        const PartDesc& px = parts[2 * l];
        const PartDesc& ph = parts[2 * l + 1];
        float* bias = blob + px.off;                         // [jg][4][RJ]
        float* xr = bias + px.bias_floats;                   // [k][jg][G][RJ]
        float* hr = blob + ph.off;
        for (int jg = 0; jg < n_jg; ++jg)
            for (int r = 0; r < RJ; ++r) {
                const int j = jg * RJ + r;
                float* bb = bias + jg * 4 * RJ + r;
                if (desc->kind == MPCDL_GRU) {
                    bb[0 * RJ] = b_ih[l][j] + b_hh[l][j];
                    bb[1 * RJ] = b_ih[l][h + j] + b_hh[l][h + j];
                    bb[2 * RJ] = b_ih[l][2 * h + j];
                    bb[3 * RJ] = b_hh[l][2 * h + j];
                } else {
                    for (int g = 0; g < 4; ++g) bb[g * RJ] = b_ih[l][g * h + j] + b_hh[l][g * h + j];
                }
                for (int g = 0; g < G; ++g) {
                    for (int k = 0; k < in; ++k) xr[(size_t)k * G * h + (jg * G + g) * RJ + r] = w_ih[l][(size_t)(g * h + j) * in + k];
                    for (int k = 0; k < h; ++k) hr[(size_t)k * G * h + (jg * G + g) * RJ + r] = w_hh[l][(size_t)(g * h + j) * h + k];
                }
            }
    }
    const PartDesc& pf = parts[2 * desc->layers];
    float* bias = blob + pf.off;
    float* rows = bias + pf.bias_floats;
    for (int j = 0; j < 6; ++j) {
        bias[j] = fc_b[j];
        for (int k = 0; k < h; ++k) rows[k * 8 + j] = fc_w[j * h + k];
    }
    return MPCDL_OK;
}


namespace {
// BF16 blob of the hidden-256 LSTM kernel (layout: lstm_tc_params.h)
int pack_lt(const mpcdl_model_desc& d, const float* const* w_ih, const float* const* w_hh, const float* const* b_ih,
            const float* const* b_hh, const float* fc_w, const float* fc_b, float* blob) {
    LtModel m;
    const uint32_t total = build_lt_offsets(d, m);
    std::fill(blob, blob + total, 0.f);
    const int h = LT_HID;
    for (int l = 0; l < d.layers; ++l)
        for (int c = 0; c < LT_NCH; ++c)
            for (int g = 0; g < 4; ++g)
                for (int j = 0; j < LT_CH; ++j) {
                    const int r = g * h + c * LT_CH + j;
                    blob[m.bias_off + (size_t)l * 4 * h + c * 256 + g * LT_CH + j] = b_ih[l][r] + b_hh[l][r];
                }
    float* fc = blob + m.fc_off;                                       // [j][8] then bias[8]
    for (int dd = 0; dd < 6; ++dd) {
        fc[h * 8 + dd] = fc_b[dd];
        for (int j = 0; j < h; ++j) fc[j * 8 + dd] = fc_w[dd * h + j];
    }
    unsigned char* wb = reinterpret_cast<unsigned char*>(blob + m.w_off);
    // one slot: [CTA rank 2][k/8][n/8][8 rows][8 k] bf16; rank r holds rows n = 128 r .. 128 r + 127 of the chunk's 256
    // (row n -> torch row (n/64)*256 + 64c + n%64), k = k0 .. k0 + nk
    auto fill_slot = [&](unsigned char* slot, const float* W, int K, int c, int k0, int nk, int kmod) {
        uint16_t* o = reinterpret_cast<uint16_t*>(slot);
        for (int kk = 0; kk < nk; ++kk)
            for (int n = 0; n < 256; ++n) {
                const int row = (n / LT_CH) * h + c * LT_CH + n % LT_CH;
                const int k = (k0 + kk) % kmod;
                const int r = n / 128, nn = n % 128;
                o[(size_t)r * 128 * nk + ((size_t)(kk / 8) * 16 + nn / 8) * 64 + (nn % 8) * 8 + kk % 8] = bf16_rne_host(W[(size_t)row * K + k]);
            }
    };
    for (int l = 0; l < d.layers; ++l)
        for (int c = 0; c < LT_NCH; ++c) {
            unsigned char* p = wb + lt_layer_off_bytes(l) + (size_t)c * lt_chunk_bytes(l);
            if (l == 0) { fill_slot(p, w_ih[0], 8, c, 0, 16, 8); p += 2 * lt_x0_slot_bytes(); }  // [W | W] against [x_hi | x_lo]
            else for (int sl = 0; sl < LT_SPC; ++sl, p += 2 * LT_SLOT_BYTES) fill_slot(p, w_ih[l], h, c, sl * LT_KS, LT_KS, h);
            for (int sl = 0; sl < LT_SPC; ++sl, p += 2 * LT_SLOT_BYTES) fill_slot(p, w_hh[l], h, c, sl * LT_KS, LT_KS, h);
        }
    return MPCDL_OK;
}
}  // namespace

size_t mpcdl_blob_floats_tc(const mpcdl_model_desc* desc) {
    if (check_desc(desc)) return 0;
    if (mc_supported(*desc)) { McModel m; return build_mc_offsets(*desc, true, m); }
    if (lt_supported(*desc)) { LtModel m; return build_lt_offsets(*desc, m); }
    if (!tc_supported(*desc)) { fail(MPCDL_ERR_UNSUPPORTED, "tensor-core blob: hidden=%d only (GRU/LSTM: window=%d or %d)", TC_HID, TC_T, TC_T_LONG); return 0; }
    TcPart parts[2 * MAX_LAYERS];
    uint32_t bo, fo;
    return build_tc_parts(*desc, parts, bo, fo);
}

int mpcdl_pack_rnn_tc(const mpcdl_model_desc* desc, const float* const* w_ih, const float* const* w_hh,
                      const float* const* b_ih, const float* const* b_hh, const float* fc_w, const float* fc_b, float* blob) {
    int rc = check_desc(desc);
    if (rc) return rc;
    if (!w_ih || !w_hh || !b_ih || !b_hh || !fc_w || !fc_b || !blob) return fail(MPCDL_ERR_ARG, "mpcdl_pack_rnn_tc: bad arguments");
    if (lt_supported(*desc)) return pack_lt(*desc, w_ih, w_hh, b_ih, b_hh, fc_w, fc_b, blob);
    if (!tc_supported(*desc)) return fail(MPCDL_ERR_UNSUPPORTED, "tensor-core blob: GRU/LSTM with hidden=%d window=%d|%d, or LSTM hidden=%d, only", TC_HID, TC_T, TC_T_LONG, LT_HID);
    TcPart parts[2 * MAX_LAYERS];
    uint32_t bias_off, fc_off;
    const uint32_t total = build_tc_parts(*desc, parts, bias_off, fc_off);
    std::fill(blob, blob + total, 0.f);
    const int h = TC_HID;
    for (int l = 0; l < desc->layers; ++l) {
        float* bb = blob + bias_off + (size_t)l * h * 4;
        for (int j = 0; j < h; ++j) {
            if (desc->kind == MPCDL_GRU) {       // b_r, b_z (ih + hh summed), b_in, b_hn (kept apart: r gates h_n + b_hn)
                bb[j * 4 + 0] = b_ih[l][j] + b_hh[l][j];
                bb[j * 4 + 1] = b_ih[l][h + j] + b_hh[l][h + j];
                bb[j * 4 + 2] = b_ih[l][2 * h + j];
                bb[j * 4 + 3] = b_hh[l][2 * h + j];
            } else {
                for (int g = 0; g < 4; ++g) bb[j * 4 + g] = b_ih[l][g * h + j] + b_hh[l][g * h + j];
            }
        }
        for (int part = 0; part < 2; ++part) {
            const TcPart& p = parts[2 * l + part];
            const float* W = part == 0 ? w_ih[l] : w_hh[l];          // torch layout [N = G*h rows][K]
            const int K = p.K, N = p.N, kc = p.kc;
            for (int n = 0; n < N; ++n)
                for (int k = 0; k < K; ++k) {
                    // GRU weight_ih rows are packed [n; r; z] (D columns [nx | r | z | nh], rollout_rnn_tc.cu issue_chunk)
                    const int src = (desc->kind == MPCDL_GRU && part == 0) ? ((n / h + 2) % 3) * h + n % h : n;
                    const float w = W[(size_t)src * K + k];
                    const float hi = tf32_rna_host(w), lo = tf32_rna_host(w - hi);
                    const int c = k / kc, kk = k % kc;
                    const size_t e = (size_t)(kk / 4) * (N / 8) * 32 + (size_t)(n / 8) * 32 + (n % 8) * 4 + (kk % 4);
                    float* chunk = blob + p.off + (size_t)c * p.chunk_floats;
                    chunk[e] = hi;
                    chunk[(size_t)kc * N + e] = lo;
                }
        }
    }
    float* fc = blob + fc_off;                                         // [j][8] then bias[8]
    for (int d = 0; d < 6; ++d) {
        fc[h * 8 + d] = fc_b[d];
        for (int j = 0; j < h; ++j) fc[j * 8 + d] = fc_w[d * h + j];
    }
    return MPCDL_OK;
}

int mpcdl_pack_mlp_tc(const mpcdl_model_desc* desc, const float* const* W, const float* const* b, float* blob) {
    int rc = check_desc(desc);
    if (rc) return rc;
    if (!mc_supported(*desc)) return fail(MPCDL_ERR_UNSUPPORTED, "tensor-core MLP blob: hidden=%d only", TC_HID);
    if (!W || !b || !blob) return fail(MPCDL_ERR_ARG, "mpcdl_pack_mlp_tc: bad arguments");
    McModel m;
    const uint32_t total = build_mc_offsets(*desc, true, m);
    std::fill(blob, blob + total, 0.f);
    const int h = TC_HID, L = desc->layers;
    for (int l = 0; l <= L; ++l) {
        const int K = l == 0 ? 8 : h;
        for (int j = 0; j < h; ++j) blob[m.bias_off + l * h + j] = b[l][j];
        float* hi = blob + m.w_off[l];
        float* lo = hi + (size_t)K * h;
        for (int n = 0; n < h; ++n)
            for (int k = 0; k < K; ++k) {
                const float w = W[l][(size_t)n * K + k];
                const float whi = tf32_rna_host(w);
                const size_t e = (size_t)(k / 4) * (h / 8) * 32 + (size_t)(n / 8) * 32 + (n % 8) * 4 + (k % 4);
                hi[e] = whi;
                lo[e] = tf32_rna_host(w - whi);
            }
    }
    float* out = blob + m.out_off;                                      // W^T [j][8] then bias[8], FP32 (CUDA-core head)
    for (int d = 0; d < 6; ++d) {
        out[h * 8 + d] = b[L + 1][d];
        for (int j = 0; j < h; ++j) out[j * 8 + d] = W[L + 1][d * h + j];
    }
    return MPCDL_OK;
}

size_t mpcdl_noise_blob_floats(int history, int size, int n_layers) {
    size_t total = 0;
    int K = 8 * history;
    for (int i = 0; i <= n_layers; ++i) {
        int out = i == n_layers ? 12 : size;
        total += (size_t)(K + 1) * out;
        K = size;
    }
    return total;
}

int mpcdl_pack_noise(int history, int size, int n_layers, const float* const* W, const float* const* b, float* blob) {
    if (!W || !b || !blob || history < 1 || size < 1 || n_layers < 0) return fail(MPCDL_ERR_ARG, "mpcdl_pack_noise: bad arguments");
    int K = 8 * history;
    float* o = blob;
    for (int i = 0; i <= n_layers; ++i) {
        int out = i == n_layers ? 12 : size;
        for (int k = 0; k < K; ++k)
            for (int j = 0; j < out; ++j) o[k * out + j] = W[i][j * K + k];
        for (int j = 0; j < out; ++j) o[K * out + j] = b[i][j];
        o += (K + 1) * out;
        K = size;
    }
    return MPCDL_OK;
}

size_t mpcdl_rollout_scratch_bytes(const mpcdl_rollout_args* a) {
    if (a && (a->precision == MPCDL_PREC_TF32X3 || a->precision == MPCDL_PREC_TF32) && tc_supported(a->model) && a->N >= 1) {
        // window 8 on the tcgen05 GRU / LSTM kernel: per-CTA layer output sequences (rollout_rnn_tc_epi.inl)
        const long long ntiles = (a->N + TC_TN - 1) / TC_TN;
        return (size_t)std::min<long long>(ntiles, planning_sms()) * tc_seq_bytes_per_cta(a->model.window);
    }
    if (!a || a->precision != MPCDL_PREC_BF16 || !lt_supported(a->model) || a->N < 1) return 0;
    LtPlan lp;
    make_lt_plan(*a, lp, planning_sms());
    return (size_t)lp.grid * lt_scratch_bytes_per_cta(a->model.window, a->noise_wblob ? 8 * a->noise_history : 0);
}

int mpcdl_rollout_plan(const mpcdl_rollout_args* a, mpcdl_plan* out) {
    int rc = check_args(a);
    if (rc) return rc;
    if (!out) return fail(MPCDL_ERR_ARG, "plan is NULL");
    const int sms = planning_sms();
    if (is_lt(*a)) {
        LtPlan lp;
        make_lt_plan(*a, lp, sms);
        memset(out, 0, sizeof(*out));
        out->rt = 1; out->rj = 16; out->n_tg = LT_TN; out->tile_traj = LT_TN; out->threads = LT_THREADS; out->grid = lp.grid;
        out->smem_bytes = lp.smem_bytes; out->resident = 0; out->nslots = LT_NSLOTS; out->slot_floats = LT_SLOT_BYTES / 4;
        out->ntiles = lp.ntiles; out->variant = 6; out->ksplit = 1;
        return MPCDL_OK;
    }
    if (a->precision != MPCDL_PREC_FP32 && a->model.kind == MPCDL_MLP) {
        McPlan mp;
        rc = make_mc_plan(*a, mp, sms);
        if (rc) return rc;
        memset(out, 0, sizeof(*out));
        out->rt = 1; out->rj = 64; out->n_tg = TC_TN; out->tile_traj = TC_TN; out->threads = mc_threads(mp.tiles_per_cta); out->grid = mp.grid;
        out->smem_bytes = mp.smem_bytes; out->resident = 1; out->nslots = 1; out->ntiles = mp.ntiles; out->variant = 5;
        out->ksplit = mp.tiles_per_cta;
        return MPCDL_OK;
    }
    if (a->precision != MPCDL_PREC_FP32) {
        TcPlan tp;
        rc = make_tc_plan(*a, tp, sms);
        if (rc) return rc;
        memset(out, 0, sizeof(*out));
        out->rt = 1; out->rj = TC_UPT; out->n_tg = TC_TN; out->tile_traj = TC_TN; out->threads = TC_THREADS; out->grid = tp.grid;
        out->smem_bytes = tp.smem_bytes; out->resident = tp.resident; out->nslots = tp.nslots; out->slot_floats = tp.slot_floats;
        out->ntiles = tp.ntiles; out->variant = 4; out->ksplit = tp.group;
        return MPCDL_OK;
    }
    Plan pl;
    rc = make_plan(*a, pl, sms);
    if (rc) return rc;
    out->rt = pl.rt; out->rj = pl.rj; out->n_tg = pl.n_tg; out->tile_traj = pl.TN; out->threads = pl.nct + (pl.variant == 2 ? 0 : 32);
    out->grid = pl.grid; out->smem_bytes = pl.smem_bytes; out->resident = pl.resident; out->nslots = pl.nslots;
    out->slot_floats = pl.slot_floats; out->ntiles = pl.ntiles; out->variant = pl.variant; out->ksplit = pl.ks;
    return MPCDL_OK;
}

int mpcdl_rollout_static_scratch(const mpcdl_rollout_args* a, void* stream) {
    if (!a) return fail(MPCDL_ERR_ARG, "args is NULL");
    if ((a->precision != MPCDL_PREC_TF32X3 && a->precision != MPCDL_PREC_TF32) || !tc_supported(a->model) || a->model.window != TC_T_LONG)
        return fail(MPCDL_ERR_UNSUPPORTED, "mpcdl_rollout_static_scratch: the GRU / LSTM tensor-core rollout with hidden=%d window=%d (TF32X3 or TF32) only; "
                    "got kind=%d hidden=%d window=%d precision=%d", TC_HID, TC_T_LONG, a->model.kind, a->model.hidden, a->model.window, a->precision);
    if (a->N > TC_STATIC_MAX_TRAJ)
        return fail(MPCDL_ERR_UNSUPPORTED, "mpcdl_rollout_static_scratch: N=%lld exceeds the static scratch (%lld trajectories = %d tiles of %d); "
                    "use mpcdl_rollout with a caller scratch", (long long)a->N, TC_STATIC_MAX_TRAJ, TC_STATIC_CTAS, TC_TN);
    mpcdl_rollout_args b = *a;                           // the caller's scratch fields are ignored: stand-ins that pass check_args
    b.scratch = &b;
    b.scratch_bytes = (uint64_t)TC_STATIC_CTAS * tc_seq_bytes_per_cta(TC_T_LONG);
    int rc = check_args(&b);
    if (rc) return rc;
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    TcPlan tp;
    rc = make_tc_plan(b, tp, di->sms);
    if (rc) return rc;
    if (tp.grid > TC_STATIC_CTAS) return fail(MPCDL_ERR_UNSUPPORTED, "mpcdl_rollout_static_scratch: grid %d exceeds %d CTAs", tp.grid, TC_STATIC_CTAS);
    TcRolloutParams q;
    memset(&q, 0, sizeof(q));
    build_tc_parts(b.model, q.m.parts, q.m.bias_off, q.m.fc_off);
    q.m.blob = b.wblob_tc;
    q.m.cell = b.model.kind == MPCDL_GRU ? 0 : 1;
    q.m.prec = b.precision; q.m.layers = b.model.layers; q.m.window = b.model.window; q.m.stateful = b.stateful != 0;
    q.m.nslots = tp.nslots; q.m.slot_floats = tp.slot_floats; q.m.resident = tp.resident; q.m.group = tp.group; q.m.noise_alias = tp.alias;
    fill_noise(q.nz, b);
    fill_io(q.io, b, tp.ntiles);
    cudaError_t e = launch_rnn_tc_static(q, tp.grid, tp.smem_bytes, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "mpcdl_rollout_static_scratch: launch failed: %s (smem=%d grid=%d)", cudaGetErrorString(e), tp.smem_bytes, tp.grid);
    return MPCDL_OK;
}

int mpcdl_rollout(const mpcdl_rollout_args* a, void* stream) {
    int rc = check_args(a);
    if (rc) return rc;
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    const int num_sms = di->sms;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (is_lt(*a)) {
        LtPlan lp;
        make_lt_plan(*a, lp, num_sms);
        LtRolloutParams q;
        memset(&q, 0, sizeof(q));
        build_lt_offsets(a->model, q.m);
        q.m.blob = a->wblob_tc;
        q.m.stateful = a->stateful != 0;
        fill_noise(q.nz, *a);
        q.scratch = static_cast<unsigned char*>(a->scratch);
#ifdef MPCDL_LT_DEBUG      // development builds only (make DEFS=-DMPCDL_LT_DEBUG, tools/lt_dbg.py): host-mapped progress words
        if (const char* e = getenv("MPCDL_LT_DBG")) q.dbg = reinterpret_cast<unsigned int*>(strtoull(e, nullptr, 0));
#endif
        fill_io(q.io, *a, lp.ntiles);
        cudaError_t e = launch_lstm_tc(q, lp.grid, lp.smem_bytes, st);
        if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "BF16 LSTM kernel launch failed: %s (smem=%d)", cudaGetErrorString(e), lp.smem_bytes);
        return MPCDL_OK;
    }
    if (a->precision != MPCDL_PREC_FP32 && a->model.kind == MPCDL_MLP) {
        McPlan mp;
        rc = make_mc_plan(*a, mp, num_sms);
        if (rc) return rc;
        McRolloutParams q;
        memset(&q, 0, sizeof(q));
        build_mc_offsets(a->model, a->precision == MPCDL_PREC_TF32X3, q.m);
        q.m.blob = a->wblob_tc; q.m.prec = a->precision; q.m.tiles_per_cta = mp.tiles_per_cta;
        fill_noise(q.nz, *a);
        fill_io(q.io, *a, mp.ntiles);
        cudaError_t e = launch_mlp_tc(q, mp.grid, mp.smem_bytes, st);
        if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "tensor-core MLP kernel launch failed: %s (smem=%d)", cudaGetErrorString(e), mp.smem_bytes);
        return MPCDL_OK;
    }
    if (a->precision != MPCDL_PREC_FP32) {
        TcPlan tp;
        rc = make_tc_plan(*a, tp, num_sms);
        if (rc) return rc;
        TcRolloutParams q;
        memset(&q, 0, sizeof(q));
        build_tc_parts(a->model, q.m.parts, q.m.bias_off, q.m.fc_off);
        q.m.blob = a->wblob_tc;
        q.m.cell = a->model.kind == MPCDL_GRU ? 0 : 1;
        q.m.prec = a->precision; q.m.layers = a->model.layers; q.m.window = a->model.window; q.m.stateful = a->stateful != 0;
        q.m.nslots = tp.nslots; q.m.slot_floats = tp.slot_floats; q.m.resident = tp.resident; q.m.group = tp.group; q.m.noise_alias = tp.alias;
        fill_noise(q.nz, *a);
        fill_io(q.io, *a, tp.ntiles);
        q.seqbuf = static_cast<float*>(a->scratch);
        cudaError_t e = launch_rnn_tc(q, tp.grid, tp.smem_bytes, st);
        if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "tensor-core kernel launch failed: %s (smem=%d nslots=%d)", cudaGetErrorString(e), tp.smem_bytes, tp.nslots);
        return MPCDL_OK;
    }
    Plan pl;
    rc = make_plan(*a, pl, num_sms);
    if (rc) return rc;

    RolloutParams p;
    memset(&p, 0, sizeof(p));
    const uint32_t total = build_parts(a->model, p.m.parts, p.m.nparts);
    fill_model(p.m, *a, pl, total);
    fill_noise(p.nz, *a);
    fill_io(p.io, *a, pl.ntiles);
    if (pl.variant != 2) set_stacking(p, *a, pl.TN, total);
    cudaError_t e;
    if (pl.variant == 2) e = launch_mlp2(p, pl.rt, pl.rj, pl.ks, pl.ng, pl.grid, pl.smem_bytes, st);
    else if (a->model.kind == MPCDL_MLP) e = launch_mlp(p, pl.rt, pl.rj, pl.grid, pl.smem_bytes, st);
    else e = launch_rnn(p, a->model.kind == MPCDL_GRU ? 0 : 1, pl.rt, pl.rj, pl.grid, pl.smem_bytes, st);
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "kernel launch failed: %s (rt=%d n_tg=%d smem=%d)", cudaGetErrorString(e), pl.rt, pl.n_tg, pl.smem_bytes);
    return MPCDL_OK;
}

int mpcdl_noise_forward(const mpcdl_noise_args* a, void* stream) {
    if (!a) return fail(MPCDL_ERR_ARG, "args is NULL");
    if (!a->noise_wblob || !a->q || !a->u || !a->out) return fail(MPCDL_ERR_ARG, "noise_wblob/q/u/out must be non-NULL device pointers");
    if (a->N < 1 || a->H < 1) return fail(MPCDL_ERR_ARG, "N=%lld H=%d must be >= 1", (long long)a->N, a->H);
    if (!a->closed_loop && a->H != 1) return fail(MPCDL_ERR_ARG, "FeedforwardNoiseModel.forward is one step: H=%d must be 1", a->H);
    if (a->closed_loop && (a->mu || a->scale)) return fail(MPCDL_ERR_ARG, "mu/scale outputs exist for the one-step forward only");
    if (a->history < 1 || a->history > 8) return fail(MPCDL_ERR_UNSUPPORTED, "noise history=%d outside [1,8]", a->history);
    if (a->size < 4 || a->size > 64 || a->size % 4) return fail(MPCDL_ERR_UNSUPPORTED, "noise size=%d must be a multiple of 4 in [4,64]", a->size);
    if (a->n_layers < 0 || a->n_layers > MAX_NOISE_LAYERS) return fail(MPCDL_ERR_UNSUPPORTED, "noise n_layers=%d outside [0,%d]", a->n_layers, MAX_NOISE_LAYERS);
    if (a->act < MPCDL_ACT_IDENTITY || a->act > MPCDL_ACT_SIGMOID) return fail(MPCDL_ERR_UNSUPPORTED, "noise activation %d unsupported", a->act);
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    AuxParams p;
    memset(&p, 0, sizeof(p));
    p.nz.blob = a->noise_wblob; p.nz.enabled = 1; p.nz.hist = a->history; p.nz.size = a->size; p.nz.nlayers = a->n_layers;
    p.nz.act = a->act; p.nz.in_dim = 8 * a->history;
    p.nz.blob_floats = (int)mpcdl_noise_blob_floats(a->history, a->size, a->n_layers);
    p.q = a->q; p.u = a->u; p.eps = a->eps; p.seed = a->seed; p.traj_offset = a->traj_offset; p.step0 = a->step0;
    p.out = a->out; p.mu = a->mu; p.scale = a->scale; p.N = a->N; p.H = a->H;
    p.ntiles = (int)((a->N + AUX_TN - 1) / AUX_TN);
    const int smem = aux_smem_bytes(p.nz.blob_floats, p.nz.in_dim, p.nz.size);
    if (smem > SMEM_LIMIT) return fail(MPCDL_ERR_UNSUPPORTED, "noise net needs %d bytes of shared memory", smem);
    cudaError_t e = launch_aux(p, a->closed_loop != 0, std::min(p.ntiles, di->sms), smem, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "noise kernel launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

size_t mpcdl_mppi_workspace_floats(int H) { return H < 1 ? 0 : (size_t)MPPI_GRID * (2 * (size_t)H + 2); }

int mpcdl_mppi_min(const float* costs, int64_t N, float* workspace, float* cmin_out, void* stream) {
    if (!costs || !workspace || !cmin_out || N < 1) return fail(MPCDL_ERR_ARG, "mpcdl_mppi_min: bad arguments");
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_mppi_min(costs, N, workspace, cmin_out, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "mppi_min launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_mppi_reduce(const mpcdl_mppi_args* a, void* stream) {
    if (!a) return fail(MPCDL_ERR_ARG, "args is NULL");
    if (!a->costs || !a->U || !a->workspace || !a->sums) return fail(MPCDL_ERR_ARG, "costs/U/workspace/sums must be non-NULL device pointers");
    if (a->N < 1 || a->H < 1 || a->H > MPPI_MAX_H) return fail(MPCDL_ERR_ARG, "N=%lld must be >= 1 and H=%d in [1,%d]", (long long)a->N, a->H, MPPI_MAX_H);
    if (!(a->lambda > 0.f)) return fail(MPCDL_ERR_ARG, "lambda must be > 0");
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    MppiParams p;
    memset(&p, 0, sizeof(p));
    p.costs = a->costs; p.N = a->N; p.traj_offset = a->traj_offset; p.H = a->H; p.U = a->U; p.ctrl_seed = a->ctrl_seed;
    for (int i = 0; i < 2; ++i) { p.sig[i] = a->ctrl_sigma[i]; p.ulo[i] = a->ctrl_lo[i]; p.uhi[i] = a->ctrl_hi[i]; }
    p.inv_lambda = 1.0f / a->lambda; p.cmin = a->cmin; p.workspace = a->workspace; p.sums = a->sums;
    cudaError_t e = cudaSuccess;
    if (!a->cmin) {
        e = launch_mppi_min(a->costs, a->N, a->workspace, a->sums + 2 * a->H + 1, st);
        p.cmin = a->sums + 2 * a->H + 1;
    }
    if (e == cudaSuccess) e = launch_mppi_reduce(p, st);
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "mppi_reduce launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_error_metrics(const float* output, const float* labels, int64_t B, float* mae, float* rmse, void* stream) {
    if (!output || !labels || !mae || !rmse || B < 1) return fail(MPCDL_ERR_ARG, "mpcdl_error_metrics: bad arguments");
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_error_metrics(output, labels, B, mae, rmse, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "error_metrics launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

namespace {
int check_dataset(const mpcdl_dataset_desc* d, DatasetParams& q) {
    if (!d) return fail(MPCDL_ERR_ARG, "dataset desc is NULL");
    if (!d->rows || !d->bag_row_off || !d->samp_off) return fail(MPCDL_ERR_ARG, "rows/bag_row_off/samp_off must be non-NULL device pointers");
    if (d->n_bags < 1 || d->n < 1) return fail(MPCDL_ERR_ARG, "n_bags=%d n=%lld must be >= 1", d->n_bags, (long long)d->n);
    if (d->window < 0 || d->window > MAX_WINDOW) return fail(MPCDL_ERR_UNSUPPORTED, "window=%d outside [0,%d]", d->window, MAX_WINDOW);
    q.rows = d->rows; q.bag_row_off = reinterpret_cast<const long long*>(d->bag_row_off);
    q.samp_off = reinterpret_cast<const long long*>(d->samp_off); q.n_bags = d->n_bags; q.window = d->window; q.n = d->n;
    return MPCDL_OK;
}
}  // namespace

int mpcdl_dataset_stats(const mpcdl_dataset_desc* d, float* stats, void* stream) {
    DatasetParams q;
    if (int rc = check_dataset(d, q)) return rc;
    if (!stats) return fail(MPCDL_ERR_ARG, "stats is NULL");
    if (d->n < 2) return fail(MPCDL_ERR_ARG, "the unbiased std needs at least 2 samples, got %lld", (long long)d->n);
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_dataset_stats(q, stats, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "dataset_stats launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_dataset_build(const mpcdl_dataset_desc* d, const float* stats, float* featrs, float* labels, void* stream) {
    DatasetParams q;
    if (int rc = check_dataset(d, q)) return rc;
    if (!featrs || !labels) return fail(MPCDL_ERR_ARG, "featrs/labels must be non-NULL device pointers");
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_dataset_build(q, stats, featrs, labels, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "dataset_build launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_rollout_inputs(const mpcdl_dataset_desc* d, int32_t H, float* init, float* U, float* target, void* stream) {
    DatasetParams q;
    if (int rc = check_dataset(d, q)) return rc;
    if (!init || !U || H < 1) return fail(MPCDL_ERR_ARG, "init/U must be non-NULL device pointers and H=%d >= 1", H);
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_rollout_inputs(q, H, init, U, target, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "rollout_inputs launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_curve_metrics(const float* pred, const float* target, const float* lstat, int32_t n_models, int64_t n, int32_t H,
                        float* mae, float* rmse, void* stream) {
    if (!pred || !target || !lstat || !mae || !rmse || n_models < 1 || n < 1 || H < 1) return fail(MPCDL_ERR_ARG, "mpcdl_curve_metrics: bad arguments");
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_curve_metrics(pred, target, lstat, n_models, n, H, mae, rmse, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "curve_metrics launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_resample(const mpcdl_resample_args* a, void* stream) {
    if (!a) return fail(MPCDL_ERR_ARG, "args is NULL");
    if (!a->t || !a->q || !a->u || !a->q_out || !a->u_out) return fail(MPCDL_ERR_ARG, "t/q/u/q_out/u_out must be non-NULL device pointers");
    if (a->M < 1 || a->step < 1 || a->n_interp < 0 || !(a->dataset_dt > 0.0)) return fail(MPCDL_ERR_ARG, "mpcdl_resample: M=%lld step=%d n_interp=%lld dt=%g", (long long)a->M, a->step, (long long)a->n_interp, a->dataset_dt);
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    ResampleParams p;
    p.t = a->t; p.q = a->q; p.u = a->u; p.M = a->M; p.dt = a->dataset_dt; p.step = a->step;
    p.n_q = (a->n_interp + a->step - 1) / a->step; p.n_u = a->n_interp / a->step;
    p.q_out = a->q_out; p.u_out = a->u_out;
    cudaError_t e = launch_resample(p, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "resample launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_debug_poison_smem(void* stream) {
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_poison_smem(di->sms, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "poison_smem launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_cdr_decode(const uint8_t* blob, const int64_t* offsets, int64_t n, double* out, int32_t* status, void* stream) {
    if (!blob || !offsets || !out || !status || n < 1) return fail(MPCDL_ERR_ARG, "mpcdl_cdr_decode: blob/offsets/out/status must be non-NULL device pointers and n >= 1");
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_cdr_decode(blob, reinterpret_cast<const long long*>(offsets), n, out, status, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "cdr_decode launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int64_t mpcdl_train_param_count(int32_t hidden, int32_t layers) {
    if (hidden < 1 || layers < 0) return 0;
    return (int64_t)hidden * 8 + hidden + (int64_t)layers * ((int64_t)hidden * hidden + hidden) + 6 * (int64_t)hidden + 6;
}

int mpcdl_train_mlp_epoch(const mpcdl_train_args* a, void* stream) {
    if (!a) return fail(MPCDL_ERR_ARG, "args is NULL");
    if (!a->theta || !a->adam_m || !a->adam_v || !a->featrs || !a->labels || !a->perm)
        return fail(MPCDL_ERR_ARG, "theta/adam_m/adam_v/featrs/labels/perm must be non-NULL device pointers");
    if (a->n_models < 1 || a->n_samples < 1 || a->batch < 1 || a->n_rows < 1 || a->step0 < 0)
        return fail(MPCDL_ERR_ARG, "n_models=%d n_samples=%d batch=%d n_rows=%lld step0=%lld must be positive", a->n_models, a->n_samples, a->batch,
                    (long long)a->n_rows, (long long)a->step0);
    if ((a->hidden != 16 && a->hidden != 32 && a->hidden != 64) || a->layers < 1 || a->layers > MAX_LAYERS)
        return fail(MPCDL_ERR_UNSUPPORTED, "training kernel: hidden=%d (16, 32, 64) layers=%d (1..%d)", a->hidden, a->layers, MAX_LAYERS);
    if (!(a->lr > 0.0) || !(a->beta1 >= 0.0 && a->beta1 < 1.0) || !(a->beta2 >= 0.0 && a->beta2 < 1.0) || !(a->eps >= 0.0))
        return fail(MPCDL_ERR_ARG, "Adam hyper-parameters out of range (lr=%g beta1=%g beta2=%g eps=%g)", a->lr, a->beta1, a->beta2, a->eps);
    const int smem = train_smem_bytes(a->hidden, a->layers);
    if (smem > SMEM_LIMIT) return fail(MPCDL_ERR_UNSUPPORTED, "training kernel: hidden=%d layers=%d need %d bytes of shared memory", a->hidden, a->layers, smem);
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    TrainParams p;
    memset(&p, 0, sizeof(p));
    p.theta = a->theta; p.adam_m = a->adam_m; p.adam_v = a->adam_v; p.featrs = a->featrs; p.labels = a->labels; p.perm = a->perm;
    p.losses = a->losses; p.n_samples = a->n_samples; p.batch = a->batch; p.hidden = a->hidden; p.layers = a->layers;
    p.step0 = a->step0; p.lr = a->lr; p.beta1 = a->beta1; p.beta2 = a->beta2; p.eps = a->eps;
    cudaError_t e = launch_train_mlp(p, a->n_models, smem, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "training kernel launch failed: %s (smem=%d)", cudaGetErrorString(e), smem);
    return MPCDL_OK;
}

int64_t mpcdl_train_rnn_param_count(int32_t kind, int32_t hidden, int32_t layers) {
    if ((kind != 1 && kind != 2) || hidden < 1 || layers < 1) return 0;
    return train_rnn_param_count(kind, hidden, layers);
}

int64_t mpcdl_train_rnn_work_floats(int32_t kind, int32_t hidden, int32_t layers, int32_t window) {
    if ((kind != 1 && kind != 2) || hidden < 1 || layers < 1 || window < 1) return 0;
    return train_rnn_work_floats(kind, hidden, layers, window);
}

int mpcdl_train_rnn_epoch(const mpcdl_train_rnn_args* a, void* stream) {
    if (!a) return fail(MPCDL_ERR_ARG, "args is NULL");
    if (!a->theta || !a->adam_m || !a->adam_v || !a->featrs || !a->labels || !a->perm || !a->work)
        return fail(MPCDL_ERR_ARG, "theta/adam_m/adam_v/featrs/labels/perm/work must be non-NULL device pointers");
    if (a->n_models < 1 || a->n_samples < 1 || a->batch < 1 || a->n_rows < 1 || a->step0 < 0)
        return fail(MPCDL_ERR_ARG, "n_models=%d n_samples=%d batch=%d n_rows=%lld step0=%lld must be positive", a->n_models, a->n_samples, a->batch,
                    (long long)a->n_rows, (long long)a->step0);
    if ((a->kind != 1 && a->kind != 2) || (a->hidden != 16 && a->hidden != 32 && a->hidden != 64 && a->hidden != 128) || a->layers < 1 ||
        a->layers > MAX_LAYERS || a->window < 1 || a->window > MAX_WINDOW)
        return fail(MPCDL_ERR_UNSUPPORTED, "recurrent training kernel: kind=%d (1 GRU, 2 LSTM) hidden=%d (16, 32, 64, 128) layers=%d (1..%d) window=%d (1..%d)",
                    a->kind, a->hidden, a->layers, MAX_LAYERS, a->window, MAX_WINDOW);
    if (!(a->lr > 0.0) || !(a->beta1 >= 0.0 && a->beta1 < 1.0) || !(a->beta2 >= 0.0 && a->beta2 < 1.0) || !(a->eps >= 0.0))
        return fail(MPCDL_ERR_ARG, "Adam hyper-parameters out of range (lr=%g beta1=%g beta2=%g eps=%g)", a->lr, a->beta1, a->beta2, a->eps);
    const long long np = train_rnn_param_count(a->kind, a->hidden, a->layers);
    if (a->param_stride < np || (a->param_stride & 3))
        return fail(MPCDL_ERR_ARG, "param_stride=%lld must be a multiple of 4 and >= %lld parameters", (long long)a->param_stride, np);
    if (((uintptr_t)a->theta | (uintptr_t)a->adam_m | (uintptr_t)a->adam_v | (uintptr_t)a->work) & 15)
        return fail(MPCDL_ERR_ARG, "theta/adam_m/adam_v/work must be 16-byte aligned");
    const int smem = train_rnn_smem_bytes(a->kind, a->hidden);
    if (smem > SMEM_LIMIT) return fail(MPCDL_ERR_UNSUPPORTED, "recurrent training kernel: hidden=%d needs %d bytes of shared memory", a->hidden, smem);
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    TrainRnnParams p;
    memset(&p, 0, sizeof(p));
    p.theta = a->theta; p.adam_m = a->adam_m; p.adam_v = a->adam_v; p.featrs = a->featrs; p.labels = a->labels; p.perm = a->perm;
    p.losses = a->losses; p.work = a->work; p.work_floats = train_rnn_work_floats(a->kind, a->hidden, a->layers, a->window);
    p.param_stride = a->param_stride; p.n_samples = a->n_samples; p.batch = a->batch; p.kind = a->kind; p.hidden = a->hidden;
    p.layers = a->layers; p.window = a->window; p.step0 = a->step0; p.lr = a->lr; p.beta1 = a->beta1; p.beta2 = a->beta2; p.eps = a->eps;
    cudaError_t e = launch_train_rnn(p, a->n_models, di->sms, smem, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "recurrent training kernel launch failed: %s (smem=%d)", cudaGetErrorString(e), smem);
    return MPCDL_OK;
}

int mpcdl_l1_batches(const float* output, const float* labels, int64_t n, int32_t batch, float* losses, void* stream) {
    if (!output || !labels || !losses || n < 1 || batch < 1) return fail(MPCDL_ERR_ARG, "mpcdl_l1_batches: bad arguments");
    const DevInfo* di = nullptr;
    if (int drc = sm100_device(di)) return drc;
    cudaError_t e = launch_l1_batches(output, labels, n, batch, losses, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "l1_batches launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_ffma_peak(float* out_device, int iters, int grid, void* stream) {
    if (!out_device || iters < 1 || grid < 1) return fail(MPCDL_ERR_ARG, "mpcdl_ffma_peak: bad arguments");
    cudaError_t e = launch_ffma_peak(out_device, iters, grid, false, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "ffma_peak launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

int mpcdl_ffma2_peak(float* out_device, int iters, int grid, void* stream) {
    if (!out_device || iters < 1 || grid < 1) return fail(MPCDL_ERR_ARG, "mpcdl_ffma2_peak: bad arguments");
    cudaError_t e = launch_ffma_peak(out_device, iters, grid, true, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return fail(MPCDL_ERR_CUDA, "ffma2_peak launch failed: %s", cudaGetErrorString(e));
    return MPCDL_OK;
}

}  // extern "C"
