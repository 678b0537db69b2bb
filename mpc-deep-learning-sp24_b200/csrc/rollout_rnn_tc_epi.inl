// Epilogue warps of rollout_rnn_tc_kernel (included inside the kernel body; warps 0 .. 4 TC_UG - 1).
// Thread <-> data: TMEM lane / trajectory row = 32*(warp%4) + lane; hidden units [UPT*uh, UPT*uh+UPT), uh = warp/4, UPT = TC_UPT.
// The layer's sequence lives in registers as seq[T][UPT] and is ROTATED by one frame after every cell, so the
// frame loop is a real loop (4x less code than unrolling it): at cell (l, f)
//   seq[0] = layer l-1 output of frame f (consumed as x, then overwritten by this cell's output)
//   seq[T-1] = layer l output of frame f-1 (the recurrent input, f > 0)
// Window 8 (T > TC_T) would need 256 registers for the sequence: there only seq[0] (x, then the output) and seq[1] (the
// recurrent input) are registers and the sequence goes through p.seqbuf [T][64 units][128 rows] (global memory, L2): a
// thread stores its UPT units of frame f after the cell and loads the next cell's input -- always words it wrote itself,
// so no barrier or fence is involved; layer l overwrites frame f only after it has consumed layer l-1's frame f.
{
    const int q = warp & 3, uh = warp >> 2;
    const int row = q * 32 + lane;
    constexpr int UPT = TC_UPT;
    const int u0 = uh * UPT;
    const bool owner = uh == 0;
    const uint32_t tlane = tbase + (static_cast<uint32_t>(q * 32) << 16);
    float* const normS = sm + L.norm;                  // f_mean[T*8] f_std[T*8] l_mean[6] l_std[6]
    float* const xwin = sm + L.xwin;                   // raw window [T][8][TN]
    float* const ypart = sm + L.ypart;                 // [TC_UG][6][TN]
    float* const nin = sm + L.nin;
    float* const nact0 = sm + L.nact0;
    const int nact_flip = L.nact1 - L.nact0;
    float* const nzb = sm + L.nzb;
    const float4* const biasS = reinterpret_cast<const float4*>(sm + L.bias);
    const float* const fcS = sm + L.fc;
    uint32_t cellpar = 0;
    constexpr bool SCR = T > TC_T;                     // sequence through the global scratch (see the header)
    constexpr int SR = SCR ? 2 : T;
    constexpr bool PRE = FAST && !SCR && T % 2 == 0;   // next cell's input operand staged under this cell's MMAs (TF32, window 4)
    float* const seqb = SCR ? p.seqbuf + (size_t)blockIdx.x * T * HID * TN : nullptr;
    constexpr float C_SIG = FAST ? 0.5f : -1.4426950408889634f;       // pre-activation scale of sigmoid gates
    constexpr float C_TANH = FAST ? 1.0f : 2.8853900817779268f;       // ... of tanh gates (biases pre-scaled alike)

    if (CELL == CELL_GRU) {                            // nh accumulates onto zero (see issue_chunk)
        uint32_t zr[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) zr[i] = 0u;
#pragma unroll
        for (int b = 0; b < UPT; b += 16) umma::tmem_st16(tlane + TC_D + 192 + u0 + b, zr);
    }

    for (int tile = blockIdx.x; tile < p.io.ntiles; tile += gridDim.x) {
        long long n = (long long)tile * TN + row;
        const bool valid = n < N;
        if (n >= N) n = N - 1;
        float cost = 0.f;
        if (owner) {                                   // initial window -> xwin[f][k][row] (own column only)
            const float4* w4 = reinterpret_cast<const float4*>(p.io.init + n * p.io.init_mul * T * 8);
#pragma unroll
            for (int f = 0; f < T; ++f) {
                float4 a = __ldg(w4 + 2 * f), b = __ldg(w4 + 2 * f + 1);
                float* d = xwin + f * 8 * TN + row;
                d[0] = a.x; d[TN] = a.y; d[2 * TN] = a.z; d[3 * TN] = a.w;
                d[4 * TN] = b.x; d[5 * TN] = b.y; d[6 * TN] = b.z; d[7 * TN] = b.w;
            }
        }
        int base = 0;                                  // physical slot of logical frame 0
        float seq[SR][UPT];
#pragma unroll
        for (int f = 0; f < SR; ++f)
#pragma unroll
            for (int j = 0; j < UPT; ++j) seq[f][j] = 0.f;
        for (int t = 0; t < H; ++t) {
            float u[2] = {0.f, 0.f};
            if (owner) {
                float2 uu = load_control(p.io, n, t);
                u[0] = uu.x; u[1] = uu.y;
                if (p.nz.enabled && !alias) {          // newest frame feeds the noise history
                    int pf = base + T - 1; if (pf >= T) pf -= T;
                    float s[6], uf[2];
#pragma unroll
                    for (int d = 0; d < 6; ++d) s[d] = xwin[(pf * 8 + d) * TN + row];
                    uf[0] = xwin[(pf * 8 + 6) * TN + row]; uf[1] = xwin[(pf * 8 + 7) * TN + row];
                    noise_push(nin, TN, row, p.nz.hist, t == 0, s, uf);
                }
            }
            for (int l = 0; l < NL; ++l) {
                float cst[UPT];                        // LSTM cell state (register resident across the frames)
#pragma unroll
                for (int j = 0; j < UPT; ++j) cst[j] = 0.f;
                const float* carry_h = p.io.carry;
                float* carry_c = p.io.carry;
                if (stateful) {
                    carry_h = p.io.carry + ((long long)l * N + n) * HID + u0;
                    if (CELL == CELL_LSTM) {
                        carry_c = p.io.carry + ((long long)(NL + l) * N + n) * HID + u0;
#pragma unroll
                        for (int j = 0; j < UPT; j += 4) {
                            float4 v = *reinterpret_cast<const float4*>(carry_c + j);
                            cst[j] = v.x; cst[j + 1] = v.y; cst[j + 2] = v.z; cst[j + 3] = v.w;
                        }
                    }
                }
#pragma unroll 1
                for (int f = 0; f < T; ++f) {
                    const bool has_h = f > 0 || stateful;
                    // ---- publish the A operands of cell (l, f)
                    // One-pass TF32 leaves the lo columns free: the input operand is double-buffered there (PRE), buffer = f & 1,
                    // and every cell but a step's first finds its x already staged by the cell before it (see below).
                    const uint32_t axo = PRE ? static_cast<uint32_t>(f & 1) * 64u : 0u;
                    auto stage_x0 = [&](int fr, uint32_t off) {        // layer 0: per-position z-score of frame fr (deep_models.py:35)
                        int pf = base + fr; if (pf >= T) pf -= T;
                        float xn[8];
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            xn[k] = __fdiv_rn(__fsub_rn(xwin[(pf * 8 + k) * TN + row], normS[fr * 8 + k]), normS[T * 8 + fr * 8 + k]);
                        stage8<X3>(tlane + TC_AX_HI + off, tlane + TC_AX_LO, xn);
                    };
                    if (!PRE || (l == 0 && f == 0)) {
                        if (l == 0) {
                            if (owner) stage_x0(f, axo);
                        } else {
#pragma unroll
                            for (int b = 0; b < UPT; b += 16) stage16<X3>(tlane + TC_AX_HI + u0 + b, tlane + TC_AX_LO + u0 + b, &seq[0][b]);
                        }
                    }
                    if (f > 0) {
#pragma unroll
                        for (int b = 0; b < UPT; b += 16) stage16<X3>(tlane + TC_AH_HI + u0 + b, tlane + TC_AH_LO + u0 + b, &seq[SR - 1][b]);
                    } else if (stateful) {
                        float hv[16];
#pragma unroll
                        for (int hb = 0; hb < UPT / 16; ++hb) {
#pragma unroll
                            for (int j = 0; j < 16; j += 4) {
                                float4 v = *reinterpret_cast<const float4*>(carry_h + hb * 16 + j);
                                hv[j] = v.x; hv[j + 1] = v.y; hv[j + 2] = v.z; hv[j + 3] = v.w;
                            }
                            stage16<X3>(tlane + TC_AH_HI + u0 + hb * 16, tlane + TC_AH_LO + u0 + hb * 16, hv);
                        }
                    }
                    umma::wait_st();
                    umma::fence_before_sync();
                    mbar_arrive(opnd_ready);
                    if (PRE && !(l == NL - 1 && f == T - 1)) {
                        // While this cell's MMAs run the warps would only wait: stage the NEXT cell's input operand into the other
                        // buffer.  It never depends on this cell: frame f + 1 of layer l - 1 (or of the window), or -- last frame
                        // of a layer -- frame 0 of layer l, computed T - 1 cells ago; before the rotation both are seq[1].
                        const uint32_t nxo = static_cast<uint32_t>((f + 1) & 1) * 64u;     // T is even: frame 0 of the next layer -> 0
                        if (l == 0 && f + 1 < T) {
                            if (owner) stage_x0(f + 1, nxo);
                        } else {
#pragma unroll
                            for (int b = 0; b < UPT; b += 16) stage16<X3>(tlane + TC_AX_HI + nxo + u0 + b, tlane + TC_AX_LO, &seq[SR > 1 ? 1 : 0][b]);
                        }
                    }
                    // ---- gate pre-activations of this cell
                    mbar_wait(acc_ready, cellpar);
                    cellpar ^= 1u;
                    umma::fence_after_sync();
                    // Software pipeline over blocks of 8 units: the tcgen05.ld of block ub + 1 is in flight while block ub goes
                    // through the MUFU chain (the warps of an SM sub-partition leave the barrier together, so without it the
                    // TMEM read port and the MUFU pipe take turns).
                    float gq[2][4][8];
                    {
                        const uint32_t c0 = tlane + TC_D + u0;
                        umma::tmem_ld8(c0, gq[0][0]); umma::tmem_ld8(c0 + 64, gq[0][1]);
                        umma::tmem_ld8(c0 + 128, gq[0][2]); umma::tmem_ld8(c0 + 192, gq[0][3]);
                    }
#pragma unroll
                    for (int ub = 0; ub < UPT / 8; ++ub) {
                        float hold[8];
                        float (&g0)[8] = gq[ub & 1][0], (&g1)[8] = gq[ub & 1][1], (&g2)[8] = gq[ub & 1][2], (&g3)[8] = gq[ub & 1][3];
                        const uint32_t c0 = tlane + TC_D + u0 + ub * 8;
                        umma::wait_ld();
                        if (ub + 1 < UPT / 8) {
                            umma::tmem_ld8(c0 + 8, gq[(ub + 1) & 1][0]); umma::tmem_ld8(c0 + 8 + 64, gq[(ub + 1) & 1][1]);
                            umma::tmem_ld8(c0 + 8 + 128, gq[(ub + 1) & 1][2]); umma::tmem_ld8(c0 + 8 + 192, gq[(ub + 1) & 1][3]);
                        }
                        if (CELL == CELL_GRU) {        // nh: re-zero for the next cell's accumulate-only recurrent part
                            const uint32_t zr[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
                            umma::tmem_st8(c0 + 192, zr);
                        }
                        if (f > 0) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) hold[i] = seq[SR - 1][ub * 8 + i];
                        } else if (stateful) {
                            float4 v0 = *reinterpret_cast<const float4*>(carry_h + ub * 8), v1 = *reinterpret_cast<const float4*>(carry_h + ub * 8 + 4);
                            hold[0] = v0.x; hold[1] = v0.y; hold[2] = v0.z; hold[3] = v0.w;
                            hold[4] = v1.x; hold[5] = v1.y; hold[6] = v1.z; hold[7] = v1.w;
                        } else {
#pragma unroll
                            for (int i = 0; i < 8; ++i) hold[i] = 0.f;
                        }
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int j = ub * 8 + i;
                            const float4 b = biasS[l * HID + u0 + j];          // pre-scaled (kernel prologue)
                            if (CELL == CELL_GRU) {
                                // D columns [nx | r | z | nh]; torch GRU cell: n = tanh(i_n + b_in + r*(h_n + b_hn)); h' = (h - n)*z + n
                                const float rg = act_sig_scaled<FAST>(fmaf(g1[i], C_SIG, b.x));
                                const float zg = act_sig_scaled<FAST>(fmaf(g2[i], C_SIG, b.y));
                                const float a = fmaf(g0[i], C_TANH, b.z), d = fmaf(g3[i], C_TANH, b.w);
                                const float ng = act_tanh_scaled<FAST>(fmaf(rg, d, a));
                                seq[0][j] = fmaf(hold[i] - ng, zg, ng);
                            } else {
                                // D columns [i | f | g | o]; torch LSTM cell: c' = f*c + i*g; h' = o*tanh(c')
                                const float ig = act_sig_scaled<FAST>(fmaf(g0[i], C_SIG, b.x));
                                const float fg = act_sig_scaled<FAST>(fmaf(g1[i], C_SIG, b.y));
                                const float gg = act_tanh_scaled<FAST>(fmaf(g2[i], C_TANH, b.z));
                                const float og = act_sig_scaled<FAST>(fmaf(g3[i], C_SIG, b.w));
                                const float cn = fmaf(fg, cst[j], ig * gg);
                                cst[j] = cn;
                                seq[0][j] = og * act_tanh_scaled<FAST>(cn * C_TANH);
                            }
                        }
                    }
                    umma::fence_before_sync();         // D reads are ordered before the next cell's MMAs
                    if (stateful && f == T - 1 && valid) {
                        float* ch = p.io.carry + ((long long)l * N + n) * HID + u0;
#pragma unroll
                        for (int j = 0; j < UPT; j += 4) {
                            *reinterpret_cast<float4*>(ch + j) = make_float4(seq[0][j], seq[0][j + 1], seq[0][j + 2], seq[0][j + 3]);
                            if (CELL == CELL_LSTM)
                                *reinterpret_cast<float4*>(carry_c + j) = make_float4(cst[j], cst[j + 1], cst[j + 2], cst[j + 3]);
                        }
                    }
                    // ---- rotate: this cell's output becomes seq[T-1], the next frame's input becomes seq[0]
                    if (!SCR) {
#pragma unroll
                        for (int j = 0; j < UPT; ++j) {
                            const float o = seq[0][j];
#pragma unroll
                            for (int k = 0; k + 1 < SR; ++k) seq[k][j] = seq[k + 1][j];
                            seq[SR - 1][j] = o;
                        }
                    } else {
                        float* const sq = seqb + u0 * TN + row;
                        if (l + 1 < NL) {
#pragma unroll
                            for (int j = 0; j < UPT; ++j) __stcg(sq + (f * HID + j) * TN, seq[0][j]);
                        }
#pragma unroll
                        for (int j = 0; j < UPT; ++j) seq[1][j] = seq[0][j];
                        const int nf = f + 1 < T ? f + 1 : 0, nl = f + 1 < T ? l : l + 1;     // the next cell of this step
                        if (nl > 0 && nl < NL) {
#pragma unroll
                            for (int j = 0; j < UPT; ++j) seq[0][j] = __ldcg(sq + (nf * HID + j) * TN);
                        }
                    }
                }
            }
            // ---- head: full_c(out[:, -1, :]) partial sums over this thread's UPT units (last frame = seq[T-1])
            {
                float y[8];
#pragma unroll
                for (int d = 0; d < 8; ++d) y[d] = 0.f;
#pragma unroll
                for (int j = 0; j < UPT; ++j) {
                    const float4 w0 = *reinterpret_cast<const float4*>(fcS + (u0 + j) * 8);
                    const float4 w1 = *reinterpret_cast<const float4*>(fcS + (u0 + j) * 8 + 4);
                    const float a = seq[SR - 1][j];
                    y[0] = fmaf(a, w0.x, y[0]); y[1] = fmaf(a, w0.y, y[1]); y[2] = fmaf(a, w0.z, y[2]); y[3] = fmaf(a, w0.w, y[3]);
                    y[4] = fmaf(a, w1.x, y[4]); y[5] = fmaf(a, w1.y, y[5]);
                }
#pragma unroll
                for (int d = 0; d < 6; ++d) ypart[(uh * 6 + d) * TN + row] = y[d];
            }
            if (alias) {
                // The noise net's buffers alias the weight ring: every MMA of this step is complete (acc_ready of the last cell)
                // and the control thread loads nothing for the next step before ring_free.  Weights are re-read (9 KB, L2) and the
                // history is REBUILT from the window buffer: entry e (0 = oldest) of noise_model.py:95-99's deques at step t is
                // window frame max(T - hist + e, T - 1 - t) -- the first push left-pads with copies of the then-newest frame.
                for (int i = tid; i < p.nz.blob_floats; i += TC_EPI_THREADS) sm[L.nblob + i] = __ldg(p.nz.blob + i);
                if (owner) {
                    const int hist = p.nz.hist;
                    for (int e = 0; e < hist; ++e) {
                        int fe = T - hist + e;
                        if (fe < T - 1 - t) fe = T - 1 - t;
                        int pf = base + fe; if (pf >= T) pf -= T;
#pragma unroll
                        for (int i = 0; i < 6; ++i) nin[(e * 6 + i) * TN + row] = xwin[(pf * 8 + q_perm(i)) * TN + row];
                        nin[(6 * hist + e * 2) * TN + row] = xwin[(pf * 8 + 6) * TN + row];
                        nin[(6 * hist + e * 2 + 1) * TN + row] = xwin[(pf * 8 + 7) * TN + row];
                    }
                }
            }
            named_bar_sync(1, TC_EPI_THREADS);
            if (p.nz.enabled) {                        // residual-noise net on the CUDA cores (noise_model.py:34-40)
                const float* W = sm + L.nblob;
                int K = p.nz.in_dim;
                for (int ph = 0; ph <= p.nz.nlayers; ++ph) {
                    const bool last = ph == p.nz.nlayers;
                    const int outw = last ? 12 : p.nz.size;
                    dense_resident<8 / TC_UG, 4>(ph == 0 ? nin : nact0 + ((ph - 1) & 1) * nact_flip, K, W, outw, outw,
                                         last ? ACT_IDENTITY : p.nz.act, last ? nzb : nact0 + (ph & 1) * nact_flip, TN, tid, TC_EPI_THREADS);
                    W += (K + 1) * outw; K = p.nz.size;
                    named_bar_sync(1, TC_EPI_THREADS);
                }
            }
            // ---- owner: next state, shift the window (overwrite the oldest physical frame)
            if (owner) {
                float yv[6], nzr[12];                  // bias, then the unit groups in ascending order
#pragma unroll
                for (int d = 0; d < 6; ++d) {
                    float a = fcS[HID * 8 + d];
#pragma unroll
                    for (int g = 0; g < TC_UG; ++g) a += ypart[(g * 6 + d) * TN + row];
                    yv[d] = a;
                }
                if (p.nz.enabled) {
#pragma unroll
                    for (int i = 0; i < 12; ++i) nzr[i] = nzb[i * TN + row];
                }
                float nxt[6];
                owner_finish(p, normS + 2 * T * 8, yv, nzr, 1, 0, n, valid, t, u, nxt, cost);
                float* d = xwin + base * 8 * TN + row;
#pragma unroll
                for (int k = 0; k < 6; ++k) d[k * TN] = nxt[k];
                d[6 * TN] = u[0]; d[7 * TN] = u[1];
                if (alias) mbar_arrive(ring_free);     // this row's noise outputs are consumed: the ring may be refilled
            }
            base = base + 1 == T ? 0 : base + 1;
        }
        if (owner && valid && p.io.costs) p.io.costs[n] = cost;
    }
}
