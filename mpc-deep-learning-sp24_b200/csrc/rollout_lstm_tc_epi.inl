        // ================================================================ epilogue warps (included by rollout_lstm_tc.cu)
        // thread <-> (trajectory row = TMEM lane, unit group g): 16 of every chunk's 64 hidden units, all four gates
        const int q = warp & 3, g = warp >> 2;
        const int row = q * 32 + lane;
        const bool owner = g == 0;                            // tid < 128: keeps the trajectory's state, cost, window
        const uint32_t tlane = tbase + (static_cast<uint32_t>(q * 32) << 16);
        const uint32_t rowoff = (row >> 3) * 128 + (row & 7) * 16;          // byte offset of this row inside a 16-B k-chunk column
        unsigned char* const scratch = p.scratch + (size_t)blockIdx.x * lt_scratch_bytes_per_cta(T, nin_dim);
        float* const nhist = xwin + T * 8 * LT_TN;             // noise history [in_dim][128], row-private like xwin
        const LtNoiseSmem NZ = lt_noise_layout(p.nz);
        float* const nzS = reinterpret_cast<float*>(ah);       // noise-net buffers: A_h is idle between the last cell of a step and the next step
        const long long Ntot = p.io.N;
        const float* const norm_l = normS + 2 * T * 8;
        constexpr int NEPI = LT_EPI_WARPS * 32;
        uint32_t dq = 0;

        for (int pt = pair; pt < npt; pt += npairs) {
            const int tile = 2 * pt + (int)rank;               // an odd tile count leaves the last pair's rank 1 a dummy tile
            long long n = (long long)tile * LT_TN + row;
            const bool valid = n < p.io.N;
            if (n >= p.io.N) n = p.io.N - 1;
            float cost = 0.f;
            int base = 0;                                      // physical slot of logical frame 0
            if (owner) {
                const float* w = p.io.init + n * p.io.init_mul * T * 8;
                for (int i = 0; i < T * 8; ++i) xwin[i * LT_TN + row] = __ldg(w + i);
            }
            // layer-0 operand tile of frame f: z-scored [hi(8) | lo(8)] bf16, written by the owner threads (A_x must be free)
            auto stage_l0 = [&](int f) {
                if (owner) {
                    int pf = base + f; if (pf >= T) pf -= T;
                    float xn[8], lo[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        xn[k] = __fdiv_rn(__fsub_rn(xwin[(pf * 8 + k) * LT_TN + row], normS[f * 8 + k]), normS[T * 8 + f * 8 + k]);
                        lo[k] = xn[k] - lt_bf16_round(xn[k]);
                    }
                    *reinterpret_cast<uint4*>(ax + rowoff) =
                        make_uint4(lt_pack(xn[0], xn[1]), lt_pack(xn[2], xn[3]), lt_pack(xn[4], xn[5]), lt_pack(xn[6], xn[7]));
                    *reinterpret_cast<uint4*>(ax + LT_A_LBO + rowoff) =
                        make_uint4(lt_pack(lo[0], lo[1]), lt_pack(lo[2], lo[3]), lt_pack(lo[4], lo[5]), lt_pack(lo[6], lo[7]));
                    umma::fence_proxy_async_smem();
                    named_bar_sync(2, LT_TN);
                    if (tid == 0) mbar_arrive(&bars->ax_full);
                }
            };
            // stateful: frame 0's recurrent operand of layer l = the carried h (FP32 plane l), rounded to BF16 like every h.
            // Every thread stages the 64 units it owns (the ones it wrote at the end of the previous step / the caller's
            // initial carry).  A_h must be free: called where every MMA of the previous cell is complete.
            auto stage_h0 = [&](int l) {
                const float* hc = p.io.carry + ((long long)l * Ntot + n) * LT_HID;
#pragma unroll
                for (int c = 0; c < LT_NCH; ++c)
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        const int u0 = c * LT_CH + 16 * g + 8 * half;
                        const float4 a = *reinterpret_cast<const float4*>(hc + u0), b = *reinterpret_cast<const float4*>(hc + u0 + 4);
                        *reinterpret_cast<uint4*>(ah + (size_t)(c * 8 + g * 2 + half) * LT_A_LBO + rowoff) =
                            make_uint4(lt_pack(a.x, a.y), lt_pack(a.z, a.w), lt_pack(b.x, b.y), lt_pack(b.z, b.w));
                    }
                umma::fence_proxy_async_smem();
                named_bar_sync(1, NEPI);
                if (tid == 0) { mbar_arrive(&bars->ah_full[0]); mbar_arrive(&bars->ah_full[1]); }
            };
            for (int t = 0; t < H; ++t) {
                float u[2] = {0.f, 0.f};
                if (owner) { const float2 uu = load_control(p.io, n, t); u[0] = uu.x; u[1] = uu.y; }
                float y[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                for (int l = 0; l < NL; ++l) {
                    named_bar_sync(1, NEPI);                                   // the previous layer's bias readers are done
                    for (int i = tid; i < 1024; i += NEPI) biasS[i] = __ldg(m.blob + m.bias_off + l * 1024 + i);
                    named_bar_sync(1, NEPI);
                    float cst[LT_NCH * 16];
#pragma unroll
                    for (int i = 0; i < LT_NCH * 16; ++i) cst[i] = 0.f;        // stateless: (h, c) = 0 at every call (deep_models.py:82-83)
                    if (stateful) {                                             // carried cell state: plane L + l
                        const float* cc = p.io.carry + ((long long)(NL + l) * Ntot + n) * LT_HID;
#pragma unroll
                        for (int c = 0; c < LT_NCH; ++c)
#pragma unroll
                            for (int hq = 0; hq < 4; ++hq) {
                                const float4 v = *reinterpret_cast<const float4*>(cc + c * LT_CH + 16 * g + 4 * hq);
                                cst[c * 16 + 4 * hq] = v.x; cst[c * 16 + 4 * hq + 1] = v.y; cst[c * 16 + 4 * hq + 2] = v.z; cst[c * 16 + 4 * hq + 3] = v.w;
                            }
                    }
                    for (int f = 0; f < T; ++f) {
                        if (l == 0 && f == 0) { stage_l0(0); if (stateful) stage_h0(0); }
                        int l2 = l, f2 = f + 1;                                  // the next cell
                        if (f2 == T) { l2 = l + 1; f2 = 0; }
                        const bool last = l == NL - 1 && f == T - 1;
                        unsigned char* const hout = scratch + (size_t)f * LT_A_BYTES;       // in place: slot f held layer l-1's frame f (this cell's x, already in A_x)
#pragma unroll
                        for (int c = 0; c < LT_NCH; ++c) {
                            const uint32_t buf = dq & 1u;
                            LT_WAIT(&bars->d_full[buf], (dq >> 1) & 1u, 5, dq);
                            ++dq;
                            if (c == LT_NCH - 1 && l2 < NL) {
                                // every MMA of this cell is complete: A_x / A_h are free.  Stage what the next cell needs and
                                // what already exists, so its x part and three quarters of its h part run under this epilogue
                                if (l2 == 0) stage_l0(f2);
                                else if (T > 1 && tid == 0) {
                                    const unsigned char* src = scratch + (size_t)f2 * LT_A_BYTES;
                                    mbar_expect_tx(&bars->ax_full, LT_A_BYTES);
                                    bulk_g2s(ax, src, LT_A_BYTES / 2, &bars->ax_full);
                                    bulk_g2s(ax + LT_A_BYTES / 2, src + LT_A_BYTES / 2, LT_A_BYTES / 2, &bars->ax_full);
                                }
                                if (f2 == 0 && stateful) stage_h0(l2);        // the next layer starts from ITS carried h
                                if (f2 > 0 && tid == 0) {        // units 0..191 of h_f (chunks 0..2: fenced + barrier below)
                                    mbar_expect_tx(&bars->ah_full[0], 3 * LT_A_BYTES / 4);
                                    bulk_g2s(ah, hout, LT_A_BYTES / 2, &bars->ah_full[0]);
                                    bulk_g2s(ah + LT_A_BYTES / 2, hout + LT_A_BYTES / 2, LT_A_BYTES / 4, &bars->ah_full[0]);
                                }
                            }
                            if (blockIdx.x == 0 && tid == 0) LT_MARK(0, dq);
                            umma::fence_after_sync();
#pragma unroll
                            for (int half = 0; half < 2; ++half) {
                                const int u0 = 16 * g + 8 * half;                  // first of this thread's 8 units inside the chunk
                                float vi[8], vf[8], vg[8], vo[8], hv[8];
                                umma::tmem_ld8(tlane + buf * 256u + 0u + u0, vi);
                                umma::tmem_ld8(tlane + buf * 256u + 64u + u0, vf);
                                umma::tmem_ld8(tlane + buf * 256u + 128u + u0, vg);
                                umma::tmem_ld8(tlane + buf * 256u + 192u + u0, vo);
                                umma::wait_ld();
                                const float* bc = biasS + c * 256 + u0;
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    const float gi = lt_sig(vi[j] + bc[j]), gf = lt_sig(vf[j] + bc[64 + j]);
                                    const float gg = lt_tanh(vg[j] + bc[128 + j]), go = lt_sig(vo[j] + bc[192 + j]);
                                    float& cc = cst[c * 16 + half * 8 + j];
                                    cc = fmaf(gf, cc, gi * gg);                    // c' = f c + i g
                                    hv[j] = go * lt_tanh(cc);                      // h' = o tanh(c')
                                }
                                *reinterpret_cast<uint4*>(hout + (size_t)(c * 8 + g * 2 + half) * LT_A_LBO + rowoff) =
                                    make_uint4(lt_pack(hv[0], hv[1]), lt_pack(hv[2], hv[3]), lt_pack(hv[4], hv[5]), lt_pack(hv[6], hv[7]));
                                if (stateful && f == T - 1 && valid) {             // (h, c) after the window's last frame -> carry (FP32)
                                    float* ch = p.io.carry + ((long long)l * Ntot + n) * LT_HID + c * LT_CH + u0;
                                    float* cc = p.io.carry + ((long long)(NL + l) * Ntot + n) * LT_HID + c * LT_CH + u0;
                                    *reinterpret_cast<float4*>(ch) = make_float4(hv[0], hv[1], hv[2], hv[3]);
                                    *reinterpret_cast<float4*>(ch + 4) = make_float4(hv[4], hv[5], hv[6], hv[7]);
                                    const float* cs = &cst[c * 16 + half * 8];
                                    *reinterpret_cast<float4*>(cc) = make_float4(cs[0], cs[1], cs[2], cs[3]);
                                    *reinterpret_cast<float4*>(cc + 4) = make_float4(cs[4], cs[5], cs[6], cs[7]);
                                }
                                if (last) {                                        // full_c head on the FP32 h of the last layer's last frame
                                    const float* wf = fcG + (c * 64 + u0) * 8;
#pragma unroll
                                    for (int j = 0; j < 8; ++j) {
                                        const float4 w0 = __ldg(reinterpret_cast<const float4*>(wf + j * 8));
                                        const float2 w1 = __ldg(reinterpret_cast<const float2*>(wf + j * 8 + 4));
                                        y[0] = fmaf(hv[j], w0.x, y[0]); y[1] = fmaf(hv[j], w0.y, y[1]); y[2] = fmaf(hv[j], w0.z, y[2]);
                                        y[3] = fmaf(hv[j], w0.w, y[3]); y[4] = fmaf(hv[j], w1.x, y[4]); y[5] = fmaf(hv[j], w1.y, y[5]);
                                    }
                                }
                            }
                            umma::fence_before_sync();
                            mbar_arrive(&bars->d_empty[buf]);
                            if (c == LT_NCH - 2 && f2 > 0 && l2 < NL) {      // chunks 0..2 of h_f are in the scratch for the early A_h load
                                __threadfence();
                                lt_fence_async_global();
                                named_bar_sync(1, NEPI);
                            }
                        }
                        // ---- end of the cell: all of h_f is in the scratch; the rest of the next cell's operands follows by TMA
                        __threadfence();
                        lt_fence_async_global();
                        named_bar_sync(1, NEPI);
                        if (tid == 0 && l2 < NL) {
                            if (l2 > 0 && T == 1) {              // window 1: the next layer's input IS this cell's output
                                const unsigned char* src = scratch + (size_t)f2 * LT_A_BYTES;
                                mbar_expect_tx(&bars->ax_full, LT_A_BYTES);
                                bulk_g2s(ax, src, LT_A_BYTES / 2, &bars->ax_full);
                                bulk_g2s(ax + LT_A_BYTES / 2, src + LT_A_BYTES / 2, LT_A_BYTES / 2, &bars->ax_full);
                            }
                            if (f2 > 0) {                        // units 192..255 of h_f
                                mbar_expect_tx(&bars->ah_full[1], LT_A_BYTES / 4);
                                bulk_g2s(ah + 3 * (LT_A_BYTES / 4), hout + 3 * (LT_A_BYTES / 4), LT_A_BYTES / 4, &bars->ah_full[1]);
                            }
                        }
                    }
                }
                // ---- head: the four unit groups of a trajectory meet in shared memory; the owner finishes the step
                if (!owner) {
#pragma unroll
                    for (int d = 0; d < 6; ++d) ypart[((g - 1) * 6 + d) * LT_TN + row] = y[d];
                }
                if (p.nz.enabled) {
                    // residual-noise net (noise_model.py:34-40) on the CUDA cores: weights + history -> the idle A_h region
                    for (int i = tid; i < p.nz.blob_floats; i += NEPI) nzS[NZ.blob + i] = __ldg(p.nz.blob + i);
                    if (owner) {                               // (q, u) of the window's newest frame joins the history (noise_model.py:95-99)
                        int pf = base + T - 1; if (pf >= T) pf -= T;
                        float s6[6], uf[2];
#pragma unroll
                        for (int d = 0; d < 6; ++d) s6[d] = xwin[(pf * 8 + d) * LT_TN + row];
                        uf[0] = xwin[(pf * 8 + 6) * LT_TN + row]; uf[1] = xwin[(pf * 8 + 7) * LT_TN + row];
                        noise_push(nhist, LT_TN, row, p.nz.hist, t == 0, s6, uf);
                        for (int k = 0; k < p.nz.in_dim; ++k) nzS[NZ.nin + k * LT_TN + row] = nhist[k * LT_TN + row];
                    }
                }
                named_bar_sync(1, NEPI);
                if (p.nz.enabled) {
                    const float* W = nzS + NZ.blob;
                    int K = p.nz.in_dim;
                    for (int ph = 0; ph <= p.nz.nlayers; ++ph) {
                        const bool lastp = ph == p.nz.nlayers;
                        const int outw = lastp ? 12 : p.nz.size;
                        const float* in = ph == 0 ? nzS + NZ.nin : nzS + (((ph - 1) & 1) ? NZ.nact1 : NZ.nact0);
                        float* out = lastp ? nzS + NZ.nzb : nzS + ((ph & 1) ? NZ.nact1 : NZ.nact0);
                        dense_resident<4, 4>(in, K, W, outw, outw, lastp ? ACT_IDENTITY : p.nz.act, out, LT_TN, tid, NEPI);
                        W += (K + 1) * outw; K = p.nz.size;
                        named_bar_sync(1, NEPI);
                    }
                }
                if (owner) {
#pragma unroll
                    for (int d = 0; d < 6; ++d)
                        y[d] = y[d] + ypart[(0 * 6 + d) * LT_TN + row] + ypart[(1 * 6 + d) * LT_TN + row] + ypart[(2 * 6 + d) * LT_TN + row] +
                               __ldg(fcG + LT_HID * 8 + d);
                    float nxt[6];
                    if (p.nz.enabled) {                        // owner_finish reads y and the noise output with one column stride
                        float* yb = nzS + NZ.ybuf;
#pragma unroll
                        for (int d = 0; d < 6; ++d) yb[d * LT_TN + row] = y[d];
                        owner_finish(p, norm_l, yb, nzS + NZ.nzb, LT_TN, row, n, valid, t, u, nxt, cost);
                    } else {
                        owner_finish(p, norm_l, y, nullptr, 1, 0, n, valid, t, u, nxt, cost);
                    }
                    // window shift: the oldest physical slot takes the new frame [s_{t+1}, u_t]
#pragma unroll
                    for (int d = 0; d < 6; ++d) xwin[(base * 8 + d) * LT_TN + row] = nxt[d];
                    xwin[(base * 8 + 6) * LT_TN + row] = u[0];
                    xwin[(base * 8 + 7) * LT_TN + row] = u[1];
                }
                if (++base == T) base = 0;
                named_bar_sync(1, NEPI);       // ypart (inside the A_x region) is dead before the next step's layer-0 tile is written
            }
            if (owner && valid && p.io.costs) p.io.costs[n] = cost;
        }
