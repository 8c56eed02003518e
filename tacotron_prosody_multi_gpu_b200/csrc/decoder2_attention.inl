        // ================= E: location-sensitive attention for (row ab, context slice asl) =================
        // (Attention.py:169-210 + _compute_attention :345-366)
        if (att) {
            float* scr = SM_F32(SO_RED);  // attention scratch aliases the GEMM reduction buffers (5120 floats)
            // E1: this row's raw LSTM output (query), bf16 pairs scattered in fragment order
            if (tid < 128) {
                const uint32_t* src = p.ring_h2raw + cur * SZ_H + frag_word(KS_H, ab, 2 * tid);
                uint32_t w = ld_word(src);
                int spins = 0;
                while (w == SENT && !spin_check(g, spins)) w = ld_word(src);
                SM_F32(SO_HQ)[2 * tid] = bf16_lo(w);
                SM_F32(SO_HQ)[2 * tid + 1] = bf16_hi(w);
            }
            __syncthreads();
            // E2: q = h2raw . Wq  (query_layer, Attention.py:187-189)
            {
                const int dp = tid & 63, k8 = tid >> 6;
                float a0 = 0.f, a1 = 0.f;
#pragma unroll 8
                for (int k = k8 * 32; k < k8 * 32 + 32; ++k) {
                    const uint32_t w = SM_U32(SO_WQ)[k * 64 + dp];
                    const float h = SM_F32(SO_HQ)[k];
                    a0 = fmaf(h, bf16_lo(w), a0);
                    a1 = fmaf(h, bf16_hi(w), a1);
                }
                scr[k8 * 128 + 2 * dp] = a0;
                scr[k8 * 128 + 2 * dp + 1] = a1;
            }
            __syncthreads();
            if (tid < ATT) {
                float q = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) q += scr[i * 128 + tid];
                SM_F32(SO_QS)[tid] = q;
            }
            __syncthreads();
            // E3: energies e[t] = sum_d v_a[d] tanh(pre[t][d] + q[d])   (Attention.py:401)
            {
                const int tl = lane & 7, qd = lane >> 3;
                const float4* q4 = reinterpret_cast<const float4*>(SM_F32(SO_QS) + qd * 32);
                const float4* v4 = reinterpret_cast<const float4*>(SM_F32(SO_VA) + qd * 32);
                for (int tb = 0; tb < alen; tb += 128) {
                    const int t = tb + warp * 8 + tl;
                    float e = 0.f;
                    if (t < alen) {
                        const float4* p4 = reinterpret_cast<const float4*>(pre + (size_t)t * pstride + qd * 32);
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float4 x = p4[i], qq = q4[i], vv = v4[i];
                            e = fmaf(vv.x, tanh_approx(x.x + qq.x), e);
                            e = fmaf(vv.y, tanh_approx(x.y + qq.y), e);
                            e = fmaf(vv.z, tanh_approx(x.z + qq.z), e);
                            e = fmaf(vv.w, tanh_approx(x.w + qq.w), e);
                        }
                    }
                    e += __shfl_xor_sync(0xffffffffu, e, 8);
                    e += __shfl_xor_sync(0xffffffffu, e, 16);
                    if (qd == 0 && t < alen) SM_F32(SO_ES)[t] = e;
                }
            }
            __syncthreads();
            // E4: masked softmax over t < len (exact zeros beyond: -inf mask), cumulate (Attention.py:203-206)
            if (warp == 0) {
                float m = -INFINITY;
                if (!p.smoothing) {
                    for (int t = lane; t < alen; t += 32) m = fmaxf(m, SM_F32(SO_ES)[t]);
                    m = warp_max(m);
                }
                float sum = 0.f;
                for (int t = lane; t < alen; t += 32) {
                    const float v = p.smoothing ? sigmoidf_(SM_F32(SO_ES)[t]) : expf(SM_F32(SO_ES)[t] - m);
                    SM_F32(SO_ES)[t] = v;
                    sum += v;
                }
                sum = warp_sum(sum);
                const float inv = 1.0f / sum;
                for (int t = lane; t < alen; t += 32) SM_F32(SO_ES)[t] *= inv;
            }
            __syncthreads();
            for (int t = tid; t < p.T; t += NT) {
                const float a = t < alen ? SM_F32(SO_ES)[t] : 0.f;
                SM_F32(SO_CPAD)[15 + t] = p.cumulative ? SM_F32(SO_CPAD)[15 + t] + a : a;
                if (asl == 0) p.hist[((size_t)s * p.B + ab) * p.T + t] = a;
            }
            // E5: context slice = sum_t a[t] values[ab][t][asl*DS .. +DS)   (bf16 values, fp32 accumulate)
            {
                const int D8 = DS >> 3;
                const int TG = NT / D8;
                const int d8 = tid % D8, tg = tid / D8;
                float c[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (tg < TG) {
                    const uint4* vp = reinterpret_cast<const uint4*>(p.values + ((size_t)ab * p.T) * MEM + asl * DS) + d8;
#pragma unroll 4
                    for (int t = tg; t < alen; t += TG) {
                        const uint4 v = __ldg(vp + (size_t)t * (MEM / 8));
                        const float a = SM_F32(SO_ES)[t];
                        c[0] = fmaf(a, bf16_lo(v.x), c[0]);
                        c[1] = fmaf(a, bf16_hi(v.x), c[1]);
                        c[2] = fmaf(a, bf16_lo(v.y), c[2]);
                        c[3] = fmaf(a, bf16_hi(v.y), c[3]);
                        c[4] = fmaf(a, bf16_lo(v.z), c[4]);
                        c[5] = fmaf(a, bf16_hi(v.z), c[5]);
                        c[6] = fmaf(a, bf16_lo(v.w), c[6]);
                        c[7] = fmaf(a, bf16_hi(v.w), c[7]);
                    }
                    float4* dst = reinterpret_cast<float4*>(scr + (size_t)tg * DS + d8 * 8);
                    dst[0] = make_float4(c[0], c[1], c[2], c[3]);
                    dst[1] = make_float4(c[4], c[5], c[6], c[7]);
                }
                __syncthreads();
                if (tid < DS / 2) {
                    float x0 = 0.f, x1 = 0.f;
                    for (int q = 0; q < TG; ++q) {
                        const float2 v = *reinterpret_cast<const float2*>(scr + (size_t)q * DS + 2 * tid);
                        x0 += v.x;
                        x1 += v.y;
                    }
                    const int col = asl * DS + 2 * tid;
                    const int idx = frag_word(KS_CTX, ab, col);
                    st_word(p.ring_ctx + cur * SZ_C + idx, pack_bf16(x0, x1));
                    st_word(p.ring_ctx + clr * SZ_C + idx, SENT);
                    *reinterpret_cast<float2*>(p.ctx_state + (size_t)ab * MEM + col) = make_float2(x0, x1);
                    DEC2_FENCE();
                }
            }
            DEC2_STAMP(8)
            // E6 (off the critical path): location pre-activations of the NEXT step from the new cumulative alignments
            __syncthreads();
            att_build_pairs(p.T, tid);
            __syncthreads();
            att_location(p, pre, pstride, ab, alen, warp, lane);
            __syncthreads();
        }
        DEC2_STAMP(9)
        // ================= F: frame + stop projection, outputs, feedback (Decoder.py:273-279; Helpers.py:36-59) =================
        if (cta < 81) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            mma_seg<4>(acc, p.ring_h2raw + cur * SZ_H, KS_H, mt, kg, KG, wo, lane, g);
            mma_seg<6>(acc, p.ring_ctx + cur * SZ_C, KS_CTX, mt, kg, KG, wo + 16 * 32, lane, g);
            DEC2_STAMP(10)
            float* redb = SM_F32(SO_RED) + (st++ & 1) * RED_FLOATS;
            store_red(redb, acc, RS, mt, kg, lane);
            __syncthreads();
            bool fin = false;
            if (ept) {
                DEC2_REDSUM(y0, redb, 20)
                const float y1 = __shfl_down_sync(0xffffffffu, y0, 1);
                if (en == 0) {
                    if (cta < 80) {
                        if (evalid)
                            *reinterpret_cast<float2*>(p.frames + ((size_t)er * p.S_cap * 2 + (size_t)s * 2) * MEL + 2 * cta) = make_float2(y0, y1);
                        if (cta >= 40) {
                            // TacoTestHelper.next_inputs: next input = last frame of this step (Helpers.py:56)
                            const float x0 = evalid ? y0 : 0.f, x1 = evalid ? y1 : 0.f;
                            if (free_run) {
                                const int idx = frag_word(KS_X, er, 2 * cta - MEL);
                                st_word(p.ring_x + ((step + 1) & 3) * SZ_X + idx, pack_bf16(x0, x1));
                                st_word(p.ring_x + ((step + 3) & 3) * SZ_X + idx, SENT);
                            }
                            if (evalid) *reinterpret_cast<float2*>(p.x_state + (size_t)er * MEL + 2 * cta - MEL) = make_float2(x0, x1);
                            DEC2_FENCE();
                        }
                    } else if (evalid) {
                        const float s0 = sigmoidf_(y0), s1 = sigmoidf_(y1);
                        *reinterpret_cast<float2*>(p.stop + (size_t)er * p.S_cap * 2 + (size_t)s * 2) =
                            p.stop_sigmoid ? make_float2(s0, s1) : make_float2(y0, y1);
                        // tf.round(sigmoid) == 1 <=> sigmoid > 0.5 (half-to-even; Helpers.py:40)
                        fin = p.stop_at_any ? (s0 > 0.5f || s1 > 0.5f) : (s0 > 0.5f && s1 > 0.5f);
                    }
                }
            }
            if (cta == 80 && free_run) {
                const int cnt = __syncthreads_count(fin ? 1 : 0);
                if (tid == 0) st_word(reinterpret_cast<uint32_t*>(p.flags + s), (p.stop_at_any ? cnt > 0 : cnt == p.B) ? 2u : 1u);
            }
        }
        DEC2_STAMP(11)
        // ================= G: next step's context / zoned-state gate contributions (shadow of the x exchange) =================
        mma_seg<4>(acc1, p.ring_ctx + cur * SZ_C, KS_CTX, mt, kg, KG, w1 + 16 * 32, lane, g);
        DEC2_STAMP(12)
        mma_seg<4>(acc1, p.ring_h1z + cur * SZ_H, KS_H, mt, kg, KG, w1 + 64 * 32, lane, g);
        DEC2_STAMP(13)
        mma_seg<4>(acc2, p.ring_h2z + cur * SZ_H, KS_H, mt, kg, KG, w2 + 16 * 32, lane, g);
        DEC2_STAMP(14)
        s_done = s + 1;
