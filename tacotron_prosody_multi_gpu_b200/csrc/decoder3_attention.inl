// Attention helpers of decoder3.cu (kept in a separate file only for size).
namespace {

// pair arrays of the hi/lo bf16 split of the zero-padded cumulative alignments (Toeplitz A operand)
// from_es: the sequence is the NEW alignments a[t] (SO_ES, valid for t < len) instead of the cumulative ones (SO_CPAD)
__device__ __forceinline__ void att3_build_pairs(int T, int tid, bool from_es = false, int len = 0) {
    const float* cpad = SMF(SO_CPAD);
    const float* es = SMF(SO_ES);
    uint32_t* P = SMU(SO_P);
    const int n = (T + 46) / 2 + 1;
    auto val = [&](int i) -> float {
        if (i >= T_MAX3 + 48) return 0.f;
        if (!from_es) return cpad[i];
        const int t = i - 15;
        return (t >= 0 && t < len) ? es[t] : 0.f;
    };
    for (int i = tid; i < n && i < PW3; i += NT) {
        const float c0 = val(2 * i), c1 = val(2 * i + 1), c2 = val(2 * i + 2);
        const float h0 = __bfloat162float(__float2bfloat16_rn(c0));
        const float h1 = __bfloat162float(__float2bfloat16_rn(c1));
        const float h2 = __bfloat162float(__float2bfloat16_rn(c2));
        P[i] = pack_bf16x2(h0, h1);
        P[PW3 + i] = pack_bf16x2(h1, h2);
        P[2 * PW3 + i] = pack_bf16x2(c0 - h0, c1 - h1);
        P[3 * PW3 + i] = pack_bf16x2(c1 - h1, c2 - h2);
    }
}

// ... the same from the window [t0 - 15, t0 + Tl + 15) of a complete alignment row (position slices: the halo belongs to the
// neighbouring slices); positions outside [0, rowlen) are 0
__device__ __forceinline__ void att3_build_pairs_row(const float* row, int t0, int rowlen, int Tl, int tid) {
    uint32_t* P = SMU(SO_P);
    const int n = (Tl + 46) / 2 + 1;
    auto val = [&](int i) -> float {
        const int t = t0 + i - 15;
        return (t >= 0 && t < rowlen) ? row[t] : 0.f;
    };
    for (int i = tid; i < n && i < PW3; i += NT) {
        const float c0 = val(2 * i), c1 = val(2 * i + 1), c2 = val(2 * i + 2);
        const float h0 = __bfloat162float(__float2bfloat16_rn(c0));
        const float h1 = __bfloat162float(__float2bfloat16_rn(c1));
        const float h2 = __bfloat162float(__float2bfloat16_rn(c2));
        P[i] = pack_bf16x2(h0, h1);
        P[PW3 + i] = pack_bf16x2(h1, h2);
        P[2 * PW3 + i] = pack_bf16x2(c0 - h0, c1 - h1);
        P[3 * PW3 + i] = pack_bf16x2(c1 - h1, c2 - h2);
    }
}

// One 16-row tile of the location term: acc[t][d] = sum_k cum[t + k - 15] G[k][d], d = 8*warp + 2*tq (+1), rows t0+gq, t0+gq+8
// (Attention.py:193-197 fused: conv1d(k=31, 32 filters) then Dense(128, no bias)); hi/lo bf16 split, 3 MMAs per k-step.
struct G4 {
    uint2 hi[2], lo[2];
};
__device__ __forceinline__ float4 att3_loc_tile_core(int t0, int lane, G4 gm) {
    const int gq = lane >> 2, tq = lane & 3;
    const uint32_t* P0h = SMU(SO_P);
    const uint32_t* P1h = SMU(SO_P + PW3);
    const uint32_t* P0l = SMU(SO_P + 2 * PW3);
    const uint32_t* P1l = SMU(SO_P + 3 * PW3);
    const int ta = t0 + gq, tb = ta + 8;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
        uint4 ah, al;
        // element (row, k) = cpad[t0 + row + k]; pair (k, k+1) at index i = t0 + row + k: even -> P0[i/2], odd -> P1[(i-1)/2]
        const int ia = ta + ks * 16 + 2 * tq, ib = tb + ks * 16 + 2 * tq;
        const bool oa = ia & 1, ob = ib & 1;
        ah.x = oa ? P1h[ia >> 1] : P0h[ia >> 1];
        ah.y = ob ? P1h[ib >> 1] : P0h[ib >> 1];
        ah.z = oa ? P1h[(ia + 8) >> 1] : P0h[(ia + 8) >> 1];
        ah.w = ob ? P1h[(ib + 8) >> 1] : P0h[(ib + 8) >> 1];
        al.x = oa ? P1l[ia >> 1] : P0l[ia >> 1];
        al.y = ob ? P1l[ib >> 1] : P0l[ib >> 1];
        al.z = oa ? P1l[(ia + 8) >> 1] : P0l[(ia + 8) >> 1];
        al.w = ob ? P1l[(ib + 8) >> 1] : P0l[(ib + 8) >> 1];
        mma16816(acc, ah, gm.hi[ks]);
        mma16816(acc, ah, gm.lo[ks]);
        mma16816(acc, al, gm.hi[ks]);
    }
    return make_float4(acc[0], acc[1], acc[2], acc[3]);
}
__device__ __forceinline__ void att3_loc_tile(float (&acc)[4], int t0, int lane, const uint2 (&ghi)[2], const uint2 (&glo)[2]) {
    G4 gm;
    gm.hi[0] = ghi[0]; gm.hi[1] = ghi[1]; gm.lo[0] = glo[0]; gm.lo[1] = glo[1];
    const float4 r = att3_loc_tile_core(t0, lane, gm);
    acc[0] = r.x; acc[1] = r.y; acc[2] = r.z; acc[3] = r.w;
}

// pre[t][d] = keys[b][t][d] + b_a[d] + conv_bias.Wl[d] + location term, for all t < len of row b.
// T <= PRE_REG_T: kept in registers preR[tile][4] (thread = rows 16*tile + gq (+8), columns 8*warp + 2*tq (+1));
// otherwise written to the row's global scratch [T][128].
// incremental: the pair arrays hold the NEW alignments and the (linear) location term is accumulated: pre += T(a).G
// -- no key reload from L2 (cumulative attention only).
__device__ __forceinline__ void att3_location(const Dec3Params& p, float (&preR)[NPT][4], float* pre_scr, int b, int len, int warp, int lane,
                                              bool incremental = false, int t_off = 0) {
    const int gq = lane >> 2, tq = lane & 3;
    uint2 ghi[2], glo[2];
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
        ghi[ks] = __ldg(reinterpret_cast<const uint2*>(p.g_hi) + (ks * 16 + warp) * 32 + lane);
        glo[ks] = __ldg(reinterpret_cast<const uint2*>(p.g_lo) + (ks * 16 + warp) * 32 + lane);
    }
    const int d0 = warp * 8 + 2 * tq;
    const float bc0 = SMF(SO_BC)[d0], bc1 = SMF(SO_BC)[d0 + 1];
    if (incremental) {
        if (p.pre_reg) {
#pragma unroll
            for (int i = 0; i < NPT; ++i) {
                if (16 * i < len) {
                    float acc[4];
                    att3_loc_tile(acc, 16 * i, lane, ghi, glo);
                    preR[i][0] += acc[0];
                    preR[i][1] += acc[1];
                    preR[i][2] += acc[2];
                    preR[i][3] += acc[3];
                }
            }
        } else {
            for (int t0 = 0; t0 < len; t0 += 16) {
                const int ta = t0 + gq, tb = ta + 8;
                float acc[4];
                att3_loc_tile(acc, t0, lane, ghi, glo);
                if (ta < len) {
                    float2* d = reinterpret_cast<float2*>(pre_scr + (size_t)ta * ATT + d0);
                    const float2 o = *d;
                    *d = make_float2(o.x + acc[0], o.y + acc[1]);
                }
                if (tb < len) {
                    float2* d = reinterpret_cast<float2*>(pre_scr + (size_t)tb * ATT + d0);
                    const float2 o = *d;
                    *d = make_float2(o.x + acc[2], o.y + acc[3]);
                }
            }
        }
        return;
    }
    if (p.pre_reg) {
#pragma unroll
        for (int i = 0; i < NPT; ++i) {
            const int t0 = 16 * i;
            if (t0 < len) {
                const int ta = t0 + gq, tb = ta + 8;
                float2 ka = make_float2(0.f, 0.f), kb = make_float2(0.f, 0.f);
                if (ta < len) ka = __ldg(reinterpret_cast<const float2*>(p.keys + ((size_t)b * p.T + t_off + ta) * ATT + d0));
                if (tb < len) kb = __ldg(reinterpret_cast<const float2*>(p.keys + ((size_t)b * p.T + t_off + tb) * ATT + d0));
                float acc[4];
                att3_loc_tile(acc, t0, lane, ghi, glo);
                preR[i][0] = acc[0] + ka.x + bc0;
                preR[i][1] = acc[1] + ka.y + bc1;
                preR[i][2] = acc[2] + kb.x + bc0;
                preR[i][3] = acc[3] + kb.y + bc1;
            }
        }
    } else {
        for (int t0 = 0; t0 < len; t0 += 16) {
            const int ta = t0 + gq, tb = ta + 8;
            float2 ka = make_float2(0.f, 0.f), kb = make_float2(0.f, 0.f);
            if (ta < len) ka = __ldg(reinterpret_cast<const float2*>(p.keys + ((size_t)b * p.T + ta) * ATT + d0));
            if (tb < len) kb = __ldg(reinterpret_cast<const float2*>(p.keys + ((size_t)b * p.T + tb) * ATT + d0));
            float acc[4];
            att3_loc_tile(acc, t0, lane, ghi, glo);
            if (ta < len) *reinterpret_cast<float2*>(pre_scr + (size_t)ta * ATT + d0) = make_float2(acc[0] + ka.x + bc0, acc[1] + ka.y + bc1);
            if (tb < len) *reinterpret_cast<float2*>(pre_scr + (size_t)tb * ATT + d0) = make_float2(acc[2] + kb.x + bc0, acc[3] + kb.y + bc1);
        }
    }
}

// thread -> (row, position slot, 16-byte column chunk) of the column-sliced context reduction (att3_context_slice), packed:
// bits 0..3 cluster row, 4..11 first position (slot), 12..19 number of slots of the row (0: no work), 20..31 row length;
// the thread takes positions slot, slot + nslot, ... of columns [48 rank + 8 chunk, + 8), chunk = tid % 6, group = tid / 6
constexpr int CTX_GROUPS = 85;   // 85 groups x 6 chunks of 8 columns = 510 threads
constexpr int CTX_BASE = CTX_GROUPS - CS;   // groups dealt in proportion to the row lengths; + 1 per row

// groups of row r = floor(CTX_BASE * len_r / sum(len)) + 1  (sum <= CTX_GROUPS for <= 16 rows)
__device__ __forceinline__ int ctx_groups_of(int len, int total) { return (CTX_BASE * len) / total + 1; }

// Owner part of one attention step of cluster row `arow` (global row b; the CTA that runs it is chosen by row length, SO_OWN):
// q from the 16 partials, energies, masked softmax, cumulate, alignment history, and the multicast of the alignment row to
// every CTA of the cluster (each of which reduces a 48-column slice of the context, att3_context_slice).
__device__ __forceinline__ void att3_owner(const Dec3Params& p, float (&preR)[NPT][4], float* pre_scr, int b, int len, int arow, int s,
                                           int tid, int warp, int lane, Guard3& g, long long* tacc, long long& tprev, bool timing, bool tracing,
                                           uint32_t q_bytes, int apart, int tb0, int Tl) {
    float* scr = SMF(SO_RED);
#ifdef TACO_DEC3_TRACE_BUILD
#define A3_TR(i)                                                        \
    if (tracing && tid == 0) {                                          \
        long long gt_;                                                  \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));         \
        p.trace[((size_t)blockIdx.x * 64 + (i)) * 2] = clock64();       \
        p.trace[((size_t)blockIdx.x * 64 + (i)) * 2 + 1] = gt_;         \
    }
#else
#define A3_TR(i)
#endif
#define A3_STAMP(i)                                                     \
    if (timing) { /* whole warp 0, uniform: a divergent stamp would push the following collectives onto their slow path */ \
        long long now_;                                                 \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));        \
        if (lane == 0) tacc[i] += now_ - tprev;                         \
        tprev = now_;                                                   \
    }
    // E1: q[d] = sum over the 16 source CTAs of their K-slice partial (query_layer, Attention.py:187-189)
    {
        // 4 threads per d, 4 sources each
        const int d = tid >> 2, part = tid & 3;
        mbar_wait(MB_Q, (uint32_t)(s & 1), g);
        if (tid == 0 && !g.dead) mbar_rearm(MB_Q, q_bytes);
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) q += SMF(SO_QRECV)[(part * 4 + i) * ATT + d];
        q += __shfl_xor_sync(0xffffffffu, q, 1);
        q += __shfl_xor_sync(0xffffffffu, q, 2);
        if (part == 0) SMF(SO_QS)[d] = g.dead ? 0.f : q;
    }
    A3_TR(34)
    __syncthreads();
    A3_STAMP(8)
    A3_TR(35)
    // E2: energies e[t] = sum_d v_a[d] tanh(pre[t][d] + q[d])   (Attention.py:401)
    if (p.pre_reg) {
        const int gq = lane >> 2, tq = lane & 3, d0 = warp * 8 + 2 * tq;
        const float q0 = SMF(SO_QS)[d0], q1 = SMF(SO_QS)[d0 + 1], v0 = SMF(SO_VA)[d0], v1 = SMF(SO_VA)[d0 + 1];
#pragma unroll
        for (int i = 0; i < NPT; ++i) {
            if (16 * i < len) {
                float ea = v0 * tanh_apx(preR[i][0] + q0) + v1 * tanh_apx(preR[i][1] + q1);
                float eb = v0 * tanh_apx(preR[i][2] + q0) + v1 * tanh_apx(preR[i][3] + q1);
                ea += __shfl_xor_sync(0xffffffffu, ea, 1);
                eb += __shfl_xor_sync(0xffffffffu, eb, 1);
                ea += __shfl_xor_sync(0xffffffffu, ea, 2);
                eb += __shfl_xor_sync(0xffffffffu, eb, 2);
                if (tq == 0) {
                    scr[warp * PRE_REG_T + 16 * i + gq] = ea;
                    scr[warp * PRE_REG_T + 16 * i + gq + 8] = eb;
                }
            }
        }
    } else {
        const int tl = lane & 7, qd = lane >> 3;
        const float4* q4 = reinterpret_cast<const float4*>(SMF(SO_QS) + qd * 32);
        const float4* v4 = reinterpret_cast<const float4*>(SMF(SO_VA) + qd * 32);
        for (int tb = 0; tb < len; tb += 128) {
            const int t = tb + warp * 8 + tl;
            float e = 0.f;
            if (t < len) {
                const float4* p4 = reinterpret_cast<const float4*>(pre_scr + (size_t)t * ATT + qd * 32);
#pragma unroll
                for (int i = 0; i < NPT; ++i) {
                    const float4 x = p4[i], qq = q4[i], vv = v4[i];
                    e = fmaf(vv.x, tanh_apx(x.x + qq.x), e);
                    e = fmaf(vv.y, tanh_apx(x.y + qq.y), e);
                    e = fmaf(vv.z, tanh_apx(x.z + qq.z), e);
                    e = fmaf(vv.w, tanh_apx(x.w + qq.w), e);
                }
            }
            e += __shfl_xor_sync(0xffffffffu, e, 8);
            e += __shfl_xor_sync(0xffffffffu, e, 16);
            if (qd == 0 && t < len) SMF(SO_ES)[t] = e;
        }
    }
    __syncthreads();
    A3_STAMP(16)
    A3_TR(36)
    // E3: masked softmax over t < len (exact zeros beyond: -inf mask), Attention.py:203-206
    if (p.pre_reg) {
        // all 512 threads: thread (position t = tid >> 2, quarter = tid & 3) adds 4 of the 16 per-warp partials of e[t]; a warp
        // covers 8 positions, so the row maximum and the normaliser are two warp reductions + one 16-entry exchange each
        float* wred = SMF(SO_QS);   // q has been consumed by E2: [0,16) warp maxima, [16,32) warp sums
        const int t = tid >> 2, part = tid & 3;
        float x = 0.f;
        if (t < len) {
#pragma unroll
            for (int i = 0; i < 4; ++i) x += scr[(part * 4 + i) * PRE_REG_T + t];
        }
        x += __shfl_xor_sync(0xffffffffu, x, 1);
        x += __shfl_xor_sync(0xffffffffu, x, 2);
        const float e = t < len ? x : -INFINITY;
        float m = 0.f;
        if (!p.smoothing) {
            m = e;
            m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
            m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 8));
            m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 16));
            if (lane == 0) wred[warp] = m;
            __syncthreads();
            const float4 a0 = *reinterpret_cast<const float4*>(wred), a1 = *reinterpret_cast<const float4*>(wred + 4);
            const float4 a2 = *reinterpret_cast<const float4*>(wred + 8), a3 = *reinterpret_cast<const float4*>(wred + 12);
            m = fmaxf(fmaxf(fmaxf(fmaxf(a0.x, a0.y), fmaxf(a0.z, a0.w)), fmaxf(fmaxf(a1.x, a1.y), fmaxf(a1.z, a1.w))),
                      fmaxf(fmaxf(fmaxf(a2.x, a2.y), fmaxf(a2.z, a2.w)), fmaxf(fmaxf(a3.x, a3.y), fmaxf(a3.z, a3.w))));
        }
        const float v = t < len ? (p.smoothing ? sigmoidf_(e) : __expf(e - m)) : 0.f;
        float sm = v;
        sm += __shfl_xor_sync(0xffffffffu, sm, 4);
        sm += __shfl_xor_sync(0xffffffffu, sm, 8);
        sm += __shfl_xor_sync(0xffffffffu, sm, 16);
        if (lane == 0) wred[16 + warp] = sm;
        __syncthreads();
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 a = *reinterpret_cast<const float4*>(wred + 16 + 4 * i);
            sum += (a.x + a.y) + (a.z + a.w);
        }
        float inv = __fdividef(1.0f, sum);
        if (p.parts > 1) {
            // position slices of one row: every slice sends (its maximum, its sum of exp) to the others; the row's softmax is
            // exp(e - m_j) * exp(m_j - M) / sum_j s_j exp(m_j - M), M = max_j m_j
            const int P = p.parts;
            if (tid == 0) {
                SMF(SO_PART)[2 * apart] = m;
                SMF(SO_PART)[2 * apart + 1] = sum;
            }
            if (tid < P && tid != apart) {
                const uint32_t peer = SMU(SO_OWN)[arow * P + tid];
                asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1,%2}, [%3];"
                             ::"r"(mapa_rank(smem_addr(SMU(SO_PART) + 2 * apart), peer)), "r"(__float_as_uint(m)), "r"(__float_as_uint(sum)),
                               "r"(mapa_rank(smem_addr(SMU(SO_MBAR) + 2 * MB_PART), peer)) : "memory");
            }
            mbar_wait(MB_PART, (uint32_t)(s & 1), g);
            __syncthreads();   // thread 0's own slot
            float M = -INFINITY;
            for (int j = 0; j < P; ++j) M = fmaxf(M, SMF(SO_PART)[2 * j]);
            float Z = 0.f;
            for (int j = 0; j < P; ++j) Z += SMF(SO_PART)[2 * j + 1] * __expf(SMF(SO_PART)[2 * j] - M);
            inv = __expf(m - M) * __fdividef(1.0f, Z);   // an empty slice has m = -inf: factor 0
            __syncthreads();   // every thread has read the slots before thread 0 re-arms the exchange for the next step
            if (tid == 0 && !g.dead) mbar_rearm(MB_PART, (uint32_t)(P - 1) * 8u);
        }
        if (part == 0) SMF(SO_ES)[t] = v * inv;   // t < PRE_REG_T always; exact zeros beyond the length
    } else if (warp == 0) {
        float m = -INFINITY;
        if (!p.smoothing) {
            for (int t = lane; t < len; t += 32) m = fmaxf(m, SMF(SO_ES)[t]);
            m = warp_max(m);
        }
        float sum = 0.f;
        for (int t = lane; t < len; t += 32) {
            const float v = p.smoothing ? sigmoidf_(SMF(SO_ES)[t]) : __expf(SMF(SO_ES)[t] - m);
            SMF(SO_ES)[t] = v;
            sum += v;
        }
        sum = warp_sum(sum);
        const float inv = 1.0f / sum;
        for (int t = lane; t < len; t += 32) SMF(SO_ES)[t] *= inv;
    }
    __syncthreads();
    A3_STAMP(10)
    A3_TR(37)
    // multicast the alignment row to every CTA of the cluster (positions < len; the context slices never read beyond)
    {
        const int nv = (len + 3) >> 2;
        const int TP = p.TP;
        for (int i = tid; i < CS * nv; i += NT) {
            const int slot = i / nv, v = i - slot * nv;
            const int dst = (slot + arow) & (CS - 1);
            const uint4 val = *reinterpret_cast<const uint4*>(SMU(SO_ES) + (v << 2));
            st_async_v4(mapa_rank(smem_addr(SMU(SO_AALL) + arow * TP + tb0 + (v << 2)), dst), val, mapa_rank(smem_addr(SMU(SO_MBAR) + 2 * MB_A), dst));
        }
    }
    A3_TR(38)
    // E4: cumulate + alignment history
    for (int t = tid; t < Tl; t += NT) {
        const float a = t < len ? SMF(SO_ES)[t] : 0.f;
        SMF(SO_CPAD)[15 + t] = p.cumulative ? SMF(SO_CPAD)[15 + t] + a : a;
        p.hist[((size_t)s * p.B + b) * p.T + tb0 + t] = a;
    }
    if (!p.cumulative) __syncthreads();
    // pair arrays of the new alignments for the location term of the next step (att3_loc_chunk, in the shadow of later exchanges);
    // position slices build them after the alignment exchange, from the complete row (att3_build_pairs_row)
    if (p.parts == 1) att3_build_pairs(Tl, tid, p.cumulative != 0, len);
    A3_STAMP(15)
    A3_TR(39)
}

// Context of ALL rows of the cluster, columns [48 rank, 48 rank + 48): ctx[r][c] = sum_t a[r][t] values[r][t][c]
// (_compute_attention, Attention.py:345-366; bf16 values, fp32 accumulate).  Every CTA reduces its slice from the alignment rows
// the owners multicast, then multicasts its three k-step blocks (B-fragment order) to the whole cluster: the reduction is spread
// over 16 SMs' L2 ports instead of one per row, and nobody waits for the CTA that owns the longest row.
#ifndef D3_UB
#define D3_UB 5
#endif
__device__ __forceinline__ void att3_context_slice(const Dec3Params& p, uint32_t cmw, int row0, int R, int rank, int s, int tid,
                                                   Guard3& g, bool tracing, uint32_t a_bytes) {
    float* scr = SMF(SO_RED);
    const int TP = p.TP;
    __syncthreads();   // the K-split partials / query slices of this step (same scratch) have been read
    A3_TR(40)
    A3_TR(41)
    const int nslot = (cmw >> 12) & 255;
    if (nslot) {
        const int crow = cmw & 15, cslot = (cmw >> 4) & 255, len = cmw >> 20;
        const int b = row0 + crow;
        const int gidx = tid / 6, chunk = tid - gidx * 6;
        const uint4* vp = reinterpret_cast<const uint4*>(p.values) + (size_t)b * p.T * (MEM / 8) + 6 * rank + chunk;
        const uint32_t* al = SMU(SO_AALL) + crow * TP;
        float c[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        constexpr int UB = D3_UB;   // value rows in flight per thread and round (8 measured slower: registers)
        for (int t0 = cslot; t0 < len; t0 += nslot * UB) {
            uint4 v[UB];
            float a[UB];
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const int t = t0 + u * nslot;
                v[u] = (t < len && !(p.dbg & 32)) ? ldg_stream_v4(vp + (size_t)t * (MEM / 8)) : make_uint4(0u, 0u, 0u, 0u);
            }
            if (t0 == cslot) mbar_wait(MB_A, (uint32_t)(s & 1), g);   // after the value loads of the first batch are in flight
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const int t = t0 + u * nslot;
                a[u] = t < len ? __uint_as_float(al[t]) : 0.f;
            }
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                c[0] = fmaf(a[u], bf_lo(v[u].x), c[0]);
                c[1] = fmaf(a[u], bf_hi(v[u].x), c[1]);
                c[2] = fmaf(a[u], bf_lo(v[u].y), c[2]);
                c[3] = fmaf(a[u], bf_hi(v[u].y), c[3]);
                c[4] = fmaf(a[u], bf_lo(v[u].z), c[4]);
                c[5] = fmaf(a[u], bf_hi(v[u].z), c[5]);
                c[6] = fmaf(a[u], bf_lo(v[u].w), c[6]);
                c[7] = fmaf(a[u], bf_hi(v[u].w), c[7]);
            }
        }
        float4* dst = reinterpret_cast<float4*>(scr + gidx * 48 + chunk * 8);
        dst[0] = make_float4(c[0], c[1], c[2], c[3]);
        dst[1] = make_float4(c[4], c[5], c[6], c[7]);
    }
    A3_TR(42)
    g.dead = __syncthreads_or(g.dead ? 1 : 0) != 0;
    A3_TR(43)
    if (tid == 0 && !g.dead) {   // (thread 0 has work whenever the cluster has a row: its wait above has returned)
        mbar_wait(MB_A, (uint32_t)(s & 1), g);
        mbar_rearm(MB_A, a_bytes);
    }
    if (tid < R * 24) {
        const int r = tid / 24, cp = tid - r * 24;
        const int g0 = (int)SMU(SO_GRP)[2 * r], ng = (int)SMU(SO_GRP)[2 * r + 1];   // groups of row r: [g0, g0 + ng)
        float x0 = 0.f, x1 = 0.f;
        for (int q = 0; q < ng; ++q) {
            const float2 v = *reinterpret_cast<const float2*>(scr + (g0 + q) * 48 + 2 * cp);
            x0 += v.x;
            x1 += v.y;
        }
        const int b = row0 + r;
        const uint32_t cw = pack_bf16x2(x0, x1);
        const int k = cp >> 3, j = cp & 7;
        SMU(SO_STG_C)[k * TILE_W + bword(r, j)] = cw;
        *reinterpret_cast<float2*>(p.ctx_state + (size_t)b * MEM + 48 * rank + 2 * cp) = make_float2(x0, x1);
        p.act_hist[((size_t)s * p.B + b) * 512 + 128 + 24 * rank + cp] = cw;   // projection input row of this step, columns 256..1023
    }
    A3_TR(44)
    __syncthreads();
    A3_TR(45)
    // multicast this CTA's three k-step blocks (3 rank .. 3 rank + 2) to every CTA of the cluster
    {
        const int dst = tid >> 5, v = tid & 31;
        const int row = ((v >> 4) << 3) + ((v & 15) >> 1);
        if (row < R) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const uint4 val = *reinterpret_cast<const uint4*>(SMU(SO_STG_C) + k * TILE_W + (v << 2));
                st_async_v4(mapa_rank(smem_addr(SMU(SO_CTX) + (3 * rank + k) * TILE_W + (v << 2)), dst), val, mapa_rank(smem_addr(SMU(SO_MBAR) + 2 * MB_CTX), dst));
            }
        }
    }
    A3_TR(46)
}

// Location pre-activations of the NEXT step from the new alignments (pair arrays built by att3_owner), tiles [i0, i1) of the
// in-register path; the scratch path (T > 128) runs all its tiles in the first call.  Warp-local: no block synchronisation,
// so the chunks can sit in the shadow of the dense_2 / LSTM-1 exchanges.
__device__ __forceinline__ void att3_loc_chunk(const Dec3Params& p, float (&preR)[NPT][4], float* pre_scr, int b, int len, int warp, int lane,
                                               int i0, int i1) {
    if (!p.cumulative) {   // not incremental: pre = keys + location(a) + const, every step
        if (i0 == 0) att3_location(p, preR, pre_scr, b, len, warp, lane, false);
        return;
    }
    if (!p.pre_reg) {
        if (i0 == 0) att3_location(p, preR, pre_scr, b, len, warp, lane, true);
        return;
    }
    uint2 ghi[2], glo[2];
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
        ghi[ks] = __ldg(reinterpret_cast<const uint2*>(p.g_hi) + (ks * 16 + warp) * 32 + lane);
        glo[ks] = __ldg(reinterpret_cast<const uint2*>(p.g_lo) + (ks * 16 + warp) * 32 + lane);
    }
#pragma unroll
    for (int i = 0; i < NPT; ++i) {
        if (i >= i0 && i < i1 && 16 * i < len) {
            float acc[4];
            att3_loc_tile(acc, 16 * i, lane, ghi, glo);
            preR[i][0] += acc[0];
            preR[i][1] += acc[1];
            preR[i][2] += acc[2];
            preR[i][3] += acc[3];
        }
    }
}

}  // namespace
