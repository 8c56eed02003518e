// Body of decoder2.cu (kept in a separate file only for size).  See decoder2.cu for the design.
namespace {

// Shared-memory map: constant word offsets from the dynamic-smem base so that no pointer lives in a
// register (the kernel is register-bound: 512 threads x 128 registers).
extern __shared__ __align__(16) uint32_t smem2[];
constexpr int RED_FLOATS = 2560;
constexpr int PW = (T_MAX2 + 48) / 2;
constexpr int SO_BLOB = 0;
constexpr int SO_WQ = SO_BLOB + BLOB_WORDS;          // [256][64] bf16 pairs
constexpr int SO_RED = SO_WQ + 256 * 64;             // [2][2560] floats (aliased as attention scratch)
constexpr int SO_ES = SO_RED + 2 * RED_FLOATS;       // [T_MAX2]
constexpr int SO_CPAD = SO_ES + T_MAX2;              // [T_MAX2 + 48]  cpad[i] = cum[i - 15]
constexpr int SO_P = SO_CPAD + T_MAX2 + 48;          // 4 x [PW] pair arrays P0hi, P1hi, P0lo, P1lo
constexpr int SO_HQ = SO_P + 4 * PW;                 // [256]
constexpr int SO_QS = SO_HQ + 256;                   // [128]
constexpr int SO_VA = SO_QS + 128;                   // [128]
constexpr int SO_BC = SO_VA + 128;                   // [128]
constexpr int SO_FLAG = SO_BC + 128;                 // [4]
constexpr int SO_TACC = SO_FLAG + 4;                 // 16 x int64 stage timers
constexpr int SO_PRE = SO_TACC + 32;                 // [T][PS] when p.pre_in_smem
constexpr size_t SMEM2_FIXED = (size_t)SO_PRE * 4;
static_assert(SO_WQ % 4 == 0 && SO_RED % 4 == 0 && SO_QS % 4 == 0 && SO_VA % 4 == 0 && SO_PRE % 4 == 0 && SO_TACC % 2 == 0, "smem alignment");
#define SM_U32(off) (smem2 + (off))
#define SM_F32(off) (reinterpret_cast<float*>(smem2 + (off)))

// pre[t][d] = keys[b][t][d] + b_a[d] + conv_bias.Wl[d] + sum_k cum[t+k-15] G[k][d]   for t < len
// (Attention.py:193-197 fused: conv1d(k=31, 32 filters) followed by Dense(128, no bias)).
// Warp w owns the 8 columns d = 8w..8w+7; the Toeplitz A fragments come from the pair arrays.
__device__ __forceinline__ void att_location(const Dec2Params& p, float* pre, int pstride, int b, int len,
                                             int warp, int lane) {
    const int g = lane >> 2, t = lane & 3;
    uint2 ghi[2], glo[2];
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
        ghi[ks] = __ldg(reinterpret_cast<const uint2*>(p.g_hi) + (ks * 16 + warp) * 32 + lane);
        glo[ks] = __ldg(reinterpret_cast<const uint2*>(p.g_lo) + (ks * 16 + warp) * 32 + lane);
    }
    const uint32_t* P0h = SM_U32(SO_P);
    const uint32_t* P1h = SM_U32(SO_P + PW);
    const uint32_t* P0l = SM_U32(SO_P + 2 * PW);
    const uint32_t* P1l = SM_U32(SO_P + 3 * PW);
    const int d0 = warp * 8 + 2 * t;
    const float bc0 = SM_F32(SO_BC)[d0], bc1 = SM_F32(SO_BC)[d0 + 1];
    for (int t0 = 0; t0 < len; t0 += 16) {
        const int ta = t0 + g, tb = t0 + g + 8;
        float2 ka = make_float2(0.f, 0.f), kb = make_float2(0.f, 0.f);
        if (ta < len) ka = __ldg(reinterpret_cast<const float2*>(p.keys + ((size_t)b * p.T + ta) * ATT + d0));
        if (tb < len) kb = __ldg(reinterpret_cast<const float2*>(p.keys + ((size_t)b * p.T + tb) * ATT + d0));
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
            uint4 ah, al;
            // element (row, k) = cpad[t0 + row + k]; pair (k, k+1) at index i = t0 + row + k: even -> P0[i/2], odd -> P1[(i-1)/2]
            const int ia = ta + ks * 16 + 2 * t, ib = tb + ks * 16 + 2 * t;
            const bool oa = ia & 1, ob = ib & 1;
            ah.x = oa ? P1h[ia >> 1] : P0h[ia >> 1];
            ah.y = ob ? P1h[ib >> 1] : P0h[ib >> 1];
            ah.z = oa ? P1h[(ia + 8) >> 1] : P0h[(ia + 8) >> 1];
            ah.w = ob ? P1h[(ib + 8) >> 1] : P0h[(ib + 8) >> 1];
            al.x = oa ? P1l[ia >> 1] : P0l[ia >> 1];
            al.y = ob ? P1l[ib >> 1] : P0l[ib >> 1];
            al.z = oa ? P1l[(ia + 8) >> 1] : P0l[(ia + 8) >> 1];
            al.w = ob ? P1l[(ib + 8) >> 1] : P0l[(ib + 8) >> 1];
            mma_bf16(acc, ah, ghi[ks]);
            mma_bf16(acc, ah, glo[ks]);
            mma_bf16(acc, al, ghi[ks]);
        }
        if (ta < len) *reinterpret_cast<float2*>(pre + (size_t)ta * pstride + d0) = make_float2(acc[0] + ka.x + bc0, acc[1] + ka.y + bc1);
        if (tb < len) *reinterpret_cast<float2*>(pre + (size_t)tb * pstride + d0) = make_float2(acc[2] + kb.x + bc0, acc[3] + kb.y + bc1);
    }
}

// pair arrays of the hi/lo bf16 split of the zero-padded cumulative alignments
__device__ __forceinline__ void att_build_pairs(int T, int tid) {
    const float* cpad = SM_F32(SO_CPAD);
    uint32_t* P = SM_U32(SO_P);
    const int n = (T + 46) / 2 + 1;
    for (int i = tid; i < n && i < PW; i += NT) {
        const float c0 = cpad[2 * i], c1 = cpad[2 * i + 1], c2 = (2 * i + 2 < T_MAX2 + 48) ? cpad[2 * i + 2] : 0.f;
        const float h0 = __bfloat162float(__float2bfloat16_rn(c0));
        const float h1 = __bfloat162float(__float2bfloat16_rn(c1));
        const float h2 = __bfloat162float(__float2bfloat16_rn(c2));
        P[i] = pack_bf16(h0, h1);
        P[PW + i] = pack_bf16(h1, h2);
        P[2 * PW + i] = pack_bf16(c0 - h0, c1 - h1);
        P[3 * PW + i] = pack_bf16(c1 - h1, c2 - h2);
    }
}

__global__ void __launch_bounds__(NT, 1) decoder2_kernel(const __grid_constant__ Dec2Params p) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, cta = blockIdx.x;
    const int MT = p.MT, KG = p.KG, RS = 16 * MT + 4, rows = 16 * MT;
    const int mt = warp % MT, kg = warp / MT;
    Guard g{p.abort_flag, false, p.dbg};

    // ---- one-time loads: this CTA's weight fragments, the shared query kernel, attention vectors ----
    {
        const uint4* src = reinterpret_cast<const uint4*>(p.blobs + (size_t)cta * BLOB_WORDS);
        uint4* dst = reinterpret_cast<uint4*>(SM_U32(SO_BLOB));
        for (int i = tid; i < BLOB_WORDS / 4; i += NT) dst[i] = src[i];
        const uint4* qsrc = reinterpret_cast<const uint4*>(p.wq);
        uint4* qdst = reinterpret_cast<uint4*>(SM_U32(SO_WQ));
        for (int i = tid; i < 256 * 64 / 4; i += NT) qdst[i] = qsrc[i];
        if (tid < ATT) {
            SM_F32(SO_VA)[tid] = p.va[tid];
            SM_F32(SO_BC)[tid] = p.bconst[tid];
        }
        for (int i = tid; i < T_MAX2 + 48; i += NT) SM_F32(SO_CPAD)[i] = 0.f;
        if (tid < 4 + 32) SM_U32(SO_FLAG)[tid] = 0;
    }
#define w1 (reinterpret_cast<const uint2*>(smem2 + SO_BLOB + OFF_W1))
#define w2 (reinterpret_cast<const uint2*>(smem2 + SO_BLOB + OFF_W2))
#define wp1 (reinterpret_cast<const uint2*>(smem2 + SO_BLOB + OFF_WP1))
#define wp2 (reinterpret_cast<const uint2*>(smem2 + SO_BLOB + OFF_WP2))
#define wo (reinterpret_cast<const uint2*>(smem2 + SO_BLOB + OFF_WO))
#define bias (reinterpret_cast<const float*>(smem2 + SO_BLOB + OFF_BIAS))  /* b1[8] b2[8] bp1[2] bp2[2] bo[2] */

    // attention work item of this CTA
    const bool att = cta < p.B * p.NS;
    const int ab = att ? cta / p.NS : 0, asl = att ? cta % p.NS : 0;
    const int alen = att ? p.lens[ab] : 0;
    const int DS = MEM / p.NS;
    float* pre = p.pre_in_smem ? SM_F32(SO_PRE) : p.pre_scratch + (size_t)cta * p.T * ATT;
    const int pstride = p.pre_in_smem ? PS : ATT;

    // epilogue threads: (batch row er = tid / 8, column en = tid % 8 of the CTA's 8 gate columns); lanes en < 2
    // additionally own unit 2*cta + en of row er: private fp32 recurrent state in registers
    const bool ept = tid < rows * 8;          // whole warps (rows is a multiple of 16)
    const int er = tid >> 3, en = tid & 7;
    const bool evalid = ept && er < p.B;
    const bool eunit = ept && en < 2;
    float c1s = 0.f, h1s = 0.f, c2s = 0.f, h2s = 0.f;
    if (evalid && eunit) {
        c1s = p.c1[(size_t)er * DEC + 2 * cta + en];
        h1s = p.h1z[(size_t)er * DEC + 2 * cta + en];
        c2s = p.c2[(size_t)er * DEC + 2 * cta + en];
        h2s = p.h2z[(size_t)er * DEC + 2 * cta + en];
    }
    __syncthreads();
    if (att) {
        for (int t = tid; t < p.T; t += NT) SM_F32(SO_CPAD)[15 + t] = p.cum[(size_t)ab * p.T + t];
        __syncthreads();
        att_build_pairs(p.T, tid);
        __syncthreads();
        att_location(p, pre, pstride, ab, alen, warp, lane);
    }
    const int SZ_X = MT * KS_X * 128, SZ_H = MT * KS_H * 128, SZ_C = MT * KS_CTX * 128;
    const float zo = p.zoneout;
    const bool free_run = p.mode == TACO_MODE_FREE;

    // contributions of step t0-1's context / zoned states to the gates of step t0 (from the state slot)
    float acc1[4] = {0.f, 0.f, 0.f, 0.f}, acc2[4] = {0.f, 0.f, 0.f, 0.f};
    {
        const int sp = (p.t0 + 3) & 3;
        mma_seg<4>(acc1, p.ring_ctx + sp * SZ_C, KS_CTX, mt, kg, KG, w1 + 16 * 32, lane, g);
        mma_seg<4>(acc1, p.ring_h1z + sp * SZ_H, KS_H, mt, kg, KG, w1 + 64 * 32, lane, g);
        mma_seg<4>(acc2, p.ring_h2z + sp * SZ_H, KS_H, mt, kg, KG, w2 + 16 * 32, lane, g);
    }
    __syncthreads();

    int s_done = 0, st = 0;
    long long* tacc = reinterpret_cast<long long*>(smem2 + SO_TACC);
    long long tprev = 0;
    const bool timing = p.timing != nullptr && tid == 0;
#define DEC2_STAMP(i)                                                   \
    if (timing) {                                                       \
        long long now_;                                                 \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));        \
        tacc[i] += now_ - tprev;                                        \
        tprev = now_;                                                   \
    }
// sum of the KG partial accumulators of (row er, column en) + bias
#define DEC2_REDSUM(z, redb, b0)                                        \
    float z = bias[(b0) + en];                                          \
    for (int k_ = 0; k_ < KG; ++k_) z += (redb)[(k_ * 8 + en) * RS + er];
    if (timing) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tprev));
    for (int s = 0; s < p.S_max; ++s) {
        const int step = p.t0 + s;
        const int cur = step & 3, clr = (step + 2) & 3;
        // ================= A: prenet dense_1 (Decoder.py:346-351) =================
        {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            if (free_run) {
                mma_seg<2>(acc, p.ring_x + cur * SZ_X, KS_X, mt, kg, KG, wp1, lane, g);
                if (s > 0 && tid == 0) {
                    int v, spins = 0;
                    while ((v = (int)ld_word(reinterpret_cast<const uint32_t*>(p.flags + (s - 1)))) == 0 && !spin_check(g, spins)) {
                    }
                    SM_U32(SO_FLAG)[0] = (uint32_t)v;
                }
            } else {
                // TacoTrainingHelper.next_inputs (Helpers.py:121-124, ratio 1.0): x(step) = targets[:, r*step - 1]
                const int fidx = step * 2 - 1;
                const int gq = lane >> 2, tq = lane & 3;
                for (int ks = kg; ks < KS_X; ks += KG) {
                    uint4 a = make_uint4(0u, 0u, 0u, 0u);
                    if (fidx >= 0 && fidx < p.F_tgt) {
                        const int ra = mt * 16 + gq, rb = ra + 8, c0 = ks * 16 + 2 * tq;
                        if (ra < p.B) {
                            const float* src = p.targets + ((size_t)ra * p.F_tgt + fidx) * MEL + c0;
                            const float2 v0 = __ldg(reinterpret_cast<const float2*>(src));
                            const float2 v1 = __ldg(reinterpret_cast<const float2*>(src + 8));
                            a.x = pack_bf16(v0.x, v0.y);
                            a.z = pack_bf16(v1.x, v1.y);
                        }
                        if (rb < p.B) {
                            const float* src = p.targets + ((size_t)rb * p.F_tgt + fidx) * MEL + c0;
                            const float2 v0 = __ldg(reinterpret_cast<const float2*>(src));
                            const float2 v1 = __ldg(reinterpret_cast<const float2*>(src + 8));
                            a.y = pack_bf16(v0.x, v0.y);
                            a.w = pack_bf16(v1.x, v1.y);
                        }
                    }
                    mma_bf16(acc, a, wp1[ks * 32 + lane]);
                }
            }
            DEC2_STAMP(0)
            float* redb = SM_F32(SO_RED) + (st++ & 1) * RED_FLOATS;
            store_red(redb, acc, RS, mt, kg, lane);
            __syncthreads();
            if (free_run && s > 0 && SM_U32(SO_FLAG)[0] == 2u) break;  // batch-global stop (Helpers.py:51-52), CTA-uniform
            if (ept) {
                DEC2_REDSUM(z, redb, 16)
                z = fmaxf(z, 0.f);
                z = (evalid && eunit && keep_bit2(p, step, er, 0, 2 * cta + en)) ? z * p.keep_scale : 0.f;
                const float zn = __shfl_down_sync(0xffffffffu, z, 1);
                if (en == 0) {
                    const int idx = frag_word(KS_H, er, 2 * cta);
                    st_word(p.ring_p1 + cur * SZ_H + idx, pack_bf16(z, zn));
                    st_word(p.ring_p1 + clr * SZ_H + idx, SENT);
                    DEC2_FENCE();
                }
            }
            DEC2_STAMP(1)
        }
        // ================= B: prenet dense_2 =================
        {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            mma_seg<4>(acc, p.ring_p1 + cur * SZ_H, KS_H, mt, kg, KG, wp2, lane, g);
            DEC2_STAMP(2)
            float* redb = SM_F32(SO_RED) + (st++ & 1) * RED_FLOATS;
            store_red(redb, acc, RS, mt, kg, lane);
            __syncthreads();
            if (ept) {
                DEC2_REDSUM(z, redb, 18)
                z = fmaxf(z, 0.f);
                z = (evalid && eunit && keep_bit2(p, step, er, 1, 2 * cta + en)) ? z * p.keep_scale : 0.f;
                const float zn = __shfl_down_sync(0xffffffffu, z, 1);
                if (en == 0) {
                    const int idx = frag_word(KS_H, er, 2 * cta);
                    st_word(p.ring_p2 + cur * SZ_H + idx, pack_bf16(z, zn));
                    st_word(p.ring_p2 + clr * SZ_H + idx, SENT);
                    DEC2_FENCE();
                }
            }
            DEC2_STAMP(3)
        }
        // ================= C: decoder LSTM layer 1 (Decoder.py:256; Encoder.py:118-142) =================
        {
            mma_seg<4>(acc1, p.ring_p2 + cur * SZ_H, KS_H, mt, kg, KG, w1, lane, g);
            DEC2_STAMP(4)
            float* redb = SM_F32(SO_RED) + (st++ & 1) * RED_FLOATS;
            store_red(redb, acc1, RS, mt, kg, lane);
            acc1[0] = acc1[1] = acc1[2] = acc1[3] = 0.f;
            __syncthreads();
            if (ept) {
                // column en = gate * 2 + unit: lanes en < 2 gather their unit's j, f, o pre-activations
                DEC2_REDSUM(zi, redb, 0)
                const float zj = __shfl_down_sync(0xffffffffu, zi, 2);
                const float zf = __shfl_down_sync(0xffffffffu, zi, 4);
                const float zo_ = __shfl_down_sync(0xffffffffu, zi, 6);
                const float c_new = sigmoid_fast(zf + 1.0f) * c1s + sigmoid_fast(zi) * tanh_fast(zj);
                const float h_new = sigmoid_fast(zo_) * tanh_fast(c_new);
                c1s = (1.f - zo) * c_new + zo * c1s;
                h1s = (1.f - zo) * h_new + zo * h1s;
                const float h_new1 = __shfl_down_sync(0xffffffffu, h_new, 1);
                const float h_z1 = __shfl_down_sync(0xffffffffu, h1s, 1);
                if (en == 0) {
                    const int idx = frag_word(KS_H, er, 2 * cta);
                    st_word(p.ring_h1raw + cur * SZ_H + idx, pack_bf16(h_new, h_new1));
                    st_word(p.ring_h1z + cur * SZ_H + idx, pack_bf16(h1s, h_z1));
                    st_word(p.ring_h1raw + clr * SZ_H + idx, SENT);
                    st_word(p.ring_h1z + clr * SZ_H + idx, SENT);
                    DEC2_FENCE();
                }
            }
            DEC2_STAMP(5)
        }
        // ================= D: decoder LSTM layer 2 =================
        {
            mma_seg<4>(acc2, p.ring_h1raw + cur * SZ_H, KS_H, mt, kg, KG, w2, lane, g);
            DEC2_STAMP(6)
            float* redb = SM_F32(SO_RED) + (st++ & 1) * RED_FLOATS;
            store_red(redb, acc2, RS, mt, kg, lane);
            acc2[0] = acc2[1] = acc2[2] = acc2[3] = 0.f;
            __syncthreads();
            if (ept) {
                DEC2_REDSUM(zi, redb, 8)
                const float zj = __shfl_down_sync(0xffffffffu, zi, 2);
                const float zf = __shfl_down_sync(0xffffffffu, zi, 4);
                const float zo_ = __shfl_down_sync(0xffffffffu, zi, 6);
                const float c_new = sigmoid_fast(zf + 1.0f) * c2s + sigmoid_fast(zi) * tanh_fast(zj);
                const float h_new = sigmoid_fast(zo_) * tanh_fast(c_new);
                c2s = (1.f - zo) * c_new + zo * c2s;
                h2s = (1.f - zo) * h_new + zo * h2s;
                const float h_new1 = __shfl_down_sync(0xffffffffu, h_new, 1);
                const float h_z1 = __shfl_down_sync(0xffffffffu, h2s, 1);
                if (en == 0) {
                    const int idx = frag_word(KS_H, er, 2 * cta);
                    st_word(p.ring_h2raw + cur * SZ_H + idx, pack_bf16(h_new, h_new1));
                    st_word(p.ring_h2z + cur * SZ_H + idx, pack_bf16(h2s, h_z1));
                    st_word(p.ring_h2raw + clr * SZ_H + idx, SENT);
                    st_word(p.ring_h2z + clr * SZ_H + idx, SENT);
                    DEC2_FENCE();
                }
            }
            DEC2_STAMP(7)
        }
#include "decoder2_attention.inl"
    }

    if (timing)
        for (int i = 0; i < 16; ++i) p.timing[(size_t)cta * 16 + i] = tacc[i];
    // ---- write back the recurrent state + step count ----
    if (evalid && eunit) {
        p.c1[(size_t)er * DEC + 2 * cta + en] = c1s;
        p.h1z[(size_t)er * DEC + 2 * cta + en] = h1s;
        p.c2[(size_t)er * DEC + 2 * cta + en] = c2s;
        p.h2z[(size_t)er * DEC + 2 * cta + en] = h2s;
    }
    if (att && asl == 0)
        for (int t = tid; t < p.T; t += NT) p.cum[(size_t)ab * p.T + t] = SM_F32(SO_CPAD)[15 + t];
    if (cta == 0 && tid == 0) *p.steps_out = (g.dead || *reinterpret_cast<volatile int*>(p.abort_flag) != 0) ? -1 : s_done;
}

}  // namespace
