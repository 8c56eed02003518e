// Body of decoder3.cu (kept in a separate file only for size).  See decoder3.cu for the design.
#include "decoder3_attention.inl"

namespace {

constexpr int STOP_LAG = 8;
constexpr int STOP_RANK = 10;   // cluster rank that computes the stop-token projection
constexpr int ACT_W = 1024;     // projection input row: [h2raw 256 | context 768] (Decoder.py:272)

#define D3_STAMP(i)                                                     \
    if (timing) { /* whole warp 0, uniform: a divergent stamp would push the following collectives onto their slow path */ \
        long long now_;                                                 \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now_));        \
        if (lane == 0) tacc[i] += now_ - tprev;                         \
        tprev = now_;                                                   \
    }

// one-step trace (build with -DTACO_DEC3_TRACE_BUILD, run with TACO_DEC3_TRACE=<step>): warp 0 lane 0 stamps (clock64, globaltimer)
// at point i of step trace_step.  Compiled out of the product build: the extra code costs 2 us per step even when it never runs.
#ifdef TACO_DEC3_TRACE_BUILD
#define D3_TR(i)                                                        \
    if (tracing && tid == 0) {                                          \
        long long gt_;                                                  \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));         \
        p.trace[((size_t)blockIdx.x * 64 + (i)) * 2] = clock64();       \
        p.trace[((size_t)blockIdx.x * 64 + (i)) * 2 + 1] = gt_;         \
    }
// wait (all threads) until every valid row of the blocks [buf, buf + KS blocks) has arrived, so that the stamp that follows
// separates "waiting for the exchange" from "computing"
#define D3_PREPOLL(buf, KS)
#else
#define D3_TR(i)
#define D3_PREPOLL(buf, KS)
#endif

// Stop-token projection of launch step s_idx for the rows of this cluster (StopProjection, Decoder.py:404-410; the batch-global
// decision of TacoTestHelper.next_inputs, Helpers.py:36-59): [h2raw ; context] . Ws[1024 x 2] as 64 k-steps of one m16 tile whose
// rows 0..1 are the two stop units.  Runs on ONE rank per cluster, off the critical path; its tiles stream from L2.
__device__ __noinline__ void dec3_stop_token(const Dec3Params& p, int s_idx, int row0, int R, int cid, bool free_run, float* red, int tid,
                                             int warp, int lane, Guard3& g) {
    float acc[2][4] = {};
    uint4 a[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = ldg_stream_v4(p.stop_tiles + (size_t)(warp + 16 * i) * 32 + lane);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const uint32_t* blk = i == 0 ? SMU(SO_H2R) + warp * TILE_W : SMU(SO_CTX) + (warp + 16 * (i - 1)) * TILE_W;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint2 b = load_b(blk, j, lane, g);
            mma16816(acc[j], a[i], b);
        }
    }
    store_red3(red, acc, warp, lane, 2);
    __syncthreads();
    bool fin = false;
    if (tid < 256) {
        const int eu = tid & 15, er = tid >> 4;
        if (eu < 2 && er < R) {
            float y = p.stop_bias[eu];
#pragma unroll
            for (int k = 0; k < 16; ++k) y += red[(k * 16 + eu) * 17 + er];
            const float sg = sigmoidf_(y);
            p.stop[(size_t)(row0 + er) * p.S_cap * 2 + (size_t)s_idx * 2 + eu] = p.stop_sigmoid ? sg : y;
            fin = sg > 0.5f;   // tf.round(sigmoid) == 1 <=> sigmoid > 0.5 (half-to-even; Helpers.py:40)
        }
    }
    const int cnt = __syncthreads_count(fin ? 1 : 0);
    if (free_run && tid == 0)
        stg_relaxed(reinterpret_cast<uint32_t*>(p.flags) + (size_t)s_idx * p.NCL + cid, (p.stop_at_any ? cnt > 0 : cnt == 2 * R) ? 2u : 1u);
}

// Prenet dropout bits (always on, Decoder.py:350-351) of `step` for this CTA's 16 units of both layers and every cluster row:
// ONE Philox block per (row, layer) -- thread i < 2R: row i >> 1, layer i & 1 -- instead of one per (row, layer, unit).
__device__ __forceinline__ void keep_bits_step(const Dec3Params& p, int step, int row0, int R, int rank, int tid) {
    if (p.masks || tid >= 2 * R) return;
    const int row = tid >> 1, layer = tid & 1;
    const uint4 c = philox_block3(p.seed, (uint32_t)step, (uint32_t)(row0 + row), (uint32_t)layer, (uint32_t)(rank >> 3));
    const int wi = (rank & 7) >> 1;
    const uint32_t w = wi == 0 ? c.x : (wi == 1 ? c.y : (wi == 2 ? c.z : c.w));
    SMU(SO_KEEP)[layer * 16 + row] = (w >> (16 * (rank & 1))) & 0xFFFFu;
}

// fill one activation buffer from a row-major fp32 state array [B][width] (rows row0 .. row0+R-1 of this cluster)
__device__ __forceinline__ void fill_from_state(uint32_t* buf, const float* state, int width, int row0, int R, int tid) {
    const int pairs = width >> 1;
    for (int i = tid; i < R * pairs; i += NT) {
        const int r = i / pairs, pi = i - r * pairs;
        const int col = 2 * pi;
        const float2 v = *reinterpret_cast<const float2*>(state + (size_t)(row0 + r) * width + col);
        buf[(col >> 4) * TILE_W + bword(r, (col & 15) >> 1)] = pack_bf16x2(v.x, v.y);
    }
}

__global__ void __launch_bounds__(NT, 1) decoder3_kernel(const __grid_constant__ Dec3Params p) {
    cg::cluster_group cluster = cg::this_cluster();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rank = (int)cluster.block_rank();
    const int cid = blockIdx.x / CS;
    const int row0 = cid * p.R;
    const int R = min(p.R, p.B - row0);
    Guard3 g{p.abort_flag, (p.dbg & 512) != 0};   // dbg 512: no CTA ever waits (pure local instruction time per step; results are garbage)
    const bool free_run = p.mode == TACO_MODE_FREE;
    const float zo = p.zoneout;

    // ---- one-time loads: resident weight tiles of this rank, attention vectors, activation buffers ----
    {
        const uint4* src = reinterpret_cast<const uint4*>(p.res_blobs + (size_t)rank * RES_WORDS);
        uint4* dst = reinterpret_cast<uint4*>(SMU(0));
        for (int i = tid; i < RES_WORDS / 4; i += NT) dst[i] = src[i];
        if (tid < ATT) {
            SMF(SO_VA)[tid] = p.va[tid];
            SMF(SO_BC)[tid] = p.bconst[tid];
        }
        for (int i = tid; i < T_MAX3 + 48; i += NT) SMF(SO_CPAD)[i] = 0.f;
        if (tid < 16 + 48) SMU(SO_FLAG)[tid] = 0;
        zero_blocks(SMU(SO_P1), 6 * KS_H + KS_CTX, tid);   // all activation buffers are contiguous
        if (tid < N_MB) mbar_init(smem_addr(SMU(SO_MBAR) + 2 * tid), 1);
    }
    __syncthreads();
    // recurrent state of step t0-1 (zeros after reset): context, zoned h1 / h2, next decoder input
    fill_from_state(SMU(SO_CTX), p.ctx_state, MEM, row0, R, tid);
    fill_from_state(SMU(SO_H1Z), p.h1z, DEC, row0, R, tid);
    fill_from_state(SMU(SO_H2Z), p.h2z, DEC, row0, R, tid);
    // folded prenet dense_1 of the first step of a continuation launch reads the raw LSTM-2 output of step t0-1
    if (free_run && p.t0 > 0) fill_from_state(SMU(SO_H2R), p.h2raw, DEC, row0, R, tid);

    // epilogue threads: (unit eu = tid & 15 of this CTA's 16 units, cluster row er = tid >> 4); fp32 recurrent state in registers
    const int eu = tid & 15, er = tid >> 4;
    const bool ept = tid < 256;
    const bool evalid = ept && er < R;
    const int eb = row0 + er;
    float c1s = 0.f, h1s = 0.f, c2s = 0.f, h2s = 0.f, h2r_last = 0.f;
    if (evalid) {
        h2r_last = p.h2raw[(size_t)eb * DEC + 16 * rank + eu];
        c1s = p.c1[(size_t)eb * DEC + 16 * rank + eu];
        h1s = p.h1z[(size_t)eb * DEC + 16 * rank + eu];
        c2s = p.c2[(size_t)eb * DEC + 16 * rank + eu];
        h2s = p.h2z[(size_t)eb * DEC + 16 * rank + eu];
    }
    // attention work items: item = row * P + part (P = 1: the whole row; P > 1: position slice [128 part, 128 part + 128) of the
    // row, so that a long row (T > 128) is spread over P CTAs and keeps its pre-activations in registers), one item per CTA,
    // dealt by (local) length to ranks 0..9, 11..15 and last to rank 10, which also computes the stop-token projection of the
    // cluster (<= 15 items leave it without one).  The host guarantees R * P <= 16.
    const int P = p.parts;
    auto item_len = [&](int item) -> int {
        const int l = p.lens[row0 + item / P];
        return P == 1 ? l : max(0, min(PART_T, l - PART_T * (item % P)));
    };
    if (tid < 16) {
        int pos = 0;
        const int items = R * P;
        if (tid < items) {
            const int mylen = item_len(tid);
            for (int r2 = 0; r2 < items; ++r2) {
                const int l2 = item_len(r2);
                pos += (l2 > mylen || (l2 == mylen && r2 < tid)) ? 1 : 0;
            }
        }
        SMU(SO_OWN)[tid] = tid < items ? (pos < STOP_RANK ? pos : (pos < CS - 1 ? pos + 1 : STOP_RANK)) : 0xFFFFu;
        SMU(SO_LEN)[tid] = tid < R ? (uint32_t)p.lens[row0 + tid] : 0u;
    }
    __syncthreads();
    if (tid == 0) {   // thread groups of the context-slice reduction, dealt in proportion to the row lengths
        int total = 0, g0 = 0;
        for (int r2 = 0; r2 < R; ++r2) total += (int)SMU(SO_LEN)[r2];
        for (int r2 = 0; r2 < R; ++r2) {
            const int ng = ctx_groups_of((int)SMU(SO_LEN)[r2], total);
            SMU(SO_GRP)[2 * r2] = g0;
            SMU(SO_GRP)[2 * r2 + 1] = ng;
            g0 += ng;
        }
        // bytes one decoder step delivers into this CTA's alignment rows: every owner multicasts ceil(len / 4) 16-byte pieces
        uint32_t ab_ = 0u;
        for (int r2 = 0; r2 < R; ++r2) ab_ += ((SMU(SO_LEN)[r2] + 3u) >> 2) * 16u;
        SMU(SO_MBAR)[2 * N_MB] = ab_;
    }
    __syncthreads();
    // expected bytes per step and exchange buffer (st.async completes them on the buffer's mbarrier)
    const uint32_t BLK_BYTES = (uint32_t)CS * R * 32u, CTX_BYTES = 3u * BLK_BYTES, Q_BYTES = (uint32_t)CS * ATT * 4u;
    const uint32_t A_BYTES = SMU(SO_MBAR)[2 * N_MB];
    uint32_t cmw = 0u;
    if (tid < 6 * CTX_GROUPS) {
        const int gidx = tid / 6;
        for (int r2 = 0; r2 < R; ++r2) {
            const int g0 = (int)SMU(SO_GRP)[2 * r2], ng = (int)SMU(SO_GRP)[2 * r2 + 1];
            if (gidx >= g0 && gidx < g0 + ng) cmw = (uint32_t)r2 | ((uint32_t)(gidx - g0) << 4) | ((uint32_t)ng << 12) | (SMU(SO_LEN)[r2] << 20);
        }
    }
    int aitem = -1;
    for (int r2 = 0; r2 < R * P; ++r2)
        if ((int)SMU(SO_OWN)[r2] == rank) aitem = r2;
    const bool att = aitem >= 0;
    const int arow = att ? aitem / P : -1, apart = att ? aitem % P : 0;
    const int ab = row0 + (att ? arow : 0);
    const int tb0 = P == 1 ? 0 : PART_T * apart;                 // first position of this item
    const int Tl = P == 1 ? p.T : max(0, min(PART_T, p.T - tb0));   // positions it covers (valid ones: alen)
    const int alen = att ? item_len(aitem) : 0;
    float* pre_scr = p.pre_scratch + ((size_t)ab * p.T + tb0) * ATT;   // [T][128]: working copy when !pre_reg, saved state otherwise
    float preR[NPT][4];
#pragma unroll
    for (int i = 0; i < NPT; ++i) preR[i][0] = preR[i][1] = preR[i][2] = preR[i][3] = 0.f;
    __syncthreads();
    if (att) {
        for (int i = tid; i < Tl + 30; i += NT) {   // cumulative alignments of the item's window, 15 positions of halo on both sides
            const int t = tb0 + i - 15;
            SMF(SO_CPAD)[i] = (t >= 0 && t < p.T) ? p.cum[(size_t)ab * p.T + t] : 0.f;
        }
        __syncthreads();
        att3_build_pairs(Tl, tid);
        __syncthreads();
        if (!p.cont || !p.cumulative) {
            att3_location(p, preR, pre_scr, ab, alen, warp, lane, false, tb0);
        } else if (p.pre_reg) {
            // continuation launch: the accumulated pre-activations were saved by the previous launch (keeps cell stepping
            // bit-identical to one multi-step launch)
            const int gq = lane >> 2, tq = lane & 3, d0 = warp * 8 + 2 * tq;
#pragma unroll
            for (int i = 0; i < NPT; ++i) {
                const int ta = 16 * i + gq, tb = ta + 8;
                if (ta < alen) {
                    const float2 v = *reinterpret_cast<const float2*>(pre_scr + (size_t)ta * ATT + d0);
                    preR[i][0] = v.x;
                    preR[i][1] = v.y;
                }
                if (tb < alen) {
                    const float2 v = *reinterpret_cast<const float2*>(pre_scr + (size_t)tb * ATT + d0);
                    preR[i][2] = v.x;
                    preR[i][3] = v.y;
                }
            }
        }
    }
    if (tid == 0) {   // arm the first phase of every exchange buffer
        for (int mb = MB_P1; mb <= MB_H2Z; ++mb) mbar_rearm(mb, BLK_BYTES);
        mbar_rearm(MB_CTX, CTX_BYTES);
        if (att) mbar_rearm(MB_Q, Q_BYTES);
        if (att && P > 1) mbar_rearm(MB_PART, (uint32_t)(P - 1) * 8u);
        mbar_rearm(MB_A, A_BYTES);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // every CTA of the cluster has armed its buffers before any peer may write into them
    cluster.sync();

    const int mt = warp & 3, kg = warp >> 2;   // 4-m-tile phases: (gate tile, k group)
    const uint4* stream = p.stream_blobs + (size_t)rank * (320 * 32);
    // streamed sets of 64 tiles [4 mt][16 ks]: LSTM-1 prenet rows, LSTM-2 h1 rows, LSTM-1 h1z rows, LSTM-2 h2z rows, LSTM-1 context k-steps 32..47
    const uint4* sW1P = stream + (size_t)(0 + mt * KS_H) * 32;
    const uint4* sW2H = stream + (size_t)(64 + mt * KS_H) * 32;
    const uint4* sW1H = stream + (size_t)(128 + mt * KS_H) * 32;
    const uint4* sW2Z = stream + (size_t)(192 + mt * KS_H) * 32;
    const uint4* sW1C = stream + (size_t)(256 + mt * KS_H) * 32;
    const uint32_t* rW1C = SMU(SO_W1C) + mt * W1C_RES * TILE_W;
    const float* bias = SMF(SO_BIAS);   // b1[64] b2[64] bp1[16] bp2[16] bo[16]
    float* red = SMF(SO_RED);

    // contributions of step t0-1's context / zoned states to the gates of step t0
    float acc1[2][4] = {}, acc2[2][4] = {};
    {
        uint4 t0[4], t1[4], t2[4];
        prefetch4(t0, sW1C, kg, lane);
        prefetch4(t1, sW1H, kg, lane);
        prefetch4(t2, sW2Z, kg, lane);
        gemm_smem4(acc1, rW1C, SMU(SO_CTX), W1C_RES, kg, 4, lane, g);
        gemm_regs(acc1, t0, SMU(SO_CTX) + W1C_RES * TILE_W, kg, lane, g);
        gemm_regs(acc1, t1, SMU(SO_H1Z), kg, lane, g);
        gemm_regs(acc2, t2, SMU(SO_H2Z), kg, lane, g);
    }
    __syncthreads();   // (the blocks are re-armed after dense_1 of the first step, which reads the context and h2raw blocks again)

    long long* tacc = reinterpret_cast<long long*>(smem3 + SO_TACC);
    long long tprev = 0;
    const bool timing = p.timing != nullptr && warp == 0;
    if (timing) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tprev));
    int s_done = 0;
    bool stopped = false;
    // prenet dropout bits (always on, Decoder.py:350-351) of the step about to run; those of step s + 1 are drawn while step s
    // waits for its attention
    bool keep0 = false, keep1 = false;
    if (p.masks) {
        keep0 = evalid && keep_bit3(p, p.t0, eb, 0, 16 * rank + eu);
        keep1 = evalid && keep_bit3(p, p.t0, eb, 1, 16 * rank + eu);
    } else {
        keep_bits_step(p, p.t0, row0, R, rank, tid);   // read after the barrier that precedes the first epilogue
    }

    for (int s = 0; s < p.S_max; ++s) {
        const int step = p.t0 + s;
#ifdef TACO_DEC3_TRACE_BUILD
        const bool tracing = p.trace != nullptr && s == p.trace_step;
#else
        constexpr bool tracing = false;
#endif
        D3_TR(0)
        // ---- batch-global stop (Helpers.py:51-52), evaluated STOP_LAG steps late so that clusters never wait for each other; speculative steps are discarded ----
        // Warp 15, lane i handles the word of cluster i.
        if (free_run && warp == 15 && !(p.dbg & 1)) {
            if (s >= STOP_LAG) {
                cp_async_wait_all();   // words of step s - STOP_LAG, requested one step ago
                uint32_t v = lane < p.NCL ? *reinterpret_cast<volatile uint32_t*>(SMU(SO_FLAG) + 8 + lane) : 1u;
                int spins = 0;
                while (v == 0u && !spin_check3(g, spins)) v = ldg_relaxed(reinterpret_cast<const uint32_t*>(p.flags) + (size_t)(s - STOP_LAG) * p.NCL + lane);
                const unsigned stopm = __ballot_sync(0xffffffffu, (v & 2u) != 0u);
                const unsigned deadm = __ballot_sync(0xffffffffu, g.dead);
                const unsigned full = (1u << p.NCL) - 1u;
                if (lane == 0) SMU(SO_FLAG)[0] = ((p.stop_at_any ? stopm != 0u : (stopm & full) == full) || deadm != 0u) ? 2u : 1u;
            }
            if (s >= STOP_LAG - 1 && lane < p.NCL) cp_async_4(SMU(SO_FLAG) + 8 + lane, p.flags + (size_t)(s - (STOP_LAG - 1)) * p.NCL + lane);
        }
        uint4 aW1C[4];
        prefetch4(aW1C, sW1C, kg, lane);   // used right after dense_1 (context rows of step s-1 -> LSTM-1 gates)
        // ================= P1: prenet dense_1 (Decoder.py:346-351) =================
        // Free-running steps > 0 (TacoTestHelper.next_inputs, Helpers.py:56: x = last frame of the previous step): the frame
        // projection of that frame and dense_1 are ONE precomposed contraction of [h2raw ; context] (K = 1024):
        //   relu(([h2raw;ctx].Wo[:,80:160] + bo[80:160]).Wp1 + bp1) = relu([h2raw;ctx].(Wo[:,80:160].Wp1) + (bo[80:160].Wp1 + bp1))
        // -- the feedback frame is never materialised, so the step has no projection phase and no frame exchange.
        const bool folded = free_run && step > 0;
        const uint32_t par = (uint32_t)(s & 1), par_prev = par ^ 1u;
        if (s > 0) {   // raw LSTM-2 output and context of step s-1 (every mode: the stop token and the next gates read them)
            mbar_wait(MB_H2R, par_prev, g);
            mbar_wait(MB_CTX, par_prev, g);
            if (tid == 0 && !g.dead) {
                mbar_rearm(MB_H2R, BLK_BYTES);
                mbar_rearm(MB_CTX, CTX_BYTES);
            }
        }
        {
            float acc[2][4] = {};
            if (folded) {
                D3_PREPOLL(SMU(SO_H2R), KS_H)
                D3_PREPOLL(SMU(SO_CTX), KS_CTX)
                D3_TR(1)
                gemm_smem4(acc, SMU(SO_WF), SMU(SO_H2R), KS_H, warp, 16, lane, g);
                if (!(p.dbg & 64)) gemm_smem4(acc, SMU(SO_WF) + KS_H * TILE_W, SMU(SO_CTX), KS_CTX, warp, 16, lane, g);
                store_red3(red, acc, warp, lane, 2);
            } else if (!free_run && warp < KS_X) {
                // TacoTrainingHelper.next_inputs (Helpers.py:121-124, ratio 1.0): x(step) = targets[:, r*step - 1]
                const int fidx = step * 2 - 1;
                const int gq = lane >> 2, tq = lane & 3;
                const uint4 a = *reinterpret_cast<const uint4*>(SMU(SO_WP1) + warp * TILE_W + (lane << 2));
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    uint2 bfrag = make_uint2(0u, 0u);
                    const int r = 8 * j + gq;
                    if (fidx >= 0 && fidx < p.F_tgt && r < R) {
                        const float* src = p.targets + ((size_t)(row0 + r) * p.F_tgt + fidx) * MEL + warp * 16 + 2 * tq;
                        const float2 v0 = __ldg(reinterpret_cast<const float2*>(src));
                        const float2 v1 = __ldg(reinterpret_cast<const float2*>(src + 8));
                        bfrag.x = pack_bf16x2(v0.x, v0.y);
                        bfrag.y = pack_bf16x2(v1.x, v1.y);
                    }
                    mma16816(acc[j], a, bfrag);
                }
                store_red3(red, acc, warp, lane, 2);
            }
            D3_STAMP(0)
            D3_TR(2)
            __syncthreads();
            D3_TR(3)
            if (free_run && s >= STOP_LAG && SMU(SO_FLAG)[0] == 2u) {
                stopped = true;
                break;   // CTA- and cluster-uniform: every CTA reads the same flag words
            }
            if (ept) {
                float z;
                if (folded) {
                    z = bias[160 + eu];
#pragma unroll
                    for (int k = 0; k < 16; ++k) z += red[(k * 16 + eu) * 17 + er];
                } else {
                    z = bias[128 + eu];   // <GO> frame of zeros (Helpers.py:131-133): dense_1 = bias
                    if (!free_run) {
#pragma unroll
                        for (int k = 0; k < KS_X; ++k) z += red[(k * 16 + eu) * 17 + er];
                    }
                }
                z = fmaxf(z, 0.f);
                if (!p.masks) {
                    keep0 = evalid && ((SMU(SO_KEEP)[er] >> eu) & 1u);
                    keep1 = evalid && ((SMU(SO_KEEP)[16 + er] >> eu) & 1u);
                }
                z = keep0 ? z * p.keep_scale : 0.f;
                const float zn = __shfl_down_sync(0xffffffffu, z, 1);
                if (!(eu & 1)) SMU(SO_STG_R)[bword(er, eu >> 1)] = pack_bf16x2(z, zn);
            }
            D3_TR(4)
            __syncthreads();
            D3_TR(5)
            multicast_block(SMU(SO_STG_R), SMU(SO_P1), MB_P1, rank, R, tid);
            D3_STAMP(1)
            D3_TR(6)
        }
        // ---- shadow of the dense_1 exchange: what step s-1 left behind ----
        if (s > 0) {
            // context of step s-1 -> LSTM-1 gates of this step
            if (!(p.dbg & 2)) {
            gemm_smem4(acc1, rW1C, SMU(SO_CTX), W1C_RES, kg, 4, lane, g);
            gemm_regs(acc1, aW1C, SMU(SO_CTX) + W1C_RES * TILE_W, kg, lane, g);
            }
            D3_STAMP(11)
            D3_TR(7)
            // stop token of step s-1 (Decoder.py:404-410; Helpers.py:36-59), on the one rank that owns no long attention item
            if (rank == STOP_RANK) dec3_stop_token(p, s - 1, row0, R, cid, free_run, red, tid, warp, lane, g);
            D3_STAMP(12)
            D3_TR(8)
        }
        uint4 aW1P[4];
        prefetch4(aW1P, sW1P, kg, lane);   // used two exchanges later (L1)
        if (rank == STOP_RANK && s > 0) __syncthreads();   // the stop token's reduction scratch is reused by dense_2 (CTA-uniform branch)
        D3_TR(9)
        D3_PREPOLL(SMU(SO_P1), KS_H)
        D3_TR(10)
        // ================= P2: prenet dense_2 =================
        {
            float acc[2][4] = {};
            mbar_wait(MB_P1, par, g);
            if (tid == 0 && !g.dead) mbar_rearm(MB_P1, BLK_BYTES);
            gemm_smem4(acc, SMU(SO_WP2), SMU(SO_P1), KS_H, warp, 16, lane, g);
            store_red3(red, acc, warp, lane, 2);
            D3_STAMP(2)
            D3_TR(11)
            __syncthreads();
            D3_TR(12)
            if (ept) {
                float z = bias[144 + eu];
#pragma unroll
                for (int k = 0; k < 16; ++k) z += red[(k * 16 + eu) * 17 + er];
                z = fmaxf(z, 0.f);
                z = keep1 ? z * p.keep_scale : 0.f;
                const float zn = __shfl_down_sync(0xffffffffu, z, 1);
                if (!(eu & 1)) SMU(SO_STG_R)[bword(er, eu >> 1)] = pack_bf16x2(z, zn);
            }
            D3_TR(13)
            __syncthreads();
            D3_TR(14)
            multicast_block(SMU(SO_STG_R), SMU(SO_P2), MB_P2, rank, R, tid);
            D3_STAMP(3)
            D3_TR(15)
        }
        // ---- shadow of the dense_2 exchange: first half of the location term of this step (alignments of step s-1) ----
        if (att && s > 0 && !(p.dbg & 8)) {
            att3_loc_chunk(p, preR, pre_scr, ab, alen, warp, lane, 0, NPT / 2);
            D3_STAMP(13)
        }
        D3_TR(16)
        D3_PREPOLL(SMU(SO_P2), KS_H)
        D3_TR(17)
        // ================= L1: decoder LSTM layer 1 (Decoder.py:256; Encoder.py:118-142) =================
        {
            mbar_wait(MB_P2, par, g);
            if (tid == 0 && !g.dead) mbar_rearm(MB_P2, BLK_BYTES);
            gemm_regs(acc1, aW1P, SMU(SO_P2), kg, lane, g);
            uint4 aW2H[4];
            prefetch4(aW2H, sW2H, kg, lane);   // in flight during the LSTM-1 epilogue and the h1 exchange
            store_red3(red, acc1, kg * 4 + mt, lane, 2);
#pragma unroll
            for (int j = 0; j < 2; ++j) acc1[j][0] = acc1[j][1] = acc1[j][2] = acc1[j][3] = 0.f;
            D3_STAMP(4)
            D3_TR(18)
            __syncthreads();
            D3_TR(19)
            if (ept) {
                float z[4];
#pragma unroll
                for (int gt = 0; gt < 4; ++gt) {
                    z[gt] = bias[gt * 16 + eu];
#pragma unroll
                    for (int k = 0; k < 4; ++k) z[gt] += red[((k * 4 + gt) * 16 + eu) * 17 + er];
                }
                // gate order i, j, f, o; forget bias 1.0
                const float c_new = sigmoid_f(z[2] + 1.0f) * c1s + sigmoid_f(z[0]) * tanh_f(z[1]);
                float h_new = sigmoid_f(z[3]) * tanh_f(c_new);
                if (evalid) {
                    c1s = (1.f - zo) * c_new + zo * c1s;
                    h1s = (1.f - zo) * h_new + zo * h1s;
                } else {
                    h_new = 0.f;
                }
                const float hn1 = __shfl_down_sync(0xffffffffu, h_new, 1);
                const float hz1 = __shfl_down_sync(0xffffffffu, h1s, 1);
                if (!(eu & 1)) {
                    SMU(SO_STG_R)[bword(er, eu >> 1)] = pack_bf16x2(h_new, hn1);
                    SMU(SO_STG_Z)[bword(er, eu >> 1)] = pack_bf16x2(h1s, hz1);
                }
            }
            D3_TR(20)
            __syncthreads();
            D3_TR(21)
            multicast_block(SMU(SO_STG_R), SMU(SO_H1R), MB_H1R, rank, R, tid);
            multicast_block(SMU(SO_STG_Z), SMU(SO_H1Z), MB_H1Z, rank, R, tid);
            D3_STAMP(5)
            D3_TR(22)
            // ---- shadow of the h1 exchange: second half of the location term ----
            if (att && s > 0 && !(p.dbg & 8)) {
                att3_loc_chunk(p, preR, pre_scr, ab, alen, warp, lane, NPT / 2, NPT);
                D3_STAMP(14)
            }
            D3_TR(23)
            D3_PREPOLL(SMU(SO_H1R), KS_H)
            D3_TR(24)
        // ================= L2: decoder LSTM layer 2 =================
            mbar_wait(MB_H1R, par, g);
            if (tid == 0 && !g.dead) mbar_rearm(MB_H1R, BLK_BYTES);
            gemm_regs(acc2, aW2H, SMU(SO_H1R), kg, lane, g);
        }
        uint4 aW1H[4];
        prefetch4(aW1H, sW1H, kg, lane);   // in flight during the LSTM-2 epilogue and the query exchange
        {
            store_red3(red, acc2, kg * 4 + mt, lane, 2);
#pragma unroll
            for (int j = 0; j < 2; ++j) acc2[j][0] = acc2[j][1] = acc2[j][2] = acc2[j][3] = 0.f;
            D3_STAMP(6)
            D3_TR(25)
            __syncthreads();
            D3_TR(26)
            if (ept) {
                float z[4];
#pragma unroll
                for (int gt = 0; gt < 4; ++gt) {
                    z[gt] = bias[64 + gt * 16 + eu];
#pragma unroll
                    for (int k = 0; k < 4; ++k) z[gt] += red[((k * 4 + gt) * 16 + eu) * 17 + er];
                }
                const float c_new = sigmoid_f(z[2] + 1.0f) * c2s + sigmoid_f(z[0]) * tanh_f(z[1]);
                float h_new = sigmoid_f(z[3]) * tanh_f(c_new);
                if (evalid) {
                    c2s = (1.f - zo) * c_new + zo * c2s;
                    h2s = (1.f - zo) * h_new + zo * h2s;
                    h2r_last = h_new;
                } else {
                    h_new = 0.f;
                }
                const float hn1 = __shfl_down_sync(0xffffffffu, h_new, 1);
                const float hz1 = __shfl_down_sync(0xffffffffu, h2s, 1);
                if (!(eu & 1)) {
                    const uint32_t hw = pack_bf16x2(h_new, hn1);
                    SMU(SO_STG_R)[bword(er, eu >> 1)] = hw;
                    SMU(SO_STG_Z)[bword(er, eu >> 1)] = pack_bf16x2(h2s, hz1);
                    // projection input row of this step, first 256 columns (the frame projection runs after the launch)
                    if (evalid) p.act_hist[((size_t)s * p.B + eb) * (ACT_W / 2) + 8 * rank + (eu >> 1)] = hw;
                }
            }
            D3_TR(27)
            __syncthreads();
            D3_TR(28)
            // K-slice of the query projection: qpart[row][d] = sum_{u in my 16 units} h2raw[row][u] Wq[u][d]
            if (warp < 8) {
                float accq[2][4] = {};
                const uint4 a = *reinterpret_cast<const uint4*>(SMU(SO_WQ) + warp * TILE_W + (lane << 2));
                const int gq = lane >> 2, tq = lane & 3;
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const uint2 bfrag = *reinterpret_cast<const uint2*>(SMU(SO_STG_R) + (j << 6) + (lane << 1));
                    mma16816(accq[j], a, bfrag);
                    float* q0 = red + (8 * j + 2 * tq) * ATT + warp * 16 + gq;   // [row][d]
                    q0[0] = accq[j][0];
                    q0[ATT] = accq[j][1];
                    q0[8] = accq[j][2];
                    q0[ATT + 8] = accq[j][3];
                }
            }
            D3_TR(29)
            __syncthreads();
            D3_TR(30)
            {
                const int row = tid >> 5, v = tid & 31;
                if (row < R) {
                    const uint4 val = *reinterpret_cast<const uint4*>(red + row * ATT + (v << 2));
                    for (int j = 0; j < P; ++j) {   // every position slice of the row needs the query
                        const uint32_t own = SMU(SO_OWN)[row * P + j];
                        st_async_v4(mapa_rank(smem_addr(SMU(SO_QRECV) + rank * ATT + (v << 2)), own), val, mapa_rank(smem_addr(SMU(SO_MBAR) + 2 * MB_Q), own));
                    }
                }
            }
            multicast_block(SMU(SO_STG_R), SMU(SO_H2R), MB_H2R, rank, R, tid);
            multicast_block(SMU(SO_STG_Z), SMU(SO_H2Z), MB_H2Z, rank, R, tid);
            D3_STAMP(7)
            D3_TR(31)
        }
        uint4 aW2Z[4];
        prefetch4(aW2Z, sW2Z, kg, lane);
        // dropout bits of the next step
        if (p.masks) {
            keep0 = evalid && keep_bit3(p, step + 1, eb, 0, 16 * rank + eu);
            keep1 = evalid && keep_bit3(p, step + 1, eb, 1, 16 * rank + eu);
        } else {
            keep_bits_step(p, step + 1, row0, R, rank, tid);   // both layers' bits of this step were read in the dense_1 epilogue
        }
        // ---- zoned states of this step -> gates of the next step (in the shadow of the query exchange) ----
        D3_PREPOLL(SMU(SO_H1Z), KS_H)
        D3_PREPOLL(SMU(SO_H2Z), KS_H)
        D3_TR(32)
        mbar_wait(MB_H1Z, par, g);
        mbar_wait(MB_H2Z, par, g);
        if (tid == 0 && !g.dead) {
            mbar_rearm(MB_H1Z, BLK_BYTES);
            mbar_rearm(MB_H2Z, BLK_BYTES);
        }
        if (!(p.dbg & 16)) {
        gemm_regs(acc1, aW1H, SMU(SO_H1Z), kg, lane, g);
        gemm_regs(acc2, aW2Z, SMU(SO_H2Z), kg, lane, g);
        }
        D3_STAMP(9)
        D3_TR(33)
        // ================= ATT: location-sensitive attention =================
        // owner of a cluster row: query, energies, softmax, alignment row -> every CTA; then every CTA: its 48 context columns
        if (att) att3_owner(p, preR, pre_scr, ab, alen, arow, s, tid, warp, lane, g, tacc, tprev, timing, tracing, Q_BYTES, apart, tb0, Tl);
        att3_context_slice(p, cmw, row0, R, rank, s, tid, g, tracing, A_BYTES);
        // a position slice needs the alignments of its neighbours (15 positions of halo) for the location term of the next step:
        // pair arrays from the complete row every CTA holds after the alignment exchange
        if (att && P > 1) att3_build_pairs_row(SMF(SO_AALL) + arow * p.TP, tb0, (int)SMU(SO_LEN)[arow], Tl, tid);
        D3_STAMP(17)
        D3_TR(47)
        s_done = s + 1;
    }
    // what the next iteration would have done for the last executed step: its stop token (the context and h2raw blocks are
    // still in place) and the location term of the following step (saved below for a continuation launch)
    if (!stopped && s_done > 0) {
        if (rank == STOP_RANK) dec3_stop_token(p, s_done - 1, row0, R, cid, free_run, red, tid, warp, lane, g);
        if (att) att3_loc_chunk(p, preR, pre_scr, ab, alen, warp, lane, 0, NPT);
    }

    if (timing && lane == 0)
        for (int i = 0; i < 24; ++i) p.timing[(size_t)blockIdx.x * 24 + i] = tacc[i];
    // ---- write back the recurrent state + step count ----
    if (evalid) {
        p.c1[(size_t)eb * DEC + 16 * rank + eu] = c1s;
        p.h1z[(size_t)eb * DEC + 16 * rank + eu] = h1s;
        p.c2[(size_t)eb * DEC + 16 * rank + eu] = c2s;
        p.h2z[(size_t)eb * DEC + 16 * rank + eu] = h2s;
        p.h2raw[(size_t)eb * DEC + 16 * rank + eu] = h2r_last;
    }
    if (att) {
        for (int t = tid; t < Tl; t += NT) p.cum[(size_t)ab * p.T + tb0 + t] = SMF(SO_CPAD)[15 + t];
        if (p.pre_reg) {
            const int gq = lane >> 2, tq = lane & 3, d0 = warp * 8 + 2 * tq;
#pragma unroll
            for (int i = 0; i < NPT; ++i) {
                const int ta = 16 * i + gq, tb = ta + 8;
                if (ta < alen) *reinterpret_cast<float2*>(pre_scr + (size_t)ta * ATT + d0) = make_float2(preR[i][0], preR[i][1]);
                if (tb < alen) *reinterpret_cast<float2*>(pre_scr + (size_t)tb * ATT + d0) = make_float2(preR[i][2], preR[i][3]);
            }
        }
    }
    if (blockIdx.x == 0 && tid == 0) {
        int steps = s_done;
        if (free_run) {
            // the stop bits of the last (up to) STOP_LAG executed steps have not been consumed yet
            const int first = max(0, s_done - STOP_LAG);
            steps = stopped ? s_done - STOP_LAG + 1 : s_done;
            if (!stopped) {
                for (int st = first; st < s_done; ++st) {
                    bool any = false, all = true;
                    for (int i = 0; i < p.NCL; ++i) {
                        uint32_t v = 0u;
                        int spins = 0;
                        while ((v = ldg_relaxed(reinterpret_cast<const uint32_t*>(p.flags) + (size_t)st * p.NCL + i)) == 0u && !spin_check3(g, spins)) {
                        }
                        any = any || (v & 2u);
                        all = all && (v & 2u);
                    }
                    if (p.stop_at_any ? any : all) {
                        steps = st + 1;
                        break;
                    }
                }
            }
        }
        *p.steps_out = (g.dead || *reinterpret_cast<volatile int*>(p.abort_flag) != 0) ? -1 : steps;
    }
    // no CTA may exit while a peer can still write into its shared memory
    cluster.sync();
}

}  // namespace
