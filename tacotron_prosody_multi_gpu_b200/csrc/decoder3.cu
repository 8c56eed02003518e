// Persistent CLUSTER decoder kernel (sm_100a): thread-block clusters of 16 CTAs, exchanges through
// distributed shared memory, bf16 mma.sync contractions, fp32 state.
//
// Computes TacotronDecoderCell.__call__ for every decoder step (reference Decoder.py:246-292;
// CustomDecoder.step / helpers, Decoder.py:117-146, Utils/Helpers.py:36-59,115-128;
// LocationSensitiveAttention, Attention.py:169-210,345-401), decomposed by what
// the measurements in tools/microbench_hop.cu say bounds a step on B200:
//   one L2 hop (st.relaxed.gpu -> polled ld.relaxed.gpu)   0.45 us,  128-CTA all-gather through L2 1.6-2.2 us
//   one DSMEM hop (st.shared::cluster -> local smem poll)   0.19 us,  16-CTA all-gather through DSMEM 0.5 us
// so the batch is split over CLUSTERS (rows r0..r0+R-1 per cluster, no inter-cluster traffic except the
// batch-global stop bit) and inside a cluster every activation vector is exchanged through DSMEM:
//
//  * CTA rank c of a cluster owns units 16c..16c+15 of every layer (64 LSTM gate columns, 16 prenet
//    columns, 16 projection columns) for ALL rows of the cluster, and the attention of cluster row c.
//  * mma.sync.m16n8k16: A = weight tile [16 output columns x 16 k] (fragment-major in shared memory for
//    the phases on the critical path, streamed fragment-major from L2 for the contributions that only
//    depend on step s-1), B = activations [16 k x 8 rows] read from the CTA's own shared memory.
//  * "Data is the flag": a producer multicasts its bf16 pairs straight into every peer's activation
//    buffer (st.shared::cluster.v4); buffers are pre-filled with the sentinel 0xFFFFFFFF (two bf16 NaNs);
//    consumers poll their LOCAL shared memory while loading the B fragments and re-arm the buffer after
//    use.  No fences, no mbarriers, no grid barrier.
//  * query projection without a resident Wq: q = h2raw.Wq is K-split -- CTA c contracts its own 16 units
//    (one k-step) for all rows and sends each row's partial [128] to the CTA that owns the row's attention.
//  * batch-global stop (Helpers.py:51-52) across clusters: one flag word per (step, cluster) in global
//    memory, consumed two steps late (speculative steps are discarded), so no step waits on L2.
#include <cooperative_groups.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "taco_internal.h"

namespace cg = cooperative_groups;

namespace {

constexpr int CS = TACO_DEC3_CLUSTER;  // 16 CTAs per cluster
constexpr int NT = TACO_THREADS;       // 512 threads = 16 warps
constexpr int MEL = 80, PRE = 256, DEC = 256, MEM = 768, ATT = 128;
constexpr uint32_t SENT = 0xFFFFFFFFu;
constexpr int T_MAX3 = TACO_DEC3_TMAX;  // 512
#ifndef D3_NPT
#define D3_NPT 8
#endif
constexpr int NPT = D3_NPT;             // 16-position tiles of location pre-activations an attention work item keeps in registers
constexpr int PRE_REG_T = 16 * NPT;     // positions per work item
constexpr int KS_X = 5, KS_H = 16, KS_CTX = 48;
constexpr int TILE_W = 128;             // words of one A tile (32 lanes x uint4) and of one B k-step block (2 n-tiles x 32 lanes x uint2)

// ---- shared-memory map (32-bit word offsets) ----
constexpr int SO_WF = 0;                          // [64 ks] folded feedback projection . prenet dense_1 (ks 0..15: h2raw, 16..63: context)
constexpr int SO_WP1 = SO_WF + 64 * TILE_W;       // [5]  prenet dense_1 itself (teacher-forced steps)
constexpr int SO_WP2 = SO_WP1 + 5 * TILE_W;       // [16]
constexpr int W1C_RES = 32;                       // k-steps of the LSTM-1 context rows kept in shared memory (of 48)
constexpr int SO_W1C = SO_WP2 + 16 * TILE_W;      // [4 mt][32 ks]  LSTM-1, context rows [256, 256 + 512)
constexpr int SO_WQ = SO_W1C + 4 * W1C_RES * TILE_W;  // [8 mt]     query kernel rows of this CTA's 16 units
constexpr int SO_BIAS = SO_WQ + 8 * TILE_W;       // b1[64] b2[64] bp1[16] bp2[16] bfold[16]
constexpr int RES_WORDS = SO_BIAS + 176;          // resident blob words per rank
constexpr int SO_P1 = RES_WORDS;                  // activation buffers (B-fragment order, one 128-word block per k-step)
constexpr int SO_P2 = SO_P1 + KS_H * TILE_W;
constexpr int SO_H1R = SO_P2 + KS_H * TILE_W;
constexpr int SO_H1Z = SO_H1R + KS_H * TILE_W;
constexpr int SO_H2R = SO_H1Z + KS_H * TILE_W;
constexpr int SO_H2Z = SO_H2R + KS_H * TILE_W;
constexpr int SO_CTX = SO_H2Z + KS_H * TILE_W;
constexpr int SO_RED = SO_CTX + KS_CTX * TILE_W;  // 4352 floats: K-split partials; aliased as attention scratch
constexpr int RED_WORDS = 16 * 16 * 17;
constexpr int SO_QRECV = SO_RED + RED_WORDS;      // [16 src][128] fp32 query partials of this CTA's attention row
constexpr int SO_STG_R = SO_QRECV + 16 * ATT;     // staging: raw block, zoned block, context row
constexpr int SO_STG_Z = SO_STG_R + TILE_W;
constexpr int SO_STG_C = SO_STG_Z + TILE_W;       // [3 k-steps][128 words]: this CTA's context slice, B-fragment order
constexpr int SO_ES = SO_STG_C + 3 * TILE_W;      // [T_MAX3]
constexpr int SO_CPAD = SO_ES + T_MAX3;           // [T_MAX3 + 48]  cpad[i] = cum[i - 15]
constexpr int PW3 = (T_MAX3 + 48) / 2;
constexpr int SO_P = SO_CPAD + T_MAX3 + 48;       // 4 x [PW3] pair arrays P0hi, P1hi, P0lo, P1lo
constexpr int SO_QS = SO_P + 4 * PW3;             // [128]
constexpr int SO_VA = SO_QS + ATT;
constexpr int SO_BC = SO_VA + ATT;
constexpr int SO_FLAG = SO_BC + ATT;              // [16]: 0 = stop decision, 8.. = prefetched stop words of the clusters
constexpr int SO_TACC = SO_FLAG + 16;              // 24 x int64 stage timers
constexpr int SO_OWN = SO_TACC + 48;             // [16]: cluster rank that owns the attention of cluster row i
constexpr int SO_LEN = SO_OWN + 16;              // [16]: memory length of cluster row i
constexpr int SO_GRP = SO_LEN + 16;              // [16][2]: first thread group / number of groups of row i in the context-slice reduction
constexpr int SO_MBAR = SO_GRP + 32;             // 10 mbarriers (64-bit each): one per exchange buffer, one phase per decoder step
constexpr int MB_P1 = 0, MB_P2 = 1, MB_H1R = 2, MB_H1Z = 3, MB_H2R = 4, MB_H2Z = 5, MB_CTX = 6, MB_Q = 7, MB_A = 8, MB_PART = 9, N_MB = 10;
constexpr int PART_T = PRE_REG_T;                // positions per attention work item when a row is split
constexpr int PART_MAX = 4;                      // T <= 512
constexpr int SO_KEEP = SO_MBAR + 2 * N_MB + 4;  // [2 prenet layers][16 rows]: 16 dropout keep bits (this CTA's units) of the step about to run
constexpr int SO_PART = SO_KEEP + 32;             // [PART_MAX][2]: (local maximum, local sum of exp) of every position slice of this CTA's row
constexpr int SO_AALL = SO_PART + 2 * PART_MAX;  // [R][TP] alignment rows of every cluster row, multicast by their owners
constexpr int A_CAP = TACO_DEC3_ACAP;            // words: R * TP <= A_CAP (the host picks the cluster count accordingly)
constexpr int SMEM3_WORDS = SO_AALL + A_CAP;
static_assert(RES_WORDS == TACO_DEC3_RES_WORDS, "resident blob size");
static_assert(SMEM3_WORDS * 4 <= 227 * 1024, "shared memory budget");
static_assert(SO_P1 % 4 == 0 && SO_RED % 4 == 0 && SO_QRECV % 4 == 0 && SO_STG_R % 4 == 0 && SO_STG_C % 4 == 0 && SO_ES % 4 == 0 &&
              SO_QS % 4 == 0 && SO_TACC % 2 == 0 && SO_AALL % 4 == 0 && SO_MBAR % 2 == 0, "smem alignment");

extern __shared__ __align__(16) uint32_t smem3[];
#define SMU(off) (smem3 + (off))
#define SMF(off) (reinterpret_cast<float*>(smem3 + (off)))

struct Guard3 {
    int* abort_flag;
    bool dead;
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa_rank(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_remote_v4(uint32_t addr, const uint4& v) {
    asm volatile("st.shared::cluster.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// ---- exchange synchronisation: every remote store is a st.async that completes bytes on the DESTINATION's mbarrier of that
// buffer; one mbarrier phase per decoder step (expected byte count armed by the consumer), consumers sleep in try_wait -- no
// sentinel values, no polling loads, no re-arming of the buffers.
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void st_async_v4(uint32_t addr, const uint4& v, uint32_t bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%2,%3,%4}, [%5];"
                 ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(bar) : "memory");
}
__device__ __forceinline__ uint2 lds_poll_v2(const uint32_t* p) {
    uint2 v;
    asm volatile("ld.volatile.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(smem_addr(p)) : "memory");
    return v;
}
__device__ __forceinline__ uint4 lds_poll_v4(const uint32_t* p) {
    uint4 v;
    asm volatile("ld.volatile.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_addr(p)) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ldg_relaxed(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void stg_relaxed(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ldg_stream_v4(const uint4* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

__device__ __forceinline__ void cp_async_4(uint32_t* smem_dst, const void* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// bounded spin: a poll that never completes (a bug, not a data condition) raises the abort flag instead of
// hanging the GPU; once raised every poll in every CTA falls through.
__device__ __forceinline__ bool spin_check3(Guard3& g, int& spins) {
    if (((++spins) & 1023) == 0) {
        if (*reinterpret_cast<volatile int*>(g.abort_flag) != 0) g.dead = true;
        if (spins > (1 << 24)) {
            atomicExch(g.abort_flag, 1);
            g.dead = true;
        }
    }
    return g.dead;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

__device__ __forceinline__ void mma16816(float (&c)[4], const uint4& a, const uint2& b) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y));
}

__device__ __forceinline__ float tanh_apx(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sigmoid_f(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_f(float x) { return 1.0f - __fdividef(2.0f, __expf(2.0f * x) + 1.0f); }

// word offset, inside a 128-word B block, of the bf16 pair (row r, k-pair j = 0..7 of the k-step)
__device__ __forceinline__ int bword(int r, int j) { return ((r >> 3) << 6) + ((((r & 7) << 2) | (j & 3)) << 1) + (j >> 2); }

// wait (bounded) until the mbarrier phase `parity` of exchange buffer `mb` has completed: all of its bytes have arrived
__device__ __forceinline__ void mbar_wait(int mb, uint32_t parity, Guard3& g) {
    const uint32_t bar = smem_addr(SMU(SO_MBAR) + 2 * mb);
    int spins = 0;
    while (!g.dead && !mbar_try_wait(bar, parity)) {
        if (spin_check3(g, spins)) break;
    }
}
// ... and, by ONE thread after its own wait returned, arm the next phase with the bytes one decoder step delivers
__device__ __forceinline__ void mbar_rearm(int mb, uint32_t bytes) { mbar_expect_tx(smem_addr(SMU(SO_MBAR) + 2 * mb), bytes); }

// B fragment of (k-step block, n-tile j) for this lane (the caller has waited for the buffer's mbarrier phase)
__device__ __forceinline__ uint2 load_b(const uint32_t* blk, int j, int lane, Guard3&) {
    return *reinterpret_cast<const uint2*>(blk + (j << 6) + (lane << 1));
}

// ---------------------------------------------------------------------------------------------
// The loop body of the kernel is executed by 16 warps that are almost always waiting on a dependency; what it must not
// do is miss the 32 KB instruction cache every step.  Everything that repeats is therefore ONE out-of-line routine with
// its operands passed and returned in registers (by-value structs).
// ---------------------------------------------------------------------------------------------
struct AccV {
    float v[2][4];
    int dead;
};
struct A4 {
    uint4 a[4];
};

// Just-in-time weight tiles: the 4 A tiles (k-steps kg, kg+4, kg+8, kg+12 of a 16-k-step set) this warp needs in a later
// phase are requested from L2 BEFORE the wait that precedes that phase, so their latency hides behind the exchange.
__device__ __forceinline__ void prefetch4(uint4 (&a)[4], const uint4* set_mt, int kg, int lane) {
#pragma unroll
#ifdef D3_NO_STREAM   // ablation build: how much of the step is the L2 stream of the non-resident weight tiles (results are garbage)
    for (int i = 0; i < 4; ++i) a[i] = make_uint4(0x3c003c00u + lane, 0x3c003c00u + kg, 0x3c003c00u, 0x3c003c00u + i);
#else
    for (int i = 0; i < 4; ++i) a[i] = ldg_stream_v4(set_mt + (size_t)(kg + 4 * i) * 32 + lane);
#endif
}
// the 8 B fragments (4 k-steps x 2 n-tiles) of blocks ks0, ks0 + kstride, ...
__device__ __forceinline__ void load_b4(uint2 (&b)[4][2], const uint32_t* bbuf, int ks0, int kstride, int nks, int lane, Guard3&) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (i < nks) {
#pragma unroll
            for (int j = 0; j < 2; ++j) b[i][j] = *reinterpret_cast<const uint2*>(bbuf + (ks0 + i * kstride) * TILE_W + (j << 6) + (lane << 1));
        }
    }
}
// acc[j] += A(regs i) . B(block kg + 4 i, n-tile j)
__device__ __forceinline__ AccV gemm_regs_core(AccV acc, A4 a, const uint32_t* bbuf, int kg, int lane, int* abort_flag) {
    Guard3 g{abort_flag, acc.dead != 0};
    uint2 b[4][2];
    load_b4(b, bbuf, kg, 4, 4, lane, g);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int j = 0; j < 2; ++j) mma16816(acc.v[j], a.a[i], b[i][j]);
    }
    acc.dead = g.dead;
    return acc;
}
// resident tiles, k-steps k0, k0 + kstride, ... < KS in chunks of 4
__device__ __forceinline__ AccV gemm_smem4_core(AccV acc, const uint32_t* a_tiles, const uint32_t* bbuf, int KS, int k0, int kstride, int lane,
                                              int* abort_flag) {
    Guard3 g{abort_flag, acc.dead != 0};
    for (int ksb = k0; ksb < KS; ksb += 4 * kstride) {
        const int nks = min(4, (KS - ksb + kstride - 1) / kstride);
        uint4 a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (i < nks) a[i] = *reinterpret_cast<const uint4*>(a_tiles + (ksb + i * kstride) * TILE_W + (lane << 2));
        uint2 b[4][2];
        load_b4(b, bbuf, ksb, kstride, nks, lane, g);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (i < nks) {
#pragma unroll
                for (int j = 0; j < 2; ++j) mma16816(acc.v[j], a[i], b[i][j]);
            }
        }
    }
    acc.dead = g.dead;
    return acc;
}
__device__ __forceinline__ void gemm_regs(float (&acc)[2][4], const uint4 (&a)[4], const uint32_t* bbuf, int kg, int lane, Guard3& g) {
    AccV x;
    A4 t;
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) x.v[j][i] = acc[j][i];
#pragma unroll
    for (int i = 0; i < 4; ++i) t.a[i] = a[i];
    x.dead = g.dead;
    x = gemm_regs_core(x, t, bbuf, kg, lane, g.abort_flag);
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[j][i] = x.v[j][i];
    g.dead = x.dead != 0;
}
__device__ __forceinline__ void gemm_smem4(float (&acc)[2][4], const uint32_t* a_tiles, const uint32_t* bbuf, int KS, int k0, int kstride,
                                           int lane, Guard3& g) {
    AccV x;
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) x.v[j][i] = acc[j][i];
    x.dead = g.dead;
    x = gemm_smem4_core(x, a_tiles, bbuf, KS, k0, kstride, lane, g.abort_flag);
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[j][i] = x.v[j][i];
    g.dead = x.dead != 0;
}

// partial accumulators of warp (mt, kg) -> red[((kg * MTN + mt) * 16 + m) * 17 + n]
__device__ __forceinline__ void store_red3_core(float* red, AccV acc, int slot, int lane) {
    const int gq = lane >> 2, tq = lane & 3;
    float* r0 = red + (slot * 16 + gq) * 17 + 2 * tq;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        r0[8 * j] = acc.v[j][0];
        r0[8 * j + 1] = acc.v[j][1];
        r0[8 * 17 + 8 * j] = acc.v[j][2];
        r0[8 * 17 + 8 * j + 1] = acc.v[j][3];
    }
}
__device__ __forceinline__ void store_red3(float* red, const float (&acc)[2][4], int slot, int lane, int) {
    AccV x;
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) x.v[j][i] = acc[j][i];
    x.dead = 0;
    store_red3_core(red, x, slot, lane);
}

// zero the activation buffers once: padding rows (>= R) have no producer and must read as 0
__device__ __forceinline__ void zero_blocks(uint32_t* buf, int KS, int tid) {
    uint4* b4 = reinterpret_cast<uint4*>(buf);
    for (int i = tid; i < KS * (TILE_W / 4); i += NT) b4[i] = make_uint4(0u, 0u, 0u, 0u);
}

// multicast one staged 128-word block (B-fragment order, k-step = this CTA's rank) to every CTA of the cluster
// (st.async: the 16 bytes complete on the destination's mbarrier `mb`; R rows x 32 bytes per source and block)
__device__ __forceinline__ void multicast_block(const uint32_t* stage, uint32_t* dst_buf, int mb, int rank, int R, int tid) {
    const int dst = tid >> 5, v = tid & 31;
    const int row = ((v >> 4) << 3) + ((v & 15) >> 1);
    if (row < R) {
        const uint4 val = *reinterpret_cast<const uint4*>(stage + (v << 2));
        st_async_v4(mapa_rank(smem_addr(dst_buf + rank * TILE_W + (v << 2)), dst), val, mapa_rank(smem_addr(SMU(SO_MBAR) + 2 * mb), dst));
    }
}

// Philox4x32-10 as in common.cuh (same counter / key layout, restated in oracle/philox.py), rolled and out of line
// one Philox block = the keep bits of 128 consecutive units of (step, row, layer)
__device__ __forceinline__ uint4 philox_block3(uint64_t seed, uint32_t step, uint32_t row, uint32_t layer, uint32_t blk) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    uint4 c = make_uint4(step, row, layer, blk);
    uint2 k = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll 1
    for (int i = 0; i < 10; ++i) {
        const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0;
        k.y += W1;
    }
    return c;
}
__device__ __forceinline__ bool philox_keep3(uint64_t seed, uint32_t step, uint32_t row, uint32_t layer, uint32_t unit) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    uint4 c = make_uint4(step, row, layer, unit >> 7);
    uint2 k = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll 1
    for (int i = 0; i < 10; ++i) {
        const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0;
        k.y += W1;
    }
    const uint32_t bit = unit & 127u;
    const uint32_t w = (bit >> 5) == 0 ? c.x : ((bit >> 5) == 1 ? c.y : ((bit >> 5) == 2 ? c.z : c.w));
    return (w >> (bit & 31u)) & 1u;
}
__device__ __forceinline__ bool keep_bit3(const Dec3Params& p, int step, int b, int layer, int unit) {
    if (p.masks) return p.masks[(((size_t)(step - p.t0) * p.B + b) * 2 + layer) * PRE + unit] != 0;
    return philox_keep3(p.seed, (uint32_t)step, (uint32_t)b, (uint32_t)layer, (uint32_t)unit);
}

}  // namespace

#include "decoder3_kernel.inl"

size_t taco_decoder3_smem_bytes() { return (size_t)SMEM3_WORDS * 4; }

cudaError_t taco_decoder3_max_clusters(int* out) {
    cudaError_t e = cudaFuncSetAttribute(decoder3_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(decoder3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)taco_decoder3_smem_bytes());
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(CS * 8);
    cfg.blockDim = dim3(NT);
    cfg.dynamicSmemBytes = taco_decoder3_smem_bytes();
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaOccupancyMaxActiveClusters(out, decoder3_kernel, &cfg);
}

cudaError_t taco_launch_decoder3(const Dec3Params& p, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(CS * p.NCL);
    cfg.blockDim = dim3(NT);
    cfg.dynamicSmemBytes = taco_decoder3_smem_bytes();
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CS;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, decoder3_kernel, p);
}
