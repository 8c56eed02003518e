// Persistent DATAFLOW decoder kernel (sm_100a), bf16 tensor-core contractions, fp32 state.
//
// Same computation as decoder.cu (TacotronDecoderCell.__call__, reference Decoder.py:246-292;
// CustomDecoder.step / helpers, Decoder.py:117-146, Utils/Helpers.py:36-59,115-128;
// LocationSensitiveAttention, Attention.py:169-210,345-401) but built around what bounds the step
// on B200: the latency of exchanging small activation vectors between the 128 weight-holding SMs.
//
//  * No grid barrier.  Every exchanged vector (x, p1, p2, h1raw, h1z, h2raw, h2z, ctx) lives in a
//    4-deep ring in L2 as bf16 pairs laid out in the mma.sync m16n8k16 A-FRAGMENT order.  A slot is
//    pre-filled with the sentinel 0xFFFFFFFF (two bf16 NaNs the kernel never produces); consumers
//    load their fragments with 16-byte ld.relaxed.gpu and simply retry while any word is still the
//    sentinel: data IS the flag, one L2 round trip per exchange instead of barrier + load.
//    The producer of step s clears slot (s+2)&3 of the words it owns, then fences.
//  * Gate / prenet / projection contractions: mma.sync.m16n8k16 bf16 -> fp32 register accumulators,
//    CTA c owns 2 units (8 gate columns); the 16 warps split (M-tile, K) and reduce through smem.
//    tcgen05 is deliberately not used here (the per-step tensor work is 0.2 us, a
//    TMEM round trip per phase would cost more than the contraction).
//  * The 80 % of LSTM-1 / 50 % of LSTM-2 MACs that only depend on step s-1 (context, zoned h) are
//    accumulated right after the projection, in the shadow of the feedback-frame exchange.
//  * Attention: CTA (row b, slice sl) keeps its cumulative alignments and the location-feature
//    pre-activations pre[t][d] = keys + b_a + conv31(cum).Wl in shared memory; the location term is
//    ONE Toeplitz contraction  T(cum)[t][k] . G[k][d],  G = conv_kernel . Wl, on tensor cores with a
//    hi/lo bf16 split (3 MMAs, ~2^-16 relative error); energies use tanh.approx; the context
//    reduction streams bf16 values with 16-byte loads.
#include <cuda_bf16.h>

#include "common.cuh"
#include "taco_internal.h"

namespace {

constexpr int NC = TACO_DEC2_CTAS;  // 128 CTAs, one per SM, co-resident (cooperative launch)
constexpr int NT = TACO_THREADS;    // 512 threads = 16 warps
constexpr int MEL = 80, PRE = 256, DEC = 256, MEM = 768, ATT = 128;
constexpr int KS_X = 5, KS_H = 16, KS_CTX = 48;
constexpr uint32_t SENT = 0xFFFFFFFFu;
constexpr int T_MAX2 = TACO_DEC2_TMAX;  // 512
constexpr int PS = 132;                 // smem row stride (floats) of pre[t][128]
// blob layout (32-bit words): fragment-major bf16 weights then fp32 biases
constexpr int OFF_W1 = 0, OFF_W2 = OFF_W1 + 80 * 64, OFF_WP1 = OFF_W2 + 32 * 64, OFF_WP2 = OFF_WP1 + 5 * 64,
              OFF_WO = OFF_WP2 + 16 * 64, OFF_BIAS = OFF_WO + 64 * 64, BLOB_WORDS = OFF_BIAS + 24;
static_assert(BLOB_WORDS == TACO_DEC2_BLOB_WORDS, "blob size");

struct Guard {
    int* abort_flag;
    bool dead;
    int dbg;
};
#define DEC2_FENCE() do { if (!(g.dbg & 1)) __threadfence(); } while (0)

__device__ __forceinline__ uint4 ld_frag(const uint32_t* p) {
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_word(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_word(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ bool has_sent(const uint4& v) { return v.x == SENT || v.y == SENT || v.z == SENT || v.w == SENT; }

// bounded spin: a poll that never completes (a bug, not a data condition) raises the abort flag
// instead of hanging the GPU; once raised every poll in every CTA falls through.
__device__ __forceinline__ bool spin_check(Guard& g, int& spins) {
    if (g.dbg & 2) __nanosleep(64);
    if (g.dbg & 4) __nanosleep(512);
    if (g.dbg & 8) __nanosleep(2000);
    if (((++spins) & 255) == 0) {
        if (*reinterpret_cast<volatile int*>(g.abort_flag) != 0) g.dead = true;
        if (spins > (1 << 21)) {
            atomicExch(g.abort_flag, 1);
            g.dead = true;
        }
    }
    return g.dead;
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint4& a, const uint2& b) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y));
}

// word index of the bf16 pair (row, col..col+1), col even, in the A-fragment-major layout [mt][ks][lane][4]
__device__ __forceinline__ int frag_word(int KS, int row, int col) {
    const int mt = row >> 4, rr = row & 15, ks = col >> 4, cc = col & 15;
    const int lane = ((rr & 7) << 2) | ((cc & 7) >> 1);
    const int reg = ((cc >> 3) << 1) | (rr >> 3);
    return ((mt * KS + ks) * 32 + lane) * 4 + reg;
}

// acc += A[slot](M-tile mt, k-steps kg, kg+KG, ...) . W[seg](k-step)   -- warp-collective
template <int CH>
__device__ __noinline__ void mma_seg(float (&acc)[4], const uint32_t* slot, int KS, int mt, int kg, int KG,
                                        const uint2* wseg, int lane, Guard& g) {
#pragma unroll 1
    for (int ks0 = kg; ks0 < KS; ks0 += CH * KG) {
        uint4 a[CH];
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            const int ks = ks0 + i * KG;
            if (ks < KS) a[i] = ld_frag(slot + ((mt * KS + ks) * 32 + lane) * 4);
        }
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            const int ks = ks0 + i * KG;
            if (ks < KS) {
                int spins = 0;
                while (!g.dead && __any_sync(0xffffffffu, has_sent(a[i]))) {
                    if (spin_check(g, spins)) break;
                    a[i] = ld_frag(slot + ((mt * KS + ks) * 32 + lane) * 4);
                }
                g.dead = __any_sync(0xffffffffu, g.dead);
                mma_bf16(acc, a[i], wseg[ks * 32 + lane]);
            }
        }
    }
}

// partial accumulators of warp (mt, kg) -> red[kg][n][RS] (conflict-free both ways, RS = 16*MT + 4)
__device__ __forceinline__ void store_red(float* red, const float (&acc)[4], int RS, int mt, int kg, int lane) {
    const int g = lane >> 2, t = lane & 3;
    float* r0 = red + (kg * 8 + 2 * t) * RS + mt * 16 + g;
    r0[0] = acc[0];
    r0[RS] = acc[1];
    r0[8] = acc[2];
    r0[RS + 8] = acc[3];
}

__device__ __forceinline__ float tanh_approx(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// fast-math gate non-linearities (ex2.approx + rcp.approx, |rel err| ~ 1e-6): the LSTM pointwise stage sits on the
// critical path of every step
__device__ __forceinline__ float sigmoid_fast(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_fast(float x) { return 1.0f - __fdividef(2.0f, __expf(2.0f * x) + 1.0f); }

__device__ __forceinline__ bool keep_bit2(const Dec2Params& p, int step, int b, int layer, int unit) {
    if (p.masks) return p.masks[(((size_t)(step - p.t0) * p.B + b) * 2 + layer) * PRE + unit] != 0;
    return philox_keep(p.seed, (uint32_t)step, (uint32_t)b, (uint32_t)layer, (uint32_t)unit);
}

}  // namespace

#include "decoder2_kernel.inl"

namespace {

// Fills the rings with the sentinel and converts the recurrent state (fp32, row-major) into the
// slot the first step reads: slot (t0-1)&3 of ctx / h1z / h2z, slot t0&3 of x.  Padding rows
// (>= B) of the state slots are zero; padding rows of the ctx ring have no producer, so they are
// zero in every slot and are never cleared.
__global__ void dec2_init_kernel(const Dec2Params p) {
    const int gsz = gridDim.x * blockDim.x, gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const int sp = (p.t0 + 3) & 3, sx = p.t0 & 3;
    auto fill = [&](uint32_t* ring, int KS, const float* state, int width, int state_slot, bool zero_pad) {
        const int SZ = p.MT * KS * 128;
        for (int i = gtid; i < 4 * SZ; i += gsz) {
            const int slot = i / SZ, w = i - slot * SZ;
            const int reg = w & 3, lane = (w >> 2) & 31, mk = w >> 7;
            const int ks = mk % KS, mt = mk / KS;
            const int row = mt * 16 + (lane >> 2) + (reg & 1) * 8;
            const int col = ks * 16 + (lane & 3) * 2 + (reg >> 1) * 8;
            uint32_t v = SENT;
            if (slot == state_slot && state)
                v = row < p.B ? pack_bf16(state[(size_t)row * width + col], state[(size_t)row * width + col + 1]) : 0u;
            else if (zero_pad && row >= p.B)
                v = 0u;
            ring[i] = v;
        }
    };
    fill(p.ring_x, KS_X, p.x_state, MEL, sx, false);
    fill(p.ring_p1, KS_H, nullptr, 0, -1, false);
    fill(p.ring_p2, KS_H, nullptr, 0, -1, false);
    fill(p.ring_h1raw, KS_H, nullptr, 0, -1, false);
    fill(p.ring_h1z, KS_H, p.h1z, DEC, sp, false);
    fill(p.ring_h2raw, KS_H, nullptr, 0, -1, false);
    fill(p.ring_h2z, KS_H, p.h2z, DEC, sp, false);
    fill(p.ring_ctx, KS_CTX, p.ctx_state, MEM, sp, true);
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, size_t n4) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 v = reinterpret_cast<const float4*>(in)[i];
    uint2 o;
    o.x = pack_bf16(v.x, v.y);
    o.y = pack_bf16(v.z, v.w);
    reinterpret_cast<uint2*>(out)[i] = o;
}

}  // namespace

size_t taco_decoder2_smem_bytes(int T, int* pre_in_smem) {
    const size_t limit = 227 * 1024;
    const size_t with_pre = SMEM2_FIXED + (size_t)T * PS * sizeof(float);
    if (with_pre <= limit) {
        *pre_in_smem = 1;
        return with_pre;
    }
    *pre_in_smem = 0;
    return SMEM2_FIXED;
}

cudaError_t taco_launch_decoder2_init(const Dec2Params& p, cudaStream_t stream) {
    dec2_init_kernel<<<64, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t taco_launch_f32_to_bf16(const float* in, void* out, size_t n, cudaStream_t stream) {
    if (n % 4 != 0) return cudaErrorInvalidValue;
    const size_t n4 = n / 4;
    f32_to_bf16_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, stream>>>(in, reinterpret_cast<__nv_bfloat16*>(out), n4);
    return cudaGetLastError();
}

cudaError_t taco_launch_decoder2(const Dec2Params& p, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(decoder2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    int dummy;
    const size_t smem = taco_decoder2_smem_bytes(p.T, &dummy);
    Dec2Params pl = p;
    void* args[] = {&pl};
    // cooperative launch only for its co-residency guarantee: the kernel synchronises through data, not grid.sync()
    return cudaLaunchCooperativeKernel((const void*)decoder2_kernel, dim3(NC), dim3(NT), args, smem, stream);
}
