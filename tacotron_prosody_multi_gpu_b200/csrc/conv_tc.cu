// tcgen05 / TMEM / TMA implicit-GEMM conv1d('same') + dense for sm_100a (bf16 operands, fp32
// accumulation in tensor memory).  Covers the tensor-core-bound stages of the path:
//   Postnet 5 x (conv1d k5 + BN + tanh|id)            (reference Postnet.py:28-36, Encoder.py:145-168)
//   postnet_projection + residual + clip               (Tacotron.py:297-309)
//   EncoderConvolution 3 x (conv1d k5 + BN + ReLU)     (Encoder.py:40-46)
//
// GEMM view: rows m = (batch item b, frame f), K = taps * Cin, N = Cout.  No im2col buffer exists:
// the activations are a 3-D TMA tensor (C, F, B) and the tile for tap j is the box at frame
// coordinate f0 + j - pad -- TMA zero-fills the out-of-range frames, which IS the 'same' padding,
// per batch item.  Weights are pre-transposed at load time to K-major [Cout][taps*Cin] bf16.
//
// One persistent CTA per SM, 10 warps:
//   warp 0      TMA producer (one elected lane), 4-stage mbarrier ring of {A 128x64, B BNx64} tiles
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer (M=128, N=BN, K=16 per instruction)
//   warps 2..9  epilogue: tcgen05.ld of the fp32 accumulator (TMEM lane quarter warp & 3, column half (warp - 2) >> 2),
//               folded BN scale/shift + tanh.approx | relu (+ residual + clip), vector stores.  The epilogue of a tile must
//               finish inside the MMA time of the next one (30.7 k cycles at K = 3840): with 4 warps, scalar scale / shift
//               loads and tanhf it took 54-61 k cycles per tile and bounded the kernel (profiles/r02e_conv_tc_*).
// Two accumulators (2 x 256 TMEM columns) so the epilogue of tile i overlaps the MMAs of tile i+1.
#include <cuda.h>
#include <cuda_bf16.h>

#include <cstring>

#include "taco_internal.h"

namespace {

constexpr int BM = 128;      // rows (frames of one batch item) per tile = UMMA M
constexpr int BK = 64;       // bf16 elements per stage along K = one 128-byte swizzle row
constexpr int STAGES = 4;
constexpr int EPI_WARPS = 8;   // two warps per TMEM lane quarter, each takes half of the accumulator columns
constexpr int TC_THREADS = 64 + 32 * EPI_WARPS;
constexpr int ACC_COLS = 256;  // TMEM column stride between the two accumulators

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    const uint32_t addr = smem_u32(bar);
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    } while (!done);
}

__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}

// K-major, 128-byte-swizzled shared-memory operand descriptor (SM100 UMMA): start address >> 4,
// stride-byte-offset = 1024 B (8 rows x 128 B swizzle atom), version 1, layout SWIZZLE_128B.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);          // bits [0,14)
    d |= (uint64_t)1 << 16;                           // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                 // stride byte offset, bits [32,46)
    d |= (uint64_t)1 << 46;                           // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                           // SWIZZLE_128B
    return d;
}

// kind::f16 instruction descriptor: D fp32, A/B bf16, both K-major, N>>3 at [17,23), M>>4 at [24,29).
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

struct TcParams {
    int B, F, Cin, taps, Cout;        // Cin is the padded channel count (multiple of 64)
    int tiles_f, tiles_n;             // ceil(F/128), Cout / BN
    const float* scale;               // [Cout] or null
    const float* shift;               // [Cout] or null
    int act;
    __nv_bfloat16* y_bf16;            // [B][F][y_stride] (EPI 0)
    float* y_f32;                     // [B][F][Cout]      (EPI 1 / optional fp32 copy in EPI 0)
    int y_stride;
    const float* residual;            // [B][F][Cout] (EPI 1)
    int clip;
    float lo, hi;
    int rows_pad;                     // wgrad: (taps - 1) / 2 of the convolution
    int wgrad;                        // weight-gradient mode: the "batch" index is the tap, its shift applies to the K coordinate
    int rows_valid;                   // wgrad: output rows per tap (Cin of the convolution); rows beyond are padding
};

template <int BN, int EPI>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // 1024-byte alignment is required by the 128B swizzle atoms
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int A_BYTES = BM * BK * 2;
    constexpr int B_BYTES = BN * BK * 2;
    constexpr int STAGE_BYTES = A_BYTES + ((B_BYTES + 1023) / 1024) * 1024;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
    uint64_t* empty = full + STAGES;
    uint64_t* tmem_full = empty + STAGES;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_tiles = p.B * p.tiles_f * p.tiles_n;
    const int kchunks = p.Cin / BK;
    const int KB = p.taps * kchunks;
    const int pad = (p.taps - 1) / 2;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&tmB)) : "memory");
        for (int i = 0; i < STAGES; ++i) {
            mbar_init(&full[i], 1);
            mbar_init(&empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], EPI_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const int mt = t / p.tiles_n, nt = t - mt * p.tiles_n;   // column tiles of one row tile run together: A is read from DRAM once
                const int b = mt / p.tiles_f, f0 = (mt - b * p.tiles_f) * BM;
                for (int kb = 0; kb < KB; ++kb) {
                    const int tap = kb / kchunks, cc = kb - tap * kchunks;
                    mbar_wait(&empty[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * STAGE_BYTES;
                    mbar_expect_tx(&full[stage], A_BYTES + B_BYTES);
                    if (p.wgrad) {
                        // dW[tap b][ci][co] = sum_k xT[b][ci][k] dyT[co][k]: both operands K-major over k = (item, frame); xT[b] is
                        // x shifted by the tap (materialised by transpose_bf16_kernel)
                        tma_load_3d(sa, &tmA, &full[stage], cc * BK, f0, b);
                        tma_load_2d(sa + A_BYTES, &tmB, &full[stage], cc * BK, nt * BN);
                    } else {
                        tma_load_3d(sa, &tmA, &full[stage], cc * BK, f0 + tap - pad, b);
                        tma_load_2d(sa + A_BYTES, &tmB, &full[stage], tap * p.Cin + cc * BK, nt * BN);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
                const int acc = it & 1;
                const uint32_t acc_phase = (it >> 1) & 1;
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * ACC_COLS;
                for (int kb = 0; kb < KB; ++kb) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
                    const uint64_t da = umma_desc_sw128(sa);
                    const uint64_t db = umma_desc_sw128(sa + A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        // advance 16 bf16 = 32 bytes inside the swizzle row: +2 in the (addr >> 4) field
                        umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
                    }
                    umma_commit(&empty[stage]);  // frees the smem slot once these MMAs have read it
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(&tmem_full[acc]);    // accumulator complete -> epilogue
            }
        }
    } else {
        // ===================== epilogue warps =====================
        const int q = warp & 3;             // TMEM lane quarter this warp may access
        const int half = (warp - 2) >> 2;   // which half of the accumulator columns
        constexpr int CHUNKS = BN / 16;
        const int ch0 = half == 0 ? 0 : (CHUNKS + 1) / 2, ch1 = half == 0 ? (CHUNKS + 1) / 2 : CHUNKS;
        int it = 0;
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
            const int acc = it & 1;
            const uint32_t acc_phase = (it >> 1) & 1;
            const int mt = t / p.tiles_n, nt = t - mt * p.tiles_n;
            const int b = mt / p.tiles_f, f0 = (mt - b * p.tiles_f) * BM;
            const int f = f0 + q * 32 + lane;
            const bool valid = p.wgrad ? f < p.rows_valid : f < p.F;
            const size_t row = p.wgrad ? (size_t)b * p.rows_valid + f : (size_t)b * p.F + f;
            mbar_wait(&tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t taddr = tmem_base + acc * ACC_COLS + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
            for (int ch = ch0; ch < ch1; ++ch) {
                const int c0 = ch * 16;
                uint32_t v[16];
                __syncwarp();
                tmem_ld16(taddr + c0, v);
                const int n = nt * BN + c0;
                // folded BN scale / shift of these 16 output channels (same addresses in every lane: broadcast 16-byte loads),
                // in flight while the accumulator load completes
                float sc[16], sh[16];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float4 a = p.scale ? __ldg(reinterpret_cast<const float4*>(p.scale + n) + j) : make_float4(1.f, 1.f, 1.f, 1.f);
                    const float4 d = p.shift ? __ldg(reinterpret_cast<const float4*>(p.shift + n) + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                    sc[4 * j] = a.x; sc[4 * j + 1] = a.y; sc[4 * j + 2] = a.z; sc[4 * j + 3] = a.w;
                    sh[4 * j] = d.x; sh[4 * j + 1] = d.y; sh[4 * j + 2] = d.z; sh[4 * j + 3] = d.w;
                }
                tmem_ld_wait();
                float o[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    float x = fmaf(__uint_as_float(v[j]), sc[j], sh[j]);
                    if (p.act == ACT_RELU) x = fmaxf(x, 0.f);
                    else if (p.act == ACT_TANH) asm("tanh.approx.f32 %0, %0;" : "+f"(x));   // |err| <= 2^-10.99, then rounded to bf16
                    o[j] = x;
                }
                if (!valid) {
                    // rows beyond F (partial last tile of a batch item) are computed on zero-filled
                    // TMA rows and dropped here
                } else if (EPI == 0) {
                    uint32_t pk[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        __nv_bfloat162 h = __floats2bfloat162_rn(o[2 * j], o[2 * j + 1]);
                        pk[j] = *reinterpret_cast<uint32_t*>(&h);
                    }
                    if (p.y_bf16) {
                        uint4* dst = reinterpret_cast<uint4*>(p.y_bf16 + row * p.y_stride + n);
                        dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                        dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                    }
                    if (p.y_f32) {
                        float4* d32 = reinterpret_cast<float4*>(p.y_f32 + row * p.Cout + n);
#pragma unroll
                        for (int j = 0; j < 4; ++j) d32[j] = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                    }
                } else {
                    const float4* res = reinterpret_cast<const float4*>(p.residual + row * p.Cout + n);
                    float4* dst = reinterpret_cast<float4*>(p.y_f32 + row * p.Cout + n);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float4 r = p.residual ? res[j] : make_float4(0.f, 0.f, 0.f, 0.f);
                        float4 w = make_float4(o[4 * j] + r.x, o[4 * j + 1] + r.y, o[4 * j + 2] + r.z, o[4 * j + 3] + r.w);
                        if (p.clip) {
                            w.x = fminf(fmaxf(w.x, p.lo), p.hi);
                            w.y = fminf(fmaxf(w.y, p.lo), p.hi);
                            w.z = fminf(fmaxf(w.z, p.lo), p.hi);
                            w.w = fminf(fmaxf(w.w, p.lo), p.hi);
                        }
                        dst[j] = w;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
    }
    // ---- teardown: everyone done with TMEM before the allocating warp frees it ----
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

// fp32 frames [B, F_stride, C] -> clipped fp32 [B, F, C] (decoder_output, Tacotron.py:282-284) and a
// channel-padded bf16 copy [B, F, Cpad] (the postnet's first TMA operand).  out32 may be null.
__global__ void prep_bf16_kernel(const float* __restrict__ in, float* __restrict__ out32, __nv_bfloat16* __restrict__ out16,
                                 int B, int F, long long in_bstride, int C, int Cpad, float lo, float hi, int clip) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t n = (size_t)B * F * Cpad;
    if (i >= n) return;
    int c = (int)(i % Cpad);
    size_t bf = i / Cpad;
    size_t b = bf / F, f = bf % F;
    float v = 0.f;
    if (c < C) {
        v = in[b * in_bstride + f * C + c];
        if (clip) v = fminf(fmaxf(v, lo), hi);
        if (out32) out32[bf * C + c] = v;
    }
    out16[i] = __float2bfloat16_rn(v);
}

// the same with 16-byte loads / 8-byte stores: a thread converts 4 adjacent channels (C, Cpad, in_bstride multiples of 4, aligned
// pointers, fewer than 2^31 channel quads)
__global__ void prep_bf16_v4_kernel(const float* __restrict__ in, float* __restrict__ out32, __nv_bfloat16* __restrict__ out16,
                                    unsigned n4, int F, long long in_bstride, int C4, int Cpad4, float lo, float hi, int clip) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const unsigned bf = i / (unsigned)Cpad4, c4 = i - bf * (unsigned)Cpad4;
    const unsigned b = bf / (unsigned)F, f = bf - b * (unsigned)F;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((int)c4 < C4) {
        v = *reinterpret_cast<const float4*>(in + (size_t)b * in_bstride + ((size_t)f * C4 + c4) * 4);
        if (clip) {
            v.x = fminf(fmaxf(v.x, lo), hi); v.y = fminf(fmaxf(v.y, lo), hi);
            v.z = fminf(fmaxf(v.z, lo), hi); v.w = fminf(fmaxf(v.w, lo), hi);
        }
        if (out32) *reinterpret_cast<float4*>(out32 + ((size_t)bf * C4 + c4) * 4) = v;
    }
    const __nv_bfloat162 lo2 = __floats2bfloat162_rn(v.x, v.y), hi2 = __floats2bfloat162_rn(v.z, v.w);
    uint2 w;
    w.x = *reinterpret_cast<const uint32_t*>(&lo2);
    w.y = *reinterpret_cast<const uint32_t*>(&hi2);
    *reinterpret_cast<uint2*>(out16 + (size_t)i * 4) = w;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

template <int BN, int EPI>
cudaError_t launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const TcParams& p, int n_sm, cudaStream_t st) {
    constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
    constexpr int STAGE_BYTES = A_BYTES + ((B_BYTES + 1023) / 1024) * 1024;
    const int smem = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
    static bool attr = false;
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    const int n_tiles = p.B * p.tiles_f * p.tiles_n;
    const int grid = n_tiles < n_sm ? n_tiles : n_sm;
    conv_tc_kernel<BN, EPI><<<grid, TC_THREADS, smem, st>>>(tmA, tmB, p);
    return cudaGetLastError();
}

}  // namespace

cudaError_t taco_launch_prep_bf16(const float* in, float* out32, void* out16, int B, int F, long long in_bstride, int C, int Cpad,
                                  float lo, float hi, int clip, cudaStream_t st) {
    size_t n = (size_t)B * F * Cpad;
    if (C % 4 == 0 && Cpad % 4 == 0 && in_bstride % 4 == 0 && n / 4 < 0x7fffffffull && (reinterpret_cast<uintptr_t>(in) & 15) == 0 &&
        (reinterpret_cast<uintptr_t>(out16) & 7) == 0 && (!out32 || (reinterpret_cast<uintptr_t>(out32) & 15) == 0)) {
        const unsigned n4 = (unsigned)(n / 4);
        prep_bf16_v4_kernel<<<(n4 + 255) / 256, 256, 0, st>>>(in, out32, reinterpret_cast<__nv_bfloat16*>(out16), n4, F, in_bstride, C / 4,
                                                              Cpad / 4, lo, hi, clip);
        return cudaGetLastError();
    }
    prep_bf16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in, out32, reinterpret_cast<__nv_bfloat16*>(out16), B, F, in_bstride, C,
                                                                  Cpad, lo, hi, clip);
    return cudaGetLastError();
}

cudaError_t taco_launch_conv_tc(const ConvTcParams& c, int n_sm, cudaStream_t st) {
    EncodeTiledFn enc = encode_fn();
    if (!enc) return cudaErrorNotSupported;
    if (c.Cin % BK != 0 || c.taps < 1) return cudaErrorInvalidValue;
    const int BN = c.out_f32_only ? c.Cout : 256;
    if (c.out_f32_only ? (c.Cout % 16 != 0 || c.Cout > 256) : (c.Cout % 256 != 0)) return cudaErrorInvalidValue;

    CUtensorMap tmA, tmB;
    {
        cuuint64_t dims[3] = {(cuuint64_t)c.Cin, (cuuint64_t)c.F, (cuuint64_t)c.B};
        cuuint64_t strides[2] = {(cuuint64_t)c.Cin * 2, (cuuint64_t)c.F * c.Cin * 2};
        cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)BM, 1};
        cuuint32_t es[3] = {1, 1, 1};
        CUresult r = enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(c.x), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)c.taps * c.Cin, (cuuint64_t)c.Cout};
        cuuint64_t strides[1] = {(cuuint64_t)c.taps * c.Cin * 2};
        cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)BN};
        cuuint32_t es[2] = {1, 1};
        CUresult r = enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(c.wt), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    }
    TcParams p;
    p.B = c.B; p.F = c.F; p.Cin = c.Cin; p.taps = c.taps; p.Cout = c.Cout;
    p.tiles_f = (c.F + BM - 1) / BM;
    p.tiles_n = c.Cout / BN;
    p.scale = c.scale; p.shift = c.shift; p.act = c.act;
    p.y_bf16 = reinterpret_cast<__nv_bfloat16*>(c.y_bf16);
    p.y_f32 = c.y_f32;
    p.y_stride = c.y_stride;
    p.residual = c.residual;
    p.clip = c.clip; p.lo = c.clip_lo; p.hi = c.clip_hi;
    p.wgrad = 0; p.rows_valid = 0; p.rows_pad = 0;
    if (c.out_f32_only) {
        if (c.Cout == 80) return launch<80, 1>(tmA, tmB, p, n_sm, st);
        if (c.Cout == 128) return launch<128, 1>(tmA, tmB, p, n_sm, st);
        return cudaErrorInvalidValue;
    }
    return launch<256, 0>(tmA, tmB, p, n_sm, st);
}

// fp32 [B][T][C] -> bf16 transposed, K-major operands of the weight-gradient contraction: out[s][c][b * Tp + tp] =
// in[b][tp - margin + s - shift0][c] (0 outside [0, T)) for s < nshift.  nshift = 1, shift0 = 0: the plain transpose with `margin`
// zero frames in front of every item; nshift = taps, shift0 = pad: one copy per tap, shifted in time -- TMA cannot start a box at an
// inner coordinate that is not a multiple of 16 bytes, so the tap shift is materialised here.  32-frame x 32-channel tiles (+ halo).
__global__ void transpose_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, int B, int T, int C, int Cpad, int Tp,
                                      int margin, int nshift, int shift0) {
    __shared__ float tile[32 + 8][33];
    const int b = blockIdx.z;
    const int t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
    for (int i = ty; i < 32 + nshift - 1; i += 8) {
        const int t = t0 + i - margin - shift0, c = c0 + tx;
        tile[i][tx] = (t >= 0 && t < T && c < C) ? in[((size_t)b * T + t) * C + c] : 0.f;
    }
    __syncthreads();
    const size_t K = (size_t)B * Tp;
    for (int s = 0; s < nshift; ++s)
        for (int i = ty; i < 32; i += 8) {
            const int c = c0 + i, tp = t0 + tx;
            if (c < Cpad && tp < Tp) out[((size_t)s * Cpad + c) * K + (size_t)b * Tp + tp] = __float2bfloat16_rn(tile[tx + s][i]);
        }
}

cudaError_t taco_launch_transpose_bf16(const float* in, void* out, int B, int T, int C, int Cpad, int Tp, int margin, int nshift, int shift0,
                                       cudaStream_t st) {
    if (nshift < 1 || nshift > 9) return cudaErrorInvalidValue;
    dim3 grid((Tp + 31) / 32, (Cpad + 31) / 32, B);
    transpose_bf16_kernel<<<grid, dim3(32, 8), 0, st>>>(in, reinterpret_cast<__nv_bfloat16*>(out), B, T, C, Cpad, Tp, margin, nshift, shift0);
    return cudaGetLastError();
}

// dW [taps][Ci][Co] (fp32) = conv1d weight gradient from xT bf16 [taps][Ci_pad][K] (one time-shifted copy per tap), dyT bf16 [Co][K], K = B * Tp (multiple of 64), Ci_pad
// a multiple of 128, Co a multiple of 256: one tile per (tap, 128 input channels, 256 output channels), K-loop over all frames.
cudaError_t taco_launch_conv_wgrad_tc(const void* xT, const void* dyT, float* dw, int K, int Ci, int Ci_pad, int Co, int taps, int n_sm,
                                      cudaStream_t st) {
    EncodeTiledFn enc = encode_fn();
    if (!enc) return cudaErrorNotSupported;
    if (K % BK != 0 || Ci_pad % BM != 0 || Co % 256 != 0) return cudaErrorInvalidValue;
    CUtensorMap tmA, tmB;
    {
        cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)Ci_pad, (cuuint64_t)taps};
        cuuint64_t strides[2] = {(cuuint64_t)K * 2, (cuuint64_t)K * Ci_pad * 2};
        cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)BM, 1};
        cuuint32_t es[3] = {1, 1, 1};
        if (enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(xT), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return cudaErrorInvalidValue;
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)Co};
        cuuint64_t strides[1] = {(cuuint64_t)K * 2};
        cuuint32_t box[2] = {(cuuint32_t)BK, 256};
        cuuint32_t es[2] = {1, 1};
        if (enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(dyT), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return cudaErrorInvalidValue;
    }
    TcParams p;
    memset(&p, 0, sizeof p);
    p.B = taps; p.F = Ci_pad; p.Cin = K; p.taps = 1; p.Cout = Co;
    p.tiles_f = Ci_pad / BM;
    p.tiles_n = Co / 256;
    p.act = ACT_NONE;
    p.y_f32 = dw; p.y_bf16 = nullptr;
    p.wgrad = 1; p.rows_valid = Ci; p.rows_pad = (taps - 1) / 2;
    return launch<256, 0>(tmA, tmB, p, n_sm, st);
}
