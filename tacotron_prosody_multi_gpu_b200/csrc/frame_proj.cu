// Frame projection of the cluster decoder, hoisted out of the recurrence (sm_100a).
//
// FrameProjection (reference Decoder.py:371-380): frames = [LSTM_output ; context] . Wo[1024 x 160] + bo, r = 2 frames of
// 80 mels per decoder step (Tacotron.py:280 reshapes [B, S, 160] -> [B, 2S, 80]).  Inside the persistent kernel nothing
// depends on the frames any more (the feedback frame is folded into prenet dense_1, decoder3_kernel.inl), so decoder3 only
// stores the projection INPUT rows of every step, bf16 [S][B][1024] -- the same bf16 operands its in-loop projection
// contracted -- and this kernel runs the projection of all executed steps as one [S*B, 1024] x [1024, 160] contraction:
// bf16 mma.sync m16n8k16, fp32 accumulate, cp.async double-buffered 64-wide K chunks.
// Bound: HBM (reads S*B*2 KB once, writes S*B*640 B); 94 MB + 29 MB at C2 (B = 64, S = 720).
#include <cuda_bf16.h>

#include "taco_internal.h"

namespace {

constexpr int FP_BM = 64, FP_N = 160, FP_K = 1024, FP_KC = 64;
constexpr int FP_LD = FP_KC / 2 + 4;   // words per smem row: 32 data + 4 pad -> fragment loads hit 32 distinct banks
constexpr int FP_THREADS = 128;
constexpr int FP_STAGE_WORDS = (FP_BM + FP_N) * FP_LD;

__device__ __forceinline__ void cp_async16(uint32_t* dst, const void* src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(FP_THREADS) frame_projection_kernel(const uint32_t* __restrict__ act, const uint32_t* __restrict__ wo_t,
                                                                       const float* __restrict__ bo, const int* __restrict__ steps,
                                                                       int S_max, int B, int S_cap, float* __restrict__ frames) {
    extern __shared__ __align__(16) uint32_t fp_smem[];
    const int n_steps = min(*steps < 0 ? 0 : *steps, S_max);
    const int M = n_steps * B;
    const int m0 = blockIdx.x * FP_BM;
    if (m0 >= M) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gq = lane >> 2, tq = lane & 3;

    auto load_stage = [&](int stage, int kc) {
        uint32_t* As = fp_smem + stage * FP_STAGE_WORDS;
        uint32_t* Bs = As + FP_BM * FP_LD;
        // 64 (+160) rows x 128 bytes = 8 x 16-byte segments per row
        for (int i = tid; i < (FP_BM + FP_N) * 8; i += FP_THREADS) {
            const int row = i >> 3, seg = i & 7;
            if (row < FP_BM) {
                const int m = min(m0 + row, M - 1);
                cp_async16(As + row * FP_LD + seg * 4, act + (size_t)m * (FP_K / 2) + kc * (FP_KC / 2) + seg * 4);
            } else {
                const int n = row - FP_BM;
                cp_async16(Bs + n * FP_LD + seg * 4, wo_t + (size_t)n * (FP_K / 2) + kc * (FP_KC / 2) + seg * 4);
            }
        }
        cp_async_commit();
    };

    float acc[FP_N / 8][4];
#pragma unroll
    for (int i = 0; i < FP_N / 8; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;

    constexpr int NKC = FP_K / FP_KC;
    load_stage(0, 0);
    for (int kc = 0; kc < NKC; ++kc) {
        if (kc + 1 < NKC) {
            load_stage((kc + 1) & 1, kc + 1);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const uint32_t* As = fp_smem + (kc & 1) * FP_STAGE_WORDS + (warp * 16) * FP_LD;
        const uint32_t* Bs = fp_smem + (kc & 1) * FP_STAGE_WORDS + FP_BM * FP_LD;
#pragma unroll
        for (int ks = 0; ks < FP_KC / 16; ++ks) {
            uint32_t a[4];
            a[0] = As[gq * FP_LD + ks * 8 + tq];
            a[1] = As[(gq + 8) * FP_LD + ks * 8 + tq];
            a[2] = As[gq * FP_LD + ks * 8 + 4 + tq];
            a[3] = As[(gq + 8) * FP_LD + ks * 8 + 4 + tq];
#pragma unroll
            for (int nt = 0; nt < FP_N / 8; ++nt) {
                const uint32_t b0 = Bs[(nt * 8 + gq) * FP_LD + ks * 8 + tq];
                const uint32_t b1 = Bs[(nt * 8 + gq) * FP_LD + ks * 8 + 4 + tq];
                mma_bf16(acc[nt], a, b0, b1);
            }
        }
        __syncthreads();
    }
    // epilogue: row m = s * B + b -> frames[b][2 s + f][c], n = 80 f + c
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int m = m0 + warp * 16 + gq + 8 * half;
        if (m >= M) continue;
        const int s = m / B, b = m - s * B;
        float* out = frames + ((size_t)b * S_cap * 2 + (size_t)s * 2) * 80;
#pragma unroll
        for (int nt = 0; nt < FP_N / 8; ++nt) {
            const int n = nt * 8 + 2 * tq;
            const float2 v = make_float2(acc[nt][2 * half] + __ldg(bo + n), acc[nt][2 * half + 1] + __ldg(bo + n + 1));
            *reinterpret_cast<float2*>(out + n) = v;   // n = 80 f + c and rows of a step are adjacent: out[80 f + c]
        }
    }
}

}  // namespace

cudaError_t taco_launch_frame_projection(const uint32_t* act_hist, const uint32_t* wo_t, const float* bo, const int* steps, int S_max, int B,
                                         int S_cap, float* frames, cudaStream_t stream) {
    const size_t smem = (size_t)2 * FP_STAGE_WORDS * sizeof(uint32_t);
    cudaError_t e = cudaFuncSetAttribute(frame_projection_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    const int M = S_max * B;
    frame_projection_kernel<<<(M + FP_BM - 1) / FP_BM, FP_THREADS, smem, stream>>>(act_hist, wo_t, bo, steps, S_max, B, S_cap, frames);
    return cudaGetLastError();
}

// fp32 -> bf16 copy of the attention values (the per-step stream of the cluster decoder), 4 elements per thread
__global__ void f32_to_bf16_kernel(const float* __restrict__ in, uint2* __restrict__ out, size_t n4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 v = reinterpret_cast<const float4*>(in)[i];
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    out[i] = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
}

cudaError_t taco_launch_f32_to_bf16(const float* in, void* out, size_t n, cudaStream_t stream) {
    if (n % 4 != 0) return cudaErrorInvalidValue;
    const size_t n4 = n / 4;
    f32_to_bf16_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, stream>>>(in, reinterpret_cast<uint2*>(out), n4);
    return cudaGetLastError();
}
