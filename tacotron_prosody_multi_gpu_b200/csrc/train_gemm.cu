// train_gemm.cu -- fp32 contractions of the training step: dense GEMM with general
// strides, and conv1d('same') forward / data-gradient / weight-gradient as implicit GEMMs (no im2col,
// padding by predicated loads).  SIMT fp32 FMA with fp32 accumulation: the training path is compared
// against fp32 autograd through the oracle, so the first version keeps full fp32 operands; the inference
// path's tcgen05 kernel (conv_tc.cu) stays the bf16 product for the forward-only convs.
//
// One templated kernel: C[M,N] = sum_k A(m,k) * B(k,n) (+ beta*C) (+ bias[n]); A and B are loader functors.
// Tiles: BM x BN x BK with a TM x TN register micro-tile per thread, shared-memory double buffering.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <cuda_bf16.h>
#include <cstdint>

#include "../../include/taco_b200_train.h"
#include "common.cuh"
#include "taco_internal.h"   // conv_tc.cu: the tcgen05 implicit-GEMM convolution shared with the synthesis path

namespace {

struct APlain {   // A(m,k) = A[m*sm + k*sk]
    const float* p; long long sm, sk; int k_contig;
    __device__ __forceinline__ float operator()(int m, int k) const { return p[m * sm + k * sk]; }
};
struct APair {    // A(m,k) = k < K1 ? A1[m*ld1 + k] : A2[m*ld2 + k - K1]   (two k-contiguous row blocks side by side)
    const float *p1, *p2; long long ld1, ld2; int K1, k_contig;
    __device__ __forceinline__ float operator()(int m, int k) const { return k < K1 ? p1[m * ld1 + k] : p2[m * ld2 + (k - K1)]; }
};
struct BPlain {   // B(k,n) = B[k*sk + n*sn]
    const float* p; long long sk, sn; int n_contig;
    __device__ __forceinline__ float operator()(int k, int n) const { return p[k * sk + n * sn]; }
};
// conv1d forward: A(m=(b,t), k=(j,ci)) = X[b, t + j - left, ci]  (zero outside [0,T))
struct AConvFwd {
    const float* x; int T, Ci, left, k_contig;
    __device__ __forceinline__ float operator()(int m, int k) const {
        int b = m / T, t = m - b * T, j = k / Ci, ci = k - j * Ci, tt = t + j - left;
        return (tt >= 0 && tt < T) ? x[((long long)b * T + tt) * Ci + ci] : 0.f;
    }
};
// conv1d data gradient: A(m=(b,t), k=(j,co)) = dY[b, t - j + left, co];  B(k=(j,co), n=ci) = W[j][ci][co]
struct AConvDgrad {
    const float* dy; int T, Co, left, k_contig;
    __device__ __forceinline__ float operator()(int m, int k) const {
        int b = m / T, t = m - b * T, j = k / Co, co = k - j * Co, tt = t - j + left;
        return (tt >= 0 && tt < T) ? dy[((long long)b * T + tt) * Co + co] : 0.f;
    }
};
struct BConvDgrad {
    const float* w; int Ci, Co, n_contig;
    __device__ __forceinline__ float operator()(int k, int n) const {
        int j = k / Co, co = k - j * Co;
        return w[((long long)j * Ci + n) * Co + co];
    }
};
// conv1d weight gradient: A(m=(j,ci), k=(b,t)) = X[b, t + j - left, ci];  B(k=(b,t), n=co) = dY[(b,t), co]
struct AConvWgrad {
    const float* x; int T, Ci, left, k_contig;
    __device__ __forceinline__ float operator()(int m, int k) const {
        int j = m / Ci, ci = m - j * Ci, b = k / T, t = k - b * T, tt = t + j - left;
        return (tt >= 0 && tt < T) ? x[((long long)b * T + tt) * Ci + ci] : 0.f;
    }
};

// conv2d(3x3, stride 2, SAME) weight gradient as an implicit GEMM: m = (tap = 3 kh + kw, ci), k = output position (b, ho, wo),
// A(m, k) = X[b, 2 ho + kh - pt, 2 wo + kw - pl, ci] (NHWC, zero outside); B = dY [positions][Co]; C = dW [3,3,Ci,Co]
struct AConv2dWgrad {
    const float* x; int H, W, Ci, Ho, Wo, pt, pl, k_contig;
    __device__ __forceinline__ float operator()(int m, int k) const {
        const int tap = m / Ci, ci = m - tap * Ci, kh = tap / 3, kw = tap - 3 * kh;
        const int r = k / Wo, wo = k - r * Wo, b = r / Ho, ho = r - b * Ho;
        const int hi = 2 * ho + kh - pt, wi = 2 * wo + kw - pl;
        return (hi >= 0 && hi < H && wi >= 0 && wi < W) ? x[(((long long)b * H + hi) * W + wi) * Ci + ci] : 0.f;
    }
};

template <int BM, int BN, int BK, int TM, int TN, class AL, class BL>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_kernel(AL A, BL Bm, float* __restrict__ C, long long ldc, int M, int N, int K, float beta, const float* __restrict__ bias,
            int kchunk /* K range per blockIdx.z; gridDim.z > 1 => split-K with atomic accumulation into C */,
            float* __restrict__ C2 = nullptr, long long ldc2 = 0, int N1 = 0 /* columns >= N1 go to C2[m*ldc2 + n - N1] */) {
    constexpr int NT = (BM / TM) * (BN / TN);
    __shared__ float As[2][BK][BM + 4];
    __shared__ float Bs[2][BK][BN + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int tx = tid % (BN / TN), ty = tid / (BN / TN);
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    constexpr int A_PER = (BM * BK + NT - 1) / NT, B_PER = (BK * BN + NT - 1) / NT;
    float ra[A_PER], rb[B_PER];
    auto gload = [&](int k0) {
#pragma unroll
        for (int i = 0; i < A_PER; ++i) {
            int e = tid + i * NT;
            int mm, kk;
            if (A.k_contig) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
            float v = 0.f;
            if (e < BM * BK && m0 + mm < M && k0 + kk < K) v = A(m0 + mm, k0 + kk);
            ra[i] = v;
        }
#pragma unroll
        for (int i = 0; i < B_PER; ++i) {
            int e = tid + i * NT;
            int nn, kk;
            if (Bm.n_contig) { nn = e % BN; kk = e / BN; } else { kk = e % BK; nn = e / BK; }
            float v = 0.f;
            if (e < BK * BN && n0 + nn < N && k0 + kk < K) v = Bm(k0 + kk, n0 + nn);
            rb[i] = v;
        }
    };
    auto sstore = [&](int buf) {
#pragma unroll
        for (int i = 0; i < A_PER; ++i) {
            int e = tid + i * NT;
            int mm, kk;
            if (A.k_contig) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
            if (e < BM * BK) As[buf][kk][mm] = ra[i];
        }
#pragma unroll
        for (int i = 0; i < B_PER; ++i) {
            int e = tid + i * NT;
            int nn, kk;
            if (Bm.n_contig) { nn = e % BN; kk = e / BN; } else { kk = e % BK; nn = e / BK; }
            if (e < BK * BN) Bs[buf][kk][nn] = rb[i];
        }
    };
    const int kbeg = blockIdx.z * kchunk;
    if (gridDim.z > 1) K = min(K, kbeg + kchunk);
    const int nk = (K - kbeg + BK - 1) / BK;
    gload(kbeg);
    sstore(0);
    __syncthreads();
    for (int it = 0; it < nk; ++it) {
        const int buf = it & 1;
        if (it + 1 < nk) gload(kbeg + (it + 1) * BK);
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; ++i) a[i] = As[buf][kk][ty * TM + i];
#pragma unroll
            for (int j = 0; j < TN; ++j) b[j] = Bs[buf][kk][tx * TN + j];
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (it + 1 < nk) sstore(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        int m = m0 + ty * TM + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            int n = n0 + tx * TN + j;
            if (n >= N) continue;
            float v = acc[i][j];
            float* c = (C2 && n >= N1) ? C2 + (long long)m * ldc2 + (n - N1) : C + (long long)m * ldc + n;
            if (gridDim.z > 1) {             // split-K: C was zeroed (beta 0) or holds the addend (beta 1)
                if (bias && blockIdx.z == 0) v += bias[n];
                atomicAdd(c, v);
                continue;
            }
            if (bias) v += bias[n];
            if (beta != 0.f) v += beta * *c;
            *c = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// bf16 tensor-core variant (TACO_TRAIN_BF16): same loader functors, operands rounded to bf16 while staged into
// shared memory, mma.sync.m16n8k16 with fp32 accumulation.  128x128x32 tiles, 8 warps (2 x 4, warp tile 64x32),
// ldmatrix fragment loads from k-contiguous [row][40-half] tiles (80-byte rows: conflict-free), split-K over
// blockIdx.z with atomic accumulation for the weight-gradient shapes (few output tiles, K = S*B rows).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned short f2bf(float f) {   // round-to-nearest-even (finite inputs)
    uint32_t u = __float_as_uint(f);
    u += 0x7FFFu + ((u >> 16) & 1u);
    return (unsigned short)(u >> 16);
}

template <class AL, class BL>
__global__ void __launch_bounds__(256)
gemm_mma_kernel(AL A, BL Bm, float* __restrict__ C, long long ldc, int M, int N, int K, float beta, const float* __restrict__ bias,
                int kchunk) {
    constexpr int BM = 128, BN = 128, BK = 32, LD = 40, NT = 256;
    __shared__ __align__(16) unsigned short As[2][BM][LD];
    __shared__ __align__(16) unsigned short Bs[2][BN][LD];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    float acc[4][4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int r = 0; r < 4; ++r) acc[i][j][r] = 0.f;
    constexpr int PER = BM * BK / NT;   // 16 elements of A and of B per thread per k-tile
    float ra[PER], rb[PER];
    auto gload = [&](int k0) {
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            int e = tid + i * NT, mm, kk;
            if (A.k_contig) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
            ra[i] = (m0 + mm < M && k0 + kk < K) ? A(m0 + mm, k0 + kk) : 0.f;
        }
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            int e = tid + i * NT, nn, kk;
            if (Bm.n_contig) { nn = e % BN; kk = e / BN; } else { kk = e % BK; nn = e / BK; }
            rb[i] = (n0 + nn < N && k0 + kk < K) ? Bm(k0 + kk, n0 + nn) : 0.f;
        }
    };
    auto sstore = [&](int buf) {
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            int e = tid + i * NT, mm, kk;
            if (A.k_contig) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
            As[buf][mm][kk] = f2bf(ra[i]);
        }
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            int e = tid + i * NT, nn, kk;
            if (Bm.n_contig) { nn = e % BN; kk = e / BN; } else { kk = e % BK; nn = e / BK; }
            Bs[buf][nn][kk] = f2bf(rb[i]);
        }
    };
    const int kbeg = blockIdx.z * kchunk;
    if (gridDim.z > 1) K = min(K, kbeg + kchunk);
    const int nk = (K - kbeg + BK - 1) / BK;
    if (nk > 0) {
        gload(kbeg);
        sstore(0);
    }
    __syncthreads();
    for (int it = 0; it < nk; ++it) {
        const int buf = it & 1;
        if (it + 1 < nk) gload(kbeg + (it + 1) * BK);
#pragma unroll
        for (int ks = 0; ks < BK; ks += 16) {
            uint32_t af[4][4], bfr[4][2];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t addr = smem_u32(&As[buf][wm * 64 + i * 16 + (lane & 15)][ks + (lane >> 4) * 8]);
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(af[i][0]), "=r"(af[i][1]), "=r"(af[i][2]), "=r"(af[i][3]) : "r"(addr));
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {   // two n-tiles per ldmatrix.x4
                const int id = lane >> 3;
                const uint32_t addr = smem_u32(&Bs[buf][wn * 32 + j * 16 + (id >> 1) * 8 + (lane & 7)][ks + (id & 1) * 8]);
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(bfr[2 * j][0]), "=r"(bfr[2 * j][1]), "=r"(bfr[2 * j + 1][0]), "=r"(bfr[2 * j + 1][1]) : "r"(addr));
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                                 : "+f"(acc[i][j][0]), "+f"(acc[i][j][1]), "+f"(acc[i][j][2]), "+f"(acc[i][j][3])
                                 : "r"(af[i][0]), "r"(af[i][1]), "r"(af[i][2]), "r"(af[i][3]), "r"(bfr[j][0]), "r"(bfr[j][1]));
        }
        if (it + 1 < nk) sstore(buf ^ 1);
        __syncthreads();
    }
    const int g = lane >> 2, tq = lane & 3;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int m = m0 + wm * 64 + i * 16 + g + (r >> 1) * 8;
                const int n = n0 + wn * 32 + j * 8 + tq * 2 + (r & 1);
                if (m >= M || n >= N) continue;
                float v = acc[i][j][r];
                float* c = C + (long long)m * ldc + n;
                if (gridDim.z > 1) {
                    if (bias && blockIdx.z == 0) v += bias[n];
                    atomicAdd(c, v);
                    continue;
                }
                if (bias) v += bias[n];
                if (beta != 0.f) v += beta * *c;
                *c = v;
            }
}

int g_train_precision = 0;   // TACO_TRAIN_FP32 / TACO_TRAIN_BF16 (taco_op_set_precision)

template <class AL, class BL>
int launch_gemm_mma(AL A, BL Bm, float* C, long long ldc, int M, int N, int K, float beta, const float* bias, cudaStream_t st) {
    dim3 grid((N + 127) / 128, (M + 127) / 128, 1);
    int kchunk = K;
    const int ctas = grid.x * grid.y;
    if ((beta == 0.f || beta == 1.f) && ctas < 148 && K >= 1024) {
        int splits = (296 + ctas - 1) / ctas;
        if (splits > K / 256) splits = K / 256;
        if (splits > 1) {
            kchunk = ((K + splits - 1) / splits + 31) / 32 * 32;
            grid.z = (K + kchunk - 1) / kchunk;
            if (grid.z > 1 && beta == 0.f &&
                cudaMemset2DAsync(C, (size_t)ldc * sizeof(float), 0, (size_t)N * sizeof(float), (size_t)M, st) != cudaSuccess)
                return -3;
        }
    }
    gemm_mma_kernel<AL, BL><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, beta, bias, kchunk);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

// ---------------------------------------------------------------------------------------------------------------
// Vectorised bf16 tensor-core kernel: operands are fetched as 8-element segments along their CONTIGUOUS axis
// (two 16-byte loads, index arithmetic once per segment) and staged in shared memory in their natural
// orientation; fragments of an operand whose contiguous axis is m / n (not k) are read with ldmatrix.trans.
// Requires every contiguous extent to be a multiple of 8 (dispatch falls back to gemm_mma_kernel otherwise).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ld8(const float* p, float v[8]) {
    if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = p[i];
    }
}
__device__ __forceinline__ void zero8(float v[8]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = 0.f;
}
// A operands: load8(m, k, v) -- K_CONTIG: v[i] = A(m, k+i);  else: v[i] = A(m+i, k).   (in-range m, k guaranteed)
struct VAPlainK { static constexpr bool K_CONTIG = true; const float* p; long long ld;
    __device__ __forceinline__ void load8(int m, int k, float v[8]) const { ld8(p + (long long)m * ld + k, v); } };
struct VAPlainM { static constexpr bool K_CONTIG = false; const float* p; long long ld;      // stored [K][M]
    __device__ __forceinline__ void load8(int m, int k, float v[8]) const { ld8(p + (long long)k * ld + m, v); } };
struct VAConvFwd { static constexpr bool K_CONTIG = true; const float* x; int T, Ci, left;
    __device__ __forceinline__ void load8(int m, int k, float v[8]) const {
        int b = m / T, t = m - b * T, j = k / Ci, ci = k - j * Ci, tt = t + j - left;
        if (tt < 0 || tt >= T) { zero8(v); return; }
        ld8(x + ((long long)b * T + tt) * Ci + ci, v);
    } };
struct VAConvDgrad { static constexpr bool K_CONTIG = true; const float* dy; int T, Co, left;
    __device__ __forceinline__ void load8(int m, int k, float v[8]) const {
        int b = m / T, t = m - b * T, j = k / Co, co = k - j * Co, tt = t - j + left;
        if (tt < 0 || tt >= T) { zero8(v); return; }
        ld8(dy + ((long long)b * T + tt) * Co + co, v);
    } };
struct VAConvWgrad { static constexpr bool K_CONTIG = false; const float* x; int T, Ci, left;   // m = (j, ci), k = (b, t)
    __device__ __forceinline__ void load8(int m, int k, float v[8]) const {
        int j = m / Ci, ci = m - j * Ci, b = k / T, t = k - b * T, tt = t + j - left;
        if (tt < 0 || tt >= T) { zero8(v); return; }
        ld8(x + ((long long)b * T + tt) * Ci + ci, v);
    } };
// B operands: load8(k, n, v) -- K_CONTIG: v[i] = B(k+i, n);  else: v[i] = B(k, n+i)
struct VBPlainN { static constexpr bool K_CONTIG = false; const float* p; long long ld;     // stored [K][N]
    __device__ __forceinline__ void load8(int k, int n, float v[8]) const { ld8(p + (long long)k * ld + n, v); } };
struct VBPlainK { static constexpr bool K_CONTIG = true; const float* p; long long ld;      // stored [N][K]
    __device__ __forceinline__ void load8(int k, int n, float v[8]) const { ld8(p + (long long)n * ld + k, v); } };
struct VBConvDgrad { static constexpr bool K_CONTIG = true; const float* w; int Ci, Co;      // k = (j, co), n = ci
    __device__ __forceinline__ void load8(int k, int n, float v[8]) const {
        int j = k / Co, co = k - j * Co;
        ld8(w + ((long long)j * Ci + n) * Co + co, v);
    } };

__device__ __forceinline__ uint4 pack8(const float v[8]) {
    return make_uint4((uint32_t)f2bf(v[0]) | ((uint32_t)f2bf(v[1]) << 16), (uint32_t)f2bf(v[2]) | ((uint32_t)f2bf(v[3]) << 16),
                      (uint32_t)f2bf(v[4]) | ((uint32_t)f2bf(v[5]) << 16), (uint32_t)f2bf(v[6]) | ((uint32_t)f2bf(v[7]) << 16));
}

template <class AL, class BL>
__global__ void __launch_bounds__(256)
gemm_mma2_kernel(AL A, BL Bm, float* __restrict__ C, long long ldc, int M, int N, int K, float beta, const float* __restrict__ bias,
                 int kchunk) {
    constexpr int BM = 128, BN = 128, BK = 32;
    constexpr int LDK = BK + 8;     // halves per row of a k-contiguous tile  [rows][LDK]   (80 B rows)
    constexpr int LDT = BM + 8;     // halves per row of a transposed tile    [BK][LDT]     (272 B rows)
    constexpr int TILE = BM * LDK > BK * LDT ? BM * LDK : BK * LDT;
    __shared__ __align__(16) unsigned short As[2][TILE];
    __shared__ __align__(16) unsigned short Bs[2][TILE];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    float acc[4][4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int r = 0; r < 4; ++r) acc[i][j][r] = 0.f;
    // segment ownership: k-contiguous tile: row = tid>>2 (+64), k offset 8*(tid&3); transposed tile: k row = tid>>4 (+16), offset 8*(tid&15)
    float ra[2][8], rb[2][8];
    auto gload = [&](int k0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (AL::K_CONTIG) {
                const int m = m0 + (tid >> 2) + 64 * h, k = k0 + 8 * (tid & 3);
                if (m < M && k < K) A.load8(m, k, ra[h]); else zero8(ra[h]);
            } else {
                const int k = k0 + (tid >> 4) + 16 * h, m = m0 + 8 * (tid & 15);
                if (m < M && k < K) A.load8(m, k, ra[h]); else zero8(ra[h]);
            }
            if (BL::K_CONTIG) {
                const int n = n0 + (tid >> 2) + 64 * h, k = k0 + 8 * (tid & 3);
                if (n < N && k < K) Bm.load8(k, n, rb[h]); else zero8(rb[h]);
            } else {
                const int k = k0 + (tid >> 4) + 16 * h, n = n0 + 8 * (tid & 15);
                if (n < N && k < K) Bm.load8(k, n, rb[h]); else zero8(rb[h]);
            }
        }
    };
    auto sstore = [&](int buf) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (AL::K_CONTIG) *reinterpret_cast<uint4*>(&As[buf][((tid >> 2) + 64 * h) * LDK + 8 * (tid & 3)]) = pack8(ra[h]);
            else *reinterpret_cast<uint4*>(&As[buf][((tid >> 4) + 16 * h) * LDT + 8 * (tid & 15)]) = pack8(ra[h]);
            if (BL::K_CONTIG) *reinterpret_cast<uint4*>(&Bs[buf][((tid >> 2) + 64 * h) * LDK + 8 * (tid & 3)]) = pack8(rb[h]);
            else *reinterpret_cast<uint4*>(&Bs[buf][((tid >> 4) + 16 * h) * LDT + 8 * (tid & 15)]) = pack8(rb[h]);
        }
    };
    const int kbeg = blockIdx.z * kchunk;
    if (gridDim.z > 1) K = min(K, kbeg + kchunk);
    const int nk = (K - kbeg + BK - 1) / BK;
    if (nk > 0) {
        gload(kbeg);
        sstore(0);
    }
    __syncthreads();
    const int id = lane >> 3, l7 = lane & 7;
    for (int it = 0; it < nk; ++it) {
        const int buf = it & 1;
        if (it + 1 < nk) gload(kbeg + (it + 1) * BK);
#pragma unroll
        for (int ks = 0; ks < BK; ks += 16) {
            uint32_t af[4][4], bfr[4][2];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int mb = wm * 64 + i * 16;
                if (AL::K_CONTIG) {
                    const uint32_t addr = smem_u32(&As[buf][(mb + (lane & 15)) * LDK + ks + (lane >> 4) * 8]);
                    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                                 : "=r"(af[i][0]), "=r"(af[i][1]), "=r"(af[i][2]), "=r"(af[i][3]) : "r"(addr));
                } else {   // stored [k][m]: matrices (m lo,k lo) (m hi,k lo) (m lo,k hi) (m hi,k hi), transposed on load
                    const uint32_t addr = smem_u32(&As[buf][(ks + (id >> 1) * 8 + l7) * LDT + mb + (id & 1) * 8]);
                    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                                 : "=r"(af[i][0]), "=r"(af[i][1]), "=r"(af[i][2]), "=r"(af[i][3]) : "r"(addr));
                }
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {   // two n-tiles per ldmatrix.x4: (nt0,k lo) (nt0,k hi) (nt1,k lo) (nt1,k hi)
                const int nb = wn * 32 + j * 16;
                if (BL::K_CONTIG) {
                    const uint32_t addr = smem_u32(&Bs[buf][(nb + (id >> 1) * 8 + l7) * LDK + ks + (id & 1) * 8]);
                    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                                 : "=r"(bfr[2 * j][0]), "=r"(bfr[2 * j][1]), "=r"(bfr[2 * j + 1][0]), "=r"(bfr[2 * j + 1][1]) : "r"(addr));
                } else {   // stored [k][n]
                    const uint32_t addr = smem_u32(&Bs[buf][(ks + (id & 1) * 8 + l7) * LDT + nb + (id >> 1) * 8]);
                    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                                 : "=r"(bfr[2 * j][0]), "=r"(bfr[2 * j][1]), "=r"(bfr[2 * j + 1][0]), "=r"(bfr[2 * j + 1][1]) : "r"(addr));
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                                 : "+f"(acc[i][j][0]), "+f"(acc[i][j][1]), "+f"(acc[i][j][2]), "+f"(acc[i][j][3])
                                 : "r"(af[i][0]), "r"(af[i][1]), "r"(af[i][2]), "r"(af[i][3]), "r"(bfr[j][0]), "r"(bfr[j][1]));
        }
        if (it + 1 < nk) sstore(buf ^ 1);
        __syncthreads();
    }
    const int g = lane >> 2, tq = lane & 3;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int m = m0 + wm * 64 + i * 16 + g + (r >> 1) * 8;
                const int n = n0 + wn * 32 + j * 8 + tq * 2 + (r & 1);
                if (m >= M || n >= N) continue;
                float v = acc[i][j][r];
                float* c = C + (long long)m * ldc + n;
                if (gridDim.z > 1) {
                    if (bias && blockIdx.z == 0) v += bias[n];
                    atomicAdd(c, v);
                    continue;
                }
                if (bias) v += bias[n];
                if (beta != 0.f) v += beta * *c;
                *c = v;
            }
}

template <class AL, class BL>
int launch_gemm_mma2(AL A, BL Bm, float* C, long long ldc, int M, int N, int K, float beta, const float* bias, cudaStream_t st) {
    dim3 grid((N + 127) / 128, (M + 127) / 128, 1);
    int kchunk = K;
    const int ctas = grid.x * grid.y;
    if ((beta == 0.f || beta == 1.f) && ctas < 148 && K >= 1024) {
        int splits = (296 + ctas - 1) / ctas;
        if (splits > K / 256) splits = K / 256;
        if (splits > 1) {
            kchunk = ((K + splits - 1) / splits + 31) / 32 * 32;
            grid.z = (K + kchunk - 1) / kchunk;
            if (grid.z > 1 && beta == 0.f &&
                cudaMemset2DAsync(C, (size_t)ldc * sizeof(float), 0, (size_t)N * sizeof(float), (size_t)M, st) != cudaSuccess)
                return -3;
        }
    }
    gemm_mma2_kernel<AL, BL><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, beta, bias, kchunk);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}
inline bool use_mma(int M, int N, int K) {
    if (g_train_precision != TACO_TRAIN_BF16) return false;
    if ((long long)M * N >= 64LL * 64 * 64 && (long long)M * N * K >= (1LL << 24)) return true;
    return M >= 128 && N >= 128 && K >= 4096;      // weight gradients of the narrow layers: few 128 x 128 tiles, split over K
}


// ---------------------------------------------------------------------------------------------
// conv1d forward / data gradient on the tcgen05 kernel of the synthesis path (conv_tc.cu) in the bf16 precision mode:
// operands are packed to bf16 per call (activations channel-padded to a multiple of 64, weights K-major [Cout][taps * Cin_pad];
// the data gradient is the same convolution of dY with the tap-reversed, channel-transposed kernel), fp32 output.
// Scratch buffers are process-wide and grow-only.
// ---------------------------------------------------------------------------------------------
struct TcScratch {
    void* p = nullptr;
    size_t bytes = 0;
};
TcScratch g_tc_x, g_tc_w, g_tc_b, g_tc_xt, g_tc_dyt, g_tc_dw;
bool g_conv_tc = getenv("TACO_TRAIN_CONV_TC") ? atoi(getenv("TACO_TRAIN_CONV_TC")) != 0 : true;   // 0: keep the mma.sync kernels (A/B)
int g_tc_sms = 0;

bool tc_reserve(TcScratch& s, size_t bytes) {
    if (s.bytes >= bytes) return true;
    if (s.p) cudaFree(s.p);
    s.p = nullptr;
    s.bytes = 0;
    if (cudaMalloc(&s.p, bytes) != cudaSuccess) return false;
    s.bytes = bytes;
    return true;
}

// w [k][Ci][Co] fp32 -> out [N][k * Kpad] bf16.  fwd: N = Co, channel c = ci, out[co][tap][ci] = w[tap][ci][co];
// dgrad: N = Ci, channel c = co, out[ci][tap][co] = w[k - 1 - tap][ci][co].
__global__ void pack_conv_w_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out, int k, int Ci, int Co, int Kpad, int dgrad) {
    const int N = dgrad ? Ci : Co, C = dgrad ? Co : Ci;
    const long long total = (long long)N * k * Kpad;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int c = (int)(i % Kpad);
    const int tap = (int)((i / Kpad) % k);
    const int n = (int)(i / ((long long)Kpad * k));
    float v = 0.f;
    if (c < C) v = dgrad ? w[((size_t)(k - 1 - tap) * Ci + n) * Co + c] : w[((size_t)tap * Ci + c) * Co + n];
    out[i] = __float2bfloat16_rn(v);
}

// returns 1 if the tcgen05 kernel took the call, 0 if the shape is not covered, < 0 on error
int conv_tc_train(const float* x, const float* w, const float* bias, float* y, int B, int T, int Cin, int Cout, int k, int dgrad,
                  cudaStream_t st) {
    static int dbg = getenv("TACO_TRAIN_CONV_DBG") ? 8 : 0;
    if (dbg > 0) {
        --dbg;
        fprintf(stderr, "[conv_tc_train] prec %d B %d T %d Cin %d Cout %d k %d dgrad %d bias %p\n", g_train_precision, B, T, Cin, Cout, k, dgrad, (const void*)bias);
    }
    if (g_train_precision != TACO_TRAIN_BF16) return 0;
    if (!(Cout % 256 == 0 || Cout == 80 || Cout == 128) || (long long)B * T < 4096) return 0;
    const int Cpad = (Cin + 63) / 64 * 64;
    if (!g_tc_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_tc_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    if (!tc_reserve(g_tc_x, (size_t)B * T * Cpad * 2) || !tc_reserve(g_tc_w, (size_t)Cout * k * Cpad * 2)) return -3;
    if (taco_launch_prep_bf16(x, nullptr, g_tc_x.p, B, T, (long long)T * Cin, Cin, Cpad, 0.f, 0.f, 0, st) != cudaSuccess) return -3;
    const long long nw = (long long)Cout * k * Cpad;
    // fwd: w is [k][Ci = Cin][Co = Cout]; dgrad: the caller's kernel is [k][Ci = Cout][Co = Cin]
    pack_conv_w_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, st>>>(w, reinterpret_cast<__nv_bfloat16*>(g_tc_w.p), k, dgrad ? Cout : Cin,
                                                                     dgrad ? Cin : Cout, Cpad, dgrad);
    // the epilogue reads the bias with 16-byte loads; parameter views of the flat training buffer are only 4-byte aligned
    const float* bias_al = nullptr;
    if (bias) {
        if (!tc_reserve(g_tc_b, 4096 * sizeof(float)) || Cout > 4096) return -3;
        if (cudaMemcpyAsync(g_tc_b.p, bias, (size_t)Cout * sizeof(float), cudaMemcpyDeviceToDevice, st) != cudaSuccess) return -3;
        bias_al = reinterpret_cast<const float*>(g_tc_b.p);
    }
    ConvTcParams c;
    memset(&c, 0, sizeof c);
    c.x = g_tc_x.p; c.wt = g_tc_w.p;
    c.B = B; c.F = T; c.Cin = Cpad; c.taps = k; c.Cout = Cout;
    c.scale = nullptr; c.shift = bias_al; c.act = ACT_NONE;
    if (reinterpret_cast<uintptr_t>(y) & 15) return 0;   // 16-byte stores in the epilogue
    c.y_bf16 = nullptr; c.y_stride = 0; c.y_f32 = y;
    c.out_f32_only = (Cout % 256 != 0) ? 1 : 0;
    c.residual = nullptr; c.clip = 0;
    if (taco_launch_conv_tc(c, g_tc_sms, st) != cudaSuccess) {
        cudaGetLastError();
        return 0;   // e.g. the driver entry point for tensor maps is unavailable: the mma.sync kernel takes the call
    }
    return 1;
}

// conv1d weight gradient on the same kernel (wgrad mode): dW[tap][ci][co] = sum_(b,t) x[b, t + tap - pad, ci] dY[b, t, co] with both
// operands transposed to K-major bf16 [channels][B * Tp] (zero margins of `pad` frames around every item).  1 = done, 0 = not covered.
int conv_wgrad_tc_train(const float* x, const float* dy, float* dw, int B, int T, int Ci, int Co, int k, cudaStream_t st) {
    if (g_train_precision != TACO_TRAIN_BF16 || Co % 256 != 0 || (long long)B * T < 4096) return 0;
    const int pad = (k - 1) / 2;
    const int Tp = (T + 2 * pad + 63) / 64 * 64;
    const int Ci_pad = (Ci + 127) / 128 * 128;
    const long long K = (long long)B * Tp;
    if (!g_tc_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_tc_sms, cudaDevAttrMultiProcessorCount, dev);
    }
    cudaError_t e;
#define TC_TRY(what, expr)                                                                          \
    if ((e = (expr)) != cudaSuccess) {                                                               \
        fprintf(stderr, "[taco_b200] conv wgrad (tcgen05): %s: %s\n", what, cudaGetErrorString(e)); \
        return -3;                                                                                   \
    }
    TC_TRY("scratch", (tc_reserve(g_tc_xt, (size_t)k * Ci_pad * K * 2) && tc_reserve(g_tc_dyt, (size_t)Co * K * 2)) ? cudaSuccess : cudaErrorMemoryAllocation)
    TC_TRY("transpose x", taco_launch_transpose_bf16(x, g_tc_xt.p, B, T, Ci, Ci_pad, Tp, pad, k, pad, st))
    TC_TRY("transpose dy", taco_launch_transpose_bf16(dy, g_tc_dyt.p, B, T, Co, Co, Tp, pad, 1, 0, st))
#undef TC_TRY
    // the epilogue stores 16 bytes at a time; gradient views of the flat training buffer are only 4-byte aligned
    float* out = dw;
    const size_t dw_bytes = (size_t)k * Ci * Co * sizeof(float);
    if (reinterpret_cast<uintptr_t>(dw) & 15) {
        if (!tc_reserve(g_tc_dw, dw_bytes)) return -3;
        out = reinterpret_cast<float*>(g_tc_dw.p);
    }
    if (taco_launch_conv_wgrad_tc(g_tc_xt.p, g_tc_dyt.p, out, (int)K, Ci, Ci_pad, Co, k, g_tc_sms, st) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    if (out != dw && cudaMemcpyAsync(dw, out, dw_bytes, cudaMemcpyDeviceToDevice, st) != cudaSuccess) return -3;
    return 1;
}

template <class AL, class BL>
int launch_gemm(AL A, BL Bm, float* C, long long ldc, int M, int N, int K, float beta, const float* bias, cudaStream_t st) {
    if (M <= 0 || N <= 0) return 0;
    const long long work = (long long)M * N;
    // bf16 tensor cores for the throughput-bound contractions; the skinny per-step ones stay fp32 (latency-bound)
    if (g_train_precision == TACO_TRAIN_BF16 && work >= 64LL * 64 * 64 && (long long)M * N * K >= (1LL << 24))
        return launch_gemm_mma(A, Bm, C, ldc, M, N, K, beta, bias, st);
    if (work >= 128LL * 128 * 96) {          // enough 128x128 tiles to fill the 148 SMs
        dim3 grid((N + 127) / 128, (M + 127) / 128);
        gemm_kernel<128, 128, 8, 8, 8, AL, BL><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, beta, bias, K);
    } else if (work >= 64LL * 64 * 64) {
        dim3 grid((N + 63) / 64, (M + 63) / 64);
        gemm_kernel<64, 64, 16, 4, 4, AL, BL><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, beta, bias, K);
    } else {
        // skinny per-step contractions (M = batch): latency-bound on the K loop -> split K over up to 8 CTAs
        dim3 grid((N + 31) / 32, (M + 31) / 32, 1);
        int kchunk = K;
        const int ctas = grid.x * grid.y;
        if ((beta == 0.f || beta == 1.f) && K >= 256 && ctas < 148) {
            // weight gradients with few outputs and K = all rows of the batch (up to 160 000): as many K slices as fill the GPU twice
            int splits = ((K >= 4096 ? 592 : 296) + ctas - 1) / ctas;
            const int cap = K >= 4096 ? 128 : 8;
            if (splits > cap) splits = cap;
            if (splits > K / 128) splits = K / 128;
            if (splits > 1) {
                kchunk = ((K + splits - 1) / splits + 31) / 32 * 32;
                grid.z = (K + kchunk - 1) / kchunk;
                if (grid.z > 1 && beta == 0.f &&
                    cudaMemset2DAsync(C, (size_t)ldc * sizeof(float), 0, (size_t)N * sizeof(float), (size_t)M, st) != cudaSuccess)
                    return -3;
            }
        }
        static const bool log_small = getenv("TACO_TRAIN_GEMM_LOG") != nullptr;   // which contractions take the small-tile SIMT path
        if (log_small) fprintf(stderr, "[taco gemm32] M=%d N=%d K=%d splits=%u\n", M, N, K, grid.z);
        gemm_kernel<32, 32, 32, 2, 2, AL, BL><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, beta, bias, kchunk);
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace

// The per-step (M = batch <= 32 rows) contractions of the recurrences are latency-bound, not throughput-bound: a CTA of the
// general kernel walks its K range in 32-wide stages, one global-load round trip per stage.  This kernel takes a 32 x 32 output
// tile and a K chunk of KC, issues ALL loads of both operand chunks at once (2 x KC/4 per thread in flight), stages them in shared
// memory with one barrier, contracts from there (2 x 4 outputs per thread) and accumulates with atomics (split-K over blockIdx.z).
template <int KC, class AL>
__global__ void __launch_bounds__(128) skinny_acc_kernel(AL A, BPlain Bm, float* __restrict__ C, long long ldc, float* __restrict__ C2,
                                                         long long ldc2, int N1, int M, int N, int K) {
    constexpr int NT = 128, PER = 32 * KC / NT;
    __shared__ float As[KC][33];
    __shared__ __align__(16) float Bs[KC][36];
    const int tid = threadIdx.x, m0 = blockIdx.y * 32, n0 = blockIdx.x * 32, k0 = blockIdx.z * KC;
    float ra[PER], rb[PER];
    // B is the weight operand (constant during a time loop): its loads go out before the programmatic-launch wait, i.e. while
    // the kernel that produces A may still be running
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int e = tid + i * NT;
        int nn, kk;
        if (Bm.n_contig) { nn = e & 31; kk = e >> 5; } else { kk = e % KC; nn = e / KC; }
        rb[i] = (n0 + nn < N && k0 + kk < K) ? Bm(k0 + kk, n0 + nn) : 0.f;
    }
    pdl_wait();
    pdl_launch_dependents();
#pragma unroll
    for (int i = 0; i < PER; ++i) {              // A is k-contiguous in both uses (activations [M, K])
        const int e = tid + i * NT, kk = e % KC, mm = e / KC;
        ra[i] = (m0 + mm < M && k0 + kk < K) ? A(m0 + mm, k0 + kk) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        const int e = tid + i * NT;
        As[e % KC][e / KC] = ra[i];
        if (Bm.n_contig) Bs[e >> 5][e & 31] = rb[i]; else Bs[e % KC][e / KC] = rb[i];
    }
    __syncthreads();
    const int ty = tid >> 3, tx = tid & 7;       // rows 2 ty, 2 ty + 1; columns 4 tx .. 4 tx + 3
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll 16
    for (int k = 0; k < KC; ++k) {
        const float a0 = As[k][2 * ty], a1 = As[k][2 * ty + 1];
        const float4 b = *reinterpret_cast<const float4*>(&Bs[k][4 * tx]);
        acc[0][0] = fmaf(a0, b.x, acc[0][0]); acc[0][1] = fmaf(a0, b.y, acc[0][1]); acc[0][2] = fmaf(a0, b.z, acc[0][2]); acc[0][3] = fmaf(a0, b.w, acc[0][3]);
        acc[1][0] = fmaf(a1, b.x, acc[1][0]); acc[1][1] = fmaf(a1, b.y, acc[1][1]); acc[1][2] = fmaf(a1, b.z, acc[1][2]); acc[1][3] = fmaf(a1, b.w, acc[1][3]);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int m = m0 + 2 * ty + i;
        if (m >= M) continue;
        const int nb = n0 + 4 * tx;
        float* c4 = (C2 && nb >= N1) ? C2 + (long long)m * ldc2 + (nb - N1) : C + (long long)m * ldc + nb;
        if (nb + 3 < N && (!C2 || nb + 3 < N1 || nb >= N1) && (reinterpret_cast<uintptr_t>(c4) & 15) == 0) {
            atomicAdd(reinterpret_cast<float4*>(c4), make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));   // one 16-byte reduction
            continue;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = nb + j;
            if (n >= N) continue;
            atomicAdd((C2 && n >= N1) ? C2 + (long long)m * ldc2 + (n - N1) : C + (long long)m * ldc + n, acc[i][j]);
        }
    }
}

// The same kernel with 16-byte operand loads for the shapes of the time loops (K % 64 == 0, N % 32 == 0, every row of both operands
// 16-byte aligned -- train/params.py aligns the variables of the flat parameter buffer for this): 4 + 4 loads per thread instead of
// 16 + 16, no per-element index arithmetic, 6 LDS.128 per 32 FMA in the contraction.  BT: B is stored [N][K] (k-contiguous).
__device__ __forceinline__ const float* a_row_ptr(const APlain& A, int m, int k0) { return A.p + (long long)m * A.sm + k0; }
__device__ __forceinline__ const float* a_row_ptr(const APair& A, int m, int k0) {
    return k0 < A.K1 ? A.p1 + (long long)m * A.ld1 + k0 : A.p2 + (long long)m * A.ld2 + (k0 - A.K1);
}
template <class AL, bool BT>
__global__ void __launch_bounds__(128) skinny_acc_vec_kernel(AL A, const float* __restrict__ Bp, long long ldb, float* __restrict__ C,
                                                             long long ldc, float* __restrict__ C2, long long ldc2, int N1, int M) {
    constexpr int KC = 64, NT = 128, AS = KC + 4;      // row stride of the k-contiguous tiles: 16-byte aligned rows
    __shared__ __align__(16) float As[32][AS];
    __shared__ __align__(16) float Bs[BT ? 32 * AS : KC * 36];   // BT: [n][k] (stride AS), else [k][n] (stride 36)
    const int tid = threadIdx.x, m0 = blockIdx.y * 32, n0 = blockIdx.x * 32, k0 = blockIdx.z * KC;
    float4 ra[4], rb[4];
    // the weight operand does not depend on the previous kernel of the chain: requested before the programmatic-launch wait
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int e = tid + i * NT;
        if (BT) rb[i] = *reinterpret_cast<const float4*>(Bp + (long long)(n0 + (e >> 4)) * ldb + k0 + 4 * (e & 15));
        else rb[i] = *reinterpret_cast<const float4*>(Bp + (long long)(k0 + (e >> 3)) * ldb + n0 + 4 * (e & 7));
    }
    pdl_wait();
    pdl_launch_dependents();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int e = tid + i * NT, mm = e >> 4;
        ra[i] = (m0 + mm < M) ? *reinterpret_cast<const float4*>(a_row_ptr(A, m0 + mm, k0) + 4 * (e & 15)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int e = tid + i * NT;
        *reinterpret_cast<float4*>(&As[e >> 4][4 * (e & 15)]) = ra[i];
        if (BT) *reinterpret_cast<float4*>(&Bs[(((e >> 4) >> 2) + 8 * ((e >> 4) & 3)) * AS + 4 * (e & 15)]) = rb[i];   // row n at (n / 4) + 8 (n % 4): conflict-free reads below
        else *reinterpret_cast<float4*>(&Bs[(e >> 3) * 36 + 4 * (e & 7)]) = rb[i];
    }
    __syncthreads();
    const int ty = tid >> 3, tx = tid & 7;       // rows 2 ty, 2 ty + 1; columns 4 tx .. 4 tx + 3
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll 4
    for (int k4 = 0; k4 < KC / 4; ++k4) {
        const float4 a0 = *reinterpret_cast<const float4*>(&As[2 * ty][4 * k4]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[2 * ty + 1][4 * k4]);
        const float a0s[4] = {a0.x, a0.y, a0.z, a0.w}, a1s[4] = {a1.x, a1.y, a1.z, a1.w};
        if (BT) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 b = *reinterpret_cast<const float4*>(&Bs[(tx + 8 * j) * AS + 4 * k4]);     // column 4 tx + j
                acc[0][j] = fmaf(a0.x, b.x, acc[0][j]); acc[0][j] = fmaf(a0.y, b.y, acc[0][j]); acc[0][j] = fmaf(a0.z, b.z, acc[0][j]); acc[0][j] = fmaf(a0.w, b.w, acc[0][j]);
                acc[1][j] = fmaf(a1.x, b.x, acc[1][j]); acc[1][j] = fmaf(a1.y, b.y, acc[1][j]); acc[1][j] = fmaf(a1.z, b.z, acc[1][j]); acc[1][j] = fmaf(a1.w, b.w, acc[1][j]);
            }
        } else {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const float4 b = *reinterpret_cast<const float4*>(&Bs[(4 * k4 + kk) * 36 + 4 * tx]);
                acc[0][0] = fmaf(a0s[kk], b.x, acc[0][0]); acc[0][1] = fmaf(a0s[kk], b.y, acc[0][1]); acc[0][2] = fmaf(a0s[kk], b.z, acc[0][2]); acc[0][3] = fmaf(a0s[kk], b.w, acc[0][3]);
                acc[1][0] = fmaf(a1s[kk], b.x, acc[1][0]); acc[1][1] = fmaf(a1s[kk], b.y, acc[1][1]); acc[1][2] = fmaf(a1s[kk], b.z, acc[1][2]); acc[1][3] = fmaf(a1s[kk], b.w, acc[1][3]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int m = m0 + 2 * ty + i;
        if (m >= M) continue;
        const int nb = n0 + 4 * tx;             // N1 % 32 == 0: a thread's 4 columns never straddle the two outputs
        float* c4 = (C2 && nb >= N1) ? C2 + (long long)m * ldc2 + (nb - N1) : C + (long long)m * ldc + nb;
        atomicAdd(reinterpret_cast<float4*>(c4), make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));   // one 16-byte reduction
    }
}

inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
inline bool vec_ok(const APlain& A, int K) { return A.sk == 1 && A.sm % 4 == 0 && al16(A.p); }
inline bool vec_ok(const APair& A, int K) { return A.K1 % 64 == 0 && A.ld1 % 4 == 0 && A.ld2 % 4 == 0 && al16(A.p1) && al16(A.p2); }

// skinny (M = batch) contraction with a two-block A and/or a two-block C, always accumulating (beta = 1), split-K
template <class AL>
int launch_skinny_acc(AL A, BPlain Bm, float* C, long long ldc, float* C2, long long ldc2, int N1, int M, int N, int K, cudaStream_t st) {
    static const bool one_shot = !getenv("TACO_TRAIN_SKINNY_OLD");
    static const bool vec = !getenv("TACO_TRAIN_SKINNY_SCALAR");
    if (one_shot && vec && K >= 128 && K % 64 == 0 && N % 32 == 0 && vec_ok(A, K) && al16(Bm.p) && al16(C) && ldc % 4 == 0 &&
        (Bm.n_contig ? (Bm.sn == 1 && Bm.sk % 4 == 0) : (Bm.sk == 1 && Bm.sn % 4 == 0)) &&
        (!C2 || (N1 % 32 == 0 && al16(C2) && ldc2 % 4 == 0))) {
        const dim3 grid(N / 32, (M + 31) / 32, K / 64);
        if (Bm.n_contig) launch_pdl(skinny_acc_vec_kernel<AL, false>, grid, dim3(128), 0, st, A, Bm.p, Bm.sk, C, ldc, C2, ldc2, N1, M);
        else launch_pdl(skinny_acc_vec_kernel<AL, true>, grid, dim3(128), 0, st, A, Bm.p, Bm.sn, C, ldc, C2, ldc2, N1, M);
        return cudaGetLastError() == cudaSuccess ? 0 : -3;
    }
    if (one_shot && K >= 128) {
        const int tiles = ((N + 31) / 32) * ((M + 31) / 32);
        static const int kc_env = getenv("TACO_SKINNY_KC") ? atoi(getenv("TACO_SKINNY_KC")) : 64;   // measured: 64 beats 128 at K = 1024
        if (kc_env == 128 && tiles * ((K + 127) / 128) >= 200) {
            launch_pdl(skinny_acc_kernel<128, AL>, dim3((N + 31) / 32, (M + 31) / 32, (K + 127) / 128), dim3(128), 0, st, A, Bm, C, ldc, C2, ldc2, N1, M, N, K);
        } else {
            launch_pdl(skinny_acc_kernel<64, AL>, dim3((N + 31) / 32, (M + 31) / 32, (K + 63) / 64), dim3(128), 0, st, A, Bm, C, ldc, C2, ldc2, N1, M, N, K);
        }
        return cudaGetLastError() == cudaSuccess ? 0 : -3;
    }
    dim3 grid((N + 31) / 32, (M + 31) / 32, 1);
    int kchunk = K;
    const int ctas = grid.x * grid.y;
    if (K >= 256 && ctas < 148) {
        int splits = (296 + ctas - 1) / ctas;
        if (splits > 8) splits = 8;
        if (splits > K / 128) splits = K / 128;
        if (splits > 1) {
            kchunk = ((K + splits - 1) / splits + 31) / 32 * 32;
            grid.z = (K + kchunk - 1) / kchunk;
        }
    }
    gemm_kernel<32, 32, 32, 2, 2, AL, BPlain><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, 1.f, nullptr, kchunk, C2, ldc2, N1);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

// Reference-encoder conv2d weight gradient as a split-K implicit GEMM (called from train_misc.cu): every (tap, ci) x co tile
// reads its share of X and dY once, instead of one CTA per (tap, ci) row re-reading all of dY.  Returns 1 when the shape
// does not qualify (the caller then runs its own kernel), 0 on success, < 0 on error.
int taco_train_conv2d_wgrad_gemm(const float* x, const float* dy, float* dw, int B, int H, int W, int Ci, int Co, int Ho, int Wo,
                                 int pt, int pl, cudaStream_t st) {
    const long long npos = (long long)B * Ho * Wo;
    const int M = 9 * Ci;
    if (Ci < 8 || npos < 256 || npos >= (1LL << 31) - 65536) return 1;
    if (cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)M * Co, st) != cudaSuccess) return -3;
    const AConv2dWgrad A{x, H, W, Ci, Ho, Wo, pt, pl, 0};
    const BPlain Bm{dy, Co, 1, 1};
    const bool big = Co >= 64 && M >= 256;
    const int tile = big ? 64 : 32;
    dim3 grid((Co + tile - 1) / tile, (M + tile - 1) / tile, 1);
    const int ctas = grid.x * grid.y;
    int splits = (592 + ctas - 1) / ctas;
    if (splits > (int)(npos / 256)) splits = (int)(npos / 256);
    if (splits < 1) splits = 1;
    const int kchunk = (int)(((npos + splits - 1) / splits + 31) / 32 * 32);
    grid.z = (unsigned)((npos + kchunk - 1) / kchunk);
    if (big) gemm_kernel<64, 64, 16, 4, 4, AConv2dWgrad, BPlain><<<grid, 256, 0, st>>>(A, Bm, dw, Co, M, Co, (int)npos, 0.f, nullptr, kchunk);
    else gemm_kernel<32, 32, 32, 2, 2, AConv2dWgrad, BPlain><<<grid, 256, 0, st>>>(A, Bm, dw, Co, M, Co, (int)npos, 0.f, nullptr, kchunk);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

extern "C" {

/* C += [A1 | A2] . B   (A1 [M,K1], A2 [M,K2] row-major, B [K1+K2, N] row-major): the x- and h-part of an LSTM gate contraction */
int taco_op_gemm_pair_acc(const float* A1, long long lda1, int K1, const float* A2, long long lda2, int K2, const float* B, long long ldb,
                          float* C, long long ldc, int M, int N, void* stream) {
    if (!A1 || !A2 || !B || !C || M < 1 || N < 1 || K1 < 1 || K2 < 1) return -1;
    return launch_skinny_acc(APair{A1, A2, lda1, lda2, K1, 1}, BPlain{B, ldb, 1, 1}, C, ldc, nullptr, 0, 0, M, N, K1 + K2, (cudaStream_t)stream);
}
/* [C1 | C2] += A . B^T   (A [M,K], B stored [N1+N2, K] row-major): the gradient of the same contraction w.r.t. its two inputs */
int taco_op_gemm_nt_split_acc(const float* A, long long lda, const float* B, long long ldb, float* C1, long long ldc1, int N1, float* C2,
                              long long ldc2, int N2, int M, int K, void* stream) {
    if (!A || !B || !C1 || !C2 || M < 1 || N1 < 1 || N2 < 1 || K < 1) return -1;
    return launch_skinny_acc(APlain{A, lda, 1, 1}, BPlain{B, 1, ldb, 0}, C1, ldc1, C2, ldc2, N1, M, N1 + N2, K, (cudaStream_t)stream);
}

/* C += A . B (tb = 0, B [K,N]) or A . B^T (tb = 1, B [N,K]) for M = batch rows: the recurrent contraction of the encoder LSTM loops */
int taco_op_gemm_acc(const float* A, long long lda, const float* B, long long ldb, float* C, long long ldc, int M, int N, int K, int tb,
                     void* stream) {
    if (!A || !B || !C || M < 1 || N < 1 || K < 1) return -1;
    return launch_skinny_acc(APlain{A, lda, 1, 1}, tb ? BPlain{B, 1, ldb, 0} : BPlain{B, ldb, 1, 1}, C, ldc, nullptr, 0, 0, M, N, K,
                             (cudaStream_t)stream);
}

int taco_op_set_precision(int precision) {
    if (precision != TACO_TRAIN_FP32 && precision != TACO_TRAIN_BF16) return -1;
    g_train_precision = precision;
    return 0;
}

int taco_op_gemm(const float* A, long long lda, const float* B, long long ldb, float* C, long long ldc, int M, int N, int K,
                 int ta, int tb, float beta, const float* bias, void* stream) {
    if (!A || !B || !C || K < 0) return -1;
    if (use_mma(M, N, K) && K % 8 == 0 && M % 8 == 0 && N % 8 == 0 && lda % 4 == 0 && ldb % 4 == 0) {
        cudaStream_t st = (cudaStream_t)stream;
        if (!ta && !tb) return launch_gemm_mma2(VAPlainK{A, lda}, VBPlainN{B, ldb}, C, ldc, M, N, K, beta, bias, st);
        if (ta && !tb) return launch_gemm_mma2(VAPlainM{A, lda}, VBPlainN{B, ldb}, C, ldc, M, N, K, beta, bias, st);
        if (!ta && tb) return launch_gemm_mma2(VAPlainK{A, lda}, VBPlainK{B, ldb}, C, ldc, M, N, K, beta, bias, st);
        return launch_gemm_mma2(VAPlainM{A, lda}, VBPlainK{B, ldb}, C, ldc, M, N, K, beta, bias, st);
    }
    APlain a{A, ta ? 1 : lda, ta ? lda : 1, ta ? 0 : 1};
    BPlain b{B, tb ? 1 : ldb, tb ? ldb : 1, tb ? 0 : 1};
    return launch_gemm(a, b, C, ldc, M, N, K, beta, bias, (cudaStream_t)stream);
}

int taco_op_conv1d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int T, int Ci, int Co, int k, void* stream) {
    if (!x || !w || !y) return -1;
    if (g_conv_tc) {
        const int r = conv_tc_train(x, w, bias, y, B, T, Ci, Co, k, 0, (cudaStream_t)stream);
        if (r != 0) return r < 0 ? r : 0;
    }
    if (use_mma(B * T, Co, k * Ci) && Ci % 8 == 0 && Co % 8 == 0)
        return launch_gemm_mma2(VAConvFwd{x, T, Ci, (k - 1) / 2}, VBPlainN{w, Co}, y, Co, B * T, Co, k * Ci, 0.f, bias, (cudaStream_t)stream);
    AConvFwd a{x, T, Ci, (k - 1) / 2, 1};
    BPlain b{w, Co, 1, 1};
    return launch_gemm(a, b, y, Co, B * T, Co, k * Ci, 0.f, bias, (cudaStream_t)stream);
}

int taco_op_conv1d_dgrad(const float* dy, const float* w, float* dx, int B, int T, int Ci, int Co, int k, void* stream) {
    if (!dy || !w || !dx) return -1;
    if (g_conv_tc) {
        const int r = conv_tc_train(dy, w, nullptr, dx, B, T, Co, Ci, k, 1, (cudaStream_t)stream);
        if (r != 0) return r < 0 ? r : 0;
    }
    if (use_mma(B * T, Ci, k * Co) && Ci % 8 == 0 && Co % 8 == 0)
        return launch_gemm_mma2(VAConvDgrad{dy, T, Co, (k - 1) / 2}, VBConvDgrad{w, Ci, Co}, dx, Ci, B * T, Ci, k * Co, 0.f, nullptr,
                                (cudaStream_t)stream);
    AConvDgrad a{dy, T, Co, (k - 1) / 2, 1};
    BConvDgrad b{w, Ci, Co, 0};
    return launch_gemm(a, b, dx, Ci, B * T, Ci, k * Co, 0.f, nullptr, (cudaStream_t)stream);
}

int taco_op_conv1d_wgrad(const float* x, const float* dy, float* dw, int B, int T, int Ci, int Co, int k, void* stream) {
    if (!x || !dy || !dw) return -1;
    if (g_conv_tc) {
        const int r = conv_wgrad_tc_train(x, dy, dw, B, T, Ci, Co, k, (cudaStream_t)stream);
        if (r != 0) return r < 0 ? r : 0;
    }
    if (use_mma(k * Ci, Co, B * T) && Ci % 8 == 0 && Co % 8 == 0 && (B * T) % 8 == 0)
        return launch_gemm_mma2(VAConvWgrad{x, T, Ci, (k - 1) / 2}, VBPlainN{dy, Co}, dw, Co, k * Ci, Co, B * T, 0.f, nullptr,
                                (cudaStream_t)stream);
    AConvWgrad a{x, T, Ci, (k - 1) / 2, 0};
    BPlain b{dy, Co, 1, 1};
    return launch_gemm(a, b, dw, Co, k * Ci, Co, B * T, 0.f, nullptr, (cudaStream_t)stream);
}

}  // extern "C"
