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

#include "../../include/taco_b200_train.h"

namespace {

struct APlain {   // A(m,k) = A[m*sm + k*sk]
    const float* p; long long sm, sk; int k_contig;
    __device__ __forceinline__ float operator()(int m, int k) const { return p[m * sm + k * sk]; }
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

template <int BM, int BN, int BK, int TM, int TN, class AL, class BL>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_kernel(AL A, BL Bm, float* __restrict__ C, long long ldc, int M, int N, int K, float beta, const float* __restrict__ bias) {
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
    const int nk = (K + BK - 1) / BK;
    gload(0);
    sstore(0);
    __syncthreads();
    for (int it = 0; it < nk; ++it) {
        const int buf = it & 1;
        if (it + 1 < nk) gload((it + 1) * BK);
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
            if (bias) v += bias[n];
            float* c = C + (long long)m * ldc + n;
            if (beta != 0.f) v += beta * *c;
            *c = v;
        }
    }
}

template <class AL, class BL>
int launch_gemm(AL A, BL Bm, float* C, long long ldc, int M, int N, int K, float beta, const float* bias, cudaStream_t st) {
    if (M <= 0 || N <= 0) return 0;
    const long long work = (long long)M * N;
    if (work >= 128LL * 128 * 96) {          // enough 128x128 tiles to fill the 148 SMs
        dim3 grid((N + 127) / 128, (M + 127) / 128);
        gemm_kernel<128, 128, 8, 8, 8, AL, BL><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, beta, bias);
    } else if (work >= 64LL * 64 * 64) {
        dim3 grid((N + 63) / 64, (M + 63) / 64);
        gemm_kernel<64, 64, 16, 4, 4, AL, BL><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, beta, bias);
    } else {
        dim3 grid((N + 31) / 32, (M + 31) / 32);
        gemm_kernel<32, 32, 32, 2, 2, AL, BL><<<grid, 256, 0, st>>>(A, Bm, C, ldc, M, N, K, beta, bias);
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace

extern "C" {

int taco_op_gemm(const float* A, long long lda, const float* B, long long ldb, float* C, long long ldc, int M, int N, int K,
                 int ta, int tb, float beta, const float* bias, void* stream) {
    if (!A || !B || !C || K < 0) return -1;
    APlain a{A, ta ? 1 : lda, ta ? lda : 1, ta ? 0 : 1};
    BPlain b{B, tb ? 1 : ldb, tb ? ldb : 1, tb ? 0 : 1};
    return launch_gemm(a, b, C, ldc, M, N, K, beta, bias, (cudaStream_t)stream);
}

int taco_op_conv1d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int T, int Ci, int Co, int k, void* stream) {
    if (!x || !w || !y) return -1;
    AConvFwd a{x, T, Ci, (k - 1) / 2, 1};
    BPlain b{w, Co, 1, 1};
    return launch_gemm(a, b, y, Co, B * T, Co, k * Ci, 0.f, bias, (cudaStream_t)stream);
}

int taco_op_conv1d_dgrad(const float* dy, const float* w, float* dx, int B, int T, int Ci, int Co, int k, void* stream) {
    if (!dy || !w || !dx) return -1;
    AConvDgrad a{dy, T, Co, (k - 1) / 2, 1};
    BConvDgrad b{w, Ci, Co, 0};
    return launch_gemm(a, b, dx, Ci, B * T, Ci, k * Co, 0.f, nullptr, (cudaStream_t)stream);
}

int taco_op_conv1d_wgrad(const float* x, const float* dy, float* dw, int B, int T, int Ci, int Co, int k, void* stream) {
    if (!x || !dy || !dw) return -1;
    AConvWgrad a{x, T, Ci, (k - 1) / 2, 0};
    BPlain b{dy, Co, 1, 1};
    return launch_gemm(a, b, dw, Co, k * Ci, Co, B * T, 0.f, nullptr, (cudaStream_t)stream);
}

}  // extern "C"
