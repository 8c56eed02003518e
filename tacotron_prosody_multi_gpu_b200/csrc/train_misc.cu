// train_misc.cu -- reference-encoder conv2d(3x3, stride 2, 'same') forward / gradients and the GST multi-head
// attention forward / gradient of the training step (Encoder.py:171-182; Tacotron.py:12-29; Attention.py:243-341).
// Small tensors (< 2 GFLOP per step in total): direct kernels, one thread per output element.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/taco_b200_train.h"
#include "common.cuh"

int taco_train_conv2d_wgrad_gemm(const float* x, const float* dy, float* dw, int B, int H, int W, int Ci, int Co, int Ho, int Wo,
                                 int pt, int pl, cudaStream_t st);   // train_gemm.cu

namespace {

__host__ __device__ inline int pad_before(int n) {   // SAME padding of a 3-tap stride-2 window
    int out = (n + 1) / 2;
    int total = (out - 1) * 2 + 3 - n;
    if (total < 0) total = 0;
    return total / 2;
}

__global__ void conv2d_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                  float* __restrict__ y, int B, int H, int W, int Ci, int Co, int Ho, int Wo, int pt, int pl) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * Ho * Wo * Co) return;
    int co = (int)(i % Co);
    long long r = i / Co;
    int wo = (int)(r % Wo); r /= Wo;
    int ho = (int)(r % Ho);
    int b = (int)(r / Ho);
    float acc = bias ? bias[co] : 0.f;
    for (int kh = 0; kh < 3; ++kh) {
        int h = 2 * ho + kh - pt;
        if (h < 0 || h >= H) continue;
        for (int kw = 0; kw < 3; ++kw) {
            int ww = 2 * wo + kw - pl;
            if (ww < 0 || ww >= W) continue;
            const float* xp = x + (((long long)b * H + h) * W + ww) * Ci;
            const float* wp = w + ((long long)(kh * 3 + kw) * Ci) * Co + co;
            for (int ci = 0; ci < Ci; ++ci) acc = fmaf(xp[ci], wp[(long long)ci * Co], acc);
        }
    }
    y[i] = acc;
}

__global__ void conv2d_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w, float* __restrict__ dx, int B, int H,
                                    int W, int Ci, int Co, int Ho, int Wo, int pt, int pl) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * H * W * Ci) return;
    int ci = (int)(i % Ci);
    long long r = i / Ci;
    int ww = (int)(r % W); r /= W;
    int h = (int)(r % H);
    int b = (int)(r / H);
    float acc = 0.f;
    for (int kh = 0; kh < 3; ++kh) {
        int hh = h + pt - kh;
        if (hh < 0 || (hh & 1)) continue;
        int ho = hh >> 1;
        if (ho >= Ho) continue;
        for (int kw = 0; kw < 3; ++kw) {
            int wq = ww + pl - kw;
            if (wq < 0 || (wq & 1)) continue;
            int wo = wq >> 1;
            if (wo >= Wo) continue;
            const float* dp = dy + (((long long)b * Ho + ho) * Wo + wo) * Co;
            const float* wp = w + ((long long)(kh * 3 + kw) * Ci + ci) * Co;
            if ((Co & 3) == 0) {                  // 16-byte loads: d(y) is a warp broadcast, the kernel row is private
                for (int co = 0; co < Co; co += 4) {
                    const float4 d4 = *reinterpret_cast<const float4*>(dp + co);
                    const float4 w4 = *reinterpret_cast<const float4*>(wp + co);
                    acc = fmaf(d4.x, w4.x, acc); acc = fmaf(d4.y, w4.y, acc); acc = fmaf(d4.z, w4.z, acc); acc = fmaf(d4.w, w4.w, acc);
                }
            } else {
                for (int co = 0; co < Co; ++co) acc = fmaf(dp[co], wp[co], acc);
            }
        }
    }
    dx[i] = acc;
}

// grid (9*Ci, splits); block 256 = Co lanes x (256/Co) position lanes
__global__ void conv2d_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* dw, int B, int H, int W, int Ci,
                                    int Co, int Ho, int Wo, int pt, int pl) {
    __shared__ float sh[256];
    const int kci = blockIdx.x, ci = kci % Ci, k = kci / Ci, kh = k / 3, kw = k % 3;
    const int co = threadIdx.x % Co, pl_lane = threadIdx.x / Co, n_pl = blockDim.x / Co;
    const int npos = B * Ho * Wo;                 // < 2^31 (checked by the launcher): 32-bit index arithmetic
    float acc = 0.f;
#pragma unroll 4
    for (int p = blockIdx.y * n_pl + pl_lane; p < npos; p += gridDim.y * n_pl) {
        const int r = p / Wo, wo = p - r * Wo;
        const int b = r / Ho, ho = r - b * Ho;
        int h = 2 * ho + kh - pt, ww = 2 * wo + kw - pl;
        const bool ok = h >= 0 && h < H && ww >= 0 && ww < W;      // predicated (no branch): loads of 4 iterations in flight
        const float xv = ok ? x[(((long long)b * H + h) * W + ww) * Ci + ci] : 0.f;
        acc = fmaf(xv, dy[(long long)p * Co + co], acc);
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    if (pl_lane == 0) {
        for (int l = 1; l < n_pl; ++l) acc += sh[l * Co + co];
        atomicAdd(dw + (long long)kci * Co + co, acc);
    }
}

// ---- im2col / col2im: the three conv2d contractions as plain GEMMs on taco_op_gemm (bf16 tensor cores in the bf16 mode) ----
// col[m = (b,ho,wo)][k = (kh*3+kw)*Ci + ci] = x[b, 2 ho + kh - pt, 2 wo + kw - pl, ci]   (0 outside the image)
template <int VEC>
__global__ void im2col_s2_kernel(const float* __restrict__ x, float* __restrict__ col, int H, int W, int Ci, int Ho, int Wo, int pt,
                                 int pl, long long n /* M * 9 * Ci / VEC */) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int cv = Ci / VEC;
    const int c = (int)(i % cv);
    long long r = i / cv;
    const int tap = (int)(r % 9); r /= 9;
    const int wo = (int)(r % Wo); r /= Wo;
    const int ho = (int)(r % Ho);
    const int b = (int)(r / Ho);
    const int h = 2 * ho + tap / 3 - pt, w = 2 * wo + tap % 3 - pl;
    const bool ok = h >= 0 && h < H && w >= 0 && w < W;
    const float* src = x + (((long long)b * H + h) * W + w) * Ci + c * VEC;
    if (VEC == 4) reinterpret_cast<float4*>(col)[i] = ok ? *reinterpret_cast<const float4*>(src) : make_float4(0.f, 0.f, 0.f, 0.f);
    else col[i] = ok ? *src : 0.f;
}
// dx[b,h,w,ci] = sum over the taps that reach (h,w): dcol[(b,ho,wo)][(kh*3+kw)*Ci + ci]
template <int VEC>
__global__ void col2im_s2_kernel(const float* __restrict__ dcol, float* __restrict__ dx, int H, int W, int Ci, int Ho, int Wo, int pt,
                                 int pl, long long n /* B*H*W*Ci / VEC */) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int cv = Ci / VEC;
    const int c = (int)(i % cv);
    long long r = i / cv;
    const int w = (int)(r % W); r /= W;
    const int h = (int)(r % H);
    const int b = (int)(r / H);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
        const int hh = h + pt - kh;
        if (hh < 0 || (hh & 1) || (hh >> 1) >= Ho) continue;
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) {
            const int ww = w + pl - kw;
            if (ww < 0 || (ww & 1) || (ww >> 1) >= Wo) continue;
            const float* src = dcol + ((((long long)b * Ho + (hh >> 1)) * Wo + (ww >> 1)) * 9 + kh * 3 + kw) * Ci + c * VEC;
            if (VEC == 4) { const float4 v = *reinterpret_cast<const float4*>(src); acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
            else acc.x += *src;
        }
    }
    if (VEC == 4) reinterpret_cast<float4*>(dx)[i] = acc; else dx[i] = acc.x;
}

struct Scratch { float* p = nullptr; size_t n = 0; };
Scratch g_col;     // process-wide, grow-only (cudaMalloc: 16-byte aligned)
float* col_scratch(size_t n) {
    if (g_col.n < n) {
        if (g_col.p) { cudaDeviceSynchronize(); cudaFree(g_col.p); }
        g_col.p = nullptr; g_col.n = 0;
        if (cudaMalloc(&g_col.p, n * sizeof(float)) != cudaSuccess) return nullptr;
        g_col.n = n;
    }
    return g_col.p;
}
const bool g_conv2d_gemm = !getenv("TACO_TRAIN_CONV2D_DIRECT");

int im2col(const float* x, float* col, int B, int H, int W, int Ci, int Ho, int Wo, cudaStream_t st) {
    const bool v4 = Ci % 4 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
    const long long n = (long long)B * Ho * Wo * 9 * Ci / (v4 ? 4 : 1);
    if (v4) im2col_s2_kernel<4><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, col, H, W, Ci, Ho, Wo, pad_before(H), pad_before(W), n);
    else im2col_s2_kernel<1><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(x, col, H, W, Ci, Ho, Wo, pad_before(H), pad_before(W), n);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

// ------------------------------------------------------------------------------------------------ GST attention
// block per batch row, one warp per head (head width U/nh == 32 lanes); N <= 32 tokens
__global__ void gst_fwd_kernel(const float* __restrict__ q, const float* __restrict__ k, const float* __restrict__ v,
                               const float* __restrict__ g, const float* __restrict__ bb, const float* __restrict__ tt,
                               float* __restrict__ wts, float* __restrict__ style, int N, int U, int D, int nh) {
    const int b = blockIdx.x, h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (h >= nh) return;
    const float vv = v[lane];
    const float nv = g[0] * vv * rsqrtf(warp_sum(vv * vv));
    const float qb = q[(long long)b * U + h * 32 + lane] + bb[lane];
    float score = -INFINITY;
    for (int n = 0; n < N; ++n) {
        float s = warp_sum(nv * tanhf(k[(long long)n * U + h * 32 + lane] + qb));
        if (lane == n) score = s;
    }
    float mx = warp_max(score);
    float e = lane < N ? expf(score - mx) : 0.f;
    float w = e / warp_sum(e);
    if (lane < N) wts[((long long)b * nh + h) * N + lane] = w;
    for (int e0 = lane; e0 < D; e0 += 32) {
        float acc = 0.f;
        for (int n = 0; n < N; ++n) acc = fmaf(__shfl_sync(0xffffffffu, w, n), tt[n * D + e0], acc);
        style[(long long)b * nh * D + h * D + e0] = acc;
    }
}

__global__ void gst_bwd_kernel(const float* __restrict__ dstyle, const float* __restrict__ q, const float* __restrict__ k,
                               const float* __restrict__ v, const float* __restrict__ g, const float* __restrict__ bb,
                               const float* __restrict__ tt, const float* __restrict__ wts, float* __restrict__ dq, float* dk, float* db,
                               float* dtt, float* dnv, int N, int U, int D, int nh) {
    const int b = blockIdx.x, h = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (h >= nh) return;
    const float vv = v[lane];
    const float nv = g[0] * vv * rsqrtf(warp_sum(vv * vv));
    const float qb = q[(long long)b * U + h * 32 + lane] + bb[lane];
    const float* ds = dstyle + (long long)b * nh * D + h * D;
    const float w = lane < N ? wts[((long long)b * nh + h) * N + lane] : 0.f;
    float dw = 0.f;
    if (lane < N)
        for (int e = 0; e < D; ++e) dw = fmaf(ds[e], tt[lane * D + e], dw);
    for (int i0 = 0; i0 < N * D; i0 += 32) {          // d(tt)[n, e] += w[n] * d(style)[e]
        const int i = i0 + lane, ic = i < N * D ? i : N * D - 1;
        const float wn = __shfl_sync(0xffffffffu, w, ic / D);
        if (i < N * D) atomicAdd(dtt + i, wn * ds[i % D]);
    }
    const float sdot = warp_sum(w * dw);
    const float dscore = w * (dw - sdot);
    float dnv_acc = 0.f, dq_acc = 0.f;
    for (int n = 0; n < N; ++n) {
        float dsn = __shfl_sync(0xffffffffu, dscore, n);
        float th = tanhf(k[(long long)n * U + h * 32 + lane] + qb);
        dnv_acc = fmaf(dsn, th, dnv_acc);
        float dpre = dsn * nv * (1.f - th * th);
        dq_acc += dpre;
        atomicAdd(dk + (long long)n * U + h * 32 + lane, dpre);
    }
    dq[(long long)b * U + h * 32 + lane] = dq_acc;
    atomicAdd(db + lane, dq_acc);
    atomicAdd(dnv + lane, dnv_acc);
}

// weight normalisation nv = g * v / ||v||: d(v), d(g) from d(nv); one warp
__global__ void gst_wn_bwd_kernel(const float* __restrict__ v, const float* __restrict__ g, const float* __restrict__ dnv,
                                  float* __restrict__ dv, float* __restrict__ dg) {
    const int lane = threadIdx.x;
    const float vv = v[lane], d = dnv[lane];
    const float rs = rsqrtf(warp_sum(vv * vv));
    const float s = warp_sum(d * vv);
    dv[lane] = g[0] * rs * d - g[0] * vv * rs * rs * rs * s;
    if (lane == 0) dg[0] = s * rs;
}

}  // namespace

#define ST ((cudaStream_t)stream)
static inline int rcm() { return cudaGetLastError() == cudaSuccess ? 0 : -3; }

extern "C" {

int taco_op_conv2d_fwd(const float* x, const float* w, const float* bias, float* y, int B, int H, int W, int Ci, int Co, void* stream) {
    if (!x || !w || !y || B < 1 || H < 1 || W < 1) return -1;
    const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
    long long n = (long long)B * Ho * Wo * Co;
    if (g_conv2d_gemm && Ci >= 8 && n / Co < (1LL << 31) / (9 * Ci)) {       // y = im2col(x) . w + bias
        const int M = B * Ho * Wo, K = 9 * Ci;
        float* col = col_scratch((size_t)M * K);
        if (!col) return -3;
        int rc = im2col(x, col, B, H, W, Ci, Ho, Wo, ST);
        return rc ? rc : taco_op_gemm(col, K, w, Co, y, Co, M, Co, K, 0, 0, 0.f, bias, stream);
    }
    conv2d_fwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ST>>>(x, w, bias, y, B, H, W, Ci, Co, Ho, Wo, pad_before(H), pad_before(W));
    return rcm();
}
int taco_op_conv2d_dgrad(const float* dy, const float* w, float* dx, int B, int H, int W, int Ci, int Co, void* stream) {
    if (!dy || !w || !dx) return -1;
    const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
    long long n = (long long)B * H * W * Ci;
    if (g_conv2d_gemm && Ci >= 8 && (long long)B * Ho * Wo < (1LL << 31) / (9 * Ci)) {   // d(col) = dy . w^T, then col2im
        const int M = B * Ho * Wo, K = 9 * Ci;
        float* col = col_scratch((size_t)M * K);
        if (!col) return -3;
        int rc = taco_op_gemm(dy, Co, w, Co, col, K, M, K, Co, 0, 1, 0.f, nullptr, stream);
        if (rc) return rc;
        const bool v4 = Ci % 4 == 0 && (reinterpret_cast<uintptr_t>(dx) & 15) == 0;
        const long long nv = n / (v4 ? 4 : 1);
        if (v4) col2im_s2_kernel<4><<<(unsigned)((nv + 255) / 256), 256, 0, ST>>>(col, dx, H, W, Ci, Ho, Wo, pad_before(H), pad_before(W), nv);
        else col2im_s2_kernel<1><<<(unsigned)((nv + 255) / 256), 256, 0, ST>>>(col, dx, H, W, Ci, Ho, Wo, pad_before(H), pad_before(W), nv);
        return rcm();
    }
    conv2d_dgrad_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ST>>>(dy, w, dx, B, H, W, Ci, Co, Ho, Wo, pad_before(H), pad_before(W));
    return rcm();
}
int taco_op_conv2d_wgrad(const float* x, const float* dy, float* dw, int B, int H, int W, int Ci, int Co, void* stream) {
    if (!x || !dy || !dw) return -1;
    if (Co < 1 || Co > 256 || 256 % Co != 0) return -1;
    const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
    if (g_conv2d_gemm && Ci >= 8 && (long long)B * Ho * Wo < (1LL << 31) / (9 * Ci)) {   // d(w) = im2col(x)^T . dy
        const int M = B * Ho * Wo, K = 9 * Ci;
        float* col = col_scratch((size_t)M * K);
        if (!col) return -3;
        int rc = im2col(x, col, B, H, W, Ci, Ho, Wo, ST);
        return rc ? rc : taco_op_gemm(col, K, dy, Co, dw, Co, K, Co, M, 1, 0, 0.f, nullptr, stream);
    }
    {   // Ci >= 8: split-K implicit GEMM of train_gemm.cu (fp32); the first layer (Ci = 1) keeps the row-per-CTA kernel below
        const int rc = taco_train_conv2d_wgrad_gemm(x, dy, dw, B, H, W, Ci, Co, Ho, Wo, pad_before(H), pad_before(W), ST);
        if (rc <= 0) return rc;
    }
    if (cudaMemsetAsync(dw, 0, sizeof(float) * 9 * (size_t)Ci * Co, ST) != cudaSuccess) return -3;
    const long long npos = (long long)B * Ho * Wo;
    if (npos >= (1LL << 31) - 65536) return -1;
    const int n_pl = 256 / Co;
    int splits = (int)((npos / n_pl + 63) / 64);          // >= 64 positions per thread
    int cap = (148 * 16 + 9 * Ci - 1) / (9 * Ci);
    if (splits > cap) splits = cap;
    if (splits < 1) splits = 1;
    conv2d_wgrad_kernel<<<dim3(9 * Ci, splits), 256, 0, ST>>>(x, dy, dw, B, H, W, Ci, Co, Ho, Wo, pad_before(H), pad_before(W));
    return rcm();
}
int taco_op_gst_fwd(const float* q, const float* k, const float* v, const float* g, const float* b, const float* tt,
                    float* wts, float* style, int B, int N, int U, int D, int nh, void* stream) {
    if (!q || !k || !v || !g || !b || !tt || !wts || !style) return -1;
    if (nh < 1 || nh > 32 || U != nh * 32 || N < 1 || N > 32) return -1;
    gst_fwd_kernel<<<B, nh * 32, 0, ST>>>(q, k, v, g, b, tt, wts, style, N, U, D, nh);
    return rcm();
}
int taco_op_gst_bwd(const float* dstyle, const float* q, const float* k, const float* v, const float* g, const float* b,
                    const float* tt, const float* wts, float* dq, float* dk, float* dv, float* dg, float* db, float* dtt,
                    float* scratch, int B, int N, int U, int D, int nh, void* stream) {
    if (!dstyle || !q || !k || !v || !g || !b || !tt || !wts || !dq || !dk || !dv || !dg || !db || !dtt || !scratch) return -1;
    if (nh < 1 || nh > 32 || U != nh * 32 || N < 1 || N > 32) return -1;
    if (cudaMemsetAsync(dk, 0, sizeof(float) * (size_t)N * U, ST) != cudaSuccess) return -3;
    if (cudaMemsetAsync(db, 0, sizeof(float) * 32, ST) != cudaSuccess) return -3;
    if (cudaMemsetAsync(dtt, 0, sizeof(float) * (size_t)N * D, ST) != cudaSuccess) return -3;
    if (cudaMemsetAsync(scratch, 0, sizeof(float) * 32, ST) != cudaSuccess) return -3;
    gst_bwd_kernel<<<B, nh * 32, 0, ST>>>(dstyle, q, k, v, g, b, tt, wts, dq, dk, db, dtt, scratch, N, U, D, nh);
    gst_wn_bwd_kernel<<<1, 32, 0, ST>>>(v, g, scratch, dv, dg);
    return rcm();
}

}  // extern "C"
