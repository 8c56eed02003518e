// train_ops.cu -- pointwise / reduction primitives of the training step (include/taco_b200_train.h).
// HBM-bound elementwise work: one pass per tensor, coalesced along the channel axis, block reductions
// for the per-channel statistics.  The attention step and the conv2d / GST kernels are in train_attn.cu.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/taco_b200_train.h"
#include "common.cuh"

namespace {

constexpr float BN_EPS = 1e-3f;   // tf.layers.batch_normalization default epsilon

__device__ __forceinline__ bool train_keep(uint64_t seed, uint32_t stream, unsigned long long idx, uint32_t thresh) {
    unsigned long long blk = idx >> 2;
    uint4 o = philox4x32_10(make_uint4((uint32_t)blk, stream, (uint32_t)(blk >> 32), 0x54524E31u),
                            make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    uint32_t l = (uint32_t)idx & 3u;
    uint32_t w = l == 0 ? o.x : (l == 1 ? o.y : (l == 2 ? o.z : o.w));
    return w < thresh;
}

inline int rc() { return cudaGetLastError() == cudaSuccess ? 0 : -3; }
inline unsigned nblk(long long n, int t = 256) { return (unsigned)((n + t - 1) / t); }

__device__ __forceinline__ float block_sum(float v, float* sh) {   // sh: >= 32 floats
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    v = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.f;
    if (w == 0) v = warp_sum(v);
    return v;   // valid in warp 0
}

// ---------------------------------------------------------------------------------------------- elementwise
__global__ void ew_kernel(int op, float* __restrict__ out, const float* __restrict__ a, const float* __restrict__ b, float alpha,
                          long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float x = a[i], r;
    switch (op) {
        case TACO_EW_COPY: r = x; break;
        case TACO_EW_ADD: r = x + alpha * b[i]; break;
        case TACO_EW_MUL: r = x * b[i]; break;
        case TACO_EW_SCALE: r = x * alpha; break;
        case TACO_EW_TANH: r = tanhf(x); break;
        default: { float y = b[i]; r = x * (1.f - y * y); } break;   // TANH_BWD: a = dy, b = y
    }
    out[i] = r;
}

__global__ void clip_fwd_kernel(const float* __restrict__ x, float lo, float hi, float* __restrict__ y, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = fminf(fmaxf(x[i], lo), hi);
}
__global__ void clip_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ x, float lo, float hi, float* __restrict__ dx,
                                long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { float v = x[i]; dx[i] = (v >= lo && v <= hi) ? dy[i] : 0.f; }
}

__global__ void mse_kernel(const float* __restrict__ p, const float* __restrict__ t, float* loss, float* __restrict__ dp, long long n) {
    __shared__ float sh[32];
    float acc = 0.f;
    const float inv = 1.0f / (float)n;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float d = p[i] - t[i];
        acc += d * d;
        dp[i] = 2.f * d * inv;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) atomicAdd(loss, acc * inv);
}
__global__ void sigmoid_ce_kernel(const float* __restrict__ lg, const float* __restrict__ y, float* loss, float* __restrict__ dl,
                                  long long n) {
    __shared__ float sh[32];
    float acc = 0.f;
    const float inv = 1.0f / (float)n;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float l = lg[i], t = y[i];
        acc += fmaxf(l, 0.f) - l * t + log1pf(expf(-fabsf(l)));
        dl[i] = (sigmoidf_(l) - t) * inv;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) atomicAdd(loss, acc * inv);
}

// ---------------------------------------------------------------------------------------------- per-channel reductions
// block (32, 8): 32 adjacent channels x 8 row lanes; rows strided by 8*gridDim.y; partials combined with atomics when
// gridDim.y > 1 (outputs pre-zeroed by the launcher).
template <class F>
__device__ __forceinline__ void col_reduce2(int M, int C, F f, float& s0, float& s1) {
    const int c = blockIdx.x * 32 + threadIdx.x;
    s0 = s1 = 0.f;
    if (c < C)
        for (int m = blockIdx.y * blockDim.y + threadIdx.y; m < M; m += gridDim.y * blockDim.y) f(m, c, s0, s1);
    __shared__ float sh0[8][33], sh1[8][33];
    sh0[threadIdx.y][threadIdx.x] = s0;
    sh1[threadIdx.y][threadIdx.x] = s1;
    __syncthreads();
    if (threadIdx.y == 0) {
        for (int i = 1; i < 8; ++i) { s0 += sh0[i][threadIdx.x]; s1 += sh1[i][threadIdx.x]; }
    }
}

__global__ void colsum_kernel(const float* __restrict__ x, long long ld, int M, int C, float* out) {
    float s0, s1;
    col_reduce2(M, C, [&](int m, int c, float& a, float&) { a += x[(long long)m * ld + c]; }, s0, s1);
    const int c = blockIdx.x * 32 + threadIdx.x;
    if (threadIdx.y == 0 && c < C) atomicAdd(out + c, s0);
}
__global__ void scale_kernel(float* x, float a, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) x[i] *= a;
}

// BN statistics: pass 1 sum -> mean; pass 2 centred sum of squares (two-pass for accuracy)
__global__ void bn_sum_kernel(const float* __restrict__ x, int M, int C, float* sum) {
    float s0, s1;
    col_reduce2(M, C, [&](int m, int c, float& a, float&) { a += x[(long long)m * C + c]; }, s0, s1);
    const int c = blockIdx.x * 32 + threadIdx.x;
    if (threadIdx.y == 0 && c < C) atomicAdd(sum + c, s0);
}
__global__ void bn_var_kernel(const float* __restrict__ x, int M, int C, const float* __restrict__ sum, float* sq) {
    const float invM = 1.0f / (float)M;
    float s0, s1;
    col_reduce2(M, C, [&](int m, int c, float& a, float&) { float d = x[(long long)m * C + c] - sum[c] * invM; a += d * d; }, s0, s1);
    const int c = blockIdx.x * 32 + threadIdx.x;
    if (threadIdx.y == 0 && c < C) atomicAdd(sq + c, s0);
}
__global__ void bn_finish_kernel(float* mean /* in: sum */, float* rstd /* in: centred sumsq */, int M, int C) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) {
        mean[c] = mean[c] / (float)M;
        rstd[c] = rsqrtf(rstd[c] / (float)M + BN_EPS);
    }
}
__device__ __forceinline__ float act_f(float a, int act) { return act == TACO_ACT_RELU ? fmaxf(a, 0.f) : (act == TACO_ACT_TANH ? tanhf(a) : a); }

__global__ void bn_apply_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                                const float* __restrict__ mean, const float* __restrict__ rstd, int act, uint32_t thresh, float keep,
                                uint64_t seed, uint32_t stream, float* __restrict__ y, long long n, int C) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c = (int)(i % C);
    float a = act_f((x[i] - mean[c]) * rstd[c] * gamma[c] + beta[c], act);
    if (keep < 1.f) a = train_keep(seed, stream, (unsigned long long)i, thresh) ? a / keep : 0.f;
    y[i] = a;
}
// dz = d(loss)/d(BN output before the activation), recomputed from dy, y and the dropout stream
__device__ __forceinline__ float bn_dz(float dy, float y, int act, uint32_t thresh, float keep, uint64_t seed, uint32_t stream,
                                       unsigned long long i) {
    float da = dy, a = y;
    if (keep < 1.f) {
        if (!train_keep(seed, stream, i, thresh)) return 0.f;
        da = dy / keep;
        a = y * keep;
    }
    if (act == TACO_ACT_RELU) return a > 0.f ? da : 0.f;
    if (act == TACO_ACT_TANH) return da * (1.f - a * a);
    return da;
}
__global__ void bn_bwd_reduce_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ y,
                                     const float* __restrict__ mean, const float* __restrict__ rstd, int act, uint32_t thresh,
                                     float keep, uint64_t seed, uint32_t stream, float* dgamma, float* dbeta, int M, int C) {
    float s0, s1;
    col_reduce2(M, C, [&](int m, int c, float& a, float& b) {
        long long i = (long long)m * C + c;
        float dz = bn_dz(dy[i], y[i], act, thresh, keep, seed, stream, (unsigned long long)i);
        a += dz;
        b += dz * (x[i] - mean[c]) * rstd[c];
    }, s0, s1);
    const int c = blockIdx.x * 32 + threadIdx.x;
    if (threadIdx.y == 0 && c < C) { atomicAdd(dbeta + c, s0); atomicAdd(dgamma + c, s1); }
}
__global__ void bn_bwd_apply_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ y,
                                    const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                                    const float* __restrict__ dgamma, const float* __restrict__ dbeta, int act, uint32_t thresh,
                                    float keep, uint64_t seed, uint32_t stream, float* __restrict__ dx, long long n, int C, float invM) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c = (int)(i % C);
    float dz = bn_dz(dy[i], y[i], act, thresh, keep, seed, stream, (unsigned long long)i);
    float xhat = (x[i] - mean[c]) * rstd[c];
    dx[i] = gamma[c] * rstd[c] * (dz - dbeta[c] * invM - xhat * dgamma[c] * invM);
}
__global__ void bn_moving_kernel(const float* mean, const float* rstd, float* mm, float* mv, float mom, int C) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) {
        float var = 1.f / (rstd[c] * rstd[c]) - BN_EPS;
        mm[c] = mom * mm[c] + (1.f - mom) * mean[c];
        mv[c] = mom * mv[c] + (1.f - mom) * var;
    }
}

// ---- 16-byte variants (C % 4 == 0, 16-byte aligned tensors): a thread owns 4 adjacent channels, one Philox block serves its 4
// dropout bits (train_keep draws element i from word i & 3 of block i >> 2), four rows in flight per thread in the reductions ----
__device__ __forceinline__ uint4 keep_block4(uint64_t seed, uint32_t stream, unsigned long long i0) {   // i0 % 4 == 0
    const unsigned long long blk = i0 >> 2;
    return philox4x32_10(make_uint4((uint32_t)blk, stream, (uint32_t)(blk >> 32), 0x54524E31u), make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
}
__device__ __forceinline__ float4 ld4(const float* p, long long i4) { return reinterpret_cast<const float4*>(p)[i4]; }
__device__ __forceinline__ float dz1(float dy, float y, int act, bool kept, float keep) {
    float da = dy, a = y;
    if (keep < 1.f) {
        if (!kept) return 0.f;
        da = dy / keep;
        a = y * keep;
    }
    if (act == TACO_ACT_RELU) return a > 0.f ? da : 0.f;
    if (act == TACO_ACT_TANH) return da * (1.f - a * a);
    return da;
}
__device__ __forceinline__ float4 bn_dz4(float4 dy, float4 y, int act, uint32_t thresh, float keep, uint64_t seed, uint32_t stream,
                                         unsigned long long i0) {
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (keep < 1.f) o = keep_block4(seed, stream, i0);
    return make_float4(dz1(dy.x, y.x, act, o.x < thresh, keep), dz1(dy.y, y.y, act, o.y < thresh, keep),
                       dz1(dy.z, y.z, act, o.z < thresh, keep), dz1(dy.w, y.w, act, o.w < thresh, keep));
}
// block (32, 8): 32 channel quads x 8 row lanes; f(m, c4, a, b) accumulates into two float4
template <class F>
__device__ __forceinline__ void col_reduce_v4(int M, int C4, F f, float4& s0, float4& s1) {
    const int c4 = blockIdx.x * 32 + threadIdx.x;
    s0 = s1 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c4 < C4) {
#pragma unroll 4
        for (int m = blockIdx.y * blockDim.y + threadIdx.y; m < M; m += gridDim.y * blockDim.y) f(m, c4, s0, s1);
    }
    __shared__ float4 sh0[8][33], sh1[8][33];
    sh0[threadIdx.y][threadIdx.x] = s0;
    sh1[threadIdx.y][threadIdx.x] = s1;
    __syncthreads();
    if (threadIdx.y == 0) {
        for (int i = 1; i < 8; ++i) {
            const float4 a = sh0[i][threadIdx.x], b = sh1[i][threadIdx.x];
            s0.x += a.x; s0.y += a.y; s0.z += a.z; s0.w += a.w;
            s1.x += b.x; s1.y += b.y; s1.z += b.z; s1.w += b.w;
        }
    }
}
__device__ __forceinline__ void atomic_add4(float* p, float4 v) { atomicAdd(p, v.x); atomicAdd(p + 1, v.y); atomicAdd(p + 2, v.z); atomicAdd(p + 3, v.w); }

__global__ void bn_sum_v4_kernel(const float* __restrict__ x, int M, int C4, float* sum) {
    float4 s0, s1;
    col_reduce_v4(M, C4, [&](int m, int c4, float4& a, float4&) {
        const float4 v = ld4(x, (long long)m * C4 + c4);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }, s0, s1);
    const int c4 = blockIdx.x * 32 + threadIdx.x;
    if (threadIdx.y == 0 && c4 < C4) atomic_add4(sum + 4 * c4, s0);
}
__global__ void bn_var_v4_kernel(const float* __restrict__ x, int M, int C4, const float* __restrict__ sum, float* sq) {
    const float invM = 1.0f / (float)M;
    const int c4q = blockIdx.x * 32 + threadIdx.x;
    float4 mu = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c4q < C4) { mu = ld4(sum, c4q); mu.x *= invM; mu.y *= invM; mu.z *= invM; mu.w *= invM; }
    float4 s0, s1;
    col_reduce_v4(M, C4, [&](int m, int c4, float4& a, float4&) {
        const float4 v = ld4(x, (long long)m * C4 + c4);
        const float d0 = v.x - mu.x, d1 = v.y - mu.y, d2 = v.z - mu.z, d3 = v.w - mu.w;
        a.x += d0 * d0; a.y += d1 * d1; a.z += d2 * d2; a.w += d3 * d3;
    }, s0, s1);
    if (threadIdx.y == 0 && c4q < C4) atomic_add4(sq + 4 * c4q, s0);
}
__global__ void bn_apply_v4_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                                   const float* __restrict__ mean, const float* __restrict__ rstd, int act, uint32_t thresh, float keep,
                                   uint64_t seed, uint32_t stream, float* __restrict__ y, long long n4, int C4) {
    const long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i4 >= n4) return;
    const int c4 = (int)(i4 % C4);
    const float4 v = ld4(x, i4), mu = ld4(mean, c4), rs = ld4(rstd, c4), ga = ld4(gamma, c4), be = ld4(beta, c4);
    float4 a = make_float4(act_f((v.x - mu.x) * rs.x * ga.x + be.x, act), act_f((v.y - mu.y) * rs.y * ga.y + be.y, act),
                           act_f((v.z - mu.z) * rs.z * ga.z + be.z, act), act_f((v.w - mu.w) * rs.w * ga.w + be.w, act));
    if (keep < 1.f) {
        const uint4 o = keep_block4(seed, stream, (unsigned long long)i4 * 4ull);
        a.x = o.x < thresh ? a.x / keep : 0.f; a.y = o.y < thresh ? a.y / keep : 0.f;
        a.z = o.z < thresh ? a.z / keep : 0.f; a.w = o.w < thresh ? a.w / keep : 0.f;
    }
    reinterpret_cast<float4*>(y)[i4] = a;
}
__global__ void bn_bwd_reduce_v4_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ y,
                                        const float* __restrict__ mean, const float* __restrict__ rstd, int act, uint32_t thresh,
                                        float keep, uint64_t seed, uint32_t stream, float* dgamma, float* dbeta, int M, int C4) {
    const int c4q = blockIdx.x * 32 + threadIdx.x;
    float4 mu = make_float4(0.f, 0.f, 0.f, 0.f), rs = mu;
    if (c4q < C4) { mu = ld4(mean, c4q); rs = ld4(rstd, c4q); }
    float4 s0, s1;
    col_reduce_v4(M, C4, [&](int m, int c4, float4& a, float4& b) {
        const long long i4 = (long long)m * C4 + c4;
        const float4 xv = ld4(x, i4);
        const float4 dz = bn_dz4(ld4(dy, i4), ld4(y, i4), act, thresh, keep, seed, stream, (unsigned long long)i4 * 4ull);
        a.x += dz.x; a.y += dz.y; a.z += dz.z; a.w += dz.w;
        b.x += dz.x * (xv.x - mu.x) * rs.x; b.y += dz.y * (xv.y - mu.y) * rs.y;
        b.z += dz.z * (xv.z - mu.z) * rs.z; b.w += dz.w * (xv.w - mu.w) * rs.w;
    }, s0, s1);
    if (threadIdx.y == 0 && c4q < C4) { atomic_add4(dbeta + 4 * c4q, s0); atomic_add4(dgamma + 4 * c4q, s1); }
}
__global__ void bn_bwd_apply_v4_kernel(const float* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ y,
                                       const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                                       const float* __restrict__ dgamma, const float* __restrict__ dbeta, int act, uint32_t thresh,
                                       float keep, uint64_t seed, uint32_t stream, float* __restrict__ dx, long long n4, int C4, float invM) {
    const long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i4 >= n4) return;
    const int c4 = (int)(i4 % C4);
    const float4 dz = bn_dz4(ld4(dy, i4), ld4(y, i4), act, thresh, keep, seed, stream, (unsigned long long)i4 * 4ull);
    const float4 xv = ld4(x, i4), mu = ld4(mean, c4), rs = ld4(rstd, c4), ga = ld4(gamma, c4), dg = ld4(dgamma, c4), db = ld4(dbeta, c4);
    float4 o;
    o.x = ga.x * rs.x * (dz.x - db.x * invM - (xv.x - mu.x) * rs.x * dg.x * invM);
    o.y = ga.y * rs.y * (dz.y - db.y * invM - (xv.y - mu.y) * rs.y * dg.y * invM);
    o.z = ga.z * rs.z * (dz.z - db.z * invM - (xv.z - mu.z) * rs.z * dg.z * invM);
    o.w = ga.w * rs.w * (dz.w - db.w * invM - (xv.w - mu.w) * rs.w * dg.w * invM);
    reinterpret_cast<float4*>(dx)[i4] = o;
}
inline bool al16p(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
const bool g_bn_vec = !getenv("TACO_TRAIN_BN_SCALAR");
inline dim3 col_grid_v4(int M, int C4) {
    int gx = (C4 + 31) / 32;
    int gy = 1;
    while (gx * gy < 592 && gy * 8 * 16 < M) gy *= 2;
    return dim3(gx, gy);
}

inline dim3 col_grid(int M, int C) {
    int gx = (C + 31) / 32;
    int gy = 1;
    while (gx * gy < 592 && gy * 8 * 16 < M) gy *= 2;   // ~4 CTAs per SM, >= 16 rows per thread
    return dim3(gx, gy);
}

}  // namespace

namespace {

// ---------------------------------------------------------------------------------------------- LSTM / zoneout
// optional sequence I/O of the encoder time loops fused into the pointwise kernels: gather of the step's input row from
// [B, T, .] at position s (or len - 1 - s) and scatter of the step's output row to it -- valid rows (s < len) only
struct SeqIO {
    const float* gsrc; long long gsb, gst;
    float* sdst; long long ssb, sst;
    int reverse;
};
__global__ void lstm_fwd_kernel(float* __restrict__ gates, const float* __restrict__ pre, const float* __restrict__ bias,
                                const float* __restrict__ c_prev, const float* __restrict__ h_prev, const int32_t* __restrict__ lens,
                                int s, float zoneout, int train, uint64_t seed, uint32_t sc, uint32_t sh, int row0, uint32_t thresh,
                                float* __restrict__ cnew, float* __restrict__ cz, float* __restrict__ hraw, float* __restrict__ hz,
                                int B, int H, SeqIO io) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * H) return;
    int b = i / H, u = i - b * H;
    const int len_b = lens ? lens[b] : 0x7fffffff;
    const bool valid = s < len_b;
    const long long pos = (io.reverse && lens) ? (long long)len_b - 1 - s : s;
    float mc, mh;                                  // zoneout masks: counter-based, computed under the predecessor's tail
    if (train) {
        unsigned long long idx = (unsigned long long)(row0 + b) * H + u;
        mc = train_keep(seed, sc, idx, thresh) ? 1.f : 0.f;
        mh = train_keep(seed, sh, idx, thresh) ? 1.f : 0.f;
    } else {
        mc = mh = 1.f - zoneout;
    }
    pdl_wait();
    pdl_launch_dependents();
    float z[4];
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        long long k = (long long)b * 4 * H + g * H + u;
        float v = gates[k];
        if (pre) v += pre[k];
        if (io.gsrc && valid) v += io.gsrc[(long long)b * io.gsb + pos * io.gst + g * H + u];   // the gathered row of the input projection
        if (bias) v += bias[g * H + u];
        z[g] = v;
    }
    float si = sigmoidf_(z[0]), tj = tanhf(z[1]), sf = sigmoidf_(z[2] + 1.0f), so = sigmoidf_(z[3]);
    long long k0 = (long long)b * 4 * H + u;
    gates[k0] = si; gates[k0 + H] = tj; gates[k0 + 2 * H] = sf; gates[k0 + 3 * H] = so;
    float cp = c_prev[i], hp = h_prev[i];
    float cn = sf * cp + si * tj;
    float hr = so * tanhf(cn);
    float czv = cp + mc * (cn - cp), hzv = hp + mh * (hr - hp);
    if (!valid) { czv = cp; hzv = hp; hr = 0.f; }
    cnew[i] = cn; cz[i] = czv; hraw[i] = hr; hz[i] = hzv;
    if (io.sdst && valid) io.sdst[(long long)b * io.ssb + pos * io.sst + u] = hr;               // scattered to its sequence position
}

__global__ void lstm_bwd_kernel(const float* __restrict__ acts, const float* __restrict__ c_prev, const float* __restrict__ h_prev,
                                const float* __restrict__ cnew, const int32_t* __restrict__ lens, int s, float zoneout, int train,
                                uint64_t seed, uint32_t sc, uint32_t sh, int row0, uint32_t thresh, const float* __restrict__ d_hraw,
                                const float* __restrict__ d_cz, const float* __restrict__ d_hz, float* __restrict__ dgates,
                                float* __restrict__ d_cprev, float* __restrict__ d_hprev, int B, int H, SeqIO io) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * H) return;
    int b = i / H, u = i - b * H;
    long long k0 = (long long)b * 4 * H + u;
    const int len_b = lens ? lens[b] : 0x7fffffff;
    const long long pos = (io.reverse && lens) ? (long long)len_b - 1 - s : s;
    float mc, mh;
    if (train) {
        unsigned long long idx = (unsigned long long)(row0 + b) * H + u;
        mc = train_keep(seed, sc, idx, thresh) ? 1.f : 0.f;
        mh = train_keep(seed, sh, idx, thresh) ? 1.f : 0.f;
    } else {
        mc = mh = 1.f - zoneout;
    }
    pdl_wait();
    pdl_launch_dependents();
    if (lens && s >= lens[b]) {
        dgates[k0] = 0.f; dgates[k0 + H] = 0.f; dgates[k0 + 2 * H] = 0.f; dgates[k0 + 3 * H] = 0.f;
        d_cprev[i] = d_cz[i];
        d_hprev[i] = d_hz[i];
        return;
    }
    float si = acts[k0], tj = acts[k0 + H], sf = acts[k0 + 2 * H], so = acts[k0 + 3 * H];
    float tc = tanhf(cnew[i]);
    float dhz = d_hz[i], dcz = d_cz[i];
    const float dhr = io.gsrc ? io.gsrc[(long long)b * io.gsb + pos * io.gst + u] : d_hraw[i];   // valid row (invalid ones returned above)
    float dh_new = dhr + mh * dhz;
    d_hprev[i] = (1.f - mh) * dhz;
    float dc_new = mc * dcz + dh_new * so * (1.f - tc * tc);
    d_cprev[i] = (1.f - mc) * dcz + dc_new * sf;
    const float g0 = dc_new * tj * si * (1.f - si), g1 = dc_new * si * (1.f - tj * tj), g2 = dc_new * c_prev[i] * sf * (1.f - sf),
                g3 = dh_new * tc * so * (1.f - so);
    dgates[k0] = g0; dgates[k0 + H] = g1; dgates[k0 + 2 * H] = g2; dgates[k0 + 3 * H] = g3;
    if (io.sdst) {                                // d(input projection) row at its sequence position
        float* o = io.sdst + (long long)b * io.ssb + pos * io.sst + u;
        o[0] = g0; o[H] = g1; o[2 * H] = g2; o[3 * H] = g3;
    }
}

__global__ void gather_rows_kernel(const float* __restrict__ src, long long sb, long long st, const int32_t* __restrict__ lens, int s,
                                   int reverse, float* __restrict__ out, int B, int N) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * N) return;
    int b = i / N, c = i - b * N, len = lens[b];
    pdl_wait();
    pdl_launch_dependents();
    float v = 0.f;
    if (s < len) v = src[(long long)b * sb + (long long)(reverse ? len - 1 - s : s) * st + c];
    out[i] = v;
}
__global__ void scatter_rows_kernel(const float* __restrict__ src, const int32_t* __restrict__ lens, int s, int reverse,
                                    float* __restrict__ dst, long long db, long long dt, int B, int N) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * N) return;
    int b = i / N, c = i - b * N, len = lens[b];
    pdl_wait();
    pdl_launch_dependents();
    if (s < len) dst[(long long)b * db + (long long)(reverse ? len - 1 - s : s) * dt + c] = src[i];
}

// ---------------------------------------------------------------------------------------------- GRU pointwise
__global__ void gru_pre_kernel(float* __restrict__ g, const float* __restrict__ h, float* __restrict__ rh, int B, int R) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * 2 * R) return;
    int b = i / (2 * R), c = i - b * 2 * R;
    float v = sigmoidf_(g[i]);
    g[i] = v;
    if (c < R) rh[b * R + c] = v * h[b * R + c];
}
__global__ void gru_out_kernel(float* __restrict__ cpre, const float* __restrict__ ru, const float* __restrict__ h,
                               float* __restrict__ hnew, int B, int R) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * R) return;
    int b = i / R, c = i - b * R;
    float cc = tanhf(cpre[i]);
    cpre[i] = cc;
    float u = ru[b * 2 * R + R + c];
    hnew[i] = u * h[i] + (1.f - u) * cc;
}
__global__ void gru_out_bwd_kernel(const float* __restrict__ dhnew, const float* __restrict__ ru, const float* __restrict__ h,
                                   const float* __restrict__ c, float* __restrict__ dcpre, float* __restrict__ dg,
                                   float* __restrict__ dh, int B, int R) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * R) return;
    int b = i / R, k = i - b * R;
    float u = ru[b * 2 * R + R + k], cc = c[i], d = dhnew[i];
    dcpre[i] = d * (1.f - u) * (1.f - cc * cc);
    dg[b * 2 * R + R + k] = d * (h[i] - cc) * u * (1.f - u);
    dh[i] = d * u;
}
__global__ void gru_pre_bwd_kernel(const float* __restrict__ drh, const float* __restrict__ ru, const float* __restrict__ h,
                                   float* __restrict__ dg, float* __restrict__ dh, int B, int R) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * R) return;
    int b = i / R, k = i - b * R;
    float r = ru[b * 2 * R + k], d = drh[i];
    dg[b * 2 * R + k] = d * h[i] * r * (1.f - r);
    dh[i] += d * r;
}

// ---------------------------------------------------------------------------------------------- embedding / prenet / masks
__global__ void emb_fwd_kernel(const int32_t* __restrict__ ids, const float* __restrict__ table, float* __restrict__ out, int n_tok,
                               int dim, int n_sym) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n_tok * dim) return;
    int t = (int)(i / dim), c = (int)(i - (long long)t * dim);
    int id = ids[t];
    id = id < 0 ? 0 : (id >= n_sym ? n_sym - 1 : id);
    out[i] = table[(long long)id * dim + c];
}
__global__ void emb_bwd_kernel(const int32_t* __restrict__ ids, const float* __restrict__ dout, float* dtable, int n_tok, int dim,
                               int n_sym) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n_tok * dim) return;
    int t = (int)(i / dim), c = (int)(i - (long long)t * dim);
    int id = ids[t];
    id = id < 0 ? 0 : (id >= n_sym ? n_sym - 1 : id);
    atomicAdd(dtable + (long long)id * dim + c, dout[i]);
}
__global__ void prenet_act_kernel(const float* __restrict__ z, const float* __restrict__ dout, int layer, uint64_t seed, int step0,
                                  float keep, float* __restrict__ out, int S, int B, int U, int bwd) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)S * B * U) return;
    int u = (int)(i % U);
    long long sb = i / U;
    int b = (int)(sb % B), s = (int)(sb / B);
    bool k = philox_keep(seed, (uint32_t)(step0 + s), (uint32_t)b, (uint32_t)layer, (uint32_t)u);
    if (!bwd) out[i] = k ? fmaxf(z[i], 0.f) / keep : 0.f;
    else out[i] = (k && z[i] > 0.f) ? dout[i] / keep : 0.f;   // z = saved forward OUTPUT (> 0 iff relu passed and kept)
}
__global__ void mask_rows_kernel(float* __restrict__ x, const int32_t* __restrict__ lens, int B, int T, int C) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * T * C) return;
    long long bt = i / C;
    int t = (int)(bt % T), b = (int)(bt / T);
    if (t >= lens[b]) x[i] = 0.f;
}
__global__ void tile_style_kernel(const float* __restrict__ style, float* __restrict__ memory, int B, int T, int SD, int MD, int col0) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * T * SD) return;
    int c = (int)(i % SD);
    long long bt = i / SD;
    int b = (int)(bt / T);
    memory[bt * MD + col0 + c] = style[b * SD + c];
}
__global__ void sum_style_kernel(const float* __restrict__ dmem, float* __restrict__ dstyle, int B, int T, int SD, int MD, int col0) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * SD) return;
    int b = i / SD, c = i - b * SD;
    float s = 0.f;
    for (int t = 0; t < T; ++t) s += dmem[((long long)b * T + t) * MD + col0 + c];
    dstyle[i] = s;
}

// ---------------------------------------------------------------------------------------------- optimizer
__global__ void reg_loss_kernel(const float* __restrict__ p, const float* __restrict__ regw, float* out, long long n) {
    __shared__ float sh[32];
    float acc = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        acc += regw[i] * p[i] * p[i];
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) atomicAdd(out, 0.5f * acc);
}
__global__ void grad_finalize_kernel(float* __restrict__ g, const float* __restrict__ p, const float* __restrict__ regw, float inv_world,
                                     float* sumsq, long long n) {
    __shared__ float sh[32];
    float acc = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        float v = g[i] * inv_world + regw[i] * p[i];
        g[i] = v;
        acc += v * v;
    }
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) atomicAdd(sumsq, acc);
}
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            const float* __restrict__ sumsq, float clip_norm, float lr_t, float b1, float b2, float eps, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float scale = 1.f;
    if (clip_norm > 0.f) scale = clip_norm / fmaxf(sqrtf(*sumsq), clip_norm);   // tf.clip_by_global_norm
    float gg = g[i] * scale;
    float mi = b1 * m[i] + (1.f - b1) * gg;
    float vi = b2 * v[i] + (1.f - b2) * gg * gg;
    m[i] = mi; v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
}

}  // namespace

#define ST ((cudaStream_t)stream)
#define ZERO(ptr, n) if (cudaMemsetAsync((ptr), 0, (size_t)(n) * sizeof(float), ST) != cudaSuccess) return -3

extern "C" {

int taco_op_ew(int op, float* out, const float* a, const float* b, float alpha, long long n, void* stream) {
    if (!out || !a || n < 0 || op < 0 || op > TACO_EW_TANH_BWD) return -1;
    if ((op == TACO_EW_ADD || op == TACO_EW_MUL || op == TACO_EW_TANH_BWD) && !b) return -1;
    if (n == 0) return 0;
    ew_kernel<<<nblk(n), 256, 0, ST>>>(op, out, a, b, alpha, n);
    return rc();
}
int taco_op_clip_fwd(const float* x, float lo, float hi, float* y, long long n, void* stream) {
    if (!x || !y) return -1;
    clip_fwd_kernel<<<nblk(n), 256, 0, ST>>>(x, lo, hi, y, n);
    return rc();
}
int taco_op_clip_bwd(const float* dy, const float* x, float lo, float hi, float* dx, long long n, void* stream) {
    if (!dy || !x || !dx) return -1;
    clip_bwd_kernel<<<nblk(n), 256, 0, ST>>>(dy, x, lo, hi, dx, n);
    return rc();
}
int taco_op_mse(const float* pred, const float* tgt, float* loss, float* dpred, long long n, void* stream) {
    if (!pred || !tgt || !loss || !dpred || n < 1) return -1;
    ZERO(loss, 1);
    unsigned g = nblk(n); if (g > 592) g = 592;
    mse_kernel<<<g, 256, 0, ST>>>(pred, tgt, loss, dpred, n);
    return rc();
}
int taco_op_sigmoid_ce(const float* logits, const float* labels, float* loss, float* dlogits, long long n, void* stream) {
    if (!logits || !labels || !loss || !dlogits || n < 1) return -1;
    ZERO(loss, 1);
    unsigned g = nblk(n); if (g > 592) g = 592;
    sigmoid_ce_kernel<<<g, 256, 0, ST>>>(logits, labels, loss, dlogits, n);
    return rc();
}
int taco_op_colsum(const float* x, long long ld, int M, int C, float* out, float beta, void* stream) {
    if (!x || !out || M < 1 || C < 1) return -1;
    if (beta == 0.f) { ZERO(out, C); }
    else if (beta != 1.f) { scale_kernel<<<nblk(C), 256, 0, ST>>>(out, beta, C); }
    if (g_bn_vec && ld == C && C % 4 == 0 && al16p(x) && al16p(out)) {      // dense rows: the 16-byte column-sum kernel of the batch norm
        bn_sum_v4_kernel<<<col_grid_v4(M, C / 4), dim3(32, 8), 0, ST>>>(x, M, C / 4, out);
        return rc();
    }
    colsum_kernel<<<col_grid(M, C), dim3(32, 8), 0, ST>>>(x, ld, M, C, out);
    return rc();
}
int taco_op_bn_fwd(const float* x, const float* gamma, const float* beta, int act, uint32_t keep_thresh, float keep,
                   uint64_t seed, uint32_t stream_id, float* y, float* mean, float* rstd, int M, int C, void* stream) {
    if (!x || !gamma || !beta || !y || !mean || !rstd || M < 1 || C < 1) return -1;
    ZERO(mean, C);
    ZERO(rstd, C);
    if (g_bn_vec && C % 4 == 0 && al16p(x) && al16p(y) && al16p(gamma) && al16p(beta) && al16p(mean) && al16p(rstd)) {
        const dim3 g4 = col_grid_v4(M, C / 4);
        bn_sum_v4_kernel<<<g4, dim3(32, 8), 0, ST>>>(x, M, C / 4, mean);
        bn_var_v4_kernel<<<g4, dim3(32, 8), 0, ST>>>(x, M, C / 4, mean, rstd);
        bn_finish_kernel<<<nblk(C), 256, 0, ST>>>(mean, rstd, M, C);
        const long long n4 = (long long)M * (C / 4);
        bn_apply_v4_kernel<<<nblk(n4), 256, 0, ST>>>(x, gamma, beta, mean, rstd, act, keep_thresh, keep, seed, stream_id, y, n4, C / 4);
        return rc();
    }
    dim3 g = col_grid(M, C);
    bn_sum_kernel<<<g, dim3(32, 8), 0, ST>>>(x, M, C, mean);
    bn_var_kernel<<<g, dim3(32, 8), 0, ST>>>(x, M, C, mean, rstd);
    bn_finish_kernel<<<nblk(C), 256, 0, ST>>>(mean, rstd, M, C);
    long long n = (long long)M * C;
    bn_apply_kernel<<<nblk(n), 256, 0, ST>>>(x, gamma, beta, mean, rstd, act, keep_thresh, keep, seed, stream_id, y, n, C);
    return rc();
}
int taco_op_bn_bwd(const float* dy, const float* x, const float* y, const float* mean, const float* rstd, const float* gamma,
                   int act, uint32_t keep_thresh, float keep, uint64_t seed, uint32_t stream_id, float* dx, float* dgamma,
                   float* dbeta, int M, int C, void* stream) {
    if (!dy || !x || !y || !mean || !rstd || !gamma || !dx || !dgamma || !dbeta || M < 1 || C < 1) return -1;
    ZERO(dgamma, C);
    ZERO(dbeta, C);
    if (g_bn_vec && C % 4 == 0 && al16p(dy) && al16p(x) && al16p(y) && al16p(dx) && al16p(gamma) && al16p(mean) && al16p(rstd) &&
        al16p(dgamma) && al16p(dbeta)) {
        bn_bwd_reduce_v4_kernel<<<col_grid_v4(M, C / 4), dim3(32, 8), 0, ST>>>(dy, x, y, mean, rstd, act, keep_thresh, keep, seed, stream_id,
                                                                               dgamma, dbeta, M, C / 4);
        const long long n4 = (long long)M * (C / 4);
        bn_bwd_apply_v4_kernel<<<nblk(n4), 256, 0, ST>>>(dy, x, y, mean, rstd, gamma, dgamma, dbeta, act, keep_thresh, keep, seed,
                                                         stream_id, dx, n4, C / 4, 1.0f / (float)M);
        return rc();
    }
    bn_bwd_reduce_kernel<<<col_grid(M, C), dim3(32, 8), 0, ST>>>(dy, x, y, mean, rstd, act, keep_thresh, keep, seed, stream_id,
                                                                  dgamma, dbeta, M, C);
    long long n = (long long)M * C;
    bn_bwd_apply_kernel<<<nblk(n), 256, 0, ST>>>(dy, x, y, mean, rstd, gamma, dgamma, dbeta, act, keep_thresh, keep, seed, stream_id,
                                                 dx, n, C, 1.0f / (float)M);
    return rc();
}
int taco_op_bn_update_moving(const float* mean, const float* rstd, float* mov_mean, float* mov_var, float momentum, int C, void* stream) {
    if (!mean || !rstd || !mov_mean || !mov_var) return -1;
    bn_moving_kernel<<<nblk(C), 256, 0, ST>>>(mean, rstd, mov_mean, mov_var, momentum, C);
    return rc();
}
int taco_op_lstm_fwd(float* gates, const float* pre, const float* bias, const float* c_prev, const float* h_prev,
                     const int32_t* lens, int s, float zoneout, int train, uint32_t zo_keep_thresh, uint64_t seed, uint32_t stream_c,
                     uint32_t stream_h, int row0, float* cnew, float* cz, float* hraw, float* hz, int B, int H, void* stream) {
    if (!gates || !c_prev || !h_prev || !cnew || !cz || !hraw || !hz || B < 1 || H < 1) return -1;
    launch_pdl(lstm_fwd_kernel, dim3(nblk((long long)B * H)), dim3(256), 0, ST, gates, pre, bias, c_prev, h_prev, lens, s, zoneout, train, seed, stream_c,
                                                            stream_h, row0, zo_keep_thresh, cnew, cz, hraw, hz, B, H, SeqIO{nullptr, 0, 0, nullptr, 0, 0, 0});
    return rc();
}
int taco_op_lstm_fwd_seq(float* gates, const float* xw, long long xw_sb, long long xw_st, const float* c_prev, const float* h_prev,
                         const int32_t* lens, int s, int reverse, float zoneout, int train, uint32_t zo_keep_thresh, uint64_t seed,
                         uint32_t stream_c, uint32_t stream_h, int row0, float* cnew, float* cz, float* hraw, float* hz, float* out,
                         long long out_sb, long long out_st, int B, int H, void* stream) {
    if (!gates || !xw || !c_prev || !h_prev || !lens || !cnew || !cz || !hraw || !hz || !out || B < 1 || H < 1) return -1;
    launch_pdl(lstm_fwd_kernel, dim3(nblk((long long)B * H)), dim3(256), 0, ST, gates, (const float*)nullptr, (const float*)nullptr, c_prev, h_prev, lens, s,
               zoneout, train, seed, stream_c, stream_h, row0, zo_keep_thresh, cnew, cz, hraw, hz, B, H, SeqIO{xw, xw_sb, xw_st, out, out_sb, out_st, reverse});
    return rc();
}
int taco_op_lstm_bwd(const float* acts, const float* c_prev, const float* h_prev, const float* cnew, const int32_t* lens, int s,
                     float zoneout, int train, uint32_t zo_keep_thresh, uint64_t seed, uint32_t stream_c, uint32_t stream_h, int row0,
                     const float* d_hraw, const float* d_cz, const float* d_hz, float* dgates, float* d_cprev, float* d_hprev,
                     int B, int H, void* stream) {
    if (!acts || !c_prev || !h_prev || !cnew || !d_hraw || !d_cz || !d_hz || !dgates || !d_cprev || !d_hprev) return -1;
    launch_pdl(lstm_bwd_kernel, dim3(nblk((long long)B * H)), dim3(256), 0, ST, acts, c_prev, h_prev, cnew, lens, s, zoneout, train, seed, stream_c, stream_h,
                                                            row0, zo_keep_thresh, d_hraw, d_cz, d_hz, dgates, d_cprev, d_hprev, B, H, SeqIO{nullptr, 0, 0, nullptr, 0, 0, 0});
    return rc();
}
int taco_op_lstm_bwd_seq(const float* acts, const float* c_prev, const float* h_prev, const float* cnew, const int32_t* lens, int s,
                         int reverse, float zoneout, int train, uint32_t zo_keep_thresh, uint64_t seed, uint32_t stream_c, uint32_t stream_h,
                         int row0, const float* dout, long long dout_sb, long long dout_st, const float* d_cz, const float* d_hz,
                         float* dgates, float* d_cprev, float* d_hprev, float* dxw, long long dxw_sb, long long dxw_st, int B, int H,
                         void* stream) {
    if (!acts || !c_prev || !h_prev || !cnew || !lens || !dout || !d_cz || !d_hz || !dgates || !d_cprev || !d_hprev || !dxw) return -1;
    launch_pdl(lstm_bwd_kernel, dim3(nblk((long long)B * H)), dim3(256), 0, ST, acts, c_prev, h_prev, cnew, lens, s, zoneout, train, seed, stream_c, stream_h,
               row0, zo_keep_thresh, (const float*)nullptr, d_cz, d_hz, dgates, d_cprev, d_hprev, B, H, SeqIO{dout, dout_sb, dout_st, dxw, dxw_sb, dxw_st, reverse});
    return rc();
}
int taco_op_gather_rows(const float* src, long long sb, long long st, const int32_t* lens, int s, int reverse, float* out,
                        int B, int N, void* stream) {
    if (!src || !lens || !out) return -1;
    launch_pdl(gather_rows_kernel, dim3(nblk((long long)B * N)), dim3(256), 0, ST, src, sb, st, lens, s, reverse, out, B, N);
    return rc();
}
int taco_op_scatter_rows(const float* src, const int32_t* lens, int s, int reverse, float* dst, long long db, long long dt,
                         int B, int N, void* stream) {
    if (!src || !lens || !dst) return -1;
    launch_pdl(scatter_rows_kernel, dim3(nblk((long long)B * N)), dim3(256), 0, ST, src, lens, s, reverse, dst, db, dt, B, N);
    return rc();
}
int taco_op_gru_pre(float* g, const float* h, float* rh, int B, int R, void* stream) {
    if (!g || !h || !rh) return -1;
    gru_pre_kernel<<<nblk((long long)B * 2 * R), 256, 0, ST>>>(g, h, rh, B, R);
    return rc();
}
int taco_op_gru_out(float* cpre, const float* ru, const float* h, float* hnew, int B, int R, void* stream) {
    if (!cpre || !ru || !h || !hnew) return -1;
    gru_out_kernel<<<nblk((long long)B * R), 256, 0, ST>>>(cpre, ru, h, hnew, B, R);
    return rc();
}
int taco_op_gru_out_bwd(const float* dhnew, const float* ru, const float* h, const float* c, float* dcpre, float* dg, float* dh,
                        int B, int R, void* stream) {
    if (!dhnew || !ru || !h || !c || !dcpre || !dg || !dh) return -1;
    gru_out_bwd_kernel<<<nblk((long long)B * R), 256, 0, ST>>>(dhnew, ru, h, c, dcpre, dg, dh, B, R);
    return rc();
}
int taco_op_gru_pre_bwd(const float* drh, const float* ru, const float* h, float* dg, float* dh, int B, int R, void* stream) {
    if (!drh || !ru || !h || !dg || !dh) return -1;
    gru_pre_bwd_kernel<<<nblk((long long)B * R), 256, 0, ST>>>(drh, ru, h, dg, dh, B, R);
    return rc();
}
int taco_op_embedding_fwd(const int32_t* ids, const float* table, float* out, int n_tok, int dim, int n_sym, void* stream) {
    if (!ids || !table || !out) return -1;
    emb_fwd_kernel<<<nblk((long long)n_tok * dim), 256, 0, ST>>>(ids, table, out, n_tok, dim, n_sym);
    return rc();
}
int taco_op_embedding_bwd(const int32_t* ids, const float* dout, float* dtable, int n_tok, int dim, int n_sym, void* stream) {
    if (!ids || !dout || !dtable) return -1;
    ZERO(dtable, (size_t)n_sym * dim);
    emb_bwd_kernel<<<nblk((long long)n_tok * dim), 256, 0, ST>>>(ids, dout, dtable, n_tok, dim, n_sym);
    return rc();
}
int taco_op_prenet_act_fwd(const float* z, int layer, uint64_t seed, int step0, float keep, float* out, int S, int B, int U, void* stream) {
    if (!z || !out) return -1;
    prenet_act_kernel<<<nblk((long long)S * B * U), 256, 0, ST>>>(z, nullptr, layer, seed, step0, keep, out, S, B, U, 0);
    return rc();
}
int taco_op_prenet_act_bwd(const float* dout, const float* out, int layer, uint64_t seed, int step0, float keep, float* dz,
                           int S, int B, int U, void* stream) {
    if (!dout || !out || !dz) return -1;
    prenet_act_kernel<<<nblk((long long)S * B * U), 256, 0, ST>>>(out, dout, layer, seed, step0, keep, dz, S, B, U, 1);
    return rc();
}
int taco_op_mask_rows(float* x, const int32_t* lens, int B, int T, int C, void* stream) {
    if (!x || !lens) return -1;
    mask_rows_kernel<<<nblk((long long)B * T * C), 256, 0, ST>>>(x, lens, B, T, C);
    return rc();
}
int taco_op_tile_style(const float* style, float* memory, int B, int T, int SD, int MD, int col0, void* stream) {
    if (!style || !memory) return -1;
    tile_style_kernel<<<nblk((long long)B * T * SD), 256, 0, ST>>>(style, memory, B, T, SD, MD, col0);
    return rc();
}
int taco_op_sum_style(const float* dmemory, float* dstyle, int B, int T, int SD, int MD, int col0, void* stream) {
    if (!dmemory || !dstyle) return -1;
    sum_style_kernel<<<nblk((long long)B * SD), 256, 0, ST>>>(dmemory, dstyle, B, T, SD, MD, col0);
    return rc();
}
int taco_op_reg_loss(const float* p, const float* regw, float* out, long long n, void* stream) {
    if (!p || !regw || !out) return -1;
    ZERO(out, 1);
    reg_loss_kernel<<<592, 256, 0, ST>>>(p, regw, out, n);
    return rc();
}
int taco_op_grad_finalize(float* g, const float* p, const float* regw, float inv_world, float* sumsq, long long n, void* stream) {
    if (!g || !p || !regw || !sumsq) return -1;
    grad_finalize_kernel<<<592, 256, 0, ST>>>(g, p, regw, inv_world, sumsq, n);
    return rc();
}
int taco_op_adam_step(float* p, const float* g, float* m, float* v, const float* sumsq, float clip_norm, float lr_t, float b1,
                      float b2, float eps, long long n, void* stream) {
    if (!p || !g || !m || !v || !sumsq) return -1;
    adam_kernel<<<nblk(n), 256, 0, ST>>>(p, g, m, v, sumsq, clip_norm, lr_t, b1, b2, eps, n);
    return rc();
}

}  // extern "C"
