// train_attn.cu -- one LocationSensitiveAttention step of the training graph, forward and gradient
// (Attention.py:169-210,345-401), fp32.  One CTA per batch row: the row's cumulative alignments, the
// location filters and the location dense layer live in shared memory; keys/values are streamed once
// with coalesced reads; the location conv, the location dense layer, tanh energies, masked softmax,
// cumulative update and context reduction are fused (no [B,T,128] temporaries in HBM).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/taco_b200_train.h"
#include "common.cuh"

namespace {

constexpr int NF = 32, AD = 128, NW = 8, WLS = AD + 1;   // filters, attention dim, warps per CTA, padded Wl row stride

struct AttnSmem {
    float *cum, *cw, *cb, *Wl, *qb, *va, *e;
    __device__ AttnSmem(float* base, int T, int KW) {
        cum = base; base += T + KW;
        cw = base; base += KW * NF;
        cb = base; base += NF;
        Wl = base; base += NF * WLS;
        qb = base; base += AD;
        va = base; base += AD;
        e = base;
    }
};
inline size_t attn_fwd_smem(int T, int KW) { return sizeof(float) * ((size_t)T + KW + KW * NF + NF + NF * WLS + 2 * AD + T + 8); }

__device__ __forceinline__ void attn_load_common(const AttnSmem& s, const float* q, const float* cum_prev, const float* conv_w,
                                                 const float* conv_b, const float* Wl, const float* va, const float* ba, int b, int T,
                                                 int KW) {
    const int left = (KW - 1) / 2;
    for (int i = threadIdx.x; i < T + KW - 1; i += blockDim.x) {
        int t = i - left;
        s.cum[i] = (t >= 0 && t < T) ? cum_prev[(long long)b * T + t] : 0.f;
    }
    for (int i = threadIdx.x; i < KW * NF; i += blockDim.x) s.cw[i] = conv_w[i];
    for (int i = threadIdx.x; i < NF * AD; i += blockDim.x) s.Wl[(i / AD) * WLS + (i % AD)] = Wl[i];
    for (int i = threadIdx.x; i < AD; i += blockDim.x) { s.qb[i] = q[(long long)b * AD + i] + ba[i]; s.va[i] = va[i]; }
    if (threadIdx.x < NF) s.cb[threadIdx.x] = conv_b[threadIdx.x];
}

// location features of position t (lane = filter) and the 4 pre-activations owned by this lane (d = lane + 32*dd)
__device__ __forceinline__ float loc_feature(const AttnSmem& s, int t, int KW, int lane) {
    float f = s.cb[lane];
    for (int j = 0; j < KW; ++j) f = fmaf(s.cum[t + j], s.cw[j * NF + lane], f);
    return f;
}
__device__ __forceinline__ void pre_act(const AttnSmem& s, const float* __restrict__ krow, float f, int lane, float acc[4]) {
#pragma unroll
    for (int dd = 0; dd < 4; ++dd) acc[dd] = krow[lane + 32 * dd] + s.qb[lane + 32 * dd];
#pragma unroll 8
    for (int ff = 0; ff < NF; ++ff) {
        float fv = __shfl_sync(0xffffffffu, f, ff);
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) acc[dd] = fmaf(fv, s.Wl[ff * WLS + lane + 32 * dd], acc[dd]);
    }
}

__global__ void __launch_bounds__(256) attn_fwd_kernel(const float* __restrict__ q, const float* __restrict__ cum_prev,
                                                       const float* __restrict__ keys, const float* __restrict__ values,
                                                       const int32_t* __restrict__ lens, const float* __restrict__ conv_w,
                                                       const float* __restrict__ conv_b, const float* __restrict__ Wl,
                                                       const float* __restrict__ va, const float* __restrict__ ba,
                                                       float* __restrict__ align, float* __restrict__ cum, float* __restrict__ ctx,
                                                       int T, int V, int KW) {
    extern __shared__ float smem[];
    __shared__ float red[32];
    AttnSmem s(smem, T, KW);
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int len = min(lens[b], T);
    attn_load_common(s, q, cum_prev, conv_w, conv_b, Wl, va, ba, b, T, KW);
    __syncthreads();
    for (int t = warp; t < len; t += NW) {
        float f = loc_feature(s, t, KW, lane);
        float acc[4];
        pre_act(s, keys + ((long long)b * T + t) * AD, f, lane, acc);
        float e = 0.f;
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) e = fmaf(s.va[lane + 32 * dd], tanhf(acc[dd]), e);
        e = warp_sum(e);
        if (lane == 0) s.e[t] = e;
    }
    __syncthreads();
    float mx = -INFINITY;
    for (int t = threadIdx.x; t < len; t += blockDim.x) mx = fmaxf(mx, s.e[t]);
    mx = warp_max(mx);
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
    for (int w = 1; w < NW; ++w) mx = fmaxf(mx, red[w]);
    __syncthreads();
    float sum = 0.f;
    for (int t = threadIdx.x; t < len; t += blockDim.x) { float v = expf(s.e[t] - mx); s.e[t] = v; sum += v; }
    sum = warp_sum(sum);
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    sum = 0.f;
    for (int w = 0; w < NW; ++w) sum += red[w];
    const float inv = 1.0f / sum;
    const int left = (KW - 1) / 2;
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        float a = t < len ? s.e[t] * inv : 0.f;
        if (t < len) s.e[t] = a;
        align[(long long)b * T + t] = a;
        cum[(long long)b * T + t] = s.cum[t + left] + a;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < V; c += blockDim.x) {
        const float* vp = values + (long long)b * T * V + c;
        float acc = 0.f;
        for (int t = 0; t < len; ++t) acc = fmaf(s.e[t], vp[(long long)t * V], acc);
        ctx[(long long)b * V + c] = acc;
    }
}

inline size_t attn_bwd_smem(int T, int KW, int V) {
    return attn_fwd_smem(T, KW) + sizeof(float) * ((size_t)T /*da*/ + (size_t)T * (NF + 1) /*dfl*/ + NW * AD /*dpre*/ + NW * NF /*fl*/ + V + 8);
}

__global__ void __launch_bounds__(256) attn_bwd_kernel(const float* __restrict__ dctx, const float* __restrict__ dcum,
                                                       const float* __restrict__ q, const float* __restrict__ cum_prev,
                                                       const float* __restrict__ align, const float* __restrict__ keys,
                                                       const float* __restrict__ values, const int32_t* __restrict__ lens,
                                                       const float* __restrict__ conv_w, const float* __restrict__ conv_b,
                                                       const float* __restrict__ Wl, const float* __restrict__ va,
                                                       const float* __restrict__ ba, float* __restrict__ dq, float* __restrict__ dcum_prev,
                                                       float* __restrict__ dkeys, float* __restrict__ p_conv_w, float* __restrict__ p_conv_b,
                                                       float* __restrict__ p_Wl, float* __restrict__ p_va, float* __restrict__ p_ba,
                                                       int T, int V, int KW) {
    extern __shared__ float smem[];
    __shared__ float red[32];
    AttnSmem s(smem, T, KW);
    float* s_da = s.e + T;                 // d(align) -> d(energy)
    float* s_dfl = s_da + T;               // [T][NF+1] d(location features)
    float* s_dpre = s_dfl + (size_t)T * (NF + 1);   // [NW][AD] tile of d(pre-activations); reused as reduction scratch
    float* s_fl = s_dpre + NW * AD;        // [NW][NF] tile of location features
    float* s_dctx = s_fl + NW * NF;        // [V]
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, tid = threadIdx.x;
    const int len = min(lens[b], T);
    const int left = (KW - 1) / 2;
    attn_load_common(s, q, cum_prev, conv_w, conv_b, Wl, va, ba, b, T, KW);
    for (int i = tid; i < T; i += blockDim.x) s.e[i] = align[(long long)b * T + i];
    for (int i = tid; i < V; i += blockDim.x) s_dctx[i] = dctx[(long long)b * V + i];
    __syncthreads();
    // d(align)[t] = values[t,:] . d(ctx) + d(cum_t)[t]      (cum_t = cum_{t-1} + align_t)
    for (int t = warp; t < T; t += NW) {
        float acc = 0.f;
        if (t < len) {
            const float* vp = values + ((long long)b * T + t) * V;
            for (int c = lane; c < V; c += 32) acc = fmaf(vp[c], s_dctx[c], acc);
            acc = warp_sum(acc);
        }
        if (lane == 0) s_da[t] = t < len ? acc + dcum[(long long)b * T + t] : 0.f;
    }
    __syncthreads();
    // softmax gradient: d(e) = a * (d(a) - sum_t a d(a))
    float dot = 0.f;
    for (int t = tid; t < len; t += blockDim.x) dot = fmaf(s.e[t], s_da[t], dot);
    dot = warp_sum(dot);
    if (lane == 0) red[warp] = dot;
    __syncthreads();
    dot = 0.f;
    for (int w = 0; w < NW; ++w) dot += red[w];
    __syncthreads();
    for (int t = tid; t < T; t += blockDim.x) s_da[t] = t < len ? s.e[t] * (s_da[t] - dot) : 0.f;
    __syncthreads();

    float dva_acc[4] = {0.f, 0.f, 0.f, 0.f}, dq_acc[4] = {0.f, 0.f, 0.f, 0.f};
    float wl_acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) wl_acc[i] = 0.f;
    const int ntile = (T + NW - 1) / NW;
    for (int tile = 0; tile < ntile; ++tile) {
        const int t = tile * NW + warp;
        float f = 0.f, dp[4] = {0.f, 0.f, 0.f, 0.f};
        if (t < len) {
            f = loc_feature(s, t, KW, lane);
            float acc[4];
            pre_act(s, keys + ((long long)b * T + t) * AD, f, lane, acc);
            const float de = s_da[t];
            float* dk = dkeys + ((long long)b * T + t) * AD;
#pragma unroll
            for (int dd = 0; dd < 4; ++dd) {
                float th = tanhf(acc[dd]);
                dp[dd] = de * s.va[lane + 32 * dd] * (1.f - th * th);
                dva_acc[dd] = fmaf(de, th, dva_acc[dd]);
                dq_acc[dd] += dp[dd];
                dk[lane + 32 * dd] += dp[dd];
            }
        }
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) s_dpre[warp * AD + lane + 32 * dd] = dp[dd];
        s_fl[warp * NF + lane] = f;
        __syncwarp();
        if (t < T) {
            float df = 0.f;   // d(floc[t, lane]) = sum_d dpre[d] * Wl[lane, d]
            if (t < len)
                for (int d = 0; d < AD; ++d) df = fmaf(s_dpre[warp * AD + d], s.Wl[lane * WLS + d], df);
            s_dfl[(size_t)t * (NF + 1) + lane] = df;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; ++i) {   // d(Wl)[f, d] += floc[t, f] * dpre[t, d] over the tile
            const int idx = tid + 256 * i, ff = idx / AD, d = idx - ff * AD;
            float a = wl_acc[i];
#pragma unroll
            for (int w = 0; w < NW; ++w) a = fmaf(s_fl[w * NF + ff], s_dpre[w * AD + d], a);
            wl_acc[i] = a;
        }
        __syncthreads();
    }
    // cross-warp reduction of dq / d(va) through the (now free) dpre tile
    float* scr = s_dpre;
#pragma unroll
    for (int dd = 0; dd < 4; ++dd) scr[warp * AD + lane + 32 * dd] = dq_acc[dd];
    __syncthreads();
    if (tid < AD) {
        float v = 0.f;
        for (int w = 0; w < NW; ++w) v += scr[w * AD + tid];
        dq[(long long)b * AD + tid] = v;
        p_ba[(long long)b * AD + tid] += v;
    }
    __syncthreads();
#pragma unroll
    for (int dd = 0; dd < 4; ++dd) scr[warp * AD + lane + 32 * dd] = dva_acc[dd];
    __syncthreads();
    if (tid < AD) {
        float v = 0.f;
        for (int w = 0; w < NW; ++w) v += scr[w * AD + tid];
        p_va[(long long)b * AD + tid] += v;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) p_Wl[(long long)b * NF * AD + tid + 256 * i] += wl_acc[i];
    // location convolution gradients
    for (int idx = tid; idx < KW * NF; idx += blockDim.x) {
        const int j = idx / NF, ff = idx - j * NF;
        float a = 0.f;
        for (int t = 0; t < len; ++t) a = fmaf(s.cum[t + j], s_dfl[(size_t)t * (NF + 1) + ff], a);
        p_conv_w[(long long)b * KW * NF + idx] += a;
    }
    if (tid < NF) {
        float a = 0.f;
        for (int t = 0; t < len; ++t) a += s_dfl[(size_t)t * (NF + 1) + tid];
        p_conv_b[(long long)b * NF + tid] += a;
    }
    for (int tp = tid; tp < T; tp += blockDim.x) {
        float a = dcum[(long long)b * T + tp];
        for (int j = 0; j < KW; ++j) {
            const int t = tp - j + left;       // floc[t] read cum[t + j - left] = cum[tp]
            if (t >= 0 && t < len) {
                const float* dfr = s_dfl + (size_t)t * (NF + 1);
                const float* cwr = s.cw + j * NF;
#pragma unroll 8
                for (int ff = 0; ff < NF; ++ff) a = fmaf(dfr[ff], cwr[ff], a);
            }
        }
        dcum_prev[(long long)b * T + tp] = a;
    }
}

}  // namespace

extern "C" {

int taco_op_attn_fwd(const float* q, const float* cum_prev, const float* keys, const float* values, const int32_t* lens,
                     const float* conv_w, const float* conv_b, const float* Wl, const float* va, const float* ba,
                     float* align, float* cum, float* ctx, int B, int T, int A, int V, int NFf, int KW, void* stream) {
    if (!q || !cum_prev || !keys || !values || !lens || !conv_w || !conv_b || !Wl || !va || !ba || !align || !cum || !ctx) return -1;
    if (A != AD || NFf != NF || KW < 1 || KW > 63 || T < 1 || T > 2048 || B < 1) return -1;
    size_t sm = attn_fwd_smem(T, KW);
    if (sm > 48 * 1024 && cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess) return -3;
    attn_fwd_kernel<<<B, 256, sm, (cudaStream_t)stream>>>(q, cum_prev, keys, values, lens, conv_w, conv_b, Wl, va, ba, align, cum, ctx, T, V, KW);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int taco_op_attn_bwd(const float* dctx, const float* dcum, const float* q, const float* cum_prev, const float* align,
                     const float* keys, const float* values, const int32_t* lens, const float* conv_w, const float* conv_b,
                     const float* Wl, const float* va, const float* ba, float* dq, float* dcum_prev, float* dkeys,
                     float* p_conv_w, float* p_conv_b, float* p_Wl, float* p_va, float* p_ba,
                     int B, int T, int A, int V, int NFf, int KW, void* stream) {
    if (!dctx || !dcum || !q || !cum_prev || !align || !keys || !values || !lens || !conv_w || !conv_b || !Wl || !va || !ba || !dq ||
        !dcum_prev || !dkeys || !p_conv_w || !p_conv_b || !p_Wl || !p_va || !p_ba)
        return -1;
    if (A != AD || NFf != NF || KW < 1 || KW > 63 || T < 1 || B < 1) return -1;
    size_t sm = attn_bwd_smem(T, KW, V);
    if (sm > 200 * 1024) return -1;
    if (sm > 48 * 1024 && cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess) return -3;
    attn_bwd_kernel<<<B, 256, sm, (cudaStream_t)stream>>>(dctx, dcum, q, cum_prev, align, keys, values, lens, conv_w, conv_b, Wl, va, ba,
                                                          dq, dcum_prev, dkeys, p_conv_w, p_conv_b, p_Wl, p_va, p_ba, T, V, KW);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // extern "C"
