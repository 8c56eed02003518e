// train_attn.cu -- one LocationSensitiveAttention step of the training graph, forward and gradient
// (Attention.py:169-210,345-401), fp32.  B = 32 rows per GPU is far fewer than 148 SMs, so every phase is split
// over (position tiles x batch rows) or (value-column blocks x batch rows) instead of one CTA per row:
//
//   forward : energy   grid (T/16, B)  location conv + location dense + tanh energies, keys streamed once
//             softmax  grid (B)        masked softmax, cumulative update
//             context  grid (V/128, B) alignments . values, 16-byte coalesced loads, t split over warps
//   backward: d(align) grid (T/8, B)   values . d(ctx) + d(cum)
//             softmax  grid (B)        d(energy)
//             pre      grid (T/16, B)  recomputed pre-activations -> d(keys), d(q), d(v_a), d(b_a), d(Wl), d(location features)
//             conv     grid (B)        d(location filters), d(cum_{t-1})
//
// No [B,T,128] temporaries reach HBM; the only scratch is d(location features) [B,T,32].
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "../../include/taco_b200_train.h"
#include "common.cuh"

namespace {

constexpr int NF = 32, AD = 128, NW = 8, WLS = AD + 1;   // filters, attention dim, warps per CTA, padded Wl row stride
constexpr int TPB = 16;                                   // positions per CTA in the tiled kernels (2 per warp)

// shared-memory working set of a position tile
struct TileSmem {
    float *cum, *cw, *cb, *Wl, *qb, *va;
    __device__ TileSmem(float* base, int KW) {
        cum = base; base += TPB + KW;        // cum_prev[t0 - left .. t0 + TPB - 1 + right]
        cw = base; base += KW * NF;
        cb = base; base += NF;
        Wl = base; base += NF * WLS;
        qb = base; base += AD;
        va = base;
    }
};
__host__ __device__ inline size_t tile_smem_floats(int KW) { return (size_t)TPB + KW + (size_t)KW * NF + NF + NF * WLS + 2 * AD; }

__device__ __forceinline__ void tile_load(const TileSmem& s, const float* q, const float* cum_prev, const float* conv_w, const float* conv_b,
                                          const float* Wl, const float* va, const float* ba, int b, int t0, int T, int KW) {
    const int left = (KW - 1) / 2;
    for (int i = threadIdx.x; i < TPB + KW - 1; i += blockDim.x) {
        int t = t0 + i - left;
        s.cum[i] = (t >= 0 && t < T) ? cum_prev[(long long)b * T + t] : 0.f;
    }
    for (int i = threadIdx.x; i < KW * NF; i += blockDim.x) s.cw[i] = conv_w[i];
    for (int i = threadIdx.x; i < NF * AD; i += blockDim.x) s.Wl[(i / AD) * WLS + (i % AD)] = Wl[i];
    for (int i = threadIdx.x; i < AD; i += blockDim.x) { s.qb[i] = q[(long long)b * AD + i] + ba[i]; s.va[i] = va[i]; }
    if (threadIdx.x < NF) s.cb[threadIdx.x] = conv_b[threadIdx.x];
}
// location feature `lane` of tile position i (= t - t0), and the 4 pre-activations owned by this lane (d = lane + 32*dd)
__device__ __forceinline__ float loc_feature(const TileSmem& s, int i, int KW, int lane) {
    float f = s.cb[lane];
    for (int j = 0; j < KW; ++j) f = fmaf(s.cum[i + j], s.cw[j * NF + lane], f);
    return f;
}
__device__ __forceinline__ void pre_act(const TileSmem& s, const float kv[4], float f, int lane, float acc[4]) {
#pragma unroll
    for (int dd = 0; dd < 4; ++dd) acc[dd] = kv[dd] + s.qb[lane + 32 * dd];
#pragma unroll 8
    for (int ff = 0; ff < NF; ++ff) {
        float fv = __shfl_sync(0xffffffffu, f, ff);
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) acc[dd] = fmaf(fv, s.Wl[ff * WLS + lane + 32 * dd], acc[dd]);
    }
}

// ------------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(256) attn_energy_kernel(const float* __restrict__ q, const float* __restrict__ cum_prev,
                                                          const float* __restrict__ keys, const int32_t* __restrict__ lens,
                                                          const float* __restrict__ conv_w, const float* __restrict__ conv_b,
                                                          const float* __restrict__ Wl, const float* __restrict__ va,
                                                          const float* __restrict__ ba, float* __restrict__ energy, int T, int KW) {
    extern __shared__ float smem[];
    TileSmem s(smem, KW);
    const int b = blockIdx.y, t0 = blockIdx.x * TPB, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int len = min(lens[b], T);
    if (t0 >= len) return;
    float kv[2][4];
#pragma unroll
    for (int p = 0; p < 2; ++p) {            // keys of both positions of this warp in flight before the smem fill
        const int t = t0 + p * NW + warp;
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) kv[p][dd] = t < len ? keys[((long long)b * T + t) * AD + lane + 32 * dd] : 0.f;
    }
    tile_load(s, q, cum_prev, conv_w, conv_b, Wl, va, ba, b, t0, T, KW);
    __syncthreads();
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int i = p * NW + warp, t = t0 + i;
        if (t >= len) continue;
        float acc[4];
        pre_act(s, kv[p], loc_feature(s, i, KW, lane), lane, acc);
        float e = 0.f;
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) e = fmaf(s.va[lane + 32 * dd], tanhf(acc[dd]), e);
        e = warp_sum(e);
        if (lane == 0) energy[(long long)b * T + t] = e;
    }
}

// masked softmax over the valid positions, cumulative update; energies are read from `align` and overwritten in place
__global__ void __launch_bounds__(256) attn_softmax_kernel(const float* __restrict__ cum_prev, const int32_t* __restrict__ lens,
                                                           float* __restrict__ align, float* __restrict__ cum, int T) {
    __shared__ float red[32];
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int len = min(lens[b], T);
    float* a = align + (long long)b * T;
    float mx = -INFINITY;
    for (int t = threadIdx.x; t < len; t += blockDim.x) mx = fmaxf(mx, a[t]);
    mx = warp_max(mx);
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
    for (int w = 1; w < NW; ++w) mx = fmaxf(mx, red[w]);
    __syncthreads();
    float sum = 0.f;
    for (int t = threadIdx.x; t < len; t += blockDim.x) sum += expf(a[t] - mx);
    sum = warp_sum(sum);
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    sum = 0.f;
    for (int w = 0; w < NW; ++w) sum += red[w];
    const float inv = 1.0f / sum;
    for (int t = threadIdx.x; t < T; t += blockDim.x) {
        const float v = t < len ? expf(a[t] - mx) * inv : 0.f;
        a[t] = v;
        cum[(long long)b * T + t] = cum_prev[(long long)b * T + t] + v;
    }
}

// context[b, c0 .. c0+127] = sum_t align[b,t] * values[b,t,c]: warp w takes t = w, w+8, ...; lane owns 4 columns (16-byte loads)
__global__ void __launch_bounds__(256) attn_context_kernel(const float* __restrict__ align, const float* __restrict__ values,
                                                           const int32_t* __restrict__ lens, float* __restrict__ ctx, int T, int V) {
    __shared__ float4 part[NW][32];
    const int b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c = blockIdx.x * 128 + lane * 4;
    const int len = min(lens[b], T);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < V) {
        const float* vp = values + (long long)b * T * V + c;
        const float* ap = align + (long long)b * T;
#pragma unroll 4
        for (int t = warp; t < len; t += NW) {
            const float a = ap[t];
            const float4 v = *reinterpret_cast<const float4*>(vp + (long long)t * V);
            acc.x = fmaf(a, v.x, acc.x); acc.y = fmaf(a, v.y, acc.y); acc.z = fmaf(a, v.z, acc.z); acc.w = fmaf(a, v.w, acc.w);
        }
    }
    part[warp][lane] = acc;
    __syncthreads();
    if (warp == 0 && c < V) {
#pragma unroll
        for (int w = 1; w < NW; ++w) { float4 o = part[w][lane]; acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w; }
        *reinterpret_cast<float4*>(ctx + (long long)b * V + c) = acc;
    }
}

// ------------------------------------------------------------------------------------------------ backward
// d(align)[t] = values[t,:] . d(ctx) + d(cum_t)[t]   (cum_t = cum_{t-1} + align_t); one warp per position
__global__ void __launch_bounds__(256) attn_bwd_dalign_kernel(const float* __restrict__ dctx, const float* __restrict__ dcum,
                                                              const float* __restrict__ values, const int32_t* __restrict__ lens,
                                                              float* __restrict__ da, int T, int V) {
    extern __shared__ float s_dctx[];
    const int b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int t = blockIdx.x * NW + warp;
    const int len = min(lens[b], T);
    if (blockIdx.x * NW >= len) {
        if (t < T && lane == 0) da[(long long)b * T + t] = 0.f;
        return;
    }
    for (int i = threadIdx.x; i < V; i += blockDim.x) s_dctx[i] = dctx[(long long)b * V + i];
    __syncthreads();
    if (t >= T) return;
    float acc = 0.f;
    if (t < len) {
        const float* vp = values + ((long long)b * T + t) * V;
        for (int c = lane * 4; c < V; c += 128) {
            const float4 v = *reinterpret_cast<const float4*>(vp + c);
            acc = fmaf(v.x, s_dctx[c], acc); acc = fmaf(v.y, s_dctx[c + 1], acc);
            acc = fmaf(v.z, s_dctx[c + 2], acc); acc = fmaf(v.w, s_dctx[c + 3], acc);
        }
        acc = warp_sum(acc) + dcum[(long long)b * T + t];
    }
    if (lane == 0) da[(long long)b * T + t] = t < len ? acc : 0.f;
}

// d(energy) = a * (d(a) - sum_t a d(a)), in place on `da`; also zeroes d(q) of the row (accumulated by the next kernel)
__global__ void __launch_bounds__(256) attn_bwd_softmax_kernel(const float* __restrict__ align, const int32_t* __restrict__ lens,
                                                               float* __restrict__ da, float* __restrict__ dq, int T) {
    __shared__ float red[32];
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int len = min(lens[b], T);
    const float* a = align + (long long)b * T;
    float* d = da + (long long)b * T;
    float dot = 0.f;
    for (int t = threadIdx.x; t < len; t += blockDim.x) dot = fmaf(a[t], d[t], dot);
    dot = warp_sum(dot);
    if (lane == 0) red[warp] = dot;
    __syncthreads();
    dot = 0.f;
    for (int w = 0; w < NW; ++w) dot += red[w];
    for (int t = threadIdx.x; t < T; t += blockDim.x) d[t] = t < len ? a[t] * (d[t] - dot) : 0.f;
    if (threadIdx.x < AD) dq[(long long)b * AD + threadIdx.x] = 0.f;
}

// recompute the pre-activations of a position tile; d(pre) -> d(keys) (+=), d(q) / d(b_a) / d(v_a) / d(Wl) (atomics into the
// row's accumulators), d(location features) -> dfl [B,T,NF]
__global__ void __launch_bounds__(256) attn_bwd_pre_kernel(const float* __restrict__ de, const float* __restrict__ q,
                                                           const float* __restrict__ cum_prev, const float* __restrict__ keys,
                                                           const int32_t* __restrict__ lens, const float* __restrict__ conv_w,
                                                           const float* __restrict__ conv_b, const float* __restrict__ Wl,
                                                           const float* __restrict__ va, const float* __restrict__ ba, float* dq,
                                                           float* __restrict__ dkeys, float* p_Wl, float* p_va, float* p_ba,
                                                           float* __restrict__ dfl, int T, int KW) {
    extern __shared__ float smem[];
    TileSmem s(smem, KW);
    float* s_dpre = smem + tile_smem_floats(KW);    // [NW][AD]
    float* s_fl = s_dpre + NW * AD;                 // [NW][NF]
    const int b = blockIdx.y, t0 = blockIdx.x * TPB, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, tid = threadIdx.x;
    const int len = min(lens[b], T);
    if (t0 >= len) {                                // nothing valid in this tile: d(location features) = 0
        for (int i = tid; i < TPB * NF; i += blockDim.x) {
            const int t = t0 + i / NF;
            if (t < T) dfl[((long long)b * T + t) * NF + (i % NF)] = 0.f;
        }
        return;
    }
    float kv[2][4];
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int t = t0 + p * NW + warp;
#pragma unroll
        for (int dd = 0; dd < 4; ++dd) kv[p][dd] = t < len ? keys[((long long)b * T + t) * AD + lane + 32 * dd] : 0.f;
    }
    tile_load(s, q, cum_prev, conv_w, conv_b, Wl, va, ba, b, t0, T, KW);
    __syncthreads();
    float dva_acc[4] = {0.f, 0.f, 0.f, 0.f}, dq_acc[4] = {0.f, 0.f, 0.f, 0.f};
    float wl_acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) wl_acc[i] = 0.f;
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int i = p * NW + warp, t = t0 + i;
        float f = 0.f, dp[4] = {0.f, 0.f, 0.f, 0.f};
        if (t < len) {
            f = loc_feature(s, i, KW, lane);
            float acc[4];
            pre_act(s, kv[p], f, lane, acc);
            const float dev = de[(long long)b * T + t];
            float* dk = dkeys + ((long long)b * T + t) * AD;
#pragma unroll
            for (int dd = 0; dd < 4; ++dd) {
                const float th = tanhf(acc[dd]);
                dp[dd] = dev * s.va[lane + 32 * dd] * (1.f - th * th);
                dva_acc[dd] = fmaf(dev, th, dva_acc[dd]);
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
            dfl[((long long)b * T + t) * NF + lane] = df;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; ++k) {   // d(Wl)[f, d] += floc[t, f] * dpre[t, d] over the 8 positions of this pass
            const int idx = tid + 256 * k, ff = idx / AD, d = idx - ff * AD;
            float a = wl_acc[k];
#pragma unroll
            for (int w = 0; w < NW; ++w) a = fmaf(s_fl[w * NF + ff], s_dpre[w * AD + d], a);
            wl_acc[k] = a;
        }
        __syncthreads();
    }
    // cross-warp reduction of d(q) / d(v_a) through the (now free) dpre tile, then one atomic per element per CTA
    float* scr = s_dpre;
#pragma unroll
    for (int dd = 0; dd < 4; ++dd) scr[warp * AD + lane + 32 * dd] = dq_acc[dd];
    __syncthreads();
    if (tid < AD) {
        float v = 0.f;
        for (int w = 0; w < NW; ++w) v += scr[w * AD + tid];
        atomicAdd(dq + (long long)b * AD + tid, v);
        atomicAdd(p_ba + (long long)b * AD + tid, v);
    }
    __syncthreads();
#pragma unroll
    for (int dd = 0; dd < 4; ++dd) scr[warp * AD + lane + 32 * dd] = dva_acc[dd];
    __syncthreads();
    if (tid < AD) {
        float v = 0.f;
        for (int w = 0; w < NW; ++w) v += scr[w * AD + tid];
        atomicAdd(p_va + (long long)b * AD + tid, v);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) atomicAdd(p_Wl + (long long)b * NF * AD + tid + 256 * k, wl_acc[k]);
}

// location convolution gradients of one row: d(conv_w), d(conv_b) (+= into the row partials), d(cum_{t-1})
__global__ void __launch_bounds__(512) attn_bwd_conv_kernel(const float* __restrict__ dcum, const float* __restrict__ cum_prev,
                                                            const float* __restrict__ dfl, const int32_t* __restrict__ lens,
                                                            const float* __restrict__ conv_w, float* __restrict__ dcum_prev,
                                                            float* __restrict__ p_conv_w, float* __restrict__ p_conv_b, int T, int KW) {
    extern __shared__ float smem[];
    float* s_cum = smem;                       // [T + KW - 1] padded
    float* s_cw = s_cum + T + KW;              // [KW][NF]
    float* s_dfl = s_cw + KW * NF;             // [T][NF+1]
    const int b = blockIdx.x, tid = threadIdx.x;
    const int len = min(lens[b], T), left = (KW - 1) / 2;
    for (int i = tid; i < T + KW - 1; i += blockDim.x) {
        int t = i - left;
        s_cum[i] = (t >= 0 && t < T) ? cum_prev[(long long)b * T + t] : 0.f;
    }
    for (int i = tid; i < KW * NF; i += blockDim.x) s_cw[i] = conv_w[i];
    for (int i = tid; i < len * NF; i += blockDim.x) s_dfl[(i / NF) * (NF + 1) + (i % NF)] = dfl[(long long)b * T * NF + i];
    __syncthreads();
    for (int idx = tid; idx < KW * NF; idx += blockDim.x) {
        const int j = idx / NF, ff = idx - j * NF;
        float a = 0.f;
        for (int t = 0; t < len; ++t) a = fmaf(s_cum[t + j], s_dfl[t * (NF + 1) + ff], a);
        p_conv_w[(long long)b * KW * NF + idx] += a;
    }
    if (tid < NF) {
        float a = 0.f;
        for (int t = 0; t < len; ++t) a += s_dfl[t * (NF + 1) + tid];
        p_conv_b[(long long)b * NF + tid] += a;
    }
    // d(cum_{t-1})[tp]: 4 adjacent lanes share one position, each takes every 4th tap; combined with two shuffles
    const int Tq = (T + 63) / 64 * 64;           // keep whole warps in the loop (shuffles)
    for (int i = tid; i < Tq * 4; i += blockDim.x) {
        const int tp = i >> 2, jq = i & 3;
        float a = 0.f;
        if (tp < T)
            for (int j = jq; j < KW; j += 4) {
                const int t = tp - j + left;       // floc[t] read cum[t + j - left] = cum[tp]
                if (t >= 0 && t < len) {
                    const float* dfr = s_dfl + t * (NF + 1);
                    const float* cwr = s_cw + j * NF;
#pragma unroll 8
                    for (int ff = 0; ff < NF; ++ff) a = fmaf(dfr[ff], cwr[ff], a);
                }
            }
        a += __shfl_xor_sync(0xffffffffu, a, 1);
        a += __shfl_xor_sync(0xffffffffu, a, 2);
        if (jq == 0 && tp < T) dcum_prev[(long long)b * T + tp] = a + dcum[(long long)b * T + tp];
    }
}

// Scratch of taco_op_attn_bwd_noalloc: d(location features) [B][T][NF] in a file-scope device array (4 MiB per GPU that loads the
// library, held until exit).  In host code the name is only a handle: the device address comes from cudaGetSymbolAddress, once per
// device with that device current (noalloc_scratch()).
static __device__ __align__(256) float g_attn_bwd_scratch[TACO_ATTN_BWD_NOALLOC_MAX_FLOATS];

constexpr int MAX_DEVICES = 64;
float* g_scratch_ptr[MAX_DEVICES];       // per device, nullptr until resolved
std::mutex g_scratch_mu;

int noalloc_scratch(float** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return -3;
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        float* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_attn_bwd_scratch);
        if (e != cudaSuccess || !ptr) {
            fprintf(stderr, "[taco_b200] taco_op_attn_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
            return -3;
        }
        g_scratch_ptr[dev] = ptr;
    }
    *out = g_scratch_ptr[dev];
    return 0;
}

template <class K>
inline bool ensure_smem(K kernel, size_t bytes) {
    return bytes <= 48 * 1024 || cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) == cudaSuccess;
}

}  // namespace

extern "C" {

int taco_op_attn_fwd(const float* q, const float* cum_prev, const float* keys, const float* values, const int32_t* lens,
                     const float* conv_w, const float* conv_b, const float* Wl, const float* va, const float* ba,
                     float* align, float* cum, float* ctx, int B, int T, int A, int V, int NFf, int KW, void* stream) {
    if (!q || !cum_prev || !keys || !values || !lens || !conv_w || !conv_b || !Wl || !va || !ba || !align || !cum || !ctx) return -1;
    if (A != AD || NFf != NF || KW < 1 || KW > 63 || T < 1 || B < 1 || V % 4 != 0) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t sm = sizeof(float) * tile_smem_floats(KW);
    if (!ensure_smem(attn_energy_kernel, sm)) return -3;
    attn_energy_kernel<<<dim3((T + TPB - 1) / TPB, B), 256, sm, st>>>(q, cum_prev, keys, lens, conv_w, conv_b, Wl, va, ba, align, T, KW);
    attn_softmax_kernel<<<B, 256, 0, st>>>(cum_prev, lens, align, cum, T);
    attn_context_kernel<<<dim3((V + 127) / 128, B), 256, 0, st>>>(align, values, lens, ctx, T, V);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int taco_op_attn_bwd(const float* dctx, const float* dcum, const float* q, const float* cum_prev, const float* align,
                     const float* keys, const float* values, const int32_t* lens, const float* conv_w, const float* conv_b,
                     const float* Wl, const float* va, const float* ba, float* dq, float* dcum_prev, float* dkeys,
                     float* p_conv_w, float* p_conv_b, float* p_Wl, float* p_va, float* p_ba, float* scratch,
                     int B, int T, int A, int V, int NFf, int KW, void* stream) {
    if (!dctx || !dcum || !q || !cum_prev || !align || !keys || !values || !lens || !conv_w || !conv_b || !Wl || !va || !ba || !dq ||
        !dcum_prev || !dkeys || !p_conv_w || !p_conv_b || !p_Wl || !p_va || !p_ba || !scratch)
        return -1;
    if (A != AD || NFf != NF || KW < 1 || KW > 63 || T < 1 || B < 1 || V % 4 != 0) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    float* de = dcum_prev;       // d(align) -> d(energy) live in the d(cum_{t-1}) output until the last kernel overwrites it
    float* dfl = scratch;        // [B][T][NF]
    attn_bwd_dalign_kernel<<<dim3((T + NW - 1) / NW, B), 256, sizeof(float) * V, st>>>(dctx, dcum, values, lens, de, T, V);
    attn_bwd_softmax_kernel<<<B, 256, 0, st>>>(align, lens, de, dq, T);
    const size_t sm = sizeof(float) * (tile_smem_floats(KW) + NW * AD + NW * NF);
    if (!ensure_smem(attn_bwd_pre_kernel, sm)) return -3;
    attn_bwd_pre_kernel<<<dim3((T + TPB - 1) / TPB, B), 256, sm, st>>>(de, q, cum_prev, keys, lens, conv_w, conv_b, Wl, va, ba, dq, dkeys,
                                                                      p_Wl, p_va, p_ba, dfl, T, KW);
    const size_t sm2 = sizeof(float) * ((size_t)T + KW + (size_t)KW * NF + (size_t)T * (NF + 1));
    if (sm2 > 200 * 1024) return -1;
    if (!ensure_smem(attn_bwd_conv_kernel, sm2)) return -3;
    attn_bwd_conv_kernel<<<B, 512, sm2, st>>>(dcum, cum_prev, dfl, lens, conv_w, dcum_prev, p_conv_w, p_conv_b, T, KW);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int taco_op_attn_bwd_noalloc(const float* dctx, const float* dcum, const float* q, const float* cum_prev, const float* align,
                             const float* keys, const float* values, const int32_t* lens, const float* conv_w, const float* conv_b,
                             const float* Wl, const float* va, const float* ba, float* dq, float* dcum_prev, float* dkeys,
                             float* p_conv_w, float* p_conv_b, float* p_Wl, float* p_va, float* p_ba,
                             int B, int T, int A, int V, int NFf, int KW, void* stream) {
    if (B < 1 || T < 1 || NFf < 1) return -1;
    const long long need = (long long)B * T * NFf;
    if (need > (long long)TACO_ATTN_BWD_NOALLOC_MAX_FLOATS) {
        fprintf(stderr, "[taco_b200] taco_op_attn_bwd_noalloc: B*T*NF = %lld floats of scratch exceed the static limit of %d "
                        "(TACO_ATTN_BWD_NOALLOC_MAX_FLOATS); use taco_op_attn_bwd with caller-provided scratch\n",
                need, (int)TACO_ATTN_BWD_NOALLOC_MAX_FLOATS);
        return TACO_OP_ERR_SCRATCH_LIMIT;
    }
    float* scratch = nullptr;
    const int rc = noalloc_scratch(&scratch);
    if (rc) return rc;
    // nothing in the scratch has to start at zero: attn_bwd_pre_kernel writes every [b][t < T][f] entry (zeros beyond the
    // length) before attn_bwd_conv_kernel reads it, so the previous call's contents are never observed
    return taco_op_attn_bwd(dctx, dcum, q, cum_prev, align, keys, values, lens, conv_w, conv_b, Wl, va, ba, dq, dcum_prev, dkeys,
                            p_conv_w, p_conv_b, p_Wl, p_va, p_ba, scratch, B, T, A, V, NFf, KW, stream);
}

}  // extern "C"
