// train_attn_fused.cu -- the LocationSensitiveAttention step of the training recurrence (Attention.py:169-210,345-401) as
// ONE launch forward and ONE launch backward per decoder step, with everything that does not feed the recurrence hoisted
// out of the BPTT loop (train_attn.cu keeps the op-by-op kernels: they remain the path for kernel sizes != 31 / T > 256).
//
//   fold     U[j,d] = sum_f conv_w[j,f] Wl[f,d], c0[d] = sum_f conv_b[f] Wl[f,d]: the location convolution and the location
//            dense layer are one 31-tap filter bank [31,128] (no nonlinearity between them), folded once per training step
//   forward  grid (4, B rows) in clusters of the 4 CTAs of a row, 512 threads: q = h2raw . Wq, tanh pre-activations (a 31-tap
//            window of the cumulative alignments per thread in registers) of this CTA's quarter of the positions, stored as
//            tanh(pre) [T,128] for the backward pass; the partial energies go to all 4 CTAs through distributed shared
//            memory; then every CTA does the masked softmax + cumulative update and its 192 columns of the context
//   backward ONE launch, grid (4, B), clusters of the 4 CTAs of a row: d(align) = values . (d(ctx) + carry) + d(cum) of a
//            quarter of the positions, d(pre) and the tap gradients of that quarter (4x4 register tiles over the shared-memory
//            d(pre) tile), d(q) -> d(h2raw) += d(q) Wq^T, d(cum_{t-1}) through the folded filter bank; saves d(energy) [T] per
//            step.  Three cluster barriers: start, ONE exchange of the partial sums of the softmax gradient and of d(q)
//            (d(q) = v_a (S1 - dot S2) with S1, S2 computable before dot is known), and one after every CTA has PUSHED its
//            tap-gradient rows into the halo buffers of the CTAs that need them
//   hoisted  after the loop, grid (T/8, B) looping over all S steps: d(keys), d(U), d(c0), d(v_a) from the saved d(energy)
//            and tanh(pre); then d(conv_w), d(conv_b), d(Wl), d(b_a) from d(U), d(c0) by the chain rule of the fold
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/taco_b200_train.h"
#include "common.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int AD = 128, NF = 32, KW = 31, LEFT = 15;   // attention dim, location filters, taps
constexpr int FT = 512;                                // threads of the forward / row kernels
constexpr int CS = 4;                                  // value-column slices (CTAs per row) of the forward kernel
constexpr int RS = AD + 4;                             // shared-memory row stride of [.,128] tiles (16-byte aligned, +4 banks per row)
constexpr int GS = 33;                                 // row stride of the [T,31] tap-gradient tile

__global__ void __launch_bounds__(AD) attn_fold_kernel(const float* __restrict__ cw, const float* __restrict__ cb,
                                                       const float* __restrict__ Wl, float* __restrict__ U, float* __restrict__ c0) {
    const int j = blockIdx.x, d = threadIdx.x;
    float a = 0.f;
    for (int f = 0; f < NF; ++f) a = fmaf(j < KW ? cw[j * NF + f] : cb[f], Wl[f * AD + d], a);
    if (j < KW) U[j * AD + d] = a; else c0[d] = a;
}

__device__ __forceinline__ float block_reduce(float v, float* red, bool is_max) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    v = is_max ? warp_max(v) : warp_sum(v);
    __syncthreads();                     // red may still be read from a previous reduction
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float r = red[0];
    for (int w = 1; w < nw; ++w) r = is_max ? fmaxf(r, red[w]) : r + red[w];
    return r;
}

// ------------------------------------------------------------------------------------------------ forward
__global__ void __cluster_dims__(CS, 1, 1) __launch_bounds__(FT, 1) attn_fwd_fused_kernel(
    const float* __restrict__ h2raw, const float* __restrict__ Wq, const float* __restrict__ U, const float* __restrict__ c0,
    const float* __restrict__ ba, const float* __restrict__ va, const float* __restrict__ cum_prev, const float* __restrict__ keys,
    const float* __restrict__ values, const int32_t* __restrict__ lens, float* __restrict__ q_out, float* __restrict__ th_out,
    float* __restrict__ align, float* __restrict__ cum, float* __restrict__ ctx, int T, int D, int V) {
    extern __shared__ __align__(16) float smem[];
    cg::cluster_group cl = cg::this_cluster();
    const int Tp = (T + 3) & ~3;
    float* s_cum = smem;                       // [Tp + 32]: cum_prev[t - LEFT], zero padded
    float* s_e = s_cum + Tp + 32;              // [Tp][4] partial energies (one per warp of a 128-thread group) -> [Tp] alignments
    float* s_q = s_e + Tp * 4;                 // [AD]
    float* s_qp = s_q + AD;                    // [4][AD] K-split partials of q
    float* s_h = s_qp + 4 * AD;                // [D]
    float* s_red = s_h + D;                    // [32]
    float4* s_part = reinterpret_cast<float4*>(s_red + 32);   // [ntg][nq] context partials
    const int cs = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int len = min(lens[b], T);
    const int d = tid & (AD - 1), g = tid >> 7;
    const int nblk = (len + 3) >> 2;
    // the 4 CTAs of a row (one cluster) split the 4-position blocks of the row: CTA cs, thread group g takes blocks
    // cs + 4 g, cs + 4 g + 16, ...; the keys of the first block are in flight before anything else
    int blk = cs + CS * g;
    float kv[4];
#pragma unroll
    for (int p = 0; p < 4; ++p) kv[p] = (blk < nblk && blk * 4 + p < len) ? keys[((long long)b * T + blk * 4 + p) * AD + d] : 0.f;

    float Ur[KW], wq0[32];                      // filter-bank column and the first 32 rows of this thread's Wq column: in flight
#pragma unroll
    for (int j = 0; j < KW; ++j) Ur[j] = U[j * AD + d];
    {
        const int k0 = g * (D / 4), k1 = (g == 3) ? D : k0 + D / 4;
#pragma unroll
        for (int k = 0; k < 32; ++k) wq0[k] = k0 + k < k1 ? Wq[(long long)(k0 + k) * AD + d] : 0.f;
    }
    pdl_wait();                                 // everything above reads weights / keys; below: what the previous kernels wrote
    pdl_launch_dependents();
    for (int i = tid; i < Tp + 32; i += FT) {
        const int t = i - LEFT;
        s_cum[i] = (t >= 0 && t < T) ? cum_prev[(long long)b * T + t] : 0.f;
    }
    for (int i = tid; i < D; i += FT) s_h[i] = h2raw[(long long)b * D + i];
    __syncthreads();
    {   // q = h2raw . Wq, K split over the 4 thread groups (every CTA of the row computes it)
        const int k0 = g * (D / 4), k1 = (g == 3) ? D : k0 + D / 4;
        float a = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) a = fmaf(k0 + k < k1 ? s_h[k0 + k] : 0.f, wq0[k], a);
#pragma unroll 32
        for (int k = k0 + 32; k < k1; ++k) a = fmaf(s_h[k], Wq[(long long)k * AD + d], a);
        s_qp[g * AD + d] = a;
    }
    cl.sync();                                 // q partials visible; every CTA of the cluster runs (remote stores below are safe)
    if (tid < AD) {
        const float qv = s_qp[tid] + s_qp[AD + tid] + s_qp[2 * AD + tid] + s_qp[3 * AD + tid];
        if (cs == 0) q_out[(long long)b * AD + tid] = qv;
        s_q[tid] = qv + ba[tid] + c0[tid];
    }
    __syncthreads();
    const float qd = s_q[d], vad = va[d];
    float* s_e_rem = cl.map_shared_rank(s_e, lane & (CS - 1));     // lanes 0..3 forward a partial energy to the 4 CTAs of the row
    for (; blk < nblk; blk += CS * 4) {
        const int t0 = blk * 4, nxt = blk + CS * 4;
        float kn[4];
#pragma unroll
        for (int p = 0; p < 4; ++p) kn[p] = (nxt < nblk && nxt * 4 + p < len) ? keys[((long long)b * T + nxt * 4 + p) * AD + d] : 0.f;
        float w[36];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const float4 x = *reinterpret_cast<const float4*>(s_cum + t0 + 4 * k);
            w[4 * k] = x.x; w[4 * k + 1] = x.y; w[4 * k + 2] = x.z; w[4 * k + 3] = x.w;
        }
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < KW; ++j)
#pragma unroll
            for (int p = 0; p < 4; ++p) acc[p] = fmaf(Ur[j], w[p + j], acc[p]);
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int t = t0 + p;
            float e = 0.f;
            if (t < len) {
                const float th = tanhf(acc[p] + qd + kv[p]);
                th_out[((long long)b * T + t) * AD + d] = th;
                e = vad * th;
            }
            e = warp_sum(e);
            if (lane < CS) s_e_rem[t * 4 + (warp & 3)] = e;
            kv[p] = kn[p];
        }
    }
    cl.sync();
    __syncthreads();
    // masked softmax over t < len (alignments beyond the length are exact zeros), cumulative update.  Every warp reduces the
    // whole row by itself (shuffles only): no block barriers.
    float mx = -INFINITY;
    for (int t = lane; t < len; t += 32) mx = fmaxf(mx, s_e[t * 4] + s_e[t * 4 + 1] + s_e[t * 4 + 2] + s_e[t * 4 + 3]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int t = lane; t < len; t += 32) sum += expf(s_e[t * 4] + s_e[t * 4 + 1] + s_e[t * 4 + 2] + s_e[t * 4 + 3] - mx);
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    float av[2];                                      // up to 2 positions per thread (T <= 1024)
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int t = tid + k * FT;
        av[k] = (t < len) ? expf(s_e[t * 4] + s_e[t * 4 + 1] + s_e[t * 4 + 2] + s_e[t * 4 + 3] - mx) * inv : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int t = tid + k * FT;
        if (t < T) {
            s_e[t] = av[k];
            if (cs == 0) {
                align[(long long)b * T + t] = av[k];
                cum[(long long)b * T + t] = s_cum[t + LEFT] + av[k];
            }
        }
    }
    __syncthreads();
    // context columns of this slice: thread = (column quad, position group)
    const int V4 = V >> 2, nq = (V4 + CS - 1) / CS, ntg = FT / nq;
    const int cq = tid % nq, tg = tid / nq, c4 = cs * nq + cq;
    if (tg < ntg) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c4 < V4) {
            const float4* vp = reinterpret_cast<const float4*>(values) + (long long)b * T * V4 + c4;
            float4 v[12];                              // 12 value rows in flight per thread (one batch at T <= 120)
            for (int tb = tg; tb < len; tb += 12 * ntg) {
#pragma unroll
                for (int i = 0; i < 12; ++i) v[i] = (tb + i * ntg < len) ? vp[(long long)(tb + i * ntg) * V4] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int i = 0; i < 12; ++i) {
                    const float a = (tb + i * ntg < len) ? s_e[tb + i * ntg] : 0.f;
                    acc.x = fmaf(a, v[i].x, acc.x); acc.y = fmaf(a, v[i].y, acc.y); acc.z = fmaf(a, v[i].z, acc.z); acc.w = fmaf(a, v[i].w, acc.w);
                }
            }
        }
        s_part[tg * nq + cq] = acc;
    }
    __syncthreads();
    if (tid < nq && c4 < V4) {
        float4 acc = s_part[tid];
        for (int k = 1; k < ntg; ++k) { const float4 o = s_part[k * nq + tid]; acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w; }
        reinterpret_cast<float4*>(ctx)[(long long)b * V4 + c4] = acc;
    }
}

// ------------------------------------------------------------------------------------------------ backward, in the loop
// everything of the attention gradient that feeds the recurrence: the 4 CTAs of a row (one cluster) split its positions into
// contiguous quarters (d(pre) rows, tap gradients G) and the units of d(h2raw); d(q) partial sums and the G rows of the
// neighbouring quarters (the 31-tap halo of d(cum_{t-1})) cross the cluster through distributed shared memory
constexpr int XT = 32;    // tanh(pre) values a thread of the row kernel keeps in flight (positions g, g+4, ... of its quarter)
__global__ void __cluster_dims__(CS, 1, 1) __launch_bounds__(FT, 1) attn_bwd_row_kernel(
    float* __restrict__ dctx, const float* __restrict__ carry, const float* __restrict__ dcum, const float* __restrict__ values,
    const float* __restrict__ align, const float* __restrict__ th, const int32_t* __restrict__ lens, const float* __restrict__ U,
    const float* __restrict__ va, const float* __restrict__ Wq, float* __restrict__ de_out, float* __restrict__ dq_out,
    float* __restrict__ dh2, float* __restrict__ dcum_prev, int T, int D, int V, int Qmax) {
    extern __shared__ __align__(16) float smem[];
    cg::cluster_group cl = cg::this_cluster();
    const int Tp = (T + 3) & ~3;
    float* s_dpre = smem;                         // [Qmax][RS]   d(pre) rows of this quarter
    float* s_U = s_dpre + (size_t)Qmax * RS;      // [32][RS], row 31 = 0
    float* s_H = s_U + 32 * RS;                   // [Qmax + 30][GS]: G[t][j] = sum_d dpre[t,d] U[j,d] of positions t_lo - 15 ... (own rows + halo)
    float* s_de = s_H + (size_t)(Qmax + 32) * GS; // [Tp]   (Qmax + 32 rows: keeps the float4 regions below 16-byte aligned)
    float* s_al = s_de + Tp;                      // [Tp] alignments of the row
    float* s_dq = s_al + Tp;                      // [2][4][AD] partial sums S1, S2 of this CTA's thread groups -> [AD] d(q)
    float* s_dqx = s_dq + 8 * AD;                 // [2][CS][AD] S1, S2 of every CTA of the row
    float* s_red = s_dqx + 2 * CS * AD;           // [32]
    float* s_dot = s_red + 32;                    // [CS] partial sums of a . d(align)
    float* s_dctx = s_dot + CS;                   // [V] d(ctx) + carry
    const int cs = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int len = min(lens[b], T);
    const int lenp = (len + 3) & ~3;
    const int Q = (((lenp >> 2) + CS - 1) / CS) * 4;               // positions per CTA (whole 4-position blocks), <= Qmax
    const int t_lo = cs * Q, nq = max(0, min(lenp, t_lo + Q) - t_lo);   // this CTA's rows: t_lo .. t_lo + nq - 1
    const int d = tid & (AD - 1), g = tid >> 7;
    // tanh(pre) of this thread's positions (rows g, g+4, ... of the quarter) in flight first: the longest chain waits for them
    const int nvalid = min(nq, len - t_lo);                        // rows of this quarter inside the length
    const int ni = nvalid > g ? (nvalid - g + 3) >> 2 : 0;
    const float* thp = th + ((long long)b * T + t_lo + g) * AD + d;
    float x[XT];
#pragma unroll
    for (int i = 0; i < XT; ++i) x[i] = i < ni ? thp[(long long)i * 4 * AD] : 0.f;
    // d(align) of this quarter's rows needs the value rows: the first row of every warp is in flight as well
    const int V4 = V >> 2;
    const float4* val4 = reinterpret_cast<const float4*>(values) + (long long)b * T * V4;
    float4 v0[6];
#pragma unroll
    for (int k = 0; k < 6; ++k)
        v0[k] = (warp < nvalid && lane + 32 * k < V4) ? val4[(long long)(t_lo + warp) * V4 + lane + 32 * k] : make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = tid; i < 32 * AD; i += FT) s_U[(i >> 7) * RS + (i & 127)] = (i >> 7) < KW ? U[i] : 0.f;
    for (int t = tid; t < Tp; t += FT) s_al[t] = t < len ? align[(long long)b * T + t] : 0.f;
    pdl_wait();                                  // everything above reads tensors of the forward pass; below: this step's gradients
    pdl_launch_dependents();
    for (int i = tid; i < V; i += FT) s_dctx[i] = dctx[(long long)b * V + i] + carry[(long long)b * V + i];
    cl.sync();                                   // every CTA of the row runs (remote stores below are safe) and has read d(ctx)
    if (cs == 0)
        for (int i = tid; i < V; i += FT) dctx[(long long)b * V + i] = s_dctx[i];      // total d(ctx_t), kept for d(values)
    // d(align)[t] = values[t,:] . d(ctx) + d(cum_t)[t], one warp per row; partial sum_t a[t] d(align)[t] of this quarter
    float part = 0.f;
    for (int r = warp; r < nq; r += FT / 32) {
        const int t = t_lo + r;
        float acc = 0.f;
        if (r < nvalid) {
            const float4* vp = val4 + (long long)t * V4;
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                const int c4 = lane + 32 * k;
                if (c4 < V4) {
                    const float4 v = (r == warp) ? v0[k] : vp[c4];
                    const float4 g4 = *reinterpret_cast<const float4*>(s_dctx + 4 * c4);
                    acc = fmaf(v.x, g4.x, acc); acc = fmaf(v.y, g4.y, acc); acc = fmaf(v.z, g4.z, acc); acc = fmaf(v.w, g4.w, acc);
                }
            }
            for (int c4 = lane + 192; c4 < V4; c4 += 32) {
                const float4 v = vp[c4];
                const float4 g4 = *reinterpret_cast<const float4*>(s_dctx + 4 * c4);
                acc = fmaf(v.x, g4.x, acc); acc = fmaf(v.y, g4.y, acc); acc = fmaf(v.z, g4.z, acc); acc = fmaf(v.w, g4.w, acc);
            }
            acc = warp_sum(acc) + dcum[(long long)b * T + t];
            part = fmaf(s_al[t], acc, part);
        }
        if (lane == 0) s_de[t] = acc;            // d(align) for now; 0 on the padding rows
    }
    part = block_reduce(lane == 0 ? part : 0.f, s_red, false);      // (its barriers also publish s_de = d(align) of the quarter)
    // ONE exchange for the softmax gradient and d(q):  d(e)[t] = a[t] (d(a)[t] - dot),  dot = sum_t a[t] d(a)[t]  and
    // d(q)[d] = v_a[d] sum_t d(e)[t] (1 - th[t,d]^2) = v_a[d] (S1[d] - dot S2[d])  with  S1 = sum_t a d(a) (1 - th^2),
    // S2 = sum_t a (1 - th^2): both sums are available BEFORE dot is known, so the partial (dot, S1, S2) of the 4 CTAs cross the
    // cluster together
    const float vad = va[d];
    {
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < XT; ++i) {
            const int r = g + 4 * i;
            if (r < nq) {
                const float w = s_al[t_lo + r] * (1.f - x[i] * x[i]);          // 0 beyond the length: a = 0
                s2 += w;
                s1 = fmaf(w, s_de[t_lo + r], s1);
            }
        }
        for (int r = g + 4 * XT; r < nvalid; r += 4) {                          // quarters longer than 128 positions
            const float xv = th[((long long)b * T + t_lo + r) * AD + d];
            const float w = s_al[t_lo + r] * (1.f - xv * xv);
            s2 += w;
            s1 = fmaf(w, s_de[t_lo + r], s1);
        }
        s_dq[g * AD + d] = s1;
        s_dq[(4 + g) * AD + d] = s2;
    }
    __syncthreads();
    if (tid < 2 * AD) {
        const int which = tid >> 7, dd = tid & (AD - 1);
        const float* src = s_dq + which * 4 * AD + dd;
        const float v = src[0] + src[AD] + src[2 * AD] + src[3 * AD];
#pragma unroll
        for (int r = 0; r < CS; ++r) cl.map_shared_rank(s_dqx, r)[(which * CS + cs) * AD + dd] = v;
    } else if (tid < 2 * AD + CS) {
        cl.map_shared_rank(s_dot, tid - 2 * AD)[cs] = part;
    }
    cl.sync();                                   // the partial sums of all 4 CTAs are everywhere
    const float dot = s_dot[0] + s_dot[1] + s_dot[2] + s_dot[3];
    if (tid < AD) {
        float S1 = 0.f, S2 = 0.f;
#pragma unroll
        for (int r = 0; r < CS; ++r) { S1 += s_dqx[r * AD + tid]; S2 += s_dqx[(CS + r) * AD + tid]; }
        const float v = va[tid] * (S1 - dot * S2);
        s_dq[tid] = v;                           // (the partials in s_dq were consumed before the cluster barrier)
        if (cs == 0) dq_out[(long long)b * AD + tid] = v;
    }
    for (int r = tid; r < nq; r += FT) {         // d(energy) = a (d(align) - sum a d(align)); 0 beyond the length
        const int t = t_lo + r;
        const float v = s_al[t] * (s_de[t] - dot);
        s_de[t] = v;
        if (t < T) de_out[(long long)b * T + t] = v;
    }
    if (cs == CS - 1)
        for (int t = CS * Q + tid; t < T; t += FT) de_out[(long long)b * T + t] = 0.f;
    // d(h2raw)[k] += sum_d d(q)[d] Wq[k,d] for this CTA's quarter of the units, one warp per k.  The loads are issued here; the
    // products are finished after the G tiles.
    const int kq = (D + CS - 1) / CS, k_end = min(D, (cs + 1) * kq), k0w = cs * kq + warp * 4;
    float wv[4][4], dh_old = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < 4; ++c) wv[i][c] = (k0w + i < k_end) ? Wq[(long long)(k0w + i) * AD + lane + 32 * c] : 0.f;
    if (lane < 4 && k0w + lane < k_end) dh_old = dh2[(long long)b * D + k0w + lane];
    __syncthreads();
    // d(pre)[t,d] = d(e)[t] v_a[d] (1 - tanh^2)
#pragma unroll
    for (int i = 0; i < XT; ++i) {
        const int r = g + 4 * i;
        if (r < nq) s_dpre[r * RS + d] = s_de[t_lo + r] * vad * (1.f - x[i] * x[i]);     // 0 beyond the length: s_de = 0
    }
    for (int r = g + 4 * XT; r < nq; r += 4) {
        const int t = t_lo + r;
        float dp = 0.f;
        if (t < len) { const float xv = th[((long long)b * T + t) * AD + d]; dp = s_de[t] * vad * (1.f - xv * xv); }
        s_dpre[r * RS + d] = dp;
    }
    __syncthreads();
    // G tiles of this quarter: 4 positions x 4 taps (taps jt, jt+8, jt+16, jt+24) x one eighth of d per work item; the 8
    // d-parts of a tile sit in 8 adjacent lanes (interleaved 16-byte columns: conflict-free loads) and combine by shuffles.
    // A finished G row goes to every CTA whose d(cum) positions reach it (its own rows +- 15): pushed, so that nobody reads
    // remote shared memory after the last cluster barrier.
    const int hc = tid & (CS - 1);               // lanes 0..3 of a tile's 8 lanes serve CTA 0..3 (p8 = tid & 7 below)
    float* Hc = cl.map_shared_rank(s_H, hc);
    const int h_lo = hc * Q - LEFT, h_hi = (hc == CS - 1) ? lenp : min(lenp, hc * Q + Q + LEFT);
    const int nitems = (nq >> 2) * 64;           // a multiple of 64: whole warps in or out
    for (int it = tid; it < nitems; it += FT) {
        const int p8 = it & 7, jt = (it >> 3) & 7, tt = it >> 6;
        const float* pr = s_dpre + (size_t)(tt * 4) * RS + p8 * 4;
        const float* ur = s_U + (size_t)jt * RS + p8 * 4;
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[i][k] = 0.f;
#pragma unroll
        for (int d4 = 0; d4 < 4; ++d4) {
            float4 p[4], u[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) p[i] = *reinterpret_cast<const float4*>(pr + i * RS + 32 * d4);
#pragma unroll
            for (int k = 0; k < 4; ++k) u[k] = *reinterpret_cast<const float4*>(ur + (size_t)(8 * k) * RS + 32 * d4);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    acc[i][k] = fmaf(p[i].x, u[k].x, acc[i][k]); acc[i][k] = fmaf(p[i].y, u[k].y, acc[i][k]);
                    acc[i][k] = fmaf(p[i].z, u[k].z, acc[i][k]); acc[i][k] = fmaf(p[i].w, u[k].w, acc[i][k]);
                }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int t = t_lo + tt * 4 + i;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float a = acc[i][k];
                a += __shfl_xor_sync(0xffffffffu, a, 1);
                a += __shfl_xor_sync(0xffffffffu, a, 2);
                a += __shfl_xor_sync(0xffffffffu, a, 4);
                if (p8 < CS && t >= h_lo && t < h_hi && jt + 8 * k < KW) Hc[(size_t)(t - h_lo) * GS + jt + 8 * k] = a;
            }
        }
    }
    {   // finish d(h2raw)
        float dqr[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) dqr[i] = s_dq[lane + 32 * i];
        float mine = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float a = wv[i][0] * dqr[0] + wv[i][1] * dqr[1] + wv[i][2] * dqr[2] + wv[i][3] * dqr[3];
            a = warp_sum(a);
            if (lane == i) mine = a;
        }
        if (lane < 4 && k0w + lane < k_end) dh2[(long long)b * D + k0w + lane] = dh_old + mine;
        for (int k0 = k0w + (FT / 32) * 4; k0 < k_end; k0 += (FT / 32) * 4) {       // D > 256
            for (int i = 0; i < 4 && k0 + i < k_end; ++i) {
                float a = 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) a = fmaf(Wq[(long long)(k0 + i) * AD + lane + 32 * c], dqr[c], a);
                a = warp_sum(a);
                if (lane == 0) dh2[(long long)b * D + k0 + i] += a;
            }
        }
    }
    cl.sync();                                   // all G rows (own and halo) of every CTA are in place; no remote access after this
    // d(cum_{t-1})[tp] = d(cum_t)[tp] + sum_j G[tp - j + LEFT][j]   (pre[t] reads cum[t + j - LEFT]).  CTA cs takes the tp of its
    // quarter, the last one also everything beyond 4 Q; one warp per tp, lane = tap, then a warp sum.
    {
        const int tp_hi = (cs == CS - 1) ? T : min(T, t_lo + Q);
        for (int tp = t_lo + warp; tp < tp_hi; tp += FT / 32) {
            const int t = tp - lane + LEFT;
            float a = 0.f;
            if (lane < KW && t >= 0 && t < len) a = s_H[(size_t)(t - (t_lo - LEFT)) * GS + lane];
            a = warp_sum(a);
            if (lane == 0) dcum_prev[(long long)b * T + tp] = a + dcum[(long long)b * T + tp];
        }
    }
}

// ------------------------------------------------------------------------------------------------ backward, hoisted
// All S steps of a tile of 8 positions of one row: d(keys)[b,t,:] += sum_s d(pre), and this CTA's partial sums of
// d(U)[j,d] = sum cum_{s-1}[t + j - LEFT] d(pre)[t,d], d(c0)[d] = sum d(pre), d(v_a)[d] = sum d(e) tanh(pre) -> acc[b][...]
constexpr int HP = 8, HSB = 4;      // positions per CTA, steps per shared-memory refill
constexpr int ACC_N = KW * AD + 2 * AD;   // d(U) | d(c0) | d(v_a)
__global__ void __launch_bounds__(256) attn_bwd_hoisted_kernel(const float* __restrict__ de, const float* __restrict__ th,
                                                               const float* __restrict__ cum, const int32_t* __restrict__ lens,
                                                               const float* __restrict__ va, float* __restrict__ dkeys,
                                                               float* __restrict__ acc_out, int S, int B, int T) {
    __shared__ __align__(16) float s_cum[HSB][HP + 32];   // cum[t0 - LEFT ..]
    __shared__ float s_de[HSB][HP];
    __shared__ float s_x[2][ACC_N / 2 + AD];              // cross-half reduction scratch
    const int b = blockIdx.y, t0 = blockIdx.x * HP, tid = threadIdx.x, d = tid & (AD - 1), hf = tid >> 7;
    const int len = min(lens[b], T);
    if (t0 >= len) return;
    const float vad = va[d];
    float dU[KW], dk[4] = {0.f, 0.f, 0.f, 0.f}, dc0 = 0.f, dva = 0.f;
#pragma unroll
    for (int j = 0; j < KW; ++j) dU[j] = 0.f;
    const int tb = t0 + hf * 4;               // this thread's 4 positions
    for (int s0 = 0; s0 < S; s0 += HSB) {
        const int ns = min(HSB, S - s0);
        float x[HSB][4];
#pragma unroll
        for (int k = 0; k < HSB; ++k)
#pragma unroll
            for (int p = 0; p < 4; ++p)
                x[k][p] = (k < ns && tb + p < len) ? th[(((long long)(s0 + k) * B + b) * T + tb + p) * AD + d] : 0.f;
        __syncthreads();
        for (int i = tid; i < HSB * (HP + 32); i += 256) {
            const int k = i / (HP + 32), o = i - k * (HP + 32), t = t0 + o - LEFT;
            s_cum[k][o] = (k < ns && t >= 0 && t < T) ? cum[((long long)(s0 + k) * B + b) * T + t] : 0.f;
        }
        if (tid < HSB * HP) {
            const int k = tid / HP, o = tid - k * HP;
            s_de[k][o] = (k < ns && t0 + o < len) ? de[((long long)(s0 + k) * B + b) * T + t0 + o] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < HSB; ++k) {
            float dp[4];
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const float dev = s_de[k][hf * 4 + p];
                dva = fmaf(dev, x[k][p], dva);
                dp[p] = dev * vad * (1.f - x[k][p] * x[k][p]);
                dk[p] += dp[p];
                dc0 += dp[p];
            }
            float w[36];
#pragma unroll
            for (int i = 0; i < 9; ++i) {
                const float4 c = *reinterpret_cast<const float4*>(&s_cum[k][hf * 4 + 4 * i]);
                w[4 * i] = c.x; w[4 * i + 1] = c.y; w[4 * i + 2] = c.z; w[4 * i + 3] = c.w;
            }
#pragma unroll
            for (int j = 0; j < KW; ++j)
#pragma unroll
                for (int p = 0; p < 4; ++p) dU[j] = fmaf(w[p + j], dp[p], dU[j]);
        }
    }
#pragma unroll
    for (int p = 0; p < 4; ++p)
        if (tb + p < len) dkeys[((long long)b * T + tb + p) * AD + d] += dk[p];
    // the two position halves of the CTA combine in shared memory; one atomic per element and CTA into the row's accumulator
    __syncthreads();
    float* sx = &s_x[0][0];
    if (hf == 1) {
#pragma unroll
        for (int j = 0; j < KW; ++j) sx[j * AD + d] = dU[j];
        sx[KW * AD + d] = dc0; sx[(KW + 1) * AD + d] = dva;
    }
    __syncthreads();
    if (hf == 0) {
        float* out = acc_out + (long long)b * ACC_N;
#pragma unroll
        for (int j = 0; j < KW; ++j) atomicAdd(out + j * AD + d, dU[j] + sx[j * AD + d]);
        atomicAdd(out + KW * AD + d, dc0 + sx[KW * AD + d]);
        atomicAdd(out + (KW + 1) * AD + d, dva + sx[(KW + 1) * AD + d]);
    }
}

// chain rule of the fold: block j < KW: d(conv_w)[j,:]; block KW + f: d(Wl)[f,:]; block KW + NF: d(conv_b), d(v_a), d(b_a).
// acc [B][ACC_N] is summed over the rows on the fly; results are ADDED to the outputs.
__global__ void __launch_bounds__(AD) attn_fold_bwd_kernel(const float* __restrict__ acc, const float* __restrict__ cw,
                                                           const float* __restrict__ cb, const float* __restrict__ Wl,
                                                           float* __restrict__ g_cw, float* __restrict__ g_cb, float* __restrict__ g_Wl,
                                                           float* __restrict__ g_va, float* __restrict__ g_ba, int B) {
    __shared__ float s_v[AD];
    const int blk = blockIdx.x, d = threadIdx.x, lane = d & 31, warp = d >> 5;
    auto rowsum = [&](int off) { float a = 0.f; for (int b = 0; b < B; ++b) a += acc[(long long)b * ACC_N + off + d]; return a; };
    if (blk < KW || blk == KW + NF) {                 // d(conv_w)[j,f] = sum_d dU[j,d] Wl[f,d]; d(conv_b)[f] = sum_d dc0[d] Wl[f,d]
        const float v = rowsum(blk < KW ? blk * AD : KW * AD);
        s_v[d] = v;
        __syncthreads();
        for (int f = warp; f < NF; f += AD / 32) {
            float a = 0.f;
            for (int k = lane; k < AD; k += 32) a = fmaf(s_v[k], Wl[f * AD + k], a);
            a = warp_sum(a);
            if (lane == 0) { if (blk < KW) g_cw[blk * NF + f] += a; else g_cb[f] += a; }
        }
        if (blk == KW + NF) { g_ba[d] += v; g_va[d] += rowsum((KW + 1) * AD); }
    } else {                                          // d(Wl)[f,d] = sum_j cw[j,f] dU[j,d] + cb[f] dc0[d]
        const int f = blk - KW;
        float a = cb[f] * rowsum(KW * AD);
        for (int j = 0; j < KW; ++j) a = fmaf(cw[j * NF + f], rowsum(j * AD), a);
        g_Wl[f * AD + d] += a;
    }
}

// d(values)[b,t,c] = sum_s align[s,b,t] d(ctx)[s,b,c]: one launch for all rows (it was one [T,S] x [S,V] GEMM per batch row).
// grid (V / 64, ceil(T / 32), B), 256 threads: 32 positions x 64 columns per CTA, 2 x 4 outputs per thread, S in chunks of 32.
__global__ void __launch_bounds__(256) attn_dvalues_kernel(const float* __restrict__ align, const float* __restrict__ dctx,
                                                           float* __restrict__ dvalues, int S, int B, int T, int V) {
    __shared__ float sA[32][33];
    __shared__ __align__(16) float sB[32][68];
    const int c0 = blockIdx.x * 64, t0 = blockIdx.y * 32, b = blockIdx.z, tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
    for (int s0 = 0; s0 < S; s0 += 32) {
        float ra[4];
        float4 rb[2];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int e = tid + 256 * i, ss = s0 + (e >> 5), t = t0 + (e & 31);
            ra[i] = (ss < S && t < T) ? align[((long long)ss * B + b) * T + t] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int e = tid + 256 * i, ss = s0 + (e >> 4);
            rb[i] = ss < S ? *reinterpret_cast<const float4*>(dctx + ((long long)ss * B + b) * V + c0 + 4 * (e & 15)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) { const int e = tid + 256 * i; sA[e >> 5][e & 31] = ra[i]; }
#pragma unroll
        for (int i = 0; i < 2; ++i) { const int e = tid + 256 * i; *reinterpret_cast<float4*>(&sB[e >> 4][4 * (e & 15)]) = rb[i]; }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < 32; ++k) {
            const float a0 = sA[k][2 * ty], a1 = sA[k][2 * ty + 1];
            const float4 v = *reinterpret_cast<const float4*>(&sB[k][4 * tx]);
            acc[0][0] = fmaf(a0, v.x, acc[0][0]); acc[0][1] = fmaf(a0, v.y, acc[0][1]); acc[0][2] = fmaf(a0, v.z, acc[0][2]); acc[0][3] = fmaf(a0, v.w, acc[0][3]);
            acc[1][0] = fmaf(a1, v.x, acc[1][0]); acc[1][1] = fmaf(a1, v.y, acc[1][1]); acc[1][2] = fmaf(a1, v.z, acc[1][2]); acc[1][3] = fmaf(a1, v.w, acc[1][3]);
        }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int t = t0 + 2 * ty + i;
        if (t < T) *reinterpret_cast<float4*>(dvalues + ((long long)b * T + t) * V + c0 + 4 * tx) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
}

template <class K>
inline bool ensure_smem(K kernel, size_t bytes) {
    return bytes <= 48 * 1024 || cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) == cudaSuccess;
}
inline size_t fwd_smem(int T, int D, int V) {
    const int Tp = (T + 3) & ~3, nq = ((V >> 2) + CS - 1) / CS, ntg = FT / nq;
    return sizeof(float) * ((size_t)Tp + 32 + (size_t)Tp * 4 + AD + 4 * AD + D + 32) + sizeof(float4) * (size_t)ntg * nq;
}
inline int row_qmax(int T) { return ((((T + 3) >> 2) + CS - 1) / CS) * 4; }     // positions per CTA of the row kernel at len = T
inline size_t row_smem(int T, int V) {
    const int Tp = (T + 3) & ~3, Q = row_qmax(T);
    return sizeof(float) * ((size_t)Q * RS + 32 * RS + (size_t)(Q + 32) * GS + 2 * Tp + 8 * AD + 2 * CS * AD + 32 + CS + V);
}

}  // namespace

extern "C" {

int taco_op_attn_fused_supported(int T, int D, int A, int V, int NFf, int KWf) {
    return A == AD && NFf == NF && KWf == KW && T >= 1 && T <= 1024 && D >= 4 && D % 4 == 0 && V % 4 == 0 &&
           ((V >> 2) + CS - 1) / CS <= FT && row_smem(T, V) <= 200 * 1024 && fwd_smem(T, D, V) <= 200 * 1024;
}

int taco_op_attn_fold(const float* conv_w, const float* conv_b, const float* Wl, float* U, float* c0, void* stream) {
    if (!conv_w || !conv_b || !Wl || !U || !c0) return -1;
    attn_fold_kernel<<<KW + 1, AD, 0, (cudaStream_t)stream>>>(conv_w, conv_b, Wl, U, c0);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int taco_op_attn_fwd_fused(const float* h2raw, const float* Wq, const float* U, const float* c0, const float* ba, const float* va,
                           const float* cum_prev, const float* keys, const float* values, const int32_t* lens, float* q, float* th,
                           float* align, float* cum, float* ctx, int B, int T, int D, int V, void* stream) {
    if (!h2raw || !Wq || !U || !c0 || !ba || !va || !cum_prev || !keys || !values || !lens || !q || !th || !align || !cum || !ctx) return -1;
    if (B < 1 || !taco_op_attn_fused_supported(T, D, AD, V, NF, KW)) return -1;
    const size_t sm = fwd_smem(T, D, V);
    if (!ensure_smem(attn_fwd_fused_kernel, sm)) return -3;
    launch_pdl(attn_fwd_fused_kernel, dim3(CS, B), dim3(FT), sm, (cudaStream_t)stream, h2raw, Wq, U, c0, ba, va, cum_prev, keys, values, lens, q,
               th, align, cum, ctx, T, D, V);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int taco_op_attn_bwd_fused(float* dctx, const float* carry, const float* dcum, const float* align, const float* th, const float* values,
                           const int32_t* lens, const float* U, const float* va, const float* Wq, float* de, float* dq, float* dh2,
                           float* dcum_prev, float* scratch, int B, int T, int D, int V, void* stream) {
    if (!dctx || !carry || !dcum || !align || !th || !values || !lens || !U || !va || !Wq || !de || !dq || !dh2 || !dcum_prev || !scratch) return -1;
    if (B < 1 || !taco_op_attn_fused_supported(T, D, AD, V, NF, KW)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t sm = row_smem(T, V);
    if (!ensure_smem(attn_bwd_row_kernel, sm)) return -3;
    launch_pdl(attn_bwd_row_kernel, dim3(CS, B), dim3(FT), sm, st, dctx, carry, dcum, values, align, th, lens, U, va, Wq, de, dq, dh2, dcum_prev,
               T, D, V, row_qmax(T));
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int taco_op_attn_dvalues(const float* align, const float* dctx, float* dvalues, int S, int B, int T, int V, void* stream) {
    if (!align || !dctx || !dvalues || S < 1 || B < 1 || T < 1 || V < 64 || V % 64 != 0) return -1;
    if ((reinterpret_cast<uintptr_t>(dctx) & 15) || (reinterpret_cast<uintptr_t>(dvalues) & 15)) return -1;
    attn_dvalues_kernel<<<dim3(V / 64, (T + 31) / 32, B), 256, 0, (cudaStream_t)stream>>>(align, dctx, dvalues, S, B, T, V);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

int taco_op_attn_bwd_hoisted(const float* de, const float* th, const float* cum, const int32_t* lens, const float* conv_w,
                             const float* conv_b, const float* Wl, const float* va, float* dkeys, float* g_conv_w, float* g_conv_b,
                             float* g_Wl, float* g_va, float* g_ba, float* work, int S, int B, int T, void* stream) {
    if (!de || !th || !cum || !lens || !conv_w || !conv_b || !Wl || !va || !dkeys || !g_conv_w || !g_conv_b || !g_Wl || !g_va || !g_ba || !work)
        return -1;
    if (S < 1 || B < 1 || T < 1) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(work, 0, sizeof(float) * (size_t)B * ACC_N, st) != cudaSuccess) return -3;
    attn_bwd_hoisted_kernel<<<dim3((T + HP - 1) / HP, B), 256, 0, st>>>(de, th, cum, lens, va, dkeys, work, S, B, T);
    attn_fold_bwd_kernel<<<KW + NF + 1, AD, 0, st>>>(work, conv_w, conv_b, Wl, g_conv_w, g_conv_b, g_Wl, g_va, g_ba, B);
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // extern "C"
