// Persistent cooperative decoder kernel (sm_100a): all S decoder steps of
// TacotronDecoderCell.__call__ (reference TacotronModel/modules/Decoder.py:246-292) +
// CustomDecoder.step / TacoTestHelper / TacoTrainingHelper (Decoder.py:117-146,
// Utils/Helpers.py:36-59,115-128) + LocationSensitiveAttention (Attention.py:169-210,345-401)
// in ONE launch.  Prenet / LSTM / query / projection weights are partitioned across the grid by
// output column and stay resident in shared memory for every step; the small activation vectors
// are exchanged through L2 in a k-major "XT" layout; phases are separated by a grid barrier.
//
// Per step (7 grid barriers):
//   A prenet dense_1 (+ReLU+dropout)            x_t[80]      -> p1[256]
//   B prenet dense_2 (+ReLU+dropout)            p1           -> p2[256]
//   C LSTM layer 1 (zoneout)                    [p2,ctx,h1z] -> h1raw, h1z', c1'
//   D LSTM layer 2 (zoneout)                    [h1raw,h2z]  -> h2raw, h2z', c2'
//   E query_layer  ||  location features        h2raw -> q ;  cum -> pre = keys + conv31(cum)*Wl + b_a
//   F energies + masked softmax + cumulate + context   (row b, d-slice) work items
//   G frame + stop projection, outputs, feedback, stop test
#include "common.cuh"
#include "taco_internal.h"

namespace {

constexpr int MEL = 80;
constexpr int R_OUT = 2;
constexpr int PRE = 256;   // prenet units
constexpr int DEC = 256;   // decoder LSTM units
constexpr int MEM = 768;   // memory width
constexpr int ATT = 128;   // attention dim
constexpr int LOCF = 32;   // location filters
constexpr int LOCK = 31;   // location kernel
constexpr int K1 = PRE + MEM + DEC;  // 1280
constexpr int K2 = DEC + DEC;        // 512
constexpr int KO = DEC + MEM;        // 1024
constexpr int NOUT = MEL * R_OUT + R_OUT;  // 162
constexpr int T_MAX = 1024;
constexpr int LOC_CHUNK = 64;

__device__ __forceinline__ bool keep_bit(const DecParams& p, int step, int b, int layer, int unit) {
    if (p.masks) return p.masks[(((size_t)(step - p.t0) * p.B + b) * 2 + layer) * PRE + unit] != 0;
    return philox_keep(p.seed, (uint32_t)step, (uint32_t)b, (uint32_t)layer, (uint32_t)unit);
}

__global__ void __launch_bounds__(TACO_THREADS, 1) decoder_persistent_kernel(DecParams p) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int cta = blockIdx.x;
    const int nCta = gridDim.x;
    const int U = p.U, NCL = 4 * p.U, RBP = p.RBP, GN = p.GN;
    const int bt = cta / GN, nt = cta % GN;
    const int PCN = PRE / GN;  // real prenet columns of this CTA (2 or 4), padded to 4
    const int QCN = ATT / GN;  // real query columns (1 or 2), padded to 4
    const bool has_proj = nt * 4 < NOUT;

    // ---- shared memory carve-up --------------------------------------------------------------
    float* W1 = smem;                       // [K1][NCL]
    float* W2 = W1 + K1 * NCL;              // [K2][NCL]
    float* Wp1 = W2 + K2 * NCL;             // [80][4]
    float* Wp2 = Wp1 + MEL * 4;             // [256][4]
    float* Wq = Wp2 + PRE * 4;              // [256][4]
    float* Wo = Wq + DEC * 4;               // [1024][4]
    float* bias = Wo + KO * 4;              // b1[NCL] b2[NCL] bp1[4] bp2[4] bo[4]
    const int blob = p.blob_floats;         // == offset of end of bias block
    float* part = smem + blob;              // [512*16]
    float* res = part + TACO_THREADS * 16;  // [32*16]
    float* cst1 = res + 32 * 16;            // [RBP][U] private cell state layer 1
    float* cst2 = cst1 + 32 * 4;            // [RBP][U]
    float* es = cst2 + 32 * 4;              // [T_MAX] energies / alignments
    float* qs = es + T_MAX;                 // [128]
    float* va_s = qs + ATT;                 // [128]
    float* red = va_s + ATT;                // [32] block reductions
    float* locw = red + 32;                 // conv [31][32] + bias[32] + Wl [32][128] + cum window [LOC_CHUNK+30] + f [32][LOC_CHUNK]
    float* conv_s = locw;
    float* convb_s = conv_s + LOCK * LOCF;
    float* Wl_s = convb_s + LOCF;
    float* cw_s = Wl_s + LOCF * ATT;
    float* f_s = cw_s + (LOC_CHUNK + 32);   // [LOCF][LOC_CHUNK]

    // ---- one-time loads -----------------------------------------------------------------------
    {
        const float4* src = reinterpret_cast<const float4*>(p.blobs + (size_t)nt * blob);
        float4* dst = reinterpret_cast<float4*>(smem);
        for (int i = tid; i < blob / 4; i += TACO_THREADS) dst[i] = src[i];
        for (int i = tid; i < LOCK * LOCF; i += TACO_THREADS) conv_s[i] = p.conv_w[i];
        if (tid < LOCF) convb_s[tid] = p.conv_b[tid];
        for (int i = tid; i < LOCF * ATT; i += TACO_THREADS) Wl_s[i] = p.Wl[i];
        if (tid < ATT) va_s[tid] = p.v_a[tid];
        for (int i = tid; i < RBP * U; i += TACO_THREADS) {
            int r = i / U, ul = i % U;
            int b = bt * RBP + r;
            cst1[i] = (b < p.B) ? p.c1[(size_t)b * DEC + nt * U + ul] : 0.f;
            cst2[i] = (b < p.B) ? p.c2[(size_t)b * DEC + nt * U + ul] : 0.f;
        }
    }
    const float* b1s = bias;
    const float* b2s = bias + NCL;
    const float* bp1s = bias + 2 * NCL;
    const float* bp2s = bp1s + 4;
    const float* bos = bp2s + 4;
    __syncthreads();

    GridBarrier gb{p.barrier, 0u, (unsigned)nCta};
    const size_t xt80 = (size_t)MEL * RBP, xt256 = (size_t)256 * RBP, xt768 = (size_t)MEM * RBP;
    // this batch tile's XT blocks
    float* xT = p.xT + bt * xt80;
    float* p1T = p.p1T + bt * xt256;
    float* p2T = p.p2T + bt * xt256;
    float* ctxT = p.ctxT + bt * xt768;
    float* h1rawT = p.h1rawT + bt * xt256;
    float* h2rawT = p.h2rawT + bt * xt256;
    const size_t hz_stride = (size_t)p.GB * xt256;  // ping-pong copies

    const float zo = p.zoneout;
    const float keep_scale = p.keep_scale;
    const int NS = p.NS;
    const int DS = MEM / NS, D4 = DS / 4;
    int TG = TACO_THREADS / D4;
    if (TG > 32) TG = 32;
    int steps_done = 0;

    for (int s = 0; s < p.S_max; ++s) {
        const int step = p.t0 + s;
        const int par = step & 1;
        float* h1z_rd = p.h1zT + par * hz_stride + bt * xt256;
        float* h1z_wr = p.h1zT + (par ^ 1) * hz_stride + bt * xt256;
        float* h2z_rd = p.h2zT + par * hz_stride + bt * xt256;
        float* h2z_wr = p.h2zT + (par ^ 1) * hz_stride + bt * xt256;

        // ================= A: prenet dense_1 ===================================================
        {
            XSeg seg[1] = {{xT, MEL}};
            dense_xt(seg, 1, Wp1, 4, RBP, part, res);
            if (tid < RBP * PCN) {
                int r = tid % RBP, j = tid / RBP;
                int b = bt * RBP + r, col = nt * PCN + j;
                float v = fmaxf(res[r * 4 + j] + bp1s[j], 0.f);
                bool kp = (b < p.B) ? keep_bit(p, step, b, 0, col) : false;
                p1T[(size_t)col * RBP + r] = kp ? v * keep_scale : 0.f;
            }
        }
        grid_sync(gb);
        // ================= B: prenet dense_2 ===================================================
        {
            XSeg seg[1] = {{p1T, PRE}};
            dense_xt(seg, 1, Wp2, 4, RBP, part, res);
            if (tid < RBP * PCN) {
                int r = tid % RBP, j = tid / RBP;
                int b = bt * RBP + r, col = nt * PCN + j;
                float v = fmaxf(res[r * 4 + j] + bp2s[j], 0.f);
                bool kp = (b < p.B) ? keep_bit(p, step, b, 1, col) : false;
                p2T[(size_t)col * RBP + r] = kp ? v * keep_scale : 0.f;
            }
        }
        grid_sync(gb);
        // ================= C: LSTM layer 1 =====================================================
        {
            XSeg seg[3] = {{p2T, PRE}, {ctxT, MEM}, {h1z_rd, DEC}};
            dense_xt(seg, 3, W1, NCL, RBP, part, res);
            if (tid < RBP * U) {
                int r = tid % RBP, ul = tid / RBP;
                int u = nt * U + ul;
                const float* z = res + r * NCL;
                float zi = z[0 * U + ul] + b1s[0 * U + ul];
                float zj = z[1 * U + ul] + b1s[1 * U + ul];
                float zf = z[2 * U + ul] + b1s[2 * U + ul];
                float zo_ = z[3 * U + ul] + b1s[3 * U + ul];
                float c_prev = cst1[r * U + ul];
                float h_prev = __ldcg(h1z_rd + (size_t)u * RBP + r);
                float c_new = sigmoidf_(zf + 1.0f) * c_prev + sigmoidf_(zi) * tanhf(zj);
                float h_new = sigmoidf_(zo_) * tanhf(c_new);
                cst1[r * U + ul] = (1.f - zo) * c_new + zo * c_prev;
                h1rawT[(size_t)u * RBP + r] = h_new;
                h1z_wr[(size_t)u * RBP + r] = (1.f - zo) * h_new + zo * h_prev;
            }
        }
        grid_sync(gb);
        // ================= D: LSTM layer 2 =====================================================
        {
            XSeg seg[2] = {{h1rawT, DEC}, {h2z_rd, DEC}};
            dense_xt(seg, 2, W2, NCL, RBP, part, res);
            if (tid < RBP * U) {
                int r = tid % RBP, ul = tid / RBP;
                int u = nt * U + ul;
                const float* z = res + r * NCL;
                float zi = z[0 * U + ul] + b2s[0 * U + ul];
                float zj = z[1 * U + ul] + b2s[1 * U + ul];
                float zf = z[2 * U + ul] + b2s[2 * U + ul];
                float zo_ = z[3 * U + ul] + b2s[3 * U + ul];
                float c_prev = cst2[r * U + ul];
                float h_prev = __ldcg(h2z_rd + (size_t)u * RBP + r);
                float c_new = sigmoidf_(zf + 1.0f) * c_prev + sigmoidf_(zi) * tanhf(zj);
                float h_new = sigmoidf_(zo_) * tanhf(c_new);
                cst2[r * U + ul] = (1.f - zo) * c_new + zo * c_prev;
                h2rawT[(size_t)u * RBP + r] = h_new;
                h2z_wr[(size_t)u * RBP + r] = (1.f - zo) * h_new + zo * h_prev;
            }
        }
        grid_sync(gb);
        // ================= E: query projection  ||  location features ==========================
        {
            XSeg seg[1] = {{h2rawT, DEC}};
            dense_xt(seg, 1, Wq, 4, RBP, part, res);
            if (tid < RBP * QCN) {
                int r = tid % RBP, j = tid / RBP;
                int b = bt * RBP + r;
                if (b < p.B) p.q[(size_t)b * ATT + nt * QCN + j] = res[r * 4 + j];
            }
            // location features: pre[b][t][:] = keys[b][t][:] + b_a + (conv31(cum[b]) + bc) . Wl
            const int nChunks = (p.T + LOC_CHUNK - 1) / LOC_CHUNK;
            for (int item = cta; item < p.B * nChunks; item += nCta) {
                const int b = item / nChunks, ch = item % nChunks;
                const int t0 = ch * LOC_CHUNK;
                const int len = p.lens[b];
                if (t0 >= len) continue;
                __syncthreads();
                if (tid < LOC_CHUNK + LOCK - 1) {
                    int t = t0 - (LOCK - 1) / 2 + tid;
                    cw_s[tid] = (t >= 0 && t < p.T) ? __ldcg(p.cum + (size_t)b * p.T + t) : 0.f;
                }
                __syncthreads();
                // f[c][tl], tl varies fastest over lanes -> window reads / f_s stores conflict-free, conv_s broadcast
                for (int i = tid; i < LOCF * LOC_CHUNK; i += TACO_THREADS) {
                    int tl = i & (LOC_CHUNK - 1), c = i / LOC_CHUNK;
                    float a = convb_s[c];
#pragma unroll
                    for (int k = 0; k < LOCK; ++k) a = fmaf(cw_s[tl + k], conv_s[k * LOCF + c], a);
                    f_s[c * LOC_CHUNK + tl] = a;
                }
                __syncthreads();
                // thread -> 4 t x 4 d tile: d4 = lane (32 x float4 = 128 d), tq = warp (16 x 4 = 64 t)
                {
                    const int d4 = lane, tq = warp;
                    float acc[4][4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
                    for (int c = 0; c < LOCF; ++c) {
                        float4 w = reinterpret_cast<const float4*>(Wl_s + c * ATT)[d4];
                        float4 x = reinterpret_cast<const float4*>(f_s + c * LOC_CHUNK)[tq];
                        TACO_FMA16(x, w)
                    }
                    float4 ba = reinterpret_cast<const float4*>(p.b_a)[d4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        int t = t0 + tq * 4 + i;
                        if (t < len) {
                            float4 k4 = reinterpret_cast<const float4*>(p.keys + ((size_t)b * p.T + t) * ATT)[d4];
                            float4 o = make_float4(k4.x + ba.x + acc[i][0], k4.y + ba.y + acc[i][1],
                                                   k4.z + ba.z + acc[i][2], k4.w + ba.w + acc[i][3]);
                            reinterpret_cast<float4*>(p.pre + ((size_t)b * p.T + t) * ATT)[d4] = o;
                        }
                    }
                }
            }
        }
        grid_sync(gb);
        // ================= F: energies, softmax, cumulate, context =============================
        for (int item = cta; item < p.B * NS; item += nCta) {
            const int b = item / NS, sl = item % NS;
            const int len = p.lens[b];
            const int ibt = b / RBP, ir = b % RBP;
            __syncthreads();
            if (tid < ATT) qs[tid] = __ldcg(p.q + (size_t)b * ATT + tid);
            __syncthreads();
            {
                const float4 q4 = reinterpret_cast<const float4*>(qs)[lane];
                const float4 v4 = reinterpret_cast<const float4*>(va_s)[lane];
                for (int t = warp; t < len; t += TACO_THREADS / 32) {
                    float4 x = __ldcg(reinterpret_cast<const float4*>(p.pre + ((size_t)b * p.T + t) * ATT) + lane);
                    float e = v4.x * tanhf(x.x + q4.x) + v4.y * tanhf(x.y + q4.y) + v4.z * tanhf(x.z + q4.z) +
                              v4.w * tanhf(x.w + q4.w);
                    e = warp_sum(e);
                    if (lane == 0) es[t] = e;
                }
            }
            __syncthreads();
            // normalisation over t < len (exact zeros beyond: -inf mask, Attention.py / App. A.10)
            float m = -INFINITY;
            if (!p.smoothing) {
                for (int t = tid; t < len; t += TACO_THREADS) m = fmaxf(m, es[t]);
                m = warp_max(m);
                if (lane == 0) red[warp] = m;
                __syncthreads();
                m = red[lane & 15];
                m = warp_max(m);
                __syncthreads();
            }
            float sum = 0.f;
            for (int t = tid; t < len; t += TACO_THREADS) {
                float v = p.smoothing ? sigmoidf_(es[t]) : expf(es[t] - m);
                es[t] = v;
                sum += v;
            }
            sum = warp_sum(sum);
            if (lane == 0) red[warp] = sum;
            __syncthreads();
            sum = warp_sum(lane < 16 ? red[lane] : 0.f);
            const float inv = 1.0f / sum;
            for (int t = tid; t < len; t += TACO_THREADS) es[t] *= inv;
            __syncthreads();
            if (sl == 0) {
                float* hist = p.hist + ((size_t)s * p.B + b) * p.T;
                float* cum = p.cum + (size_t)b * p.T;
                for (int t = tid; t < p.T; t += TACO_THREADS) {
                    float a = t < len ? es[t] : 0.f;
                    hist[t] = a;
                    cum[t] = p.cumulative ? (__ldcg(cum + t) + a) : a;
                }
            }
            // context slice [sl*DS, (sl+1)*DS)
            {
                const int d4 = tid % D4, tg = tid / D4;
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                if (tg < TG) {
                    const float4* vp = reinterpret_cast<const float4*>(p.values + (size_t)b * p.T * MEM) + sl * D4 + d4;
                    for (int t = tg; t < len; t += TG) {
                        float a = es[t];
                        float4 v = __ldg(vp + (size_t)t * (MEM / 4));
                        acc.x = fmaf(a, v.x, acc.x);
                        acc.y = fmaf(a, v.y, acc.y);
                        acc.z = fmaf(a, v.z, acc.z);
                        acc.w = fmaf(a, v.w, acc.w);
                    }
                    reinterpret_cast<float4*>(part)[tg * D4 + d4] = acc;
                }
                __syncthreads();
                if (tid < D4) {
                    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int g = 0; g < TG; ++g) {
                        float4 v = reinterpret_cast<const float4*>(part)[g * D4 + tid];
                        c.x += v.x; c.y += v.y; c.z += v.z; c.w += v.w;
                    }
                    float* dst = p.ctxT + ibt * xt768 + (size_t)(sl * DS + tid * 4) * RBP + ir;
                    dst[0] = c.x;
                    dst[RBP] = c.y;
                    dst[2 * RBP] = c.z;
                    dst[3 * RBP] = c.w;
                }
            }
        }
        grid_sync(gb);
        // ================= G: frame + stop projection, outputs, feedback ========================
        if (has_proj) {
            XSeg seg[2] = {{h2rawT, DEC}, {ctxT, MEM}};
            dense_xt(seg, 2, Wo, 4, RBP, part, res);
            if (tid < RBP * 4) {
                int r = tid % RBP, j = tid / RBP;
                int b = bt * RBP + r, col = nt * 4 + j;
                if (b < p.B && col < NOUT) {
                    float y = res[r * 4 + j] + bos[j];
                    if (col < MEL * R_OUT) {
                        p.frames[((size_t)b * p.S_cap * R_OUT + (size_t)s * R_OUT) * MEL + col] = y;
                        if (p.mode == TACO_MODE_FREE && col >= MEL * (R_OUT - 1))
                            xT[(size_t)(col - MEL * (R_OUT - 1)) * RBP + r] = y;
                    } else {
                        int rr = col - MEL * R_OUT;
                        float sv = p.stop_sigmoid ? sigmoidf_(y) : y;
                        p.stop[(size_t)b * p.S_cap * R_OUT + (size_t)s * R_OUT + rr] = sv;
                        // tf.round(sigmoid) == 1  <=>  sigmoid > 0.5 (half-to-even; Helpers.py:40)
                        if (sigmoidf_(y) > 0.5f) atomicAdd(p.stop_count + s, 1);
                    }
                }
            }
        }
        if (p.mode == TACO_MODE_TEACHER && nt == 0) {
            // TacoTrainingHelper.next_inputs (Helpers.py:121-124, ratio 1.0): frame r*(t+1)-1 of the targets
            const int fidx = step * R_OUT + (R_OUT - 1);
            for (int i = tid; i < RBP * MEL; i += TACO_THREADS) {
                int r = i % RBP, m_ = i / RBP;
                int b = bt * RBP + r;
                float v = 0.f;
                if (b < p.B && fidx < p.F_tgt) v = p.targets[((size_t)b * p.F_tgt + fidx) * MEL + m_];
                xT[(size_t)m_ * RBP + r] = v;
            }
        }
        grid_sync(gb);
        steps_done = s + 1;
        if (p.mode == TACO_MODE_FREE) {
            int cnt = __ldcg(p.stop_count + s);
            bool fin = p.stop_at_any ? (cnt > 0) : (cnt == p.B * R_OUT);
            if (fin) break;
        }
    }

    // ---- write back private cell state + step count -------------------------------------------
    for (int i = tid; i < RBP * U; i += TACO_THREADS) {
        int r = i / U, ul = i % U;
        int b = bt * RBP + r;
        if (b < p.B) {
            p.c1[(size_t)b * DEC + nt * U + ul] = cst1[i];
            p.c2[(size_t)b * DEC + nt * U + ul] = cst2[i];
        }
    }
    if (cta == 0 && tid == 0) *p.steps_out = steps_done;
}

// [S][B][T] -> [B][T][S] (Tacotron.py:312) via 32x32 shared tiles
__global__ void transpose_alignments_kernel(const float* __restrict__ hist, float* __restrict__ out, int S, int BT) {
    __shared__ float tile[32][33];
    int x0 = blockIdx.x * 32, y0 = blockIdx.y * 32;  // x over BT, y over S
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int s = y0 + j, x = x0 + threadIdx.x;
        tile[j][threadIdx.x] = (s < S && x < BT) ? hist[(size_t)s * BT + x] : 0.f;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        int x = x0 + j, s = y0 + threadIdx.x;
        if (x < BT && s < S) out[(size_t)x * S + s] = tile[threadIdx.x][j];
    }
}

}  // namespace

size_t taco_decoder_smem_bytes(int U) {
    size_t blob = taco_decoder_blob_floats(U);
    size_t extra = TACO_THREADS * 16 + 32 * 16 + 32 * 4 * 2 + T_MAX + ATT + ATT + 32 + LOCK * LOCF + LOCF + LOCF * ATT +
                   (LOC_CHUNK + 32) + LOCF * LOC_CHUNK;
    return (blob + extra) * sizeof(float);
}

int taco_decoder_blob_floats(int U) {
    const int NCL = 4 * U;
    return K1 * NCL + K2 * NCL + MEL * 4 + PRE * 4 + DEC * 4 + KO * 4 + 2 * NCL + 12;
}

cudaError_t taco_launch_decoder(const DecParams& p, cudaStream_t stream) {
    static bool attr_set[2] = {false, false};
    size_t smem = taco_decoder_smem_bytes(p.U);
    int idx = p.U == 2 ? 0 : 1;
    if (!attr_set[idx]) {
        cudaError_t e = cudaFuncSetAttribute(decoder_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024));
        if (e != cudaSuccess) return e;
        attr_set[idx] = true;
    }
    DecParams pl = p;
    void* args[] = {&pl};
    return cudaLaunchCooperativeKernel((const void*)decoder_persistent_kernel, dim3(p.GB * p.GN), dim3(TACO_THREADS), args, smem, stream);
}

cudaError_t taco_launch_transpose_alignments(const float* hist, float* out, int S, int B, int T, cudaStream_t stream) {
    int BT = B * T;
    dim3 grid((BT + 31) / 32, (S + 31) / 32), block(32, 8);
    transpose_alignments_kernel<<<grid, block, 0, stream>>>(hist, out, S, BT);
    return cudaGetLastError();
}
