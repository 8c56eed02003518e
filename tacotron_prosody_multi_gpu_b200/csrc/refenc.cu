// Prosody reference encoder + global-style-token attention.
//   reference_encoder (Tacotron.py:12-29): 6 x [conv2d 3x3 stride 2 'same' + BN + ReLU] (Encoder.py:171-182)
//   -> GRUCell(128) over ceil(Tm/64) steps, no length mask -> last output -> dense(128, tanh)
//   MultiheadAttention.multi_head_attention, mlp_attention, normalize=True (Attention.py:243-341)
//   random-weight GST branch (Tacotron.py:216-222)
// These run once per utterance batch and are tiny next to the decoder loop; plain fp32 SIMT.
#include "common.cuh"
#include "taco_internal.h"

namespace {

// TF 'same' stride-2 padding: pad_before = max((out-1)*2 + 3 - in, 0) / 2
__device__ __host__ inline int same_pad_before(int in) {
    int out = (in + 1) / 2;
    int total = (out - 1) * 2 + 3 - in;
    if (total < 0) total = 0;
    return total / 2;
}

__global__ void conv2d_s2_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ scale,
                                 const float* __restrict__ shift, float* __restrict__ y, int B, int H, int W, int Cin,
                                 int Cout, int Ho, int Wo, int pt, int pl) {
    size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t total = (size_t)B * Ho * Wo * Cout;
    if (idx >= total) return;
    int co = idx % Cout;
    size_t rest = idx / Cout;
    int j = rest % Wo;
    rest /= Wo;
    int i = rest % Ho;
    int b = rest / Ho;
    float acc = 0.f;
    for (int di = 0; di < 3; ++di) {
        int yy = i * 2 + di - pt;
        if (yy < 0 || yy >= H) continue;
        for (int dj = 0; dj < 3; ++dj) {
            int xx = j * 2 + dj - pl;
            if (xx < 0 || xx >= W) continue;
            const float* xp = x + (((size_t)b * H + yy) * W + xx) * Cin;
            const float* wp = w + ((size_t)(di * 3 + dj) * Cin) * Cout + co;
            for (int ci = 0; ci < Cin; ++ci) acc = fmaf(xp[ci], wp[(size_t)ci * Cout], acc);
        }
    }
    y[idx] = fmaxf(acc * scale[co] + shift[co], 0.f);
}

// One CTA per batch row.  GRUCell (App. A.9): [r,u] = sigmoid([x,h].Wg + bg); c = tanh([x, r*h].Wc + bc);
// h' = u*h + (1-u)*c; then reference_state = tanh(h_last . Wd + bd).
__global__ void __launch_bounds__(256) ref_gru_kernel(const float* __restrict__ x, int Tp, int In, const float* __restrict__ Wg,
                                                     const float* __restrict__ bg, const float* __restrict__ Wc,
                                                     const float* __restrict__ bc, const float* __restrict__ Wd,
                                                     const float* __restrict__ bd, float* __restrict__ ref_out) {
    constexpr int R = 128;
    extern __shared__ float sm[];
    float* xh = sm;             // [In + R]
    float* xr = xh + In + R;    // [In + R]
    float* gate = xr + In + R;  // [2R]
    const int b = blockIdx.x, tid = threadIdx.x;
    const int K = In + R;
    if (tid < R) xh[In + tid] = 0.f;
    for (int t = 0; t < Tp; ++t) {
        for (int k = tid; k < In; k += blockDim.x) {
            float v = x[((size_t)b * Tp + t) * In + k];
            xh[k] = v;
            xr[k] = v;
        }
        __syncthreads();
        {
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
            int k = 0;
            for (; k + 3 < K; k += 4) {
                a0 = fmaf(xh[k], Wg[(size_t)k * 2 * R + tid], a0);
                a1 = fmaf(xh[k + 1], Wg[(size_t)(k + 1) * 2 * R + tid], a1);
                a2 = fmaf(xh[k + 2], Wg[(size_t)(k + 2) * 2 * R + tid], a2);
                a3 = fmaf(xh[k + 3], Wg[(size_t)(k + 3) * 2 * R + tid], a3);
            }
            for (; k < K; ++k) a0 = fmaf(xh[k], Wg[(size_t)k * 2 * R + tid], a0);
            gate[tid] = sigmoidf_((a0 + a1) + (a2 + a3) + bg[tid]);
        }
        __syncthreads();
        if (tid < R) xr[In + tid] = gate[tid] * xh[In + tid];
        __syncthreads();
        float hn = 0.f;
        if (tid < R) {
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
            int k = 0;
            for (; k + 3 < K; k += 4) {
                a0 = fmaf(xr[k], Wc[(size_t)k * R + tid], a0);
                a1 = fmaf(xr[k + 1], Wc[(size_t)(k + 1) * R + tid], a1);
                a2 = fmaf(xr[k + 2], Wc[(size_t)(k + 2) * R + tid], a2);
                a3 = fmaf(xr[k + 3], Wc[(size_t)(k + 3) * R + tid], a3);
            }
            for (; k < K; ++k) a0 = fmaf(xr[k], Wc[(size_t)k * R + tid], a0);
            float c = tanhf((a0 + a1) + (a2 + a3) + bc[tid]);
            float u = gate[R + tid];
            hn = u * xh[In + tid] + (1.f - u) * c;
        }
        __syncthreads();
        if (tid < R) xh[In + tid] = hn;
        __syncthreads();
    }
    if (tid < R) {
        float a = bd[tid];
        for (int k = 0; k < R; ++k) a = fmaf(xh[In + k], Wd[(size_t)k * R + tid], a);
        ref_out[(size_t)b * R + tid] = tanhf(a);
    }
}

// One CTA per batch row; dims: ref 128, units U (128), heads Hh (4), tokens N (10), token dim D (64).
__global__ void __launch_bounds__(128) gst_attention_kernel(const float* __restrict__ ref, const float* __restrict__ tokens,
                                                           const float* __restrict__ Wq, const float* __restrict__ bq,
                                                           const float* __restrict__ Wk, const float* __restrict__ bk,
                                                           const float* __restrict__ v, const float* __restrict__ g,
                                                           const float* __restrict__ bb, float* __restrict__ style, int N,
                                                           int D, int Hh, int U) {
    extern __shared__ float sm[];
    float* tok = sm;             // [N][D] tanh(tokens)
    float* q = tok + N * D;      // [U]
    float* k = q + U;            // [N][U]
    float* sc = k + N * U;       // [Hh][N]
    float* refs = sc + Hh * N;   // [128]
    __shared__ float vnorm;
    const int b = blockIdx.x, tid = threadIdx.x;
    const int hu = U / Hh;
    for (int i = tid; i < N * D; i += blockDim.x) tok[i] = tanhf(tokens[i]);
    for (int i = tid; i < 128; i += blockDim.x) refs[i] = ref[(size_t)b * 128 + i];
    if (tid == 0) {
        float s = 0.f;
        for (int i = 0; i < hu; ++i) s += v[i] * v[i];
        vnorm = g[0] * rsqrtf(s);
    }
    __syncthreads();
    for (int u = tid; u < U; u += blockDim.x) {
        float a = bq[u];
        for (int i = 0; i < 128; ++i) a = fmaf(refs[i], Wq[(size_t)i * U + u], a);
        q[u] = a;
        for (int n = 0; n < N; ++n) {
            float c = bk[u];
            for (int j = 0; j < D; ++j) c = fmaf(tok[n * D + j], Wk[(size_t)j * U + u], c);
            k[n * U + u] = c;
        }
    }
    __syncthreads();
    for (int i = tid; i < Hh * N; i += blockDim.x) {
        int h = i / N, n = i % N;
        float a = 0.f;
        for (int e = 0; e < hu; ++e) a += vnorm * v[e] * tanhf(k[n * U + h * hu + e] + q[h * hu + e] + bb[e]);
        sc[i] = a;
    }
    __syncthreads();
    if (tid < Hh) {
        float m = -INFINITY;
        for (int n = 0; n < N; ++n) m = fmaxf(m, sc[tid * N + n]);
        float s = 0.f;
        for (int n = 0; n < N; ++n) {
            float e = expf(sc[tid * N + n] - m);
            sc[tid * N + n] = e;
            s += e;
        }
        for (int n = 0; n < N; ++n) sc[tid * N + n] /= s;
    }
    __syncthreads();
    for (int i = tid; i < Hh * D; i += blockDim.x) {
        int h = i / D, j = i % D;
        float a = 0.f;
        for (int n = 0; n < N; ++n) a = fmaf(sc[h * N + n], tok[n * D + j], a);
        style[(size_t)b * Hh * D + i] = a;
    }
}

__global__ void style_random_kernel(const float* __restrict__ gw, const float* __restrict__ tokens, float* __restrict__ style,
                                    int Hh, int N, int D) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= Hh * D) return;
    int h = i / D, j = i % D;
    float a = 0.f;
    for (int n = 0; n < N; ++n) a = fmaf(gw[h * N + n], tanhf(tokens[n * D + j]), a);
    style[i] = a;
}

}  // namespace

cudaError_t taco_launch_conv2d_s2(const float* x, const float* w, const float* scale, const float* shift, float* y, int B,
                                  int H, int W, int Cin, int Cout, cudaStream_t stream) {
    int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
    size_t total = (size_t)B * Ho * Wo * Cout;
    conv2d_s2_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(x, w, scale, shift, y, B, H, W, Cin, Cout, Ho, Wo,
                                                                          same_pad_before(H), same_pad_before(W));
    return cudaGetLastError();
}

cudaError_t taco_launch_ref_gru(const float* x, int B, int Tp, int In, const float* Wg, const float* bg, const float* Wc,
                                const float* bc, const float* Wd, const float* bd, float* ref_out, cudaStream_t stream) {
    size_t smem = (size_t)(2 * (In + 128) + 256) * sizeof(float);
    ref_gru_kernel<<<B, 256, smem, stream>>>(x, Tp, In, Wg, bg, Wc, bc, Wd, bd, ref_out);
    return cudaGetLastError();
}

cudaError_t taco_launch_gst_attention(const float* ref, const float* tokens, const float* Wq, const float* bq,
                                      const float* Wk, const float* bk, const float* v, const float* g, const float* b,
                                      float* style, int B, int n_tok, int tok_dim, int heads, int units,
                                      cudaStream_t stream) {
    size_t smem = (size_t)(n_tok * tok_dim + units + n_tok * units + heads * n_tok + 128) * sizeof(float);
    gst_attention_kernel<<<B, 128, smem, stream>>>(ref, tokens, Wq, bq, Wk, bk, v, g, b, style, n_tok, tok_dim, heads, units);
    return cudaGetLastError();
}

cudaError_t taco_launch_style_random(const float* gst_weights, const float* tokens, float* style, int heads, int n_tok,
                                     int tok_dim, cudaStream_t stream) {
    int n = heads * tok_dim;
    style_random_kernel<<<(n + 127) / 128, 128, 0, stream>>>(gst_weights, tokens, style, heads, n_tok, tok_dim);
    return cudaGetLastError();
}
