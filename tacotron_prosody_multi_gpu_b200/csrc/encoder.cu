// Encoder-side kernels: embedding gather (Tacotron.py:179-181), persistent bidirectional
// zoneout-LSTM (Encoder.py:68-77,118-142; bidirectional_dynamic_rnn with sequence_length,
// SURVEY App. A.2-3), style-embedding tiling + concat (Tacotron.py:215,221-223).
#include "common.cuh"
#include "taco_internal.h"

namespace {

__global__ void embedding_kernel(const int32_t* __restrict__ ids, const float* __restrict__ table, float* __restrict__ out,
                                 int n_tok, int dim4, int n_symbols) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tok * dim4) return;
    int tok = i / dim4, d = i % dim4;
    int id = ids[tok];
    id = id < 0 ? 0 : (id >= n_symbols ? n_symbols - 1 : id);
    reinterpret_cast<float4*>(out)[i] = reinterpret_cast<const float4*>(table)[(size_t)id * dim4 + d];
}

// One CTA = (batch tile bt, unit tile nt), both directions.  U units per CTA: 2 / 4 (lowest latency: 128 CTAs at B = 64) or 16
// (32 CTAs at B = 64: fits beside a running cluster decoder, which leaves 36 SMs free -- engine.EnginePool selects it).  Recurrent kernels [256][4U] of both
// directions resident in shared memory for all T steps; h exchanged through L2 (XT layout,
// ping-pong), c private.  The input half of the gate pre-activation (x.W_x + b) is hoisted out
// of the loop into one GEMM over all tokens (xw).
__global__ void __launch_bounds__(TACO_THREADS, 1) bilstm_persistent_kernel(BiLstmParams p) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x;
    const int U = p.U, NCL = 4 * U, RBP = p.RBP, GN = p.GN;
    const int cta = blockIdx.x;
    const int bt = cta / GN, nt = cta % GN;
    const int H = 256;
    float* Wh[2] = {smem, smem + H * NCL};
    float* part = smem + 2 * H * NCL;
    float* res = part + TACO_THREADS * 16;   // [RBP][NCL]
    float* cst = res + 32 * NCL;             // [2][32][U]
    for (int d = 0; d < 2; ++d) {
        const float4* src = reinterpret_cast<const float4*>(p.blobs + ((size_t)d * GN + nt) * p.blob_floats);
        float4* dst = reinterpret_cast<float4*>(Wh[d]);
        for (int i = tid; i < H * NCL / 4; i += TACO_THREADS) dst[i] = src[i];
    }
    for (int i = tid; i < 2 * 32 * U; i += TACO_THREADS) cst[i] = 0.f;
    __syncthreads();

    GridBarrier gb{p.barrier, 0u, gridDim.x};
    const size_t xt = (size_t)H * RBP;
    const size_t par_stride = (size_t)p.GB * xt;
    const size_t dir_stride = 2 * par_stride;
    const float zo = p.zoneout;

    for (int s = 0; s < p.T; ++s) {
        const int par = s & 1;
        for (int d = 0; d < 2; ++d) {
            const float* h_rd = p.hT + d * dir_stride + par * par_stride + bt * xt;
            float* h_wr = p.hT + d * dir_stride + (par ^ 1) * par_stride + bt * xt;
            XSeg seg[1] = {{h_rd, H}};
            dense_xt(seg, 1, Wh[d], NCL, RBP, part, res);
            if (tid < RBP * U) {
                int r = tid % RBP, ul = tid / RBP;
                int b = bt * RBP + r;
                int u = nt * U + ul;
                float h_prev = __ldcg(h_rd + (size_t)u * RBP + r);
                float h_out = h_prev;
                if (b < p.B) {
                    int len = p.lens[b];
                    if (s < len) {
                        int pos = d == 0 ? s : (len - 1 - s);
                        const float* xw = p.xw + ((size_t)d * p.B * p.T + (size_t)b * p.T + pos) * (4 * H);
                        const float* z = res + r * NCL;
                        float zi = z[0 * U + ul] + xw[0 * H + u];
                        float zj = z[1 * U + ul] + xw[1 * H + u];
                        float zf = z[2 * U + ul] + xw[2 * H + u];
                        float zo_ = z[3 * U + ul] + xw[3 * H + u];
                        float* cp = cst + (d * 32 + r) * U + ul;
                        float c_prev = *cp;
                        float c_new = sigmoidf_(zf + 1.0f) * c_prev + sigmoidf_(zi) * tanhf(zj);
                        float h_new = sigmoidf_(zo_) * tanhf(c_new);
                        *cp = (1.f - zo) * c_new + zo * c_prev;
                        h_out = (1.f - zo) * h_new + zo * h_prev;
                        p.out[((size_t)b * p.T + pos) * p.out_stride + d * H + u] = h_new;  // RAW h (App. A.2)
                    }
                }
                h_wr[(size_t)u * RBP + r] = h_out;
            }
        }
        grid_sync(gb);
    }
}

__global__ void style_fill_kernel(float* __restrict__ memory, const float* __restrict__ style, int B, int T, int sd,
                                  int md, int style_bstride) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * T * sd) return;
    int d = i % sd;
    int bt = i / sd;
    int b = bt / T;
    memory[(size_t)bt * md + (md - sd) + d] = style[(size_t)b * style_bstride + d];
}

}  // namespace

int taco_bilstm_blob_floats(int U) { return 256 * 4 * U; }

cudaError_t taco_launch_embedding(const int32_t* ids, const float* table, float* out, int n_tok, int dim, int n_symbols,
                                  cudaStream_t stream) {
    int n = n_tok * (dim / 4);
    embedding_kernel<<<(n + 255) / 256, 256, 0, stream>>>(ids, table, out, n_tok, dim / 4, n_symbols);
    return cudaGetLastError();
}

cudaError_t taco_launch_bilstm(const BiLstmParams& p, cudaStream_t stream) {
    size_t smem = (size_t)(2 * 256 * 4 * p.U + TACO_THREADS * 16 + 32 * 4 * p.U + 2 * 32 * p.U) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(bilstm_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    BiLstmParams pl = p;
    void* args[] = {&pl};
    return cudaLaunchCooperativeKernel((const void*)bilstm_persistent_kernel, dim3(p.GB * p.GN), dim3(TACO_THREADS), args, smem, stream);
}

cudaError_t taco_launch_style_fill(float* memory, const float* style, int B, int T, int style_dim, int mem_dim,
                                   int style_bstride, cudaStream_t stream) {
    int n = B * T * style_dim;
    style_fill_kernel<<<(n + 255) / 256, 256, 0, stream>>>(memory, style, B, T, style_dim, mem_dim, style_bstride);
    return cudaGetLastError();
}
