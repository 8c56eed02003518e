// fp32 SIMT implicit-GEMM conv1d(padding='same') / dense with fused epilogue.
// Reference ops: tf.layers.conv1d + batch_normalization + activation (Encoder.py:145-168),
// tf.layers.dense (Decoder.py:371; Attention.py memory_layer), residual add + clip
// (Tacotron.py:303-309).  BN (inference) and the conv bias are pre-folded into scale/shift.
// This is the full-precision path (TACO_PRECISION_FP32) and the numerical reference for the
// tcgen05 bf16 kernel in conv_tc.cu.
#include "common.cuh"
#include "taco_internal.h"

namespace {

constexpr int BM = 128, BN = 64, BK = 16;
constexpr int BMP = BM + 4;  // padded k-major A tile row

template <int ACT>
__global__ void __launch_bounds__(256) conv1d_f32_kernel(ConvParams p) {
    __shared__ __align__(16) float As[BK][BMP];
    __shared__ __align__(16) float Bs[BK][BN];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int M = p.B * p.T;
    const int pad = (p.k - 1) / 2;

    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    // A-load mapping: 4 lanes cover 16 consecutive channels of one row; 2 row groups per thread
    const int a_kq = tid & 3;
    const int a_row = tid >> 2;  // 0..63, second at +64
    int a_b[2], a_t[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        int m = m0 + a_row + i * 64;
        a_b[i] = m < M ? m / p.T : -1;
        a_t[i] = m < M ? m % p.T : 0;
    }
    // B-load mapping: 16 rows(k) x 64 cols = 256 float4 -> one per thread
    const int b_k = tid >> 4, b_n4 = tid & 15;

    const int cchunks = p.Cin / BK;
    for (int tap = 0; tap < p.k; ++tap) {
        for (int cc = 0; cc < cchunks; ++cc) {
            const int ci0 = cc * BK;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                int ts = a_t[i] + tap - pad;
                if (a_b[i] >= 0 && ts >= 0 && ts < p.T)
                    v = *reinterpret_cast<const float4*>(p.x + (size_t)a_b[i] * p.x_bstride + (size_t)ts * p.x_stride + ci0 + a_kq * 4);
                int r = a_row + i * 64;
                As[a_kq * 4 + 0][r] = v.x;
                As[a_kq * 4 + 1][r] = v.y;
                As[a_kq * 4 + 2][r] = v.z;
                As[a_kq * 4 + 3][r] = v.w;
            }
            {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                int n = n0 + b_n4 * 4;
                if (n < p.Cout) v = *reinterpret_cast<const float4*>(p.w + ((size_t)tap * p.Cin + ci0 + b_k) * p.Cout + n);
                *reinterpret_cast<float4*>(&Bs[b_k][b_n4 * 4]) = v;
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < BK; ++k) {
                float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 8]);
                float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * 8 + 4]);
                float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
                float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    acc[i][0] = fmaf(a[i], b.x, acc[i][0]);
                    acc[i][1] = fmaf(a[i], b.y, acc[i][1]);
                    acc[i][2] = fmaf(a[i], b.z, acc[i][2]);
                    acc[i][3] = fmaf(a[i], b.w, acc[i][3]);
                }
            }
            __syncthreads();
        }
    }
    const int n = n0 + tx * 4;
    if (n >= p.Cout) return;
    float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.scale) sc = *reinterpret_cast<const float4*>(p.scale + n);
    if (p.shift) sh = *reinterpret_cast<const float4*>(p.shift + n);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int m = m0 + ty * 8 + i;
        if (m >= M) continue;
        int b = m / p.T, t = m % p.T;
        float v[4] = {acc[i][0] * sc.x + sh.x, acc[i][1] * sc.y + sh.y, acc[i][2] * sc.z + sh.z, acc[i][3] * sc.w + sh.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (ACT == ACT_RELU) v[j] = fmaxf(v[j], 0.f);
            if (ACT == ACT_TANH) v[j] = tanhf(v[j]);
        }
        if (p.residual) {
            float4 r = *reinterpret_cast<const float4*>(p.residual + (size_t)m * p.Cout + n);
            v[0] += r.x; v[1] += r.y; v[2] += r.z; v[3] += r.w;
        }
        if (p.clip) {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = fminf(fmaxf(v[j], p.clip_lo), p.clip_hi);
        }
        *reinterpret_cast<float4*>(p.y + (size_t)b * p.y_bstride + (size_t)t * p.y_stride + n) = make_float4(v[0], v[1], v[2], v[3]);
    }
}

}  // namespace

cudaError_t taco_launch_conv1d_f32(const ConvParams& p, cudaStream_t stream) {
    if (p.Cin % BK != 0 || p.Cout % 4 != 0) return cudaErrorInvalidValue;
    dim3 grid((p.B * p.T + BM - 1) / BM, (p.Cout + BN - 1) / BN);
    switch (p.act) {
        case ACT_RELU: conv1d_f32_kernel<ACT_RELU><<<grid, 256, 0, stream>>>(p); break;
        case ACT_TANH: conv1d_f32_kernel<ACT_TANH><<<grid, 256, 0, stream>>>(p); break;
        default: conv1d_f32_kernel<ACT_NONE><<<grid, 256, 0, stream>>>(p); break;
    }
    return cudaGetLastError();
}
