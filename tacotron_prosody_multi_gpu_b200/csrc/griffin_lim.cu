// Mel -> waveform (Griffin-Lim) for sm_100a: the C ABI of include/taco_b200_audio.h.
//
// Reference: Utils/AudioProcessing/AudioPreprocess.py:152-186 (mel_to_audio_serie) and the librosa stft / istft it calls
// (center=True reflect padding, periodic Hann of win_size zero-padded to n_fft, window-sum-of-squares normalisation).
//
// One Griffin-Lim iteration is ONE kernel.  A CTA owns two adjacent frames of one utterance:
//   1. gathers the overlap-added, window-sum-normalised signal under its two windows from the previous iteration's
//      windowed frames (each sample: <= ceil(win/hop) frames, fixed order => deterministic, no atomics),
//   2. packs frame 0 into the real and frame 1 into the imaginary part of one 2048-point complex FFT held in shared
//      memory (Stockham autosort, radix 8,8,8,4 in registers, padded against bank conflicts, twiddles from an L1-resident
//      table computed in double precision),
//   3. splits the two Hermitian spectra, replaces the moduli by the target magnitudes (phase kept),
//   4. packs the two projected spectra into one inverse FFT, windows the two real outputs and writes them as the next
//      iteration's frames.
// HBM/L2 traffic per frame and iteration: n_bins magnitudes + ~(1 + hop/win) * 4 * win gathered samples read, win
// samples written -- the spectra never leave shared memory.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/taco_b200_audio.h"
#include "common.cuh"

namespace {

constexpr int N = 2048;          // n_fft
constexpr int NB = N / 2 + 1;    // bins
constexpr int GL_THREADS = 256;
constexpr uint32_t GLIM_TAG = 0x474C494D;

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// Shared-memory index of FFT point i: one pad slot per 16 points keeps the radix-8 scatter of the first pass conflict-free.
__device__ __forceinline__ int pad(int i) { return i + (i >> 4); }
constexpr int NP = N + N / 16;
// Twiddle table (global, L1-resident): pass p with radix R and Ns = product of the earlier radices holds
// exp(-2 pi i r k / (Ns R)) at TW_OFF_p + k (R - 1) + (r - 1); the first pass (Ns = 1) needs none.
constexpr int TW_OFF1 = 0, TW_OFF2 = TW_OFF1 + 8 * 7, TW_OFF3 = TW_OFF2 + 64 * 7, TW_TOTAL = TW_OFF3 + 512 * 3;

template <bool INV>
__device__ __forceinline__ void dft4(float2& x0, float2& x1, float2& x2, float2& x3) {
    const float2 a0 = make_float2(x0.x + x2.x, x0.y + x2.y);
    const float2 a1 = make_float2(x0.x - x2.x, x0.y - x2.y);
    const float2 a2 = make_float2(x1.x + x3.x, x1.y + x3.y);
    const float2 d = make_float2(x1.x - x3.x, x1.y - x3.y);
    const float2 a3 = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);   // * (+i) / (-i)
    x0 = make_float2(a0.x + a2.x, a0.y + a2.y);
    x1 = make_float2(a1.x + a3.x, a1.y + a3.y);
    x2 = make_float2(a0.x - a2.x, a0.y - a2.y);
    x3 = make_float2(a1.x - a3.x, a1.y - a3.y);
}

// 8-point DFT in registers, natural order out: X[2m] = DFT4(v[n] + v[n+4]), X[2m+1] = DFT4((v[n] - v[n+4]) W8^n)
template <bool INV>
__device__ __forceinline__ void dft8(float2 (&v)[8]) {
    constexpr float S = 0.70710678118654752f;
    float2 e0 = make_float2(v[0].x + v[4].x, v[0].y + v[4].y), o0 = make_float2(v[0].x - v[4].x, v[0].y - v[4].y);
    float2 e1 = make_float2(v[1].x + v[5].x, v[1].y + v[5].y), d1 = make_float2(v[1].x - v[5].x, v[1].y - v[5].y);
    float2 e2 = make_float2(v[2].x + v[6].x, v[2].y + v[6].y), d2 = make_float2(v[2].x - v[6].x, v[2].y - v[6].y);
    float2 e3 = make_float2(v[3].x + v[7].x, v[3].y + v[7].y), d3 = make_float2(v[3].x - v[7].x, v[3].y - v[7].y);
    float2 o1 = INV ? make_float2(S * (d1.x - d1.y), S * (d1.x + d1.y)) : make_float2(S * (d1.x + d1.y), S * (d1.y - d1.x));
    float2 o2 = INV ? make_float2(-d2.y, d2.x) : make_float2(d2.y, -d2.x);
    float2 o3 = INV ? make_float2(-S * (d3.x + d3.y), S * (d3.x - d3.y)) : make_float2(S * (d3.y - d3.x), -S * (d3.x + d3.y));
    dft4<INV>(e0, e1, e2, e3);
    dft4<INV>(o0, o1, o2, o3);
    v[0] = e0; v[1] = o0; v[2] = e1; v[3] = o1; v[4] = e2; v[5] = o2; v[6] = e3; v[7] = o3;
}

// One Stockham autosort pass of radix R (Ns = product of the earlier radices): in -> out, both padded.
// ZIN: the input is known to be zero outside [N/4, 3N/4) (a window of at most N/2 taps centred in the frame): the first
// radix-8 pass reads only the 4 of its 8 inputs that can be non-zero.  HOUT: only outputs in [N/4, 3N/4) are used (the
// windowed part of the inverse transform): the last radix-4 pass stores 2 of its 4 outputs.
template <int R, int Ns, bool INV, bool ZIN = false, bool HOUT = false>
__device__ __forceinline__ void fft_pass(const float2* __restrict__ in, float2* __restrict__ out, const float2* __restrict__ twp) {
    static_assert(!ZIN || (R == 8 && Ns == 1), "ZIN is a property of the first pass");
    static_assert(!HOUT || (R == 4 && Ns == N / 4), "HOUT is a property of the last pass");
    constexpr int T = N / R;
#pragma unroll
    for (int j = threadIdx.x; j < T; j += GL_THREADS) {
        const int k = j & (Ns - 1), pj = pad(j);
        float2 v[R];
#pragma unroll
        for (int r = 0; r < R; ++r)                                        // pad(j + r T): T is a multiple of 16
            v[r] = (ZIN && (r < R / 4 || r >= 3 * R / 4)) ? make_float2(0.f, 0.f) : in[pj + r * (T + T / 16)];
        if (Ns > 1) {
#pragma unroll
            for (int r = 1; r < R; ++r) {
                float2 w = __ldg(twp + k * (R - 1) + (r - 1));
                if (INV) w.y = -w.y;
                v[r] = cmul(v[r], w);
            }
        }
        const int pj0 = pad((j - k) * R + k);
        if constexpr (R == 8) {
            dft8<INV>(v);
        } else {
            dft4<INV>(v[0], v[1], v[2], v[3]);
        }
#pragma unroll
        for (int r = 0; r < R; ++r)                                        // pad(j0 + r Ns): no carry out of the low 4 bits
            if (!HOUT || r == 1 || r == 2) out[pj0 + r * Ns + ((r * Ns) >> 4)] = v[r];
    }
    __syncthreads();
}

// 2048-point FFT, a -> a (b is scratch): radix 8, 8, 8, 4 -- the data ends in the buffer it started in.
template <bool INV, bool ZIN = false, bool HOUT = false>
__device__ __forceinline__ void fft2048(float2* a, float2* b, const float2* tw) {
    fft_pass<8, 1, INV, ZIN>(a, b, tw);
    fft_pass<8, 8, INV>(b, a, tw + TW_OFF1);
    fft_pass<8, 64, INV>(a, b, tw + TW_OFF2);
    fft_pass<4, 512, INV, false, HOUT>(b, a, tw + TW_OFF3);
}

struct GlGeom {
    int hop, win, lpad;      // lpad = (N - win) / 2: first non-zero tap of the padded window
    int cover;               // ceil(win / hop): frames over one sample
};

// Overlap-added, normalised signal sample n (0 <= n < hop * (F - 1)) from the windowed frames yf [F][win]
// (librosa.istft: y[i*hop : i*hop + n_fft] += w * irfft(S_i); y /= sum w^2 where > tiny; y = y[n_fft/2 : -n_fft/2]).
__device__ __forceinline__ float ola_sample(const float* __restrict__ yf, const float* __restrict__ w, const GlGeom g, int F, int n) {
    const int q = n + N / 2 - g.lpad;                       // position relative to frame 0's first tap
    int i_hi = q / g.hop;
    if (i_hi > F - 1) i_hi = F - 1;
    int i_lo = (q - g.win + g.hop) / g.hop;                 // ceil((q - win + 1) / hop) for q - win + 1 >= 0
    if (q - g.win + 1 <= 0) i_lo = 0;
    float acc = 0.f, ws = 0.f;
    for (int i = i_lo; i <= i_hi; ++i) {
        const int j = q - i * g.hop;
        const float wj = __ldg(w + j);
        acc += yf[(size_t)i * g.win + j];
        ws = fmaf(wj, wj, ws);
    }
    return ws > 1.17549435e-38f ? acc / ws : acc;
}

__device__ __forceinline__ float glim_turns(uint64_t seed, uint32_t utt, uint64_t idx) {
    const uint64_t blk = idx >> 2;
    const uint4 o = philox4x32_10(make_uint4((uint32_t)blk, utt, (uint32_t)(blk >> 32), GLIM_TAG),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const uint32_t lane = (uint32_t)idx & 3u;
    const uint32_t word = lane == 0 ? o.x : lane == 1 ? o.y : lane == 2 ? o.z : o.w;
    return (float)(word >> 8) * (1.0f / 16777216.0f);
}

}  // namespace

namespace {

enum { GL_FIRST = 0, GL_ITER = 1, GL_MEASURE = 2 };

struct GlArgs {
    const float* mag;        // [B,F,NB]
    const float* turns;      // [B,F,NB] or null (GL_FIRST)
    const float* yf_in;      // [B,F,win] windowed frames of the previous iteration (GL_ITER)
    const float* sig;        // [B,sig_stride] signal (GL_MEASURE)
    long long sig_stride;
    float* yf_out;           // [B,F,win]
    float* dist_acc;         // [B,2] (GL_MEASURE)
    const int32_t* n_frames; // [B] or null
    const float2* tw;        // [TW_TOTAL] per-pass twiddles
    const float* w;          // [win]
    const float* wsum;       // [win + hop] window sum of squares under an interior frame pair
    GlGeom g;
    int F;
    uint64_t seed;
    int utt0;
};

template <int MODE, bool CENTRED>
__global__ void __launch_bounds__(GL_THREADS) gl_kernel(const GlArgs a) {
    extern __shared__ float2 gl_smem[];
    float2* A = gl_smem;
    float2* Bf = gl_smem + NP;
    const float2* tw = a.tw;
    const int b = blockIdx.y, t0 = blockIdx.x * 2, tid = threadIdx.x;
    const int Fb = a.n_frames ? a.n_frames[b] : a.F;
    if (t0 >= Fb) return;
    const bool has1 = t0 + 1 < Fb;
    const GlGeom g = a.g;
    const float* mag0 = a.mag + ((size_t)b * a.F + t0) * NB;

    if (MODE != GL_FIRST) {
        float* seg = reinterpret_cast<float*>(Bf);
        const int L = g.hop * (Fb - 1);
        const float* yf = a.yf_in ? a.yf_in + (size_t)b * a.F * g.win : nullptr;
        const int SEG = g.win + g.hop;
        const int n0 = t0 * g.hop + g.lpad - N / 2;              // signal sample under the first tap of frame t0
        const int d_lo = -((g.win - 1) / g.hop), d_hi = (SEG - 1) / g.hop;
        if (MODE == GL_ITER && n0 >= 0 && n0 + SEG <= L && t0 + d_lo >= 0 && t0 + d_hi <= Fb - 1) {
            // interior: no reflection, every covering frame exists => the window sum depends on the offset only (a.wsum)
            const float* base = yf + (size_t)t0 * g.win;
            for (int i = tid; i < SEG; i += GL_THREADS) {
                float acc = 0.f;
                for (int d = d_lo; d <= d_hi; ++d) {
                    const int j = i - d * g.hop;
                    if (j >= 0 && j < g.win) acc += base[d * g.win + j];
                }
                seg[i] = acc / __ldg(a.wsum + i);
            }
        } else {
            for (int i = tid; i < SEG; i += GL_THREADS) {
                int n = n0 + i;
                if (n < 0) n = -n;                       // np.pad(mode='reflect')
                if (n >= L) n = 2 * (L - 1) - n;
                n = min(max(n, 0), L - 1);
                seg[i] = MODE == GL_MEASURE ? a.sig[(size_t)b * a.sig_stride + n] : ola_sample(yf, a.w, g, Fb, n);
            }
        }
        __syncthreads();
        // CENTRED: window inside [N/4, 3N/4): skip the known-zero half of the transforms
        for (int j = CENTRED ? N / 4 + tid : tid; j < (CENTRED ? 3 * N / 4 : N); j += GL_THREADS) {
            const int jj = j - g.lpad;
            float2 z = make_float2(0.f, 0.f);
            if (jj >= 0 && jj < g.win) {
                const float wj = __ldg(a.w + jj);
                z.x = wj * seg[jj];
                z.y = has1 ? wj * seg[jj + g.hop] : 0.f;
            }
            A[pad(j)] = z;
        }
        __syncthreads();
        fft2048<false, CENTRED, false>(A, Bf, tw);
    }

    float err = 0.f, ref = 0.f;
    for (int k = tid; k <= N / 2; k += GL_THREADS) {
        const float m0 = mag0[k];
        const float m1 = has1 ? mag0[NB + k] : 0.f;
        float2 P0, P1;
        if (MODE == GL_FIRST) {
            float u0, u1 = 0.f;
            if (a.turns) {
                const float* tr = a.turns + ((size_t)b * a.F + t0) * NB;
                u0 = tr[k];
                if (has1) u1 = tr[NB + k];
            } else {
                u0 = glim_turns(a.seed, (uint32_t)(b + a.utt0), (uint64_t)k * Fb + t0);
                if (has1) u1 = glim_turns(a.seed, (uint32_t)(b + a.utt0), (uint64_t)k * Fb + t0 + 1);
            }
            float s, c;
            sincospif(2.f * u0, &s, &c);
            P0 = make_float2(m0 * c, m0 * s);
            sincospif(2.f * u1, &s, &c);
            P1 = make_float2(m1 * c, m1 * s);
        } else {
            const float2 Zk = A[pad(k)], Zn = A[pad((N - k) & (N - 1))];
            const float2 X0 = make_float2(0.5f * (Zk.x + Zn.x), 0.5f * (Zk.y - Zn.y));   // (Z[k] + conj Z[N-k]) / 2
            const float2 X1 = make_float2(0.5f * (Zk.y + Zn.y), 0.5f * (Zn.x - Zk.x));   // (Z[k] - conj Z[N-k]) / 2i
            const float q0 = X0.x * X0.x + X0.y * X0.y, q1 = X1.x * X1.x + X1.y * X1.y;
            if (MODE == GL_MEASURE) {
                const float r0 = sqrtf(q0), r1 = sqrtf(q1);
                err += (r0 - m0) * (r0 - m0) + (has1 ? (r1 - m1) * (r1 - m1) : 0.f);
                ref += m0 * m0 + m1 * m1;
                continue;
            }
            const float s0 = m0 * rsqrtf(q0), s1 = m1 * rsqrtf(q1);                     // m exp(i angle(X)); angle(0) = 0
            P0 = q0 > 0.f ? make_float2(X0.x * s0, X0.y * s0) : make_float2(m0, 0.f);
            P1 = q1 > 0.f ? make_float2(X1.x * s1, X1.y * s1) : make_float2(m1, 0.f);
        }
        if (k == 0 || k == N / 2) P0.y = P1.y = 0.f;       // irfft ignores the imaginary part of DC / Nyquist
        Bf[pad(k)] = make_float2(P0.x - P1.y, P0.y + P1.x);      // P0 + i P1
        if (k > 0 && k < N / 2) Bf[pad(N - k)] = make_float2(P0.x + P1.y, P1.x - P0.y);   // conj P0 + i conj P1
    }
    if (MODE == GL_MEASURE) {
        __shared__ float red[2][GL_THREADS / 32];
        for (int o = 16; o; o >>= 1) {
            err += __shfl_xor_sync(0xffffffffu, err, o);
            ref += __shfl_xor_sync(0xffffffffu, ref, o);
        }
        if ((tid & 31) == 0) { red[0][tid >> 5] = err; red[1][tid >> 5] = ref; }
        __syncthreads();
        if (tid == 0) {
            float e = 0.f, r = 0.f;
            for (int i = 0; i < GL_THREADS / 32; ++i) { e += red[0][i]; r += red[1][i]; }
            atomicAdd(a.dist_acc + 2 * b, e);
            atomicAdd(a.dist_acc + 2 * b + 1, r);
        }
        return;
    }
    __syncthreads();
    fft2048<true, false, CENTRED>(Bf, A, tw);
    float* out0 = a.yf_out + ((size_t)b * a.F + t0) * g.win;
    for (int j = tid; j < g.win; j += GL_THREADS) {
        const float wj = __ldg(a.w + j) * (1.0f / N);
        const float2 z = Bf[pad(g.lpad + j)];
        out0[j] = wj * z.x;
        if (has1) out0[g.win + j] = wj * z.y;
    }
}

__global__ void gl_dist_final_kernel(const float* acc, float* out, int B) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) out[b] = acc[2 * b + 1] > 0.f ? sqrtf(acc[2 * b] / acc[2 * b + 1]) : 0.f;
}

// final overlap-add of the last iteration's frames into the waveform
__global__ void gl_ola_kernel(const float* __restrict__ yf, const float* __restrict__ w, GlGeom g, const int32_t* n_frames, int F,
                              float* __restrict__ audio, long long stride) {
    const int b = blockIdx.y;
    const int Fb = n_frames ? n_frames[b] : F;
    const long long L = (long long)g.hop * (Fb - 1);
    for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < stride; n += (long long)gridDim.x * blockDim.x)
        audio[b * stride + n] = n < L ? ola_sample(yf + (size_t)b * F * g.win, w, g, Fb, (int)n) : 0.f;
}

// y[n] = x[n] + k y[n-1] (scipy.signal.lfilter([1], [1, -k])), in place; one CTA per utterance, chunked scan
__global__ void __launch_bounds__(1024) gl_deemphasis_kernel(float* audio, long long stride, GlGeom g, const int32_t* n_frames, int F, float k) {
    __shared__ float tail[1024];
    __shared__ float carry[1024];
    const int b = blockIdx.x, tid = threadIdx.x;
    const int Fb = n_frames ? n_frames[b] : F;
    const int L = g.hop * (Fb - 1);
    const int C = (L + 1023) / 1024;
    float* x = audio + b * stride;
    const int lo = min(tid * C, L), hi = min(lo + C, L);
    float acc = 0.f;
    for (int i = lo; i < hi; ++i) { acc = fmaf(k, acc, x[i]); x[i] = acc; }
    tail[tid] = acc;
    __syncthreads();
    if (tid == 0) {
        float c = 0.f;
        for (int t = 0; t < 1024; ++t) {
            carry[t] = c;                                   // true y just before chunk t
            const int len = min(max(L - t * C, 0), C);
            c = tail[t] + powf(k, (float)len) * c;
        }
    }
    __syncthreads();
    float pw = k * carry[tid];
    for (int i = lo; i < hi; ++i) { x[i] += pw; pw *= k; }
}

struct MelNormDev { int sn, clip, sym; float m, lo, ref, power; };

// mag[b,t,k] = max(1e-10, sum_m inv[k,m] * 10^((denorm(mel[b,t,m]) + ref) / 20)) ^ power; 8 frames per CTA share the basis reads
constexpr int MM_FRAMES = 8;
__global__ void __launch_bounds__(256) gl_mel_to_mag_kernel(const float* __restrict__ mel, const int32_t* n_frames, int F, int n_mels,
                                                             const float* __restrict__ invT /*[n_mels][NB]*/, MelNormDev nm,
                                                             float* __restrict__ mag) {
    extern __shared__ float amp[];   // [MM_FRAMES][n_mels]
    const int b = blockIdx.y, t0 = blockIdx.x * MM_FRAMES;
    const int Fb = n_frames ? n_frames[b] : F;
    for (int i = threadIdx.x; i < MM_FRAMES * n_mels; i += blockDim.x) {
        const int f = i / n_mels, m = i - f * n_mels, t = t0 + f;
        float v = 0.f;
        if (t < Fb) {
            float d = mel[((size_t)b * F + t) * n_mels + m];
            if (nm.sn) {
                if (nm.sym) {
                    if (nm.clip) d = fminf(fmaxf(d, -nm.m), nm.m);
                    d = (d + nm.m) * -nm.lo / (2.f * nm.m) + nm.lo;
                } else {
                    if (nm.clip) d = fminf(fmaxf(d, 0.f), nm.m);
                    d = d * -nm.lo / nm.m + nm.lo;
                }
            }
            v = exp10f((d + nm.ref) * 0.05f);
        }
        amp[i] = v;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < NB; k += blockDim.x) {
        float acc[MM_FRAMES];
#pragma unroll
        for (int f = 0; f < MM_FRAMES; ++f) acc[f] = 0.f;
        for (int m = 0; m < n_mels; ++m) {
            const float bv = __ldg(invT + (size_t)m * NB + k);
#pragma unroll
            for (int f = 0; f < MM_FRAMES; ++f) acc[f] = fmaf(bv, amp[f * n_mels + m], acc[f]);
        }
#pragma unroll
        for (int f = 0; f < MM_FRAMES; ++f) {
            const int t = t0 + f;
            if (t < F) mag[((size_t)b * F + t) * NB + k] = t < Fb ? powf(fmaxf(acc[f], 1e-10f), nm.power) : 0.f;
        }
    }
}

}  // namespace

struct TacoGl {
    int device = 0;
    GlGeom g{};
    int n_mels = 0;
    float2* tw = nullptr;      // [TW_TOTAL]
    float* w = nullptr;        // [win]
    float* wsum = nullptr;     // [win + hop]
    float* invT = nullptr;     // [n_mels][NB]
    float* yf[2] = {nullptr, nullptr};
    size_t yf_floats = 0;
    float* dist_acc = nullptr;
    size_t dist_cap = 0;
    // grow-only staging of the host-buffer entry point: device mel / mag / audio / n_frames, pinned bounce buffer
    void* stage[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t stage_bytes[4] = {0, 0, 0, 0};
    void* pinned = nullptr;
    size_t pinned_bytes = 0;
    std::string err;
};

namespace {

constexpr size_t GL_SMEM = 2 * NP * sizeof(float2);

int gl_fail(TacoGl* g, int code, const char* what, cudaError_t e = cudaSuccess) {
    if (g) {
        g->err = what;
        if (e != cudaSuccess) { g->err += ": "; g->err += cudaGetErrorString(e); }
    }
    return code;
}

#define GL_CUDA(call)                                                        \
    do {                                                                     \
        cudaError_t e_ = (call);                                             \
        if (e_ != cudaSuccess) return gl_fail(g, e_ == cudaErrorMemoryAllocation ? -2 : -3, #call, e_); \
    } while (0)

int gl_reserve(TacoGl* g, int B, int F) {
    const size_t need = (size_t)B * F * g->g.win;
    if (need > g->yf_floats) {
        for (int i = 0; i < 2; ++i) {
            if (g->yf[i]) cudaFree(g->yf[i]);
            g->yf[i] = nullptr;
        }
        g->yf_floats = 0;
        GL_CUDA(cudaMalloc(&g->yf[0], need * sizeof(float)));
        GL_CUDA(cudaMalloc(&g->yf[1], need * sizeof(float)));
        g->yf_floats = need;
    }
    if ((size_t)B > g->dist_cap) {
        if (g->dist_acc) cudaFree(g->dist_acc);
        g->dist_acc = nullptr;
        g->dist_cap = 0;
        GL_CUDA(cudaMalloc(&g->dist_acc, (size_t)B * 2 * sizeof(float)));
        g->dist_cap = B;
    }
    return 0;
}

template <int MODE, bool CENTRED>
int gl_launch2(TacoGl* g, const GlArgs& a, int B, cudaStream_t st) {
    static bool attr_done = false;
    if (!attr_done) {
        GL_CUDA(cudaFuncSetAttribute(gl_kernel<MODE, CENTRED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GL_SMEM));
        attr_done = true;
    }
    dim3 grid((a.F + 1) / 2, B);
    gl_kernel<MODE, CENTRED><<<grid, GL_THREADS, GL_SMEM, st>>>(a);
    GL_CUDA(cudaGetLastError());
    return 0;
}
template <int MODE>
int gl_launch(TacoGl* g, const GlArgs& a, int B, cudaStream_t st) {
    return 2 * a.g.win <= N ? gl_launch2<MODE, true>(g, a, B, st) : gl_launch2<MODE, false>(g, a, B, st);
}

}  // namespace

extern "C" {

int taco_gl_create(int device, int n_fft, int hop, int win, int n_mels, const float* inv_mel_basis, TacoGl** out) {
    if (!out) return -1;
    *out = nullptr;
    if (n_fft != N || hop < 1 || win < hop || win > N || (win & 1) || n_mels < 1 || n_mels > 512 || !inv_mel_basis) return -1;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return -3;
    TacoGl* g = new TacoGl();
    g->device = device;
    g->g.hop = hop;
    g->g.win = win;
    g->g.lpad = (N - win) / 2;
    g->g.cover = (win + hop - 1) / hop;
    g->n_mels = n_mels;
    std::vector<float2> tw(TW_TOTAL);
    {
        const int radix[3] = {8, 8, 4}, ns[3] = {8, 64, 512}, off[3] = {TW_OFF1, TW_OFF2, TW_OFF3};
        for (int p = 0; p < 3; ++p)
            for (int k = 0; k < ns[p]; ++k)
                for (int r = 1; r < radix[p]; ++r) {
                    const double ang = -2.0 * 3.14159265358979323846 * r * k / ((double)ns[p] * radix[p]);
                    tw[off[p] + k * (radix[p] - 1) + (r - 1)] = make_float2((float)cos(ang), (float)sin(ang));
                }
    }
    std::vector<float> w(win);
    for (int j = 0; j < win; ++j) w[j] = (float)(0.5 - 0.5 * cos(2.0 * 3.14159265358979323846 * j / win));   // periodic Hann
    std::vector<float> wsum(win + hop);
    for (int i = 0; i < win + hop; ++i) {
        float ws = 0.f;
        for (int d = -((win - 1) / hop); d <= (win + hop - 1) / hop; ++d) {
            const int j = i - d * hop;
            if (j >= 0 && j < win) ws = fmaf(w[j], w[j], ws);
        }
        wsum[i] = ws > 1.17549435e-38f ? ws : 1.f;
    }
    std::vector<float> invT((size_t)n_mels * NB);
    for (int k = 0; k < NB; ++k)
        for (int m = 0; m < n_mels; ++m) invT[(size_t)m * NB + k] = inv_mel_basis[(size_t)k * n_mels + m];
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaMalloc(&g->tw, TW_TOTAL * sizeof(float2));
    if (e == cudaSuccess) e = cudaMalloc(&g->w, win * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&g->invT, invT.size() * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&g->wsum, wsum.size() * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(g->wsum, wsum.data(), wsum.size() * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(g->tw, tw.data(), TW_TOTAL * sizeof(float2), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(g->w, w.data(), win * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(g->invT, invT.data(), invT.size() * sizeof(float), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        taco_gl_destroy(g);
        return e == cudaErrorMemoryAllocation ? -2 : -3;
    }
    *out = g;
    return 0;
}

void taco_gl_destroy(TacoGl* g) {
    if (!g) return;
    cudaFree(g->tw);
    cudaFree(g->w);
    cudaFree(g->invT);
    cudaFree(g->wsum);
    cudaFree(g->yf[0]);
    cudaFree(g->yf[1]);
    cudaFree(g->dist_acc);
    for (int i = 0; i < 4; ++i) cudaFree(g->stage[i]);
    if (g->pinned) cudaFreeHost(g->pinned);
    delete g;
}

const char* taco_gl_last_error(const TacoGl* g) { return g ? g->err.c_str() : "null plan"; }

int taco_gl_mel_to_mag(TacoGl* g, const float* mel_dev, const int32_t* n_frames_dev, int B, int F, const TacoMelNorm* norm,
                       float* mag_dev, void* stream) {
    if (!g) return -1;
    if (!mel_dev || !mag_dev || !norm || B < 1 || F < 1) return gl_fail(g, -1, "taco_gl_mel_to_mag: bad argument");
    MelNormDev nm{norm->signal_normalization, norm->allow_clipping_in_normalization, norm->symmetric_mels, norm->max_abs_value,
                  norm->min_level_db, norm->ref_level_db, norm->power};
    dim3 grid((F + MM_FRAMES - 1) / MM_FRAMES, B);
    gl_mel_to_mag_kernel<<<grid, 256, MM_FRAMES * g->n_mels * sizeof(float), (cudaStream_t)stream>>>(mel_dev, n_frames_dev, F, g->n_mels,
                                                                                                  g->invT, nm, mag_dev);
    GL_CUDA(cudaGetLastError());
    return 0;
}

int taco_gl_griffin_lim(TacoGl* g, const float* mag_dev, const float* turns_dev, const int32_t* n_frames_dev, uint64_t seed,
                        int utterance0, int B, int F, int iters, float preemphasis_k, float* audio_dev, long long audio_stride,
                        void* stream) {
    if (!g) return -1;
    if (!mag_dev || !audio_dev || B < 1 || F < 2 || iters < 0 || audio_stride < (long long)g->g.hop * (F - 1))
        return gl_fail(g, -1, "taco_gl_griffin_lim: bad argument");
    if ((long long)g->g.hop * (F - 1) <= N / 2 && !n_frames_dev)
        return gl_fail(g, -1, "taco_gl_griffin_lim: hop * (frames - 1) must exceed n_fft / 2 (reflect padding)");
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = gl_reserve(g, B, F)) return rc;
    GlArgs a{};
    a.mag = mag_dev;
    a.turns = turns_dev;
    a.n_frames = n_frames_dev;
    a.tw = g->tw;
    a.w = g->w;
    a.wsum = g->wsum;
    a.g = g->g;
    a.F = F;
    a.seed = seed;
    a.utt0 = utterance0;
    a.yf_out = g->yf[0];
    if (int rc = gl_launch<GL_FIRST>(g, a, B, st)) return rc;
    int cur = 0;
    for (int it = 0; it < iters; ++it) {
        a.yf_in = g->yf[cur];
        a.yf_out = g->yf[cur ^ 1];
        if (int rc = gl_launch<GL_ITER>(g, a, B, st)) return rc;
        cur ^= 1;
    }
    const int gx = (int)((audio_stride + 255) / 256 < 4096 ? (audio_stride + 255) / 256 : 4096);
    gl_ola_kernel<<<dim3(gx, B), 256, 0, st>>>(g->yf[cur], g->w, g->g, n_frames_dev, F, audio_dev, audio_stride);
    GL_CUDA(cudaGetLastError());
    if (preemphasis_k != 0.f) {
        gl_deemphasis_kernel<<<B, 1024, 0, st>>>(audio_dev, audio_stride, g->g, n_frames_dev, F, preemphasis_k);
        GL_CUDA(cudaGetLastError());
    }
    return 0;
}

int taco_gl_spectral_distance(TacoGl* g, const float* audio_dev, long long audio_stride, const float* mag_dev,
                              const int32_t* n_frames_dev, int B, int F, float* out_dev, void* stream) {
    if (!g) return -1;
    if (!audio_dev || !mag_dev || !out_dev || B < 1 || F < 2) return gl_fail(g, -1, "taco_gl_spectral_distance: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    if (int rc = gl_reserve(g, B, F)) return rc;
    GL_CUDA(cudaMemsetAsync(g->dist_acc, 0, (size_t)B * 2 * sizeof(float), st));
    GlArgs a{};
    a.mag = mag_dev;
    a.sig = audio_dev;
    a.sig_stride = audio_stride;
    a.dist_acc = g->dist_acc;
    a.n_frames = n_frames_dev;
    a.tw = g->tw;
    a.w = g->w;
    a.wsum = g->wsum;
    a.g = g->g;
    a.F = F;
    if (int rc = gl_launch<GL_MEASURE>(g, a, B, st)) return rc;
    gl_dist_final_kernel<<<(B + 127) / 128, 128, 0, st>>>(g->dist_acc, out_dev, B);
    GL_CUDA(cudaGetLastError());
    return 0;
}

int taco_mel_to_audio(TacoGl* g, const float* mel_host, const int32_t* n_frames_host, int B, int F, const TacoMelNorm* norm,
                      uint64_t seed, int utterance0, int iters, float preemphasis_k, float* audio_host, long long audio_stride) {
    if (!g) return -1;
    if (!mel_host || !audio_host || !norm || B < 1 || F < 2 || audio_stride < (long long)g->g.hop * (F - 1))
        return gl_fail(g, -1, "taco_mel_to_audio: bad argument");
    for (int b = 0; b < B; ++b) {
        const int fb = n_frames_host ? n_frames_host[b] : F;
        if (fb > F || (long long)g->g.hop * (fb - 1) <= N / 2)
            return gl_fail(g, -1, "taco_mel_to_audio: every utterance needs hop * (frames - 1) > n_fft / 2 and frames <= F");
    }
    GL_CUDA(cudaSetDevice(g->device));
    const size_t n_mel = (size_t)B * F * g->n_mels, n_mag = (size_t)B * F * NB, n_audio = (size_t)B * audio_stride;
    const size_t need[4] = {n_mel * sizeof(float), n_mag * sizeof(float), n_audio * sizeof(float), (size_t)B * sizeof(int32_t)};
    for (int i = 0; i < 4; ++i)
        if (need[i] > g->stage_bytes[i]) {
            cudaFree(g->stage[i]);
            g->stage[i] = nullptr;
            g->stage_bytes[i] = 0;
            GL_CUDA(cudaMalloc(&g->stage[i], need[i]));
            g->stage_bytes[i] = need[i];
        }
    // pinned bounce buffer: the copies run at full PCIe rate and in chunks that overlap the host memcpy of the next chunk
    const size_t CHUNK = (size_t)16 << 20;
    if (g->pinned_bytes < 2 * CHUNK) {
        if (g->pinned) cudaFreeHost(g->pinned);
        g->pinned = nullptr;
        g->pinned_bytes = 0;
        GL_CUDA(cudaHostAlloc(&g->pinned, 2 * CHUNK, cudaHostAllocDefault));
        g->pinned_bytes = 2 * CHUNK;
    }
    float* mel = static_cast<float*>(g->stage[0]);
    float* mag = static_cast<float*>(g->stage[1]);
    float* audio = static_cast<float*>(g->stage[2]);
    int32_t* nf = n_frames_host ? static_cast<int32_t*>(g->stage[3]) : nullptr;
    cudaStream_t st = nullptr;
    cudaEvent_t ev[2];
    GL_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
    GL_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
    int rc = 0;
    auto fail_cuda = [&](const char* what, cudaError_t e) { if (!rc) rc = gl_fail(g, -3, what, e); };
    {   // host -> device through the two halves of the bounce buffer
        const char* src = reinterpret_cast<const char*>(mel_host);
        size_t off = 0;
        for (int k = 0; off < need[0] && !rc; ++k, off += CHUNK) {
            const size_t n = need[0] - off < CHUNK ? need[0] - off : CHUNK;
            char* half = static_cast<char*>(g->pinned) + (k & 1) * CHUNK;
            cudaError_t e = k >= 2 ? cudaEventSynchronize(ev[k & 1]) : cudaSuccess;
            if (e != cudaSuccess) { fail_cuda("taco_mel_to_audio: staging", e); break; }
            memcpy(half, src + off, n);
            e = cudaMemcpyAsync(reinterpret_cast<char*>(mel) + off, half, n, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) e = cudaEventRecord(ev[k & 1], st);
            if (e != cudaSuccess) fail_cuda("taco_mel_to_audio: staging", e);
        }
        if (!rc && nf) {
            cudaError_t e = cudaMemcpyAsync(nf, n_frames_host, need[3], cudaMemcpyHostToDevice, st);
            if (e != cudaSuccess) fail_cuda("taco_mel_to_audio: staging", e);
        }
    }
    if (!rc) rc = taco_gl_mel_to_mag(g, mel, nf, B, F, norm, mag, st);
    if (!rc) rc = taco_gl_griffin_lim(g, mag, nullptr, nf, seed, utterance0, B, F, iters, preemphasis_k, audio, audio_stride, st);
    if (!rc) {   // device -> host, chunk k copied out of the bounce buffer while chunk k + 1 is in flight
        char* dst = reinterpret_cast<char*>(audio_host);
        const size_t total = need[2];
        const size_t nchunks = (total + CHUNK - 1) / CHUNK;
        for (size_t k = 0; k <= nchunks && !rc; ++k) {
            if (k < nchunks) {
                const size_t off = k * CHUNK, n = total - off < CHUNK ? total - off : CHUNK;
                cudaError_t e = cudaMemcpyAsync(static_cast<char*>(g->pinned) + (k & 1) * CHUNK, reinterpret_cast<char*>(audio) + off, n,
                                                cudaMemcpyDeviceToHost, st);
                if (e == cudaSuccess) e = cudaEventRecord(ev[k & 1], st);
                if (e != cudaSuccess) { fail_cuda("taco_mel_to_audio: read back", e); break; }
            }
            if (k >= 1) {
                const size_t j = k - 1, off = j * CHUNK, n = total - off < CHUNK ? total - off : CHUNK;
                cudaError_t e = cudaEventSynchronize(ev[j & 1]);
                if (e != cudaSuccess) { fail_cuda("taco_mel_to_audio: read back", e); break; }
                memcpy(dst + off, static_cast<char*>(g->pinned) + (j & 1) * CHUNK, n);
            }
        }
    }
    cudaStreamSynchronize(st);
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
    return rc;
}

}  // extern "C"
