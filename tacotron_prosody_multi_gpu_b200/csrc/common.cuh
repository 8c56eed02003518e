// Shared device helpers for the sm_100a Tacotron kernels: grid-wide barrier for the
// persistent cooperative kernels, Philox4x32-10 (dropout bits), activations, and the
// weight-stationary fp32 dense micro-kernel used by every "column-partitioned" phase.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define TACO_THREADS 512

// ---------------------------------------------------------------------------------------------
// Grid barrier.  All CTAs are co-resident (cudaLaunchCooperativeKernel).  One monotonically
// increasing counter per launch; no reset between barriers.  Thread 0 publishes the CTA's
// writes (fence), arrives, and spins with an acquire load; data exchanged across CTAs is
// always read with ld.global.cg (L1 is not coherent).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add_u32(unsigned* p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

struct GridBarrier {
    unsigned* counter;
    unsigned target;
    unsigned n;
};

__device__ __forceinline__ void grid_sync(GridBarrier& gb) {
    gb.target += gb.n;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        red_release_add_u32(gb.counter, 1u);
        while (ld_acquire_u32(gb.counter) < gb.target) {
        }
        __threadfence();
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al.), counter = (step, batch_row, layer, unit>>7), key = seed.
// keep(unit) = bit (unit & 127) of the 128-bit block.  Restated in oracle/philox.py.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
        uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += W0;
        k.y += W1;
    }
    return c;
}

__device__ __forceinline__ bool philox_keep(uint64_t seed, uint32_t step, uint32_t row, uint32_t layer, uint32_t unit) {
    uint4 o = philox4x32_10(make_uint4(step, row, layer, unit >> 7), make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    uint32_t bit = unit & 127u;
    uint32_t w = (bit >> 5) == 0 ? o.x : ((bit >> 5) == 1 ? o.y : ((bit >> 5) == 2 ? o.z : o.w));
    return (w >> (bit & 31u)) & 1u;
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch (the time loops of the training step are chains of small dependent kernels): a kernel
// launched with launch_pdl() may be scheduled while its predecessor on the stream still runs.  It must execute pdl_wait()
// before it reads anything the predecessor wrote or writes anything at all (before that: index arithmetic, Philox masks,
// loads of weights and of tensors older than the predecessor); pdl_launch_dependents() after the wait lets the NEXT kernel
// of the chain start its own prologue under this kernel's body.  Without the launch attribute both are no-ops.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#ifdef __CUDACC__
#include <stdlib.h>
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    static const bool on = !(getenv("TACO_TRAIN_PDL") && atoi(getenv("TACO_TRAIN_PDL")) == 0);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = on ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---------------------------------------------------------------------------------------------
// Weight-stationary dense micro-kernel (fp32 SIMT).
//   out[r][c] = sum_k X[k][r] * W[k][c],  r < RBP (rows = batch rows of this CTA's batch tile),
//   c < NC (this CTA's columns), k over up to 3 concatenated segments.
// X segments live in global memory in "XT" layout [K][RBP] (k-major, batch contiguous) written by
// other CTAs in the previous phase -> read with ld.global.cg.  W lives in shared memory [K][NC].
// Thread tile 4 rows x 4 cols; the (RBP/4)*(NC/4) tiles of a k-slice are adjacent lanes so the
// W / X loads broadcast; K is split TACO_THREADS/tiles ways and reduced through shared memory.
// Result: res[r*NC + c] in shared memory (valid after the trailing __syncthreads()).
// ---------------------------------------------------------------------------------------------
struct XSeg {
    const float* ptr;  // [K][RBP]
    int K;
};

__device__ __forceinline__ void dense_xt(const XSeg* segs, int nseg, const float* __restrict__ Ws, int NC, int RBP,
                                         float* part /* smem, TACO_THREADS*16 floats */, float* res /* smem RBP*NC */) {
    const int tid = threadIdx.x;
    const int tiles_c = NC >> 2;
    const int tiles = (RBP >> 2) * tiles_c;
    const int KS = TACO_THREADS / tiles;
    const int tile = tid % tiles;
    const int ks = tid / tiles;
    const int tr = tile / tiles_c;
    const int tc = tile - tr * tiles_c;
    const int rb4 = RBP >> 2;

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    int kbase = 0;
    for (int s = 0; s < nseg; ++s) {
        const float4* xp = reinterpret_cast<const float4*>(segs[s].ptr) + tr;
        const float4* wp = reinterpret_cast<const float4*>(Ws + (size_t)kbase * NC) + tc;
        const int K = segs[s].K;
        int k = ks;
        // 4-deep software pipeline of the L2 loads
        for (; k + 3 * KS < K; k += 4 * KS) {
            float4 x0 = __ldcg(xp + (size_t)(k) * rb4);
            float4 x1 = __ldcg(xp + (size_t)(k + KS) * rb4);
            float4 x2 = __ldcg(xp + (size_t)(k + 2 * KS) * rb4);
            float4 x3 = __ldcg(xp + (size_t)(k + 3 * KS) * rb4);
            float4 w0 = wp[(size_t)(k) * tiles_c];
            float4 w1 = wp[(size_t)(k + KS) * tiles_c];
            float4 w2 = wp[(size_t)(k + 2 * KS) * tiles_c];
            float4 w3 = wp[(size_t)(k + 3 * KS) * tiles_c];
#define TACO_FMA16(XV, WV)                                                                   \
    acc[0][0] = fmaf(XV.x, WV.x, acc[0][0]); acc[0][1] = fmaf(XV.x, WV.y, acc[0][1]);        \
    acc[0][2] = fmaf(XV.x, WV.z, acc[0][2]); acc[0][3] = fmaf(XV.x, WV.w, acc[0][3]);        \
    acc[1][0] = fmaf(XV.y, WV.x, acc[1][0]); acc[1][1] = fmaf(XV.y, WV.y, acc[1][1]);        \
    acc[1][2] = fmaf(XV.y, WV.z, acc[1][2]); acc[1][3] = fmaf(XV.y, WV.w, acc[1][3]);        \
    acc[2][0] = fmaf(XV.z, WV.x, acc[2][0]); acc[2][1] = fmaf(XV.z, WV.y, acc[2][1]);        \
    acc[2][2] = fmaf(XV.z, WV.z, acc[2][2]); acc[2][3] = fmaf(XV.z, WV.w, acc[2][3]);        \
    acc[3][0] = fmaf(XV.w, WV.x, acc[3][0]); acc[3][1] = fmaf(XV.w, WV.y, acc[3][1]);        \
    acc[3][2] = fmaf(XV.w, WV.z, acc[3][2]); acc[3][3] = fmaf(XV.w, WV.w, acc[3][3]);
            TACO_FMA16(x0, w0)
            TACO_FMA16(x1, w1)
            TACO_FMA16(x2, w2)
            TACO_FMA16(x3, w3)
        }
        for (; k < K; k += KS) {
            float4 x0 = __ldcg(xp + (size_t)k * rb4);
            float4 w0 = wp[(size_t)k * tiles_c];
            TACO_FMA16(x0, w0)
        }
        kbase += K;
    }
    // partials: part[(ks*tiles + tile)*16 + i*4 + j]
    float4* pp = reinterpret_cast<float4*>(part) + (size_t)(ks * tiles + tile) * 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) pp[i] = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    __syncthreads();
    // reduction: nOut = tiles*16 outputs, KS = 16*G partials each (G = TACO_THREADS / nOut)
    const int nOut = tiles * 16;
    if (nOut > TACO_THREADS) {   // wide tiles (32 rows x 64 columns): KS = TACO_THREADS / tiles partials per output, several outputs per thread
        for (int o2 = tid; o2 < nOut; o2 += TACO_THREADS) {
            float sum2 = 0.f;
            for (int q = 0; q < KS; ++q) sum2 += part[(size_t)q * nOut + o2];
            const int t = o2 >> 4, i = (o2 >> 2) & 3, j = o2 & 3;
            res[((t / tiles_c) * 4 + i) * NC + (t % tiles_c) * 4 + j] = sum2;
        }
        __syncthreads();
        return;
    }
    const int G = TACO_THREADS / nOut;  // >= 1 because tiles <= 32
    const int o = tid % nOut;
    const int g = tid / nOut;
    float sum = 0.f;
    if (g < G) {
#pragma unroll 4
        for (int q = 0; q < 16; ++q) sum += part[(size_t)((g * 16 + q) * tiles) * 16 + o];
    }
    if (G > 1) {
        __syncthreads();
        if (g < G) part[g * nOut + o] = sum;
        __syncthreads();
        if (g == 0) {
            sum = 0.f;
            for (int q = 0; q < G; ++q) sum += part[q * nOut + o];
        }
    }
    if (g == 0) {
        // o = tile*16 + i*4 + j  ->  (row, col)
        const int t = o >> 4, i = (o >> 2) & 3, j = o & 3;
        const int r = (t / tiles_c) * 4 + i;
        const int c = (t % tiles_c) * 4 + j;
        res[r * NC + c] = sum;
    }
    __syncthreads();
}
