// L2 -> SM bandwidth available to the decoder's context phase: 112 CTAs x 512 threads (the decoder's footprint), each CTA
// re-reading its own slice of an L2-resident buffer with 16-byte loads, UB loads in flight per thread (the kernel uses 5).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/probes/l2_probe scripts/probes/l2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int UB>
__global__ void __launch_bounds__(512) probe(const uint4* __restrict__ buf, size_t slice16, int iters, uint4* out) {
    const uint4* p = buf + (size_t)blockIdx.x * slice16;
    uint4 acc = make_uint4(0, 0, 0, 0);
    for (int it = 0; it < iters; ++it) {
        for (size_t i = threadIdx.x; i < slice16; i += 512 * UB) {
            uint4 v[UB];
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const size_t j = i + (size_t)u * 512;
                v[u] = make_uint4(0, 0, 0, 0);
                if (j < slice16) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(p + j));
            }
#pragma unroll
            for (int u = 0; u < UB; ++u) { acc.x ^= v[u].x; acc.y += v[u].y; acc.z ^= v[u].z; acc.w += v[u].w; }
        }
        __syncthreads();   // one dependent round per "step", like the kernel
    }
    if (acc.x == 0x12345678u) out[blockIdx.x] = acc;
}

template <int UB>
void run(const uint4* buf, size_t slice_bytes, int ctas, int iters, uint4* out) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<UB><<<ctas, 512>>>(buf, slice_bytes / 16, 20, out);
    cudaEventRecord(e0);
    probe<UB><<<ctas, 512>>>(buf, slice_bytes / 16, iters, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = (double)slice_bytes * ctas * iters;
    printf("ctas %3d  slice %6zu B  UB %2d: %8.1f GB/s   %.2f us per pass of %.2f MB\n", ctas, slice_bytes, UB, bytes / ms * 1e-6,
           1e3 * ms / iters, slice_bytes * ctas * 1e-6);
}

int main() {
    const size_t total = 64u << 20;
    uint4 *buf, *out;
    cudaMalloc(&buf, total); cudaMemset(buf, 1, total); cudaMalloc(&out, 4096 * 16);
    for (int ctas : {112, 148}) {
        for (size_t slice : {(size_t)96 * 1024, (size_t)24 * 1024}) {
            run<2>(buf, slice, ctas, 2000, out);
            run<5>(buf, slice, ctas, 2000, out);
            run<8>(buf, slice, ctas, 2000, out);
            run<16>(buf, slice, ctas, 2000, out);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
