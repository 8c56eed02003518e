// Microbenchmark (not product code): one-hop latencies that bound a decoder step on B200.
//  (1) global ping-pong between two CTAs (st.relaxed.gpu / ld.relaxed.gpu through L2)
//  (2) DSMEM ping-pong inside a cluster (st.shared::cluster + local poll; mbarrier remote arrive)
//  (3) cluster-wide all-gather of W bytes per CTA through DSMEM stores + remote mbarrier arrive,
//      with all clusters of the grid running concurrently and ~200 KB of dynamic smem per CTA
//      (the footprint of the weight-resident decoder), for cluster sizes 8 and 16.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o microbench_hop microbench_hop.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); } } while (0)

__device__ __forceinline__ unsigned ld_relaxed(const unsigned* p) { unsigned v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_relaxed(unsigned* p, unsigned v) { asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) { uint32_t r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r; }
__device__ __forceinline__ void st_cluster_u32(uint32_t addr, uint32_t v) { asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint4 v) { asm volatile("st.shared::cluster.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
__device__ __forceinline__ uint32_t ld_shared_volatile(uint32_t addr) { uint32_t v; asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory"); return v; }
__device__ __forceinline__ void mbar_init(uint32_t addr, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(addr), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) { asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t addr, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(addr), "r"(parity) : "memory");
}

// (1) global ping-pong: CTA a and CTA b of the grid, one thread each
__global__ void k_pp_global(unsigned* flags, int a, int b, int iters, long long* out) {
    if (threadIdx.x != 0) return;
    const int cta = blockIdx.x;
    if (cta != a && cta != b) return;
    unsigned* fa = flags;        // written by a
    unsigned* fb = flags + 64;   // written by b (different line)
    long long t0 = clock64();
    for (int it = 1; it <= iters; ++it) {
        if (cta == a) {
            st_relaxed(fa, it);
            while (ld_relaxed(fb) != (unsigned)it) {}
        } else {
            while (ld_relaxed(fa) != (unsigned)it) {}
            st_relaxed(fb, it);
        }
    }
    if (cta == a) out[0] = clock64() - t0;
}

// (2) DSMEM ping-pong between rank 0 and rank `peer` of one cluster
__global__ void k_pp_dsmem(int peer, int iters, int use_mbar, long long* out) {
    __shared__ __align__(8) unsigned long long bar;
    __shared__ unsigned flag;
    cg::cluster_group cl = cg::this_cluster();
    const unsigned rank = cl.block_rank();
    if (threadIdx.x == 0) { flag = 0; mbar_init(smem_u32(&bar), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    cl.sync();
    if (threadIdx.x == 0 && (rank == 0 || rank == (unsigned)peer)) {
        const unsigned other = rank == 0 ? peer : 0;
        const uint32_t rflag = mapa(smem_u32(&flag), other), rbar = mapa(smem_u32(&bar), other);
        const uint32_t lflag = smem_u32(&flag), lbar = smem_u32(&bar);
        long long t0 = clock64();
        for (int it = 1; it <= iters; ++it) {
            if (rank == 0) {
                if (use_mbar) { mbar_arrive_remote(rbar); mbar_wait(lbar, (it - 1) & 1); }
                else { st_cluster_u32(rflag, it); while (ld_shared_volatile(lflag) != (unsigned)it) {} }
            } else {
                if (use_mbar) { mbar_wait(lbar, (it - 1) & 1); mbar_arrive_remote(rbar); }
                else { while (ld_shared_volatile(lflag) != (unsigned)it) {} st_cluster_u32(rflag, it); }
            }
        }
        if (rank == 0) out[0] = clock64() - t0;
    }
    cl.sync();
}

// (3) cluster all-gather: every CTA pushes W bytes to every CTA of its cluster (incl. itself), then one
// remote arrive per (src, dst); consumers wait for CS arrivals, then sum the gathered buffer.
__global__ void __launch_bounds__(512, 1) k_cluster_allgather(int W, int iters, float* sink, long long* out) {
    extern __shared__ __align__(16) unsigned char dsm[];
    __shared__ __align__(8) unsigned long long bar[2];
    cg::cluster_group cl = cg::this_cluster();
    const unsigned rank = cl.block_rank(), CS = cl.num_blocks();
    const int tid = threadIdx.x;
    if (tid == 0) { mbar_init(smem_u32(&bar[0]), CS); mbar_init(smem_u32(&bar[1]), CS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    cl.sync();
    // gathered buffer: 2 slots x [CS][W] bytes at dsm
    const int nvec = W / 16;                  // uint4 per CTA payload
    float acc = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const int slot = it & 1;
        const uint32_t base = smem_u32(dsm) + slot * CS * W + rank * W;
        // push: thread -> (dst, vec)
        for (int i = tid; i < (int)CS * nvec; i += 512) {
            const int dst = i / nvec, v = i - dst * nvec;
            uint4 val = make_uint4(it + v, __float_as_uint(acc), rank, dst);
            st_cluster_v4(mapa(base + v * 16, dst), val);
        }
        // release our stores cluster-wide, then one elected thread per destination signals
        asm volatile("fence.acq_rel.cluster;" ::: "memory");
        __syncthreads();
        if (tid < (int)CS) mbar_arrive_remote(mapa(smem_u32(&bar[slot]), tid));
        mbar_wait(smem_u32(&bar[slot]), (it >> 1) & 1);
        // consume
        const uint4* g = reinterpret_cast<const uint4*>(dsm + slot * CS * W);
        float s = 0.f;
        for (int i = tid; i < (int)CS * nvec; i += 512) { uint4 v = g[i]; s += (float)(v.x & 3); }
        acc += s * 1e-20f;
        __syncthreads();
    }
    if (tid == 0 && blockIdx.x == 0) out[0] = clock64() - t0;
    if (acc == 123.f) sink[0] = acc;
    cl.sync();
}


// (4) cluster all-gather, "data is the flag": st.shared::cluster.v4 multicast, consumers poll LOCAL smem for the
// iteration tag (no fence, no mbarrier).
__global__ void __launch_bounds__(512, 1) k_ag_poll(int W, int iters, float* sink, long long* out) {
    extern __shared__ __align__(16) unsigned char dsm[];
    cg::cluster_group cl = cg::this_cluster();
    const unsigned rank = cl.block_rank(), CS = cl.num_blocks();
    const int tid = threadIdx.x, nvec = W / 16;
    for (int i = tid; i < 2 * (int)CS * W / 4; i += 512) reinterpret_cast<unsigned*>(dsm)[i] = 0xFFFFFFFFu;
    cl.sync();
    float acc = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const int slot = it & 1;
        const uint32_t base = smem_u32(dsm) + slot * CS * W + rank * W;
        for (int i = tid; i < (int)CS * nvec; i += 512) {
            const int dst = i / nvec, v = i - dst * nvec;
            st_cluster_v4(mapa(base + v * 16, dst), make_uint4(it, it, it, it));
        }
        const uint32_t gbase = smem_u32(dsm) + slot * CS * W;
        float s = 0.f;
        for (int i = tid; i < (int)CS * nvec; i += 512) {
            uint4 v;
            do {
                asm volatile("ld.volatile.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(gbase + i * 16) : "memory");
            } while (v.x != (unsigned)it || v.y != (unsigned)it || v.z != (unsigned)it || v.w != (unsigned)it);
            s += (float)(v.x & 3);
        }
        acc += s * 1e-20f;
        __syncthreads();
    }
    if (tid == 0 && blockIdx.x == 0) out[0] = clock64() - t0;
    if (acc == 123.f) sink[0] = acc;
    cl.sync();
}

// (5) cluster all-gather through the bulk-copy engine: local staging -> cp.async.bulk smem -> remote smem with
// complete_tx on the destination's mbarrier.
__global__ void __launch_bounds__(512, 1) k_ag_bulk(int W, int iters, float* sink, long long* out) {
    extern __shared__ __align__(128) unsigned char dsm[];
    __shared__ __align__(8) unsigned long long bar[2];
    cg::cluster_group cl = cg::this_cluster();
    const unsigned rank = cl.block_rank(), CS = cl.num_blocks();
    const int tid = threadIdx.x, nvec = W / 16;
    unsigned char* stage = dsm + 2 * CS * W;   // local staging block [W]
    if (tid == 0) { mbar_init(smem_u32(&bar[0]), 1); mbar_init(smem_u32(&bar[1]), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    if (tid == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[0])), "r"((unsigned)(CS * W)) : "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[1])), "r"((unsigned)(CS * W)) : "memory");
    }
    cl.sync();
    float acc = 0.f;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const int slot = it & 1;
        for (int i = tid; i < nvec; i += 512) reinterpret_cast<uint4*>(stage)[i] = make_uint4(it, it, it, it);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid < (int)CS) {
            const uint32_t dst = mapa(smem_u32(dsm) + slot * CS * W + rank * W, tid);
            const uint32_t rb = mapa(smem_u32(&bar[slot]), tid);
            asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst), "r"(smem_u32(stage)), "r"((unsigned)W), "r"(rb) : "memory");
        }
        mbar_wait(smem_u32(&bar[slot]), (it >> 1) & 1);
        const uint4* g = reinterpret_cast<const uint4*>(dsm + slot * CS * W);
        float s = 0.f;
        for (int i = tid; i < (int)CS * nvec; i += 512) { uint4 v = g[i]; s += (float)(v.x & 3); }
        acc += s * 1e-20f;
        __syncthreads();   // all reads of this slot done, staging reusable
        if (tid == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[slot])), "r"((unsigned)(CS * W)) : "memory");
    }
    if (tid == 0 && blockIdx.x == 0) out[0] = clock64() - t0;
    if (acc == 123.f) sink[0] = acc;
    cl.sync();
}

template <class K, class... A>
cudaError_t launch_cluster(K kern, int grid, int block, int cluster, size_t smem, A... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = 0;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, args...);
}

int main() {
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    const double ghz = clk_khz * 1e-6;
    printf("SM clock (max) %.3f GHz\n", ghz);
    long long* out; CK(cudaMallocManaged(&out, 64));
    unsigned* flags; CK(cudaMalloc(&flags, 4096)); CK(cudaMemset(flags, 0, 4096));
    float* sink; CK(cudaMalloc(&sink, 4));
    const int iters = 20000;
    for (int b : {1, 2, 37, 74, 100, 147}) {
        CK(cudaMemset(flags, 0, 4096));
        k_pp_global<<<148, 32>>>(flags, 0, b, iters, out);
        CK(cudaDeviceSynchronize());
        printf("global ping-pong cta 0 <-> %3d: one-way %.0f cyc = %.3f us\n", b, out[0] / (2.0 * iters), out[0] / (2.0 * iters) / ghz * 1e-3);
    }
    CK(cudaFuncSetAttribute(k_pp_dsmem, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    for (int cs : {2, 8, 16}) for (int mb = 0; mb < 2; ++mb) {
        cudaError_t e = launch_cluster(k_pp_dsmem, cs, 32, cs, 0, cs - 1, iters, mb, out);
        if (e != cudaSuccess) { printf("dsmem ping-pong cluster %d: launch failed %s\n", cs, cudaGetErrorString(e)); cudaGetLastError(); continue; }
        CK(cudaDeviceSynchronize());
        printf("dsmem ping-pong cluster %2d rank 0 <-> %2d (%s): one-way %.0f cyc = %.3f us\n", cs, cs - 1, mb ? "mbarrier arrive/try_wait" : "st + poll", out[0] / (2.0 * iters), out[0] / (2.0 * iters) / ghz * 1e-3);
    }
    CK(cudaFuncSetAttribute(k_cluster_allgather, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CK(cudaFuncSetAttribute(k_cluster_allgather, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (int cs : {8, 16}) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(128); cfg.blockDim = dim3(512); cfg.dynamicSmemBytes = 200 * 1024;
        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int ncl = -1; cudaError_t e = cudaOccupancyMaxActiveClusters(&ncl, k_cluster_allgather, &cfg);
        printf("cluster size %2d, 512 thr, 200 KB smem: max active clusters = %d (%s)\n", cs, ncl, cudaGetErrorString(e));
        cudaGetLastError();
        for (int grid : {cs, 128}) for (int W : {256, 512, 1024, 2048}) {
            if (2 * cs * W > 200 * 1024) continue;
            cudaError_t e2 = launch_cluster(k_cluster_allgather, grid, 512, cs, (size_t)200 * 1024, W, 2000, sink, out);
            if (e2 != cudaSuccess) { printf("allgather cs=%d grid=%d: launch failed %s\n", cs, grid, cudaGetErrorString(e2)); cudaGetLastError(); continue; }
            e2 = cudaDeviceSynchronize();
            if (e2 != cudaSuccess) { printf("allgather cs=%d grid=%d: run failed %s\n", cs, grid, cudaGetErrorString(e2)); return 1; }
            printf("cluster all-gather cs=%2d grid=%3d W=%4d B/CTA: %.0f cyc = %.3f us per exchange\n", cs, grid, W, out[0] / 2000.0, out[0] / 2000.0 / ghz * 1e-3);
        }
    }
    CK(cudaFuncSetAttribute(k_ag_poll, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CK(cudaFuncSetAttribute(k_ag_poll, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_ag_bulk, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CK(cudaFuncSetAttribute(k_ag_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (int cs : {8, 16}) for (int grid : {cs, 7 * cs}) for (int W : {256, 512, 1536, 3072}) {
        cudaError_t e2 = launch_cluster(k_ag_poll, grid, 512, cs, (size_t)200 * 1024, W, 2000, sink, out);
        if (e2 == cudaSuccess) e2 = cudaDeviceSynchronize();
        if (e2 != cudaSuccess) { printf("ag_poll cs=%d grid=%d failed %s\n", cs, grid, cudaGetErrorString(e2)); return 1; }
        const double a = out[0] / 2000.0;
        e2 = launch_cluster(k_ag_bulk, grid, 512, cs, (size_t)200 * 1024, W, 2000, sink, out);
        if (e2 == cudaSuccess) e2 = cudaDeviceSynchronize();
        if (e2 != cudaSuccess) { printf("ag_bulk cs=%d grid=%d failed %s\n", cs, grid, cudaGetErrorString(e2)); return 1; }
        const double b = out[0] / 2000.0;
        printf("all-gather cs=%2d grid=%3d W=%4d B/CTA: st.v4+local poll %.0f cyc (%.3f us) | bulk copy+mbarrier %.0f cyc (%.3f us)\n",
               cs, grid, W, a, a / ghz * 1e-3, b, b / ghz * 1e-3);
    }
    return 0;
}
