// Distributed-shared-memory throughput per SM, as the decode kernel's reduction tail and the cluster split-K exchange
// use it: every CTA of a cluster sends `bytes` of its shared memory, split evenly over the other CTAs, and the
// measurement ends when everything addressed to a CTA has landed (its mbarrier completes).
// Methods:
//   0  one cp.async.bulk.shared::cluster per destination (what the kernels do today)
//   1  st.async.b32 from 128 threads, lanes on consecutive words (128 B per warp instruction)
//   2  st.async.v4.b32 from 128 threads (512 B per warp instruction)
//   3  st.async.v4.b32 from 256 threads
//   4  the receiver pulls: ld.shared::cluster.v4 by 256 threads after a cluster barrier
// Cluster sizes 2, 4, 8; one or two CTAs per SM (the decode kernel runs two).  Prints bytes / clock / SM (sent = received).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 --expt-relaxed-constexpr -o dsmem_throughput dsmem_throughput.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../deepgemm_ascend_b200/csrc/ptx.cuh"

using namespace dga;

__device__ __forceinline__ void st_async_b32(uint32_t cluster_addr, uint32_t v, uint32_t cluster_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
                 :: "r"(cluster_addr), "r"(v), "r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void st_async_v4(uint32_t cluster_addr, uint4 v, uint32_t cluster_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];"
                 :: "r"(cluster_addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(cluster_bar) : "memory");
}
__device__ __forceinline__ uint4 ld_cluster_v4(uint32_t cluster_addr) {
    uint4 v;
    asm volatile("ld.shared::cluster.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(cluster_addr) : "memory");
    return v;
}

// shared memory: [send: bytes][recv: bytes][barrier]
__global__ void __launch_bounds__(256, 2) probe(uint32_t method, uint32_t bytes, uint32_t reps,
                                                 unsigned long long* out_clk, uint32_t* sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* send = smem;
    uint8_t* recv = smem + bytes;
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 2 * bytes);
    uint32_t csize;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(csize));
    const uint32_t rank = ptx::cluster_ctarank();
    const uint32_t per = bytes / (csize - 1) & ~127u;   // bytes per destination
    const uint32_t tid = threadIdx.x;
    for (uint32_t i = tid; i < bytes / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(send)[i] = i * 2654435761u + rank;
    if (tid == 0) { ptx::mbar_init(bar, 1); ptx::fence_mbar_init(); }
    __syncthreads();
    ptx::cluster_sync();
    unsigned long long total = 0;
    uint32_t acc = 0;
    for (uint32_t r = 0; r < reps; ++r) {
        if (method != 4 && tid == 0) ptx::mbar_arrive_expect_tx(bar, per * (csize - 1));
        ptx::cluster_sync();   // every receiver is armed; everyone starts together
        const unsigned long long t0 = clock64();
        const uint32_t send_a = ptx::smem_u32(send), recv_a = ptx::smem_u32(recv), bar_a = ptx::smem_u32(bar);
        if (method == 0) {
            ptx::fence_async_smem();
            if (tid < csize - 1) {
                const uint32_t dst = (rank + 1 + tid) % csize;
                // slot of this sender inside the destination's receive area: ((rank - dst - 1) mod csize)
                const uint32_t slot = (rank + csize - dst - 1) % csize;
                ptx::dsmem_bulk_copy(recv_a + slot * per, send_a + tid * per, per, bar, dst);
            }
        } else if (method <= 3) {
            const uint32_t nthr = method == 3 ? 256u : 128u;
            if (tid < nthr) {
                for (uint32_t d = 0; d < csize - 1; ++d) {
                    const uint32_t dst = (rank + 1 + d) % csize;
                    const uint32_t slot = (rank + csize - dst - 1) % csize;
                    const uint32_t rbase = ptx::mapa(recv_a + slot * per, dst), rbar = ptx::mapa(bar_a, dst);
                    if (method == 1) {
                        const uint32_t* src = reinterpret_cast<const uint32_t*>(send + d * per);
                        for (uint32_t i = tid; i < per / 4; i += nthr) st_async_b32(rbase + i * 4, src[i], rbar);
                    } else {
                        const uint4* src = reinterpret_cast<const uint4*>(send + d * per);
                        for (uint32_t i = tid; i < per / 16; i += nthr) st_async_v4(rbase + i * 16, src[i], rbar);
                    }
                }
            }
        } else {
            for (uint32_t d = 0; d < csize - 1; ++d) {
                const uint32_t src_cta = (rank + 1 + d) % csize;
                const uint32_t slot = (src_cta + csize - rank - 1) % csize;   // my part of the peer's send area
                const uint32_t sbase = ptx::mapa(send_a + slot * per, src_cta);
                uint4* dstp = reinterpret_cast<uint4*>(recv + d * per);
                for (uint32_t i = tid; i < per / 16; i += blockDim.x) {
                    const uint4 v = ld_cluster_v4(sbase + i * 16);
                    dstp[i] = v;
                    acc ^= v.x;
                }
            }
        }
        if (method != 4) ptx::mbar_wait(bar, r & 1);
        else __syncthreads();
        const unsigned long long t1 = clock64();
        total += t1 - t0;
        ptx::cluster_sync();   // nobody re-arms or overwrites while a peer is still receiving / reading
    }
    if (tid == 0) out_clk[blockIdx.x] = total;
    acc ^= reinterpret_cast<uint32_t*>(recv)[tid];
    if (acc == 0x12345678u) sink[0] = acc;
}

static void run(uint32_t method, uint32_t csize, uint32_t bytes, uint32_t ctas_per_sm, unsigned long long* d_clk, uint32_t* d_sink) {
    const uint32_t reps = 50;
    const int smem = int(2 * bytes + 64);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg{};
    // clusters that fit: leave the placement to the driver, ask for at most 128 (x ctas_per_sm) CTAs
    const uint32_t grid = (128 * ctas_per_sm / csize) * csize;
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(256);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = csize; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    for (int rep = 0; rep < 2; ++rep) {
        cudaMemset(d_clk, 0, 1024 * 8);
        cudaError_t e = cudaLaunchKernelEx(&cfg, probe, method, bytes, reps, d_clk, d_sink);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("method %u cluster %u: %s\n", method, csize, cudaGetErrorString(e)); return; }
    }
    static unsigned long long h[1024];
    cudaMemcpy(h, d_clk, grid * 8, cudaMemcpyDeviceToHost);
    double worst = 0, sum = 0;
    for (uint32_t b = 0; b < grid; ++b) { worst = h[b] > worst ? double(h[b]) : worst; sum += double(h[b]); }
    const uint32_t per = bytes / (csize - 1) & ~127u;
    const double sent = double(per) * (csize - 1) * ctas_per_sm;   // per SM and repetition
    static const char* names[] = {"bulk copy", "st.async.b32 x128", "st.async.v4 x128", "st.async.v4 x256", "ld.cluster.v4 x256"};
    printf("%-20s cluster %u, %u CTA/SM, %5.1f KB per CTA: %6.0f clk (mean), %6.0f (slowest CTA) -> %5.1f B/clk/SM (mean), %5.1f (slowest)\n",
           names[method], csize, ctas_per_sm, per * (csize - 1) / 1024.0, sum / grid / reps, worst / reps,
           sent / (sum / grid / reps), sent / (worst / reps));
}

int main() {
    unsigned long long* d_clk; uint32_t* d_sink;
    cudaMalloc(&d_clk, 1024 * 8); cudaMalloc(&d_sink, 4);
    for (uint32_t ctas_per_sm : {1u, 2u})
        for (uint32_t csize : {2u, 4u, 8u})
            for (uint32_t bytes : {28u * 1024, 48u * 1024})
                for (uint32_t method = 0; method < 5; ++method) {
                    if (ctas_per_sm == 2 && 2 * bytes + 64 > 110 * 1024) continue;
                    run(method, csize, bytes, ctas_per_sm, d_clk, d_sink);
                }
    return 0;
}
