// tcgen05.ld throughput per SM, with and without tensor-core work running beside it.
// The claim under test: "tcgen05.ld moves 64 B/clk per SM", which caps the fp32-scale fp8 path (every
// accumulator element leaves TMEM once per 128-wide K block) at 25 % of the tensor peak.
// One CTA per SM; 4 or 8 reader warps (warp w reads lanes 32*(w%4)..+32, the only window it may touch) sweep the
// 512 allocated columns with .32x32b.x{32,64,128} or .16x256b.x{8,16} loads; `--mma` adds one thread issuing
// 128x256x32 kind::f8f6f4 MMAs back to back into columns 0..255 from (uninitialised) shared memory.
// Prints bytes / clock / SM for every variant.   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_ld_throughput
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../deepgemm_ascend_b200/csrc/gemm_sm100.cuh"

using namespace dga;

template <int kShape>   // 0: 32x32b.x32, 1: 32x32b.x64, 2: 32x32b.x128, 3: 16x256b.x8 (16 lanes x 64 cols), 4: 16x256b.x16
__device__ __forceinline__ uint32_t ld_once(uint32_t taddr) {
    uint32_t acc = 0;
    if constexpr (kShape == 0) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(taddr, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 8) acc ^= v[i];
    } else if constexpr (kShape == 1) {
        uint32_t v[64];
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
            "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
            "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
              "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
              "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]),
              "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]),
              "=r"(v[40]), "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]),
              "=r"(v[48]), "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]),
              "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
            : "r"(taddr) : "memory");
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 64; i += 8) acc ^= v[i];
    } else if constexpr (kShape == 2) {
        // two x64 loads in flight before one wait: the largest footprint that stays in registers beside the loop
        uint32_t a[32], b[32], c[32], d[32];
        ptx::tmem_ld_32x32(taddr, a);
        ptx::tmem_ld_32x32(taddr + 32, b);
        ptx::tmem_ld_32x32(taddr + 64, c);
        ptx::tmem_ld_32x32(taddr + 96, d);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 8) acc ^= a[i] ^ b[i] ^ c[i] ^ d[i];
    } else if constexpr (kShape == 3) {
        uint32_t v[32];   // 16x256b.x8: 16 lanes x 8 x 256 bit = 64 columns of 16 lanes; 32 registers per thread
        asm volatile(
            "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
              "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
              "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr) : "memory");
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 8) acc ^= v[i];
    }
    return acc;
}
template <int kShape> constexpr uint32_t cols_per_ld() { return kShape == 0 ? 32 : kShape == 1 ? 64 : kShape == 2 ? 128 : 64; }
template <int kShape> constexpr uint32_t bytes_per_ld() { return kShape == 3 ? 16 * 64 * 4 : 32 * cols_per_ld<kShape>() * 4; }

template <int kShape>
__global__ void __launch_bounds__(320, 1) probe(uint32_t reader_warps, uint32_t iters, uint32_t with_mma,
                                                unsigned long long* out_clk, uint32_t* sink) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint32_t tmem_ptr;
    __shared__ uint64_t bar;
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 8) ptx::tmem_alloc<1>(&tmem_ptr, 512);
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = tmem_ptr;
    unsigned long long t0 = 0, t1 = 0;
    uint32_t acc = 0;
    if (warp < reader_warps) {
        const uint32_t lane_base = (warp & 3) * 32;
        // readers sweep columns 256..511 (the MMAs, when on, write 0..255); two warps on one quadrant take alternate strides
        const uint32_t span = 256, step = cols_per_ld<kShape>();
        const uint32_t taddr0 = tmem_base + (lane_base << 16) + 256;
        uint32_t col = (warp >> 2) * step % span;
        __syncwarp();
        t0 = clock64();
        for (uint32_t i = 0; i < iters; ++i) {
            acc ^= ld_once<kShape>(taddr0 + col);
            col += step * (reader_warps > 4 ? 2 : 1);
            if (col >= span) col -= span;
        }
        t1 = clock64();
    } else if (warp == 9 && with_mma) {
        if (lane == 0) {
            const uint32_t hi = static_cast<uint32_t>(make_smem_desc_hi(1024, kLayoutSw128) >> 32);
            const uint32_t a_lo = (smem_base & 0x3FFFF) >> 4, b_lo = ((smem_base + 16384) & 0x3FFFF) >> 4;
            const uint32_t idesc = make_idesc(128, 256, 0u, 0, 0);
            // keep the tensor pipe busy for at least as long as the readers run: iters * (cycles per ld at 64 B/clk) / 128 clk per MMA
            const uint32_t n_mma = iters * (bytes_per_ld<kShape>() * reader_warps / 64) / 128 * 2;
            for (uint32_t i = 0; i < n_mma; ++i) {
                const uint64_t a_desc = (static_cast<uint64_t>(hi) << 32) | (a_lo + (i & 3) * 2);
                const uint64_t b_desc = (static_cast<uint64_t>(hi) << 32) | (b_lo + (i & 3) * 2);
                ptx::mma_f8<1>(tmem_base, a_desc, b_desc, idesc, i != 0);
            }
            ptx::mma_commit<1>(&bar);
            ptx::mbar_wait(&bar, 0);
        }
        __syncwarp();
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp < reader_warps && lane == 0) out_clk[blockIdx.x * 8 + warp] = t1 - t0;
    if (acc == 0x12345678u) sink[0] = acc;
    if (warp == 8) ptx::tmem_dealloc<1>(tmem_base, 512);
}

template <int kShape>
static void run(const char* name, uint32_t reader_warps, uint32_t with_mma, unsigned long long* d_clk, uint32_t* d_sink) {
    const uint32_t iters = 4096, grid = 148;
    const int smem = 64 * 1024;
    cudaFuncSetAttribute(probe<kShape>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int rep = 0; rep < 2; ++rep) {
        cudaMemset(d_clk, 0, grid * 8 * 8);
        probe<kShape><<<grid, 320, smem>>>(reader_warps, iters, with_mma, d_clk, d_sink);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    }
    static unsigned long long h[148 * 8];
    cudaMemcpy(h, d_clk, sizeof(h), cudaMemcpyDeviceToHost);
    double worst = 0, sum = 0; int n = 0;
    for (uint32_t b = 0; b < grid; ++b)
        for (uint32_t w = 0; w < reader_warps; ++w) { double c = double(h[b * 8 + w]); worst = c > worst ? c : worst; sum += c; ++n; }
    const double bytes_sm = double(iters) * bytes_per_ld<kShape>() * reader_warps;
    printf("%-14s readers=%u mma=%u: %.1f B/clk/SM (slowest warp), %.1f (mean warp); %.0f clk per ld per warp\n", name,
           reader_warps, with_mma, bytes_sm / worst, bytes_sm / (sum / n), (sum / n) / iters);
}

int main() {
    unsigned long long* d_clk; uint32_t* d_sink;
    cudaMalloc(&d_clk, 148 * 8 * 8); cudaMalloc(&d_sink, 4);
    for (uint32_t mma = 0; mma < 2; ++mma)
        for (uint32_t readers : {4u, 8u}) {
            run<0>("32x32b.x32", readers, mma, d_clk, d_sink);
            run<1>("32x32b.x64", readers, mma, d_clk, d_sink);
            run<2>("4 x 32x32b.x32", readers, mma, d_clk, d_sink);
            run<3>("16x256b.x8", readers, mma, d_clk, d_sink);
        }
    return 0;
}
