// Small-M ("decode") GEMM for sm_100a:  C[M,N] = X[M,K] * W[N,K]^T with M <= 256.
//
// HBM-bound: the job is to stream W once at full bandwidth from every SM.
// Swap-AB: the 128 weight rows of a tile sit on the UMMA M axis (TMEM lanes),
// the (padded) M activations on the UMMA N axis (TMEM columns), so no tensor
// work is wasted on padding M up to 128.  K is split across the CTAs of a
// thread-block cluster; their fp32 partials are reduced through distributed
// shared memory in a fixed order (deterministic), so no workspace, no atomics
// and no second kernel are needed.
// This is the B200 counterpart of what the reference reaches for with its
// stream-K variant and (declared, never dispatched) multicore split-K kernel
// type 3 (op_kernel/catlass_dynamic_matmul_tiling_key.h:117-130,
// kernel/padding_streamk_matmul_kernel.h:33-108) when tiles << cores.
#pragma once
#include <type_traits>

#include "gemm_sm100.cuh"

namespace dga {

struct SmallMParams {
    uint32_t m, n, k;
    uint32_t m_pad;      // UMMA N: m rounded up to 16
    uint32_t split;      // cluster size along K
    uint32_t num_stages;
    uint32_t k_blocks;   // total 64-element K blocks
    uint32_t ldc;        // elements
    uint32_t idesc;
    uint32_t tmem_cols;  // power of two >= max(32, m_pad)
    uint32_t prefetch_kb; // K blocks of this CTA's weight slice pulled into L2 before the grid-dependency wait
    void* out;
};

struct SmallMSmem {
    uint32_t w_bytes, x_bytes, stage_bytes, off_bar, total;
};
__host__ __device__ inline SmallMSmem make_smallm_smem(uint32_t m_pad, uint32_t stages) {
    SmallMSmem l;
    l.w_bytes = 128 * 128;
    l.x_bytes = m_pad * 128;
    l.stage_bytes = l.w_bytes + l.x_bytes;
    l.off_bar = stages * l.stage_bytes;
    l.total = l.off_bar + (2 * stages + 2) * 8 + 16;
    return l;
}

__device__ __forceinline__ float ld_dsmem_f32(uint32_t cluster_addr) {
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(cluster_addr));
    return v;
}

template <typename OutT>
__device__ __forceinline__ void store_out(OutT* p, float v) {
    if constexpr (sizeof(OutT) == 4) *p = v;
    else if constexpr (std::is_same<OutT, __half>::value) *p = __float2half_rn(v);
    else *p = __float2bfloat16_rn(v);
}

template <typename OutT>
__global__ void __launch_bounds__(kNumThreads, 2)
gemm_smallm_sm100_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
                         const SmallMParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
    ptx::grid_dep_launch();

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t split = p.split;
    const uint32_t rank = split > 1 ? ptx::cluster_ctarank() : 0u;
    const uint32_t tile = blockIdx.x / split;
    const uint32_t n0 = tile * 128;
    const uint32_t stages = p.num_stages;
    // this CTA's K blocks: [kb0, kb1)
    const uint32_t kb0 = static_cast<uint32_t>((static_cast<uint64_t>(rank) * p.k_blocks) / split);
    const uint32_t kb1 = static_cast<uint32_t>((static_cast<uint64_t>(rank + 1) * p.k_blocks) / split);

    const SmallMSmem L = make_smallm_smem(p.m_pad, stages);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.off_bar);
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* tmem_full_bar = empty_bar + stages;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 2);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tm_w);
        ptx::prefetch_tensormap(&tm_x);
        // Back-to-back launches (decode: one GEMM after another): this grid becomes resident while its
        // predecessor is still draining, and HBM would sit idle through that tail and through our own first
        // round trip.  Pull the head of our weight slice into L2 now; the real loads below hit it there.
        const uint32_t pf = min(p.prefetch_kb, kb1 - kb0);
        for (uint32_t i = 0; i < pf; ++i)
            ptx::tma_prefetch_2d(&tm_w, static_cast<int32_t>((kb0 + i) * 64), static_cast<int32_t>(n0));
    }
    if (warp == 1 && lane == 0) {
        for (uint32_t s = 0; s < stages; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        ptx::mbar_init(tmem_full_bar, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 2) ptx::tmem_alloc<1>(tmem_ptr_smem, p.tmem_cols);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    ptx::grid_dep_wait();   // the predecessor on the stream may still have been running up to here

    if (warp == 0) {
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (uint32_t kb = kb0; kb < kb1; ++kb) {
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* sw = smem + stage * L.stage_bytes;
                // W is read exactly once: do not let it push X out of L2
                ptx::tma_load_2d(&tm_w, &full_bar[stage], sw, static_cast<int32_t>(kb * 64),
                                 static_cast<int32_t>(n0), ptx::kEvictFirst);
                ptx::tma_load_2d(&tm_x, &full_bar[stage], sw + L.w_bytes, static_cast<int32_t>(kb * 64), 0,
                                 ptx::kEvictLast);
                ptx::mbar_arrive_expect_tx(&full_bar[stage], L.stage_bytes);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint64_t hi = make_smem_desc_hi(1024, kLayoutSw128);
            const uint32_t lo0 = (smem_base & 0x3FFFF) >> 4;
            uint32_t stage = 0, phase = 0;
            for (uint32_t kb = kb0; kb < kb1; ++kb) {
                ptx::mbar_wait(&full_bar[stage], phase);
                ptx::tc_fence_after();
                const uint64_t w_desc = hi | (lo0 + stage * (L.stage_bytes >> 4));
                const uint64_t x_desc = w_desc + (L.w_bytes >> 4);
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k)
                    ptx::mma_f16<1>(tmem_base, w_desc + k * 2, x_desc + k * 2, p.idesc, (kb != kb0) || (k != 0));
                ptx::mma_commit<1>(&empty_bar[stage]);
                if (kb + 1 == kb1) ptx::mma_commit<1>(tmem_full_bar);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    }

    OutT* out = reinterpret_cast<OutT*>(p.out);
    // Partial tile -> registers.  After tmem_full fires every MMA (hence every
    // smem read of the ring) has completed, so the ring is reused as the
    // reduction buffer red[m][128] (fp32).
    if (warp >= 4) {
        const uint32_t row = threadIdx.x - 128;  // weight row inside the tile == TMEM lane
        const uint32_t lane_base = (warp & 3) * 32;
        ptx::mbar_wait(tmem_full_bar, 0);
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_base + (lane_base << 16);
        const bool n_ok = n0 + row < p.n;
        for (uint32_t c = 0; c < p.m_pad; c += 16) {
            uint32_t v[16];
            ptx::tmem_ld_32x16(taddr + c, v);
            ptx::tmem_ld_wait();
            if (split == 1) {
#pragma unroll
                for (uint32_t j = 0; j < 16; ++j)
                    if (n_ok && c + j < p.m)
                        store_out<OutT>(out + static_cast<uint64_t>(c + j) * p.ldc + n0 + row, __uint_as_float(v[j]));
            } else {
                float* red = reinterpret_cast<float*>(smem);
#pragma unroll
                for (uint32_t j = 0; j < 16; ++j) red[(c + j) * 128 + row] = __uint_as_float(v[j]);
            }
        }
        ptx::tc_fence_before();
    }

    if (split > 1) {
        ptx::cluster_sync();  // all partials visible cluster-wide
        // CTA `rank` finalises activations rank, rank + split, ...; 256 threads = 2 rows of 128 columns
        const uint32_t col = threadIdx.x & 127;
        const uint32_t half = threadIdx.x >> 7;
        const uint32_t local = smem_base + col * 4;
        const bool n_ok = n0 + col < p.n;
        constexpr uint32_t kU = 4;
        for (uint32_t mi = rank + split * half; mi < p.m; mi += split * 2 * kU) {
            float acc[kU];
#pragma unroll
            for (uint32_t u = 0; u < kU; ++u) acc[u] = 0.f;
            for (uint32_t s = 0; s < split; ++s) {
#pragma unroll
                for (uint32_t u = 0; u < kU; ++u) {
                    const uint32_t mm = mi + u * split * 2;
                    if (mm < p.m) acc[u] += ld_dsmem_f32(ptx::mapa(local + mm * 512, s));
                }
            }
#pragma unroll
            for (uint32_t u = 0; u < kU; ++u) {
                const uint32_t mm = mi + u * split * 2;
                if (mm < p.m && n_ok) store_out<OutT>(out + static_cast<uint64_t>(mm) * p.ldc + n0 + col, acc[u]);
            }
        }
        ptx::cluster_sync();  // nobody leaves while a peer still reads its smem
    } else {
        __syncthreads();
    }
    if (warp == 2) ptx::tmem_dealloc<1>(tmem_base, p.tmem_cols);
}

}  // namespace dga
