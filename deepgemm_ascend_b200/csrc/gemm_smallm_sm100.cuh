// Small-M ("decode") GEMM for sm_100a:  C[M,N] = X[M,K] * W[N,K]^T with M <= 256.
//
// HBM-bound: the job is to stream W once at full bandwidth from every SM.
// Swap-AB: the 128 weight rows of a tile sit on the UMMA M axis (TMEM lanes),
// the (padded) M activations on the UMMA N axis (TMEM columns), so no tensor
// work is wasted on padding M up to 128.  K is split across the CTAs of a
// thread-block cluster.  Each CTA owns a block of activation rows: every CTA
// stages its fp32 partial tile in its own shared memory (over the drained ring)
// and pushes each owner's rows there with one bulk DSMEM copy that completes on
// the owner's mbarrier; the owner adds its `split` slots in a fixed order
// (deterministic) and stores.  No workspace, no atomics, no second kernel, and
// no cluster-wide barrier between the last MMA and the stores.
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
    uint32_t issue_batch; // K blocks whose loads the producer issues back to back (<= num_stages)
    unsigned long long* trace; // debug: per CTA 12 words (see tools/trace_smallm.py); nullptr = off
    uint32_t prefetch_kb; // K blocks of this CTA's weight slice pulled into L2 before the grid-dependency wait
    void* out;
    // Workspace mode (DGA_KERNEL_DECODE_WS): no cluster.  CTA (tile, slice) = (blockIdx / split, blockIdx % split) leaves
    // its fp32 partial in slab [slice] of a [split][m][ws_pitch] workspace (coalesced: lanes = consecutive weight rows)
    // and EXITS; splitk_reduce_kernel adds the slabs.  Nothing of the launch stays on an SM through a reduction
    // tail, so the CTAs of the next launch on the stream move in as ours retire and start on their weights at once.
    float* partials;
    uint32_t ws_pitch;    // floats per workspace row (>= n)
    // Pair launch (gemm_smallm_pair_sm100_kernel): two GEMMs that share X.  Tiles [0, tiles_first) belong to the first
    // weight (n rows, output out / ldc), the rest to the second (n2 rows, output out2 / ldc2).
    uint32_t tiles_first;
    uint32_t n2;
    uint32_t ldc2;
    void* out2;
};

struct SmallMSmem {
    uint32_t w_bytes, x_bytes, stage_bytes, off_red, off_bar, total;
};
// ring (reused as the staging tile [m][128] fp32 once drained) | reduction slots red[src CTA][owned row][128] fp32
// (split > 1 only; written by the peers while this CTA may still be streaming, hence not aliased) | barriers
__host__ __device__ inline uint32_t smallm_red_bytes(uint32_t m_pad, uint32_t split) {
    return split > 1 ? split * ((m_pad + split - 1) / split) * 512u : 0u;
}
__host__ __device__ inline SmallMSmem make_smallm_smem(uint32_t m_pad, uint32_t stages, uint32_t split) {
    SmallMSmem l;
    l.w_bytes = 128 * 128;
    l.x_bytes = m_pad * 128;
    l.stage_bytes = l.w_bytes + l.x_bytes;
    l.off_red = stages * l.stage_bytes;
    l.off_bar = l.off_red + smallm_red_bytes(m_pad, split);
    l.total = l.off_bar + (2 * stages + 4) * 8 + 16;
    return l;
}

template <typename OutT>
__device__ __forceinline__ void store_out(OutT* p, float v) {
    if constexpr (sizeof(OutT) == 4) *p = v;
    else if constexpr (std::is_same<OutT, __half>::value) *p = __float2half_rn(v);
    else *p = __float2bfloat16_rn(v);
}

template <typename OutT, bool kPair>
__device__ __forceinline__ void smallm_body(const CUtensorMap& tm_w_first, const CUtensorMap& tm_w_second,
                                            const CUtensorMap& tm_x, const SmallMParams& p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
    ptx::grid_dep_launch();
    unsigned long long* tr = p.trace != nullptr ? p.trace + blockIdx.x * 12 : nullptr;
    if (tr != nullptr && threadIdx.x == 0) { tr[0] = ptx::globaltimer(); tr[2] = clock64(); }

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t split = p.split;
    const bool ws_mode = p.partials != nullptr;
    const uint32_t tile = blockIdx.x / split;
    const uint32_t rank = ws_mode ? blockIdx.x - tile * split : (split > 1 ? ptx::cluster_ctarank() : 0u);
    // which of the two GEMMs of a pair launch this cluster works for (uniform over the cluster)
    const bool second = kPair && tile >= p.tiles_first;
    const CUtensorMap& tm_w = second ? tm_w_second : tm_w_first;
    const uint32_t n0 = (second ? tile - p.tiles_first : tile) * 128;
    const uint32_t n_rows = second ? p.n2 : p.n;       // weight rows = output columns of this GEMM
    const uint32_t ldc = second ? p.ldc2 : p.ldc;
    const uint32_t stages = p.num_stages;
    // this CTA's K blocks: [kb0, kb1)
    const uint32_t kb0 = static_cast<uint32_t>((static_cast<uint64_t>(rank) * p.k_blocks) / split);
    const uint32_t kb1 = static_cast<uint32_t>((static_cast<uint64_t>(rank + 1) * p.k_blocks) / split);

    const SmallMSmem L = make_smallm_smem(p.m_pad, stages, ws_mode ? 1u : split);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.off_bar);
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* tmem_full_bar = empty_bar + stages;
    uint64_t* red_full_bar = tmem_full_bar + 1;   // completes when every peer's partial of this CTA's rows has landed
    uint64_t* ack_bar = tmem_full_bar + 2;        // one arrival per owner that has received this CTA's partial
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 4);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tm_w);
        ptx::prefetch_tensormap(&tm_x);
        // Back-to-back launches (decode: one GEMM after another): this grid becomes resident while its
        // predecessor is still draining, and HBM would sit idle through that tail and through our own first
        // round trip.  Pull the head of our weight slice into L2 now; the real loads below hit it there.
        // (workspace mode loads the head of the slice into the ring itself, below, and prefetches what follows it)
        const uint32_t skip = ws_mode ? min(stages, kb1 - kb0) : 0u;
        const uint32_t pf = min(p.prefetch_kb, kb1 - kb0 - skip);
        for (uint32_t i = 0; i < pf; ++i)
            ptx::tma_prefetch_2d(&tm_w, static_cast<int32_t>((kb0 + skip + i) * 64), static_cast<int32_t>(n0));
    }
    if (warp == 1 && lane == 0) {
        for (uint32_t s = 0; s < stages; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        ptx::mbar_init(tmem_full_bar, 1);
        ptx::mbar_init(red_full_bar, 1);
        // owners = CTAs that hold at least one activation row
        ptx::mbar_init(ack_bar, min(split, (p.m + (p.m + split - 1) / split - 1) / ((p.m + split - 1) / split)));
        ptx::fence_mbar_init();
    }
    if (warp == 2) ptx::tmem_alloc<1>(tmem_ptr_smem, p.tmem_cols);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    // Peers will write into this CTA's reduction slots and signal its mbarrier: they may only do so once it is known
    // to be running with the barrier initialised (fence_mbar_init above).  Arrive now, wait (phase 1 of the cluster
    // barrier) just before the first remote access, a whole K slice later.
    if (split > 1 && !ws_mode) ptx::cluster_arrive_relaxed();
    if (tr != nullptr && threadIdx.x == 0) tr[3] = clock64();
    // Workspace mode: the weights never depend on the preceding grid, so the first ring-full of weight boxes is
    // requested BEFORE the dependency wait (real loads into shared memory, not an L2 hint); only the activations,
    // which the predecessor may have produced, and our own stores come after it.  The bytes may complete on the
    // full barriers before their expect_tx is posted (transiently negative transaction count).
    const uint32_t preloaded = ws_mode ? min(stages, kb1 - kb0) : 0u;
    if (warp == 0 && lane == 0) {
        for (uint32_t i = 0; i < preloaded; ++i)
            ptx::tma_load_2d(&tm_w, &full_bar[i], smem + i * L.stage_bytes, static_cast<int32_t>((kb0 + i) * 64),
                             static_cast<int32_t>(n0), ptx::kEvictFirst);
    }
    ptx::grid_dep_wait();   // the predecessor on the stream may still have been running up to here
    if (tr != nullptr && threadIdx.x == 0) { tr[4] = clock64(); tr[11] = ptx::globaltimer(); }

    if (warp == 0) {
        if (lane == 0) {
            // Loads go out `issue_batch` K blocks at a time: the 128-byte pieces of one weight row that belong to
            // consecutive K blocks then reach the memory controller together (longer runs per DRAM page).
            uint32_t stage = 0, phase = 0;
            const uint32_t batch = p.issue_batch;
            for (uint32_t kb = kb0; kb < kb1;) {
                const uint32_t nb = min(batch, kb1 - kb);
                uint32_t s2 = stage, ph2 = phase;
                for (uint32_t j = 0; j < nb; ++j) {
                    ptx::mbar_wait(&empty_bar[s2], ph2 ^ 1);
                    if (++s2 == stages) { s2 = 0; ph2 ^= 1; }
                }
                for (uint32_t j = 0; j < nb; ++j, ++kb) {
                    uint8_t* sw = smem + stage * L.stage_bytes;
                    // W is read exactly once: do not let it push X out of L2
                    if (kb - kb0 >= preloaded)
                        ptx::tma_load_2d(&tm_w, &full_bar[stage], sw, static_cast<int32_t>(kb * 64),
                                         static_cast<int32_t>(n0), ptx::kEvictFirst);
                    ptx::tma_load_2d(&tm_x, &full_bar[stage], sw + L.w_bytes, static_cast<int32_t>(kb * 64), 0,
                                     ptx::kEvictLast);
                    ptx::mbar_arrive_expect_tx(&full_bar[stage], L.stage_bytes);
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
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
                if (tr != nullptr && kb == kb0) tr[5] = clock64();
                const uint64_t w_desc = hi | (lo0 + stage * (L.stage_bytes >> 4));
                const uint64_t x_desc = w_desc + (L.w_bytes >> 4);
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k)
                    ptx::mma_f16<1>(tmem_base, w_desc + k * 2, x_desc + k * 2, p.idesc, (kb != kb0) || (k != 0));
                ptx::mma_commit<1>(&empty_bar[stage]);
                if (kb + 1 == kb1) ptx::mma_commit<1>(tmem_full_bar);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
            if (tr != nullptr) tr[6] = clock64();
        }
        __syncwarp();
    }

    OutT* out = reinterpret_cast<OutT*>(second ? p.out2 : p.out);
    // Ownership of the activation rows inside a cluster: CTA d reduces rows [d * rows_per, (d + 1) * rows_per) ∩ [0, m)
    const uint32_t rows_per = (p.m + split - 1) / split;
    auto rows_of = [&](uint32_t d) { return d * rows_per >= p.m ? 0u : min(rows_per, p.m - d * rows_per); };
    const uint32_t my_rows = rows_of(rank);
    if (split > 1 && !ws_mode) {
        ptx::cluster_wait();   // every CTA of the cluster is running (see cluster_arrive_relaxed above)
        if (threadIdx.x == 0 && my_rows != 0) ptx::mbar_arrive_expect_tx(red_full_bar, split * my_rows * 512);
    }
    // Partial tile -> registers -> staging tile -> the owners' reduction slots (or straight to C when K is not split).
    if (warp >= 4) {
        const uint32_t row = threadIdx.x - 128;  // weight row inside the tile == TMEM lane
        const uint32_t lane_base = (warp & 3) * 32;
        ptx::mbar_wait(tmem_full_bar, 0);
        ptx::tc_fence_after();
        if (tr != nullptr && threadIdx.x == 128) tr[7] = clock64();
        const uint32_t taddr = tmem_base + (lane_base << 16);
        const bool n_ok = n0 + row < n_rows;
        // tmem_full fired: every MMA, hence every read of the ring, has completed; the ring is the staging tile now
        float* stage_out = reinterpret_cast<float*>(smem);
        // up to 64 activation columns per round: the TMEM loads go out back to back and share one wait
        // (columns beyond m are padding: loaded at most up to the allocation, never used)
        for (uint32_t c = 0; c < p.m; c += 64) {
            uint32_t v[64];
            const uint32_t left = p.m - c;
            if (left <= 16) {
                ptx::tmem_ld_32x16(taddr + c, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
            } else {
                ptx::tmem_ld_32x32(taddr + c, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                if (left > 32) ptx::tmem_ld_32x32(taddr + c + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
            }
            ptx::tmem_ld_wait();
            if (ws_mode) {
                float* slab = p.partials + (static_cast<uint64_t>(rank) * p.m + c) * p.ws_pitch + n0 + row;
#pragma unroll
                for (uint32_t j = 0; j < 64; ++j)
                    if (n_ok && j < left) __stcg(slab + static_cast<uint64_t>(j) * p.ws_pitch, __uint_as_float(v[j]));
            } else if (split == 1) {
#pragma unroll
                for (uint32_t j = 0; j < 64; ++j)
                    if (n_ok && j < left)
                        store_out<OutT>(out + static_cast<uint64_t>(c + j) * ldc + n0 + row, __uint_as_float(v[j]));
            } else {
#pragma unroll
                for (uint32_t j = 0; j < 64; ++j)
                    if (j < left) stage_out[(c + j) * 128 + row] = __uint_as_float(v[j]);
            }
        }
        ptx::tc_fence_before();
        if (split > 1 && !ws_mode) {
            ptx::fence_async_smem();          // generic-proxy writes -> visible to the bulk-copy engine
            ptx::named_bar_sync(1, 128);
            if (row < split) {
                // thread d ships owner d's rows: stage_out[d * rows_per ...] -> red[rank][0 ...] of CTA d
                const uint32_t cnt = rows_of(row);
                if (cnt != 0)
                    ptx::dsmem_bulk_copy(smem_base + L.off_red + rank * rows_per * 512,
                                         smem_base + row * rows_per * 512, cnt * 512, red_full_bar, row);
            }
        }
    }

    if (split > 1 && !ws_mode) {
        if (my_rows != 0) {
            ptx::mbar_wait(red_full_bar, 0);
            if (tr != nullptr && threadIdx.x == 0) tr[8] = clock64();
            // tell every source that its staging tile has been read (it may leave); then reduce
            if (threadIdx.x < split) ptx::mbar_arrive_remote(ack_bar, threadIdx.x);
            // 8 rows x 32 four-column groups at a time; slots are added in CTA order (deterministic)
            const uint32_t cg = (threadIdx.x & 31) * 4;
            const float* red = reinterpret_cast<const float*>(smem + L.off_red);
            const bool vec_ok = (reinterpret_cast<uintptr_t>(out) % (4 * sizeof(OutT)) == 0) &&
                                (ldc % 4 == 0) && (n0 + 128 <= n_rows);
            for (uint32_t r = threadIdx.x >> 5; r < my_rows; r += 8) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                for (uint32_t s = 0; s < split; ++s) {
                    const float4 v = *reinterpret_cast<const float4*>(red + (s * rows_per + r) * 128 + cg);
                    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
                }
                OutT* dst = out + static_cast<uint64_t>(rank * rows_per + r) * ldc + n0 + cg;
                if (vec_ok) {
                    if constexpr (sizeof(OutT) == 4) {
                        *reinterpret_cast<float4*>(dst) = acc;
                    } else {
                        uint2 w;
                        w.x = OutPack<OutT>::pack(acc.x, acc.y);
                        w.y = OutPack<OutT>::pack(acc.z, acc.w);
                        *reinterpret_cast<uint2*>(dst) = w;
                    }
                } else {
                    const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
                    for (uint32_t j = 0; j < 4; ++j)
                        if (n0 + cg + j < n_rows) store_out<OutT>(dst + j, a4[j]);
                }
            }
        }
        if (tr != nullptr && threadIdx.x == 0) tr[9] = clock64();
        // A CTA's outgoing copies read its staging tile until the owners have seen them complete
        ptx::mbar_wait(ack_bar, 0);
        __syncthreads();
    } else {
        __syncthreads();
    }
    if (warp == 2) ptx::tmem_dealloc<1>(tmem_base, p.tmem_cols);
    if (tr != nullptr && threadIdx.x == 0) { tr[10] = clock64(); tr[1] = ptx::globaltimer(); }
}

template <typename OutT>
__global__ void __launch_bounds__(kNumThreads, 2)
gemm_smallm_sm100_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_x,
                         const SmallMParams p) {
    smallm_body<OutT, false>(tm_w, tm_w, tm_x, p);
}

// Two decode GEMMs over the same activations (gate / up, q / kv held as separate weights) in ONE launch: the fixed
// cost of a launch (first round trip, reduction tail) is paid once for both weights' bytes.
template <typename OutT>
__global__ void __launch_bounds__(kNumThreads, 2)
gemm_smallm_pair_sm100_kernel(const __grid_constant__ CUtensorMap tm_w_first,
                              const __grid_constant__ CUtensorMap tm_w_second,
                              const __grid_constant__ CUtensorMap tm_x, const SmallMParams p) {
    smallm_body<OutT, true>(tm_w_first, tm_w_second, tm_x, p);
}

}  // namespace dga
