// fp8 e4m3 block-scaled GEMM for sm_100a, DeepGEMM semantics:
//   D[m,n] (bf16) = sum_kb  sfa[m,kb] * sfb[n/128,kb] * ( sum_{k in block kb} A[m,k] * B[n,k] )
// A [M,K] e4m3 with one fp32 scale per 1x128 strip, B [N,K] e4m3 with one fp32 scale per 128x128 block.
// Not in the reference (BASELINE.json north_star item 1).  The scales are arbitrary fp32 values, so they
// cannot ride in the instruction (kind::mxf8f6f4 block scaling wants UE8M0): every 128-wide K block is a
// separate tcgen05 accumulation (kind::f8f6f4, fp32) into TMEM, and the promotion warps pull it out
// (tcgen05.ld), multiply by sfa*sfb and add it to fp32 accumulators held in REGISTERS.
//
//   warp 0        : TMA producer          (A/B tiles -> smem ring, SWIZZLE_128B, 128 fp8 = 128 B per row)
//   warp 1        : MMA issuer            (per K block and column half: 4 x K=32 MMAs, commit -> that half's barrier)
//   warp 2        : TMEM alloc / dealloc
//   warps 4..11   : promotion + epilogue  (2 warp groups; group g owns the column half g of the tile)
//
// What paces it (measured, profiles/r02_tmem_ld_throughput.log and r02_fp8_promo_*): tcgen05.ld delivers 350-630
// B/clk/SM (4-8 reader warps, MMAs running), so reading every accumulator once per K block (512 N bytes per
// CTA) costs ~0.9 N clocks against 2 N clocks of fp8 MMAs - the TMEM read is NOT the ceiling round 1 took it for.
// The chain MMA done -> barrier -> load -> release -> barrier -> next MMA into that buffer is: with two whole-tile
// buffers it is ~600 clocks against 448 (N = 224) of MMAs.  So the tile is computed in two column HALVES with
// their own TMEM buffers and barriers: four half buffers in flight, a half may be promoted while the tensor core
// works through the three that follow it.  To make a half a contiguous run of output columns, B is staged per
// CTA as two boxes of N/4 rows (columns [h N/2 + rank N/4, +N/4) for h = 0, 1): the half-h MMA (UMMA N = N/2)
// then reads box h of both CTAs of the pair.
#pragma once
#include "gemm_sm100.cuh"

namespace dga {

constexpr uint32_t kFp8Threads = 384;
constexpr uint32_t kFp8BlockK = 128;   // elements == bytes == the scale granularity

struct Fp8Params {
    uint32_t m, n, k;
    uint32_t block_n;
    uint32_t num_stages;
    uint32_t k_blocks;
    uint32_t num_m_blocks, num_n_blocks;
    uint32_t raster_group, raster_along_m;
    uint32_t idesc;            // UMMA N = block_n / 2 (one column half per instruction)
    uint32_t sfa_ld, sfb_ld;   // row pitch of the scale tensors, in floats
    uint32_t sfb_rows;         // ceil(n / 128)
    const float* sfa;          // [m, ceil(k/128)]
    const float* sfb;          // [ceil(n/128), ceil(k/128)]
    const int32_t* gate;       // run only when *gate == gate_expect (nullptr: always)
    int32_t gate_expect;
    uint64_t hint_a, hint_b;   // L2 eviction hints of the operand loads (0 = normal), see GemmParams
};

// Bottleneck isolation, compile time only (tools/fp8_isolate.sh builds side libraries with -DDGA_FP8_ISOLATE=<flags>;
// results are wrong with any of them): 32 the promotion warps only shake hands (no TMEM loads, no FMAs), 64 the MMA
// warp does not wait for the promotion (nor the promotion for it), 128 TMEM loads but no FMAs, 256 no MMA instructions
// (commits only).  A run-time flag in these loops costs 25 %.
#ifndef DGA_FP8_ISOLATE
#define DGA_FP8_ISOLATE 0
#endif
constexpr uint32_t kFp8Isolate = DGA_FP8_ISOLATE;

struct Fp8Smem {
    uint32_t d_stage_bytes, a_stage_bytes, b_stage_bytes, sfb_pitch, off_d, off_a, off_b, off_sfb, off_bar, total;
};
// k_blocks: the weight scales of a tile (its columns touch at most three 128-wide scale rows) sit in shared memory,
// two tiles deep: [2][3][sfb_pitch] floats
__host__ __device__ inline Fp8Smem make_fp8_smem(uint32_t block_n, uint32_t cta, uint32_t stages, uint32_t k_blocks) {
    Fp8Smem l;
    l.d_stage_bytes = kBlockM * kEpiCols * 2;   // bf16 staging chunk, 2 per warp group
    l.a_stage_bytes = kBlockM * 128;
    l.b_stage_bytes = (block_n / cta) * 128;
    l.sfb_pitch = (k_blocks + 3u) & ~3u;
    l.off_d = 0;
    l.off_a = 4 * l.d_stage_bytes;
    l.off_b = l.off_a + stages * l.a_stage_bytes;
    l.off_sfb = l.off_b + stages * l.b_stage_bytes;
    l.off_bar = l.off_sfb + 2 * 3 * l.sfb_pitch * 4;
    // full[stages] empty[stages] part_full[<= 4 buffers][2 halves] part_empty[<= 4][2] + tmem ptr
    l.total = l.off_bar + (2 * stages + 16) * 8 + 16;
    return l;
}

// Scales of four consecutive K blocks of one scale row: one 16-byte load when the row allows it.  Issued a whole
// group (4 K blocks of MMAs, ~2000 clocks) before first use, so the L2 round trip never sits in the promotion
// chain.  (Round 1 fetched one float per K block two blocks ahead and rotated it through registers: the rotation
// waited on the youngest load, every K block paid an L2 latency, and the kernel ran at 2500 clocks per K block
// against 450 of MMAs.)
__device__ __forceinline__ float4 load_scales4(const float* row, uint32_t kb, uint32_t k_blocks, bool vec) {
    if (vec && kb + 4 <= k_blocks) return __ldg(reinterpret_cast<const float4*>(row + kb));
    float4 v;
    v.x = __ldg(row + min(kb, k_blocks - 1));
    v.y = __ldg(row + min(kb + 1, k_blocks - 1));
    v.z = __ldg(row + min(kb + 2, k_blocks - 1));
    v.w = __ldg(row + min(kb + 3, k_blocks - 1));
    return v;
}
__device__ __forceinline__ float pick4(const float4& v, uint32_t i) {
    return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w;
}

__device__ __forceinline__ void stage_row_bf16(uint32_t stage_base, uint32_t row, const float* v) {
    const uint32_t row_base = stage_base + row * 64;
    const uint32_t x = (row >> 1) & 3;
#pragma unroll
    for (uint32_t j = 0; j < 4; ++j) {
        const uint32_t a = OutPack<__nv_bfloat16>::pack(v[8 * j + 0], v[8 * j + 1]);
        const uint32_t b = OutPack<__nv_bfloat16>::pack(v[8 * j + 2], v[8 * j + 3]);
        const uint32_t c = OutPack<__nv_bfloat16>::pack(v[8 * j + 4], v[8 * j + 5]);
        const uint32_t d = OutPack<__nv_bfloat16>::pack(v[8 * j + 6], v[8 * j + 7]);
        ptx::st_shared_v4(row_base + ((j ^ x) << 4), a, b, c, d);
    }
}

template <int kCta, uint32_t kBN>
__global__ void __launch_bounds__(kFp8Threads, 1)
gemm_fp8_sm100_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                      const __grid_constant__ CUtensorMap tm_d, const Fp8Params p) {
    constexpr uint32_t kHalf = kBN / 2;              // columns per warp group, per MMA and per TMEM half buffer
    constexpr uint32_t kBoxN = kBN / (2 * kCta);     // B rows per TMA box (two boxes per CTA and stage)
    constexpr uint32_t kUnits = kHalf / 16;          // 16-column units of a half (a unit never straddles a scale block)
    constexpr uint32_t kLoads = (kHalf + 31) / 32;   // TMEM loads per half: 32 columns each, the last one 16 when odd
    static_assert(kBN <= 256 && kHalf % 16 == 0 && kBoxN % 8 == 0, "unsupported block_n");
    ptx::grid_dep_launch();
    if (p.gate != nullptr) {
        ptx::grid_dep_wait();     // the flag was written by a previous grid
        if (*p.gate != p.gate_expect) return;   // grid-uniform
    }
    // TMEM buffers of kBN columns: as many as the 512 columns hold (2 for 256..192, 3 for 160, 4 for 128).  Each is two
    // half buffers with their own barriers; the more are in flight, the more of the MMA <-> promotion handshake
    // (~1500 clocks from "MMAs of a half done" to "next MMAs into that half issued") hides behind other halves' MMAs.
    constexpr uint32_t kNumBuf = (512 / kBN) > 4 ? 4 : (512 / kBN);
    constexpr uint32_t kTmemCols = 512;

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (smem_base - ptx::smem_u32(smem_raw));

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t cta_rank = (kCta == 2) ? ptx::cluster_ctarank() : 0u;
    const bool is_leader = cta_rank == 0;
    const uint32_t num_clusters = gridDim.x / kCta;
    const uint32_t cluster_id = blockIdx.x / kCta;
    const uint32_t num_tiles = p.num_m_blocks * p.num_n_blocks;
    const uint32_t stages = p.num_stages;
    const uint32_t k_blocks = p.k_blocks;

    const Fp8Smem L = make_fp8_smem(kBN, kCta, stages, k_blocks);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.off_bar);
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* part_full_bar = empty_bar + stages;    // [buffer * 2 + half] partial product of one K block is in TMEM
    uint64_t* part_empty_bar = part_full_bar + 8;    // [buffer * 2 + half] that half's promotion warps have read it
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(part_empty_bar + 8);

    if constexpr (kCta == 2) ptx::cluster_sync();

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tm_a);
        ptx::prefetch_tensormap(&tm_b);
        ptx::prefetch_tensormap(&tm_d);
    }
    if (warp == 1 && lane == 0) {
        for (uint32_t s = 0; s < stages; ++s) {
            ptx::mbar_init(&full_bar[s], kCta);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (uint32_t s = 0; s < 2 * kNumBuf; ++s) {
            ptx::mbar_init(&part_full_bar[s], 1);
            ptx::mbar_init(&part_empty_bar[s], kCta * 4);   // one elected arrive per warp of the owning group, both CTAs
        }
        ptx::fence_mbar_init();
    }
    if (warp == 2) ptx::tmem_alloc<kCta>(tmem_ptr_smem, kTmemCols);
    ptx::tc_fence_before();
    if constexpr (kCta == 2) ptx::cluster_sync(); else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    ptx::grid_dep_wait();

    if (warp < 4) {
        ptx::setmaxnreg_dec<56>();
        if (warp == 0) {
            // ------------------------------------------------ TMA producer
            const uint64_t hint_a = p.hint_a != 0 ? p.hint_a : ptx::kEvictNormal;
            const uint64_t hint_b = p.hint_b != 0 ? p.hint_b : ptx::kEvictNormal;
            uint32_t stage = 0, phase = 0;
            for (uint32_t t = cluster_id; t < num_tiles; t += num_clusters) {
                uint32_t mb, nb;
                tile_coords(p, t, mb, nb);
                const int32_t m_idx = static_cast<int32_t>(mb * (kBlockM * kCta) + cta_rank * kBlockM);
                const int32_t n_idx = static_cast<int32_t>(nb * kBN + cta_rank * kBoxN);   // box of half 0; half 1: + kHalf
                for (uint32_t kb = 0; kb < k_blocks; ++kb) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (ptx::elect_one()) {
                        uint8_t* sa = smem + L.off_a + stage * L.a_stage_bytes;
                        uint8_t* sb = smem + L.off_b + stage * L.b_stage_bytes;
                        const int32_t k_idx = static_cast<int32_t>(kb * kFp8BlockK);
                        if constexpr (kCta == 2) {
                            ptx::tma_load_2d_pair(&tm_a, &full_bar[stage], sa, k_idx, m_idx, hint_a);
                            ptx::tma_load_2d_pair(&tm_b, &full_bar[stage], sb, k_idx, n_idx, hint_b);
                            ptx::tma_load_2d_pair(&tm_b, &full_bar[stage], sb + kBoxN * 128, k_idx,
                                                  n_idx + static_cast<int32_t>(kHalf), hint_b);
                            if (is_leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * (L.a_stage_bytes + L.b_stage_bytes));
                            else ptx::mbar_arrive_remote(&full_bar[stage], 0);
                        } else {
                            ptx::tma_load_2d(&tm_a, &full_bar[stage], sa, k_idx, m_idx, hint_a);
                            ptx::tma_load_2d(&tm_b, &full_bar[stage], sb, k_idx, n_idx, hint_b);
                            ptx::tma_load_2d(&tm_b, &full_bar[stage], sb + kBoxN * 128, k_idx,
                                             n_idx + static_cast<int32_t>(kHalf), hint_b);
                            ptx::mbar_arrive_expect_tx(&full_bar[stage], L.a_stage_bytes + L.b_stage_bytes);
                        }
                    }
                    __syncwarp();
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
        } else if (warp == 1 && is_leader) {
            // -------------------------------------------------- MMA issuer
            const uint32_t hi = static_cast<uint32_t>(make_smem_desc_hi(1024, kLayoutSw128) >> 32);
            const uint32_t a_lo0 = ((smem_base + L.off_a) & 0x3FFFF) >> 4;
            const uint32_t b_lo0 = ((smem_base + L.off_b) & 0x3FFFF) >> 4;
            const uint32_t a_step = L.a_stage_bytes >> 4, b_step = L.b_stage_bytes >> 4;
            const uint32_t idesc = p.idesc;
            uint32_t stage = 0, phase = 0, pc = 0;   // pc: running K-block counter (selects the TMEM buffer)
            for (uint32_t t = cluster_id; t < num_tiles; t += num_clusters) {
                for (uint32_t kb = 0; kb < k_blocks; ++kb, ++pc) {
                    const uint32_t buf = pc % kNumBuf, par = (pc / kNumBuf) & 1;
                    ptx::mbar_wait(&full_bar[stage], phase);
#pragma unroll
                    for (uint32_t h = 0; h < 2; ++h) {
                        if constexpr (!(kFp8Isolate & 64)) ptx::mbar_wait(&part_empty_bar[buf * 2 + h], par ^ 1);
                        ptx::tc_fence_after();
                        if (ptx::elect_one()) {
                            const uint32_t a_lo = a_lo0 + stage * a_step;
                            const uint32_t b_lo = b_lo0 + stage * b_step + h * ((kBoxN * 128) >> 4);
                            const uint32_t tmem_d = tmem_base + buf * kBN + h * kHalf;
#pragma unroll
                            for (uint32_t k = 0; k < 4; ++k) {   // 4 x (K = 32 fp8 = 32 bytes) inside the 128 B atom
                                const uint64_t a_desc = (static_cast<uint64_t>(hi) << 32) | (a_lo + k * 2);
                                const uint64_t b_desc = (static_cast<uint64_t>(hi) << 32) | (b_lo + k * 2);
                                if constexpr (!(kFp8Isolate & 256)) ptx::mma_f8<kCta>(tmem_d, a_desc, b_desc, idesc, k != 0);
                            }
                            ptx::mma_commit<kCta>(&part_full_bar[buf * 2 + h]);
                            if (h == 1) ptx::mma_commit<kCta>(&empty_bar[stage]);
                        }
                        __syncwarp();
                    }
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else {
        // ------------------------------------------------ promotion + epilogue (warps 4..11)
        ptx::setmaxnreg_inc<224>();
        const uint32_t group = (warp - 4) >> 2;                  // warp group == column half it owns
        const uint32_t quad = warp & 3;                          // TMEM lane quadrant this warp may read
        const uint32_t row = quad * 32 + lane;                   // accumulator row inside the CTA tile
        const uint32_t gtid = (threadIdx.x - 128) & 127;         // thread index inside the warp group (== row)
        const uint32_t bar_id = 1 + group;
        const uint32_t sbase0 = smem_base + L.off_d + group * 2 * L.d_stage_bytes;
        uint8_t* sptr0 = smem + L.off_d + group * 2 * L.d_stage_bytes;
        uint32_t pc = 0, sbuf = 0, tile_par = 0;
        float acc[kHalf];
        // Weight scales of a tile: 3 rows x k_blocks floats, copied to shared memory by the 256 promotion threads;
        // the copy for tile i + 1 is issued before the stores of tile i and lands behind them.
        const uint32_t ptid = threadIdx.x - 128;
        float* sfb_s = reinterpret_cast<float*>(smem + L.off_sfb);
        const uint32_t sfb_total = 3 * k_blocks;
        constexpr uint32_t kSfbPer = 6;   // elements per thread and pass
        auto sfb_src = [&](uint32_t tile, uint32_t idx) {
            uint32_t mb2, nb2;
            tile_coords(p, tile, mb2, nb2);
            const uint32_t r = idx / k_blocks, kb = idx - r * k_blocks;
            return p.sfb + static_cast<uint64_t>(min(((nb2 * kBN) >> 7) + r, p.sfb_rows - 1)) * p.sfb_ld + kb;
        };
        auto sfb_load = [&](uint32_t tile, uint32_t base, float (&v)[kSfbPer]) {
#pragma unroll
            for (uint32_t i = 0; i < kSfbPer; ++i) {
                const uint32_t idx = base + ptid + i * 256;
                if (idx < sfb_total) v[i] = __ldg(sfb_src(tile, idx));
            }
        };
        auto sfb_store = [&](uint32_t par, uint32_t base, const float (&v)[kSfbPer]) {
#pragma unroll
            for (uint32_t i = 0; i < kSfbPer; ++i) {
                const uint32_t idx = base + ptid + i * 256;
                if (idx < sfb_total) {
                    const uint32_t r = idx / k_blocks, kb = idx - r * k_blocks;
                    sfb_s[(par * 3 + r) * L.sfb_pitch + kb] = v[i];
                }
            }
        };
        if (cluster_id < num_tiles) {
            for (uint32_t base = 0; base < sfb_total; base += 256 * kSfbPer) {
                float v[kSfbPer];
                sfb_load(cluster_id, base, v);
                sfb_store(0, base, v);
            }
            ptx::named_bar_sync(3, 256);
        }

        // loop invariants of the promotion chain: barrier addresses (the arrival goes to the leader CTA's barrier) and the
        // TMEM address of this group's half, per buffer by plain arithmetic (a dynamically indexed array would go to
        // local memory)
        const uint32_t pf_addr0 = ptx::smem_u32(&part_full_bar[group]);
        uint32_t pe_addr0 = ptx::smem_u32(&part_empty_bar[group]);
        if constexpr (kCta == 2) pe_addr0 = ptx::mapa(pe_addr0, 0);
        const uint32_t taddr0 = tmem_base + ((quad * 32) << 16) + group * kHalf;

        for (uint32_t t = cluster_id; t < num_tiles; t += num_clusters, tile_par ^= 1) {
            uint32_t mb, nb;
            tile_coords(p, t, mb, nb);
            const uint32_t m0 = mb * (kBlockM * kCta) + cta_rank * kBlockM;
            const uint32_t n0 = nb * kBN;
            const uint32_t c0 = n0 + group * kHalf;              // first output column of this group's half
            const float* sfa_row = p.sfa + static_cast<uint64_t>(min(m0 + row, p.m - 1)) * p.sfa_ld;
            // A half is at most 128 columns wide and starts on a multiple of 16: it touches at most two 128-column
            // scale blocks, A (its first column's) and B (the next).  16-column unit u takes A's scale while
            // u < units_a, B's after; as a bit mask per unit so that the choice is one LOP3 per K block.
            const float* sfb_a = sfb_s + (tile_par * 3 + ((c0 >> 7) - (n0 >> 7))) * L.sfb_pitch;
            const float* sfb_b = sfb_a + ((((c0 + kHalf - 1) >> 7) != (c0 >> 7)) ? L.sfb_pitch : 0u);
            const uint32_t units_a = (128u - (c0 & 127u)) >> 4;
            uint32_t take_a[kUnits];
#pragma unroll
            for (uint32_t u = 0; u < kUnits; ++u) take_a[u] = u < units_a ? 0xFFFFFFFFu : 0u;
#pragma unroll
            for (uint32_t j = 0; j < kHalf; ++j) acc[j] = 0.f;

            // One K block: fold this group's half of the partial product into the running sums.  kBuf is a literal
            // in every call below (the loop is unrolled by four), sa the row's activation scale of this block.
            auto promote = [&](uint32_t kb, uint32_t buf, uint32_t par, float sa) __attribute__((always_inline)) {
                const uint32_t pf_addr = pf_addr0 + buf * 16, pe_addr = pe_addr0 + buf * 16, taddr = taddr0 + buf * kBN;
                const float pa = sa * sfb_a[kb], pb = sa * sfb_b[kb];
                float s[kUnits];
#pragma unroll
                for (uint32_t u = 0; u < kUnits; ++u)
                    s[u] = __uint_as_float((take_a[u] & __float_as_uint(pa)) | (~take_a[u] & __float_as_uint(pb)));
                if constexpr (!(kFp8Isolate & 64)) ptx::mbar_wait_addr(pf_addr, par);   // (64: both sides run free)
                ptx::tc_fence_after();
                // tcgen05.wait::ld waits for EVERY outstanding load of the thread, so loads cannot be pipelined one behind
                // the other: each wait exposes a full TMEM round trip (~150 clocks with MMAs running).  Hence batches of
                // 64 columns (two loads back to back, one wait, 32 packed FMAs) instead of one 32-column load per wait:
                // half as many round trips per K block (profiles/r02_fp8_tile_sweep.log: the time per K block was
                // ~550 clocks + 1.5 per column whatever the tile width and however many TMEM buffers were in flight).
                constexpr uint32_t kBatches = (kHalf + 63) / 64;
#pragma unroll
                for (uint32_t b = 0; b < kBatches; ++b) {
                    const uint32_t c0 = 64 * b;                       // first column of the batch inside the half
                    uint32_t v[64];
                    if constexpr ((kFp8Isolate & 32) != 0) {
#pragma unroll
                        for (uint32_t j = 0; j < 64; ++j) v[j] = b;
                    } else {
                        if (c0 + 32 <= kHalf) ptx::tmem_ld_32x32(taddr + c0, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
                        else ptx::tmem_ld_32x16(taddr + c0, *reinterpret_cast<uint32_t(*)[16]>(&v[0]));
                        if (c0 + 64 <= kHalf) ptx::tmem_ld_32x32(taddr + c0 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
                        else if (c0 + 48 <= kHalf) ptx::tmem_ld_32x16(taddr + c0 + 32, *reinterpret_cast<uint32_t(*)[16]>(&v[32]));
                        ptx::tmem_ld_wait();
                    }
                    if (b + 1 == kBatches) {
                        // the whole half is in registers: hand its TMEM buffer back to the tensor core
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (ptx::elect_one()) {
                            if constexpr (kCta == 2) ptx::mbar_arrive_cluster_addr(pe_addr);
                            else ptx::mbar_arrive_addr(pe_addr);
                        }
                    }
                    if constexpr ((kFp8Isolate & (32 | 128)) != 0) {
                        acc[c0] += __uint_as_float(v[0] ^ v[17] ^ v[34] ^ v[63]) * s[0];   // keeps the loads alive
                    } else {
#pragma unroll
                        for (uint32_t j = 0; j < 64; j += 2) {
                            if (c0 + j < kHalf)
                                ptx::fma2(acc[c0 + j], acc[c0 + j + 1], __uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                          s[(c0 + j) / 16]);
                        }
                    }
                }
            };

            // this row's activation scales of K blocks [4g, 4g + 4) are in registers while group g runs; group g + 1 is
            // in flight.  pc counts K blocks over the whole kernel: buffer = pc % kNumBuf, barrier parity = (pc / kNumBuf) & 1.
            const bool vec = (reinterpret_cast<uintptr_t>(p.sfa) & 15) == 0 && (p.sfa_ld & 3) == 0;
            float4 sa4 = load_scales4(sfa_row, 0, k_blocks, vec);
            float4 sa4n = sa4;
            auto next_scales = [&](uint32_t kb) __attribute__((always_inline)) {   // at every fourth K block
                if (kb != 0) sa4 = sa4n;
                if (kb + 4 < k_blocks) sa4n = load_scales4(sfa_row, kb + 4, k_blocks, vec);
            };
            for (uint32_t kb = 0; kb < k_blocks; ++kb, ++pc) {
                const uint32_t ki = kb & 3;
                if (ki == 0) next_scales(kb);
                promote(kb, pc % kNumBuf, (pc / kNumBuf) & 1, pick4(sa4, ki));
            }

            // ---- tile done: registers -> bf16 -> swizzled smem -> TMA store (per warp group, 32-column boxes; when
            // the half is an odd multiple of 16 wide the last box is moved back by 16 columns and rewrites them)
            const uint32_t t_next = t + num_clusters;
            float sfb_next[kSfbPer];
            if (t_next < num_tiles) sfb_load(t_next, 0, sfb_next);
#pragma unroll
            for (uint32_t i = 0; i < kLoads; ++i) {
                const uint32_t col = (32 * (i + 1) <= kHalf) ? 32 * i : kHalf - 32;   // inside the half
                if (c0 + col < p.n) {
                    if (gtid == 0) ptx::tma_store_wait_read<1>();
                    ptx::named_bar_sync(bar_id, 128);
                    stage_row_bf16(sbase0 + sbuf * L.d_stage_bytes, row, &acc[col]);
                    ptx::fence_async_smem();
                    ptx::named_bar_sync(bar_id, 128);
                    if (gtid == 0) {
                        if (m0 < p.m)
                            ptx::tma_store_2d(&tm_d, sptr0 + sbuf * L.d_stage_bytes,
                                              static_cast<int32_t>(c0 + col), static_cast<int32_t>(m0));
                        ptx::tma_store_commit();
                    }
                    sbuf ^= 1;
                }
            }
            if (t_next < num_tiles) {
                // nobody reads buffer tile_par ^ 1 during this tile; the barrier below orders these stores before
                // the first read of the next tile
                sfb_store(tile_par ^ 1, 0, sfb_next);
                for (uint32_t base = 256 * kSfbPer; base < sfb_total; base += 256 * kSfbPer) {
                    float v[kSfbPer];
                    sfb_load(t_next, base, v);
                    sfb_store(tile_par ^ 1, base, v);
                }
                ptx::named_bar_sync(3, 256);
            }
        }
        if (gtid == 0) ptx::tma_store_wait_all<0>();
    }

    // ---------------------------------------------------------------- teardown
    ptx::tc_fence_before();
    if constexpr (kCta == 2) ptx::cluster_sync(); else __syncthreads();
    if (warp == 2) ptx::tmem_dealloc<kCta>(tmem_base, kTmemCols);
}

}  // namespace dga
