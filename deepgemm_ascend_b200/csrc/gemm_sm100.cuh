// Persistent, warp-specialised dense GEMM for sm_100a:  D[M,N] = A[M,K] * B[N,K]^T
//
// Replaces the reference's CommonMatmulKernel / SmallMatmulKernel device path
// (aclnn_catlass_dynamic_matmul/op_kernel/kernel/common_matmul_kernel.h:28-78,
// small_matmul_kernel.h:28-74) and, with kBMn = true and OutT = float, the
// framework's row-major x row-major fp32-output MMAD kernel
// (deep_gemm_ascend/framework/csrc/jit/generate_code.hpp:123-369).
//
// Shape of the machine:
//   warp 0 lane 0 : TMA producer   (global -> smem ring, SWIZZLE_128B boxes)
//   warp 1 lane 0 : MMA issuer     (tcgen05.mma, fp32 accumulators in TMEM; leader CTA only)
//   warp 2        : TMEM allocator / deallocator
//   warps 4..7    : epilogue       (tcgen05.ld -> cvt -> swizzled smem -> TMA store)
// kCta = 2 pairs two CTAs (one cluster) on a 256 x BLOCK_N tile with
// cta_group::2 MMAs: each CTA loads its own 128 rows of A and half of the
// B rows, so per-SM smem traffic for B is halved.
// Two TMEM accumulator stages let the epilogue of tile i overlap the mainloop
// of tile i+1.  BLOCK_N, stage count and rasterisation are runtime values
// chosen by the host tile selector (tile_selector.cpp).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cstdint>

#include "ptx.cuh"

namespace dga {

constexpr uint32_t kBlockM = 128;      // rows per CTA (UMMA M = 128 * kCta)
constexpr uint32_t kBlockKBytes = 128; // one SWIZZLE_128B atom along K per stage
constexpr uint32_t kEpiCols = 32;      // columns per epilogue chunk / TMA store box
constexpr uint32_t kNumThreads = 256;
constexpr uint32_t kMaxStages = 12;

struct GemmParams {
    uint32_t m, n, k;
    uint32_t block_n;        // cluster-tile N (multiple of 16, <= 256; odd multiples only for whole-tile K-major plans)
    uint32_t num_stages;     // smem ring depth
    uint32_t k_blocks;       // ceil(k / elements-per-128B)
    uint32_t num_m_blocks;   // cluster tiles along M
    uint32_t num_n_blocks;   // cluster tiles along N
    uint32_t raster_group;   // tiles per rasterisation group (role of swizzleOffset)
    uint32_t raster_along_m; // 1: consecutive tiles walk M inside a group of N blocks (role of swizzleDirection)
    uint32_t idesc;          // tcgen05 instruction descriptor
    uint32_t b_stage_bytes;  // bytes of B per CTA per stage
    uint32_t debug;          // bottleneck isolation (DGA_DEBUG_FLAGS): 1 no TMA loads, 2 no MMAs, 4 no stores
    uint32_t k_block_elems;  // elements per 128-byte K block: 64 (16-bit) or 128 (fp8)
    // kMx (hardware block-scaled fp8) only:
    uint32_t sf_n_chunks;    // 128-column scale chunks per N tile
    uint32_t sf_m_chunks;    // ceil(m / 128)
    const int32_t* gate;     // device flag; the kernel runs only when *gate == gate_expect (nullptr: always)
    int32_t gate_expect;
    // stream-K (role of PaddingStreamkMatmulKernel): the first dp_tiles tiles are walked whole, one per cluster
    // and wave; each of the remaining sk_tiles tiles (the ragged last wave) is cut into split_k K slices with
    // aligned boundaries, and the sk_tiles * split_k slices are dealt out round-robin (tile-major) to all clusters.
    uint32_t dp_tiles, sk_tiles;
    uint32_t num_peers;      // extra destinations of every output chunk (PeerMaps)
    uint32_t epi_cols;       // columns per TMA store: 32, or 64 for 16-bit outputs when block_n % 64 == 0 (128-byte rows)
    unsigned long long* trace; // debug & 16: per CTA {t_start, t_first_data, t_last_mma_issued, t_end (ns), clk_start, clk_end}
    // kMode 1 (batched: A [B,M,K], B [B,K,N] or [B,N,K], D [B,M,N], 3-D tensor maps): tiles per batch entry
    // kMode 2 (M-grouped contiguous: A [M_total,K], B [G,N,K], D [M_total,N]): group of every row, -1 = skip
    // kMode 4 (M-grouped masked: batched addressing over groups, group_index[g] = valid rows of group g; 128-row
    //         blocks at or past it are skipped by every role and left unwritten)
    // kMode 3 (split-K: work item = K slice `i % split_k` of tile `i / split_k`; fp32 partial tiles go to slab
    //         [slice] of a [split_k, M, N] workspace behind a 3-D tensor map, splitk_reduce_kernel adds them up)
    uint32_t tiles_per_batch;
    uint32_t split_k;
    const int32_t* group_index;
    float* sk_partials;      // [slice item][cta rank][128][block_n] fp32
    int32_t* sk_sems;        // [slice item][cta rank] "partial is in its slot" flags: the giver stores sk_nonce, the
                             // owner waits for it and stores 0 back.  A per-launch nonce instead of a counter so
                             // that a workspace the caller recycled (arbitrary bytes) needs no clearing launch and
                             // cannot stall the owner; the write-back of 0 keeps CUDA-graph replays (same nonce) right
    int32_t sk_nonce;        // non-zero
    // L2 eviction-priority hints of the operand loads (ptx::kEvict*; 0 = normal).  A panel that the tiles in flight
    // will not touch again soon goes evict-first, the one the following waves re-read evict-last, so that the
    // resident set of a long-K problem stays under the L2 capacity (fp8 config 4: profiles/r02_fp8_raster_sweep.log)
    uint64_t hint_a, hint_b;
};

// N-sharded execution (BASELINE config 5): besides its own C, a rank's epilogue stores every finished chunk
// straight into the same column slice of each peer's C over NVLink (TMA stores on peer-mapped addresses), so
// the all-gather of C rides on the GEMM tile by tile instead of following it.
constexpr uint32_t kMaxPeers = 7;
struct PeerMaps {
    CUtensorMap m[kMaxPeers];
};

// One unit of work of a cluster: K blocks [kb0, kb1) of `tile`.  Every warp role walks the same sequence.
struct WorkItem {
    uint32_t tile, kb0, kb1;
    uint32_t item;           // stream-K: index of this slice among all slices (its partial slot); unused otherwise
};
// kSk = false (every variant except stream-K): whole tiles only; the slice arithmetic and, in the epilogue, the
// whole partial-tile path drop out of the kernel (measured: 1.2 us of a 30 us launch at M=512, 0.9 of 8.6 us at
// 1024^3, profiles/r01_jit_keep_rt.log).
// Why aligned slices and not an even deal of K blocks: clusters that work on tiles of one row / column panel at the
// same K offset share the panel through L2 in time; with an even deal every cluster sits at its own K offset, every
// panel block is fetched once per tile, and the variant ran load-bound (profiles/r01_streamk_probe.log).
template <bool kSk>
struct WorkIter {
    uint32_t clusters, dp_tiles, kb_total;
    uint32_t next_dp;
    uint32_t next_sk, sk_items, split;
    __device__ __forceinline__ WorkIter(const GemmParams& p, uint32_t cluster_id, uint32_t num_clusters)
        : clusters(num_clusters), dp_tiles(p.dp_tiles), kb_total(p.k_blocks), next_dp(cluster_id),
          next_sk(cluster_id), sk_items(kSk ? p.sk_tiles * p.split_k : 0u), split(kSk ? p.split_k : 1u) {}
    __device__ __forceinline__ bool next(WorkItem& w) {
        if (next_dp < dp_tiles) {
            w.tile = next_dp; w.kb0 = 0; w.kb1 = kb_total; w.item = 0;
            next_dp += clusters;
            return true;
        }
        if constexpr (!kSk) {
            return false;
        } else {
            if (next_sk >= sk_items) return false;
            const uint32_t local = next_sk / split, slice = next_sk - local * split;
            w.tile = dp_tiles + local;
            w.kb0 = slice * kb_total / split;
            w.kb1 = (slice + 1) * kb_total / split;
            w.item = next_sk;
            next_sk += clusters;
            return true;
        }
    }
};

// --- descriptors -----------------------------------------------------------
// tcgen05 shared-memory matrix descriptor (sm_100 layout): start>>4 [0,14),
// LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48), layout type [61,64).
__host__ __device__ constexpr uint64_t make_smem_desc_hi(uint32_t sbo_bytes, uint32_t layout_type) {
    return (static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46) |
           (static_cast<uint64_t>(layout_type) << 61);
}
constexpr uint32_t kLayoutSw128 = 2;

// Instruction descriptor for kind::f16 / kind::f8f6f4 with fp32 accumulation.
// a_fmt/b_fmt: f16 kind: 0 = fp16, 1 = bf16;  f8f6f4 kind: 0 = e4m3, 1 = e5m2.
__host__ __device__ constexpr uint32_t make_idesc(uint32_t umma_m, uint32_t umma_n, uint32_t ab_fmt,
                                                  uint32_t a_mn_major, uint32_t b_mn_major) {
    return (1u << 4) | (ab_fmt << 7) | (ab_fmt << 10) | (a_mn_major << 15) | (b_mn_major << 16) |
           ((umma_n >> 3) << 17) | ((umma_m >> 4) << 24);
}

// --- tile order --------------------------------------------------------------
// Grouped rasterisation: tiles are walked in groups of `raster_group` blocks of
// one dimension so that the ~74 tiles in flight cover a compact rectangle and
// share A/B panels in L2 (what DynamicGemmIdentityBlockSwizzle(offset,
// direction) does for the reference, tiling_params.h:63-64).
template <typename P>
__device__ __forceinline__ void tile_coords(const P& p, uint32_t t, uint32_t& mb, uint32_t& nb) {
    const uint32_t g = p.raster_group;
    if (p.raster_along_m) {
        // groups of g N-blocks; inside a group walk N fastest, then M
        const uint32_t per_group = g * p.num_m_blocks;
        const uint32_t gi = t / per_group, r = t - gi * per_group;
        const uint32_t first = gi * g;
        const uint32_t size = min(g, p.num_n_blocks - first);
        mb = r / size;
        nb = first + (r - mb * size);
    } else {
        const uint32_t per_group = g * p.num_n_blocks;
        const uint32_t gi = t / per_group, r = t - gi * per_group;
        const uint32_t first = gi * g;
        const uint32_t size = min(g, p.num_m_blocks - first);
        nb = r / size;
        mb = first + (r - nb * size);
    }
}

template <typename OutT> struct OutPack;
template <> struct OutPack<__nv_bfloat16> {
    static __device__ __forceinline__ uint32_t pack(float a, float b) {
        __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&v);
    }
};
template <> struct OutPack<__half> {
    static __device__ __forceinline__ uint32_t pack(float a, float b) {
        __half2 v = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&v);
    }
};

// Write this thread's 32 fp32 accumulator columns as one row of the staging
// tile, in the swizzle the TMA store descriptor expects (SWIZZLE_64B for
// 2-byte outputs: 64 B rows; SWIZZLE_128B for fp32: 128 B rows).
template <typename OutT>
__device__ __forceinline__ void stage_row(uint32_t stage_base, uint32_t row, const uint32_t (&v)[32], uint32_t half = 0,
                                          bool wide = false) {
    if constexpr (sizeof(OutT) == 2) {
        // narrow: 64-byte rows, SWIZZLE_64B; wide: 128-byte rows (two 32-column halves), SWIZZLE_128B
        const uint32_t row_base = stage_base + row * (wide ? 128 : 64);
        const uint32_t x = wide ? (row & 7) : ((row >> 1) & 3);
        const uint32_t j0 = half * 4;
#pragma unroll
        for (uint32_t jj = 0; jj < 4; ++jj) {
            const uint32_t j = jj;
            const uint32_t a = OutPack<OutT>::pack(__uint_as_float(v[8 * j + 0]), __uint_as_float(v[8 * j + 1]));
            const uint32_t b = OutPack<OutT>::pack(__uint_as_float(v[8 * j + 2]), __uint_as_float(v[8 * j + 3]));
            const uint32_t c = OutPack<OutT>::pack(__uint_as_float(v[8 * j + 4]), __uint_as_float(v[8 * j + 5]));
            const uint32_t d = OutPack<OutT>::pack(__uint_as_float(v[8 * j + 6]), __uint_as_float(v[8 * j + 7]));
            ptx::st_shared_v4(row_base + (((j0 + j) ^ x) << 4), a, b, c, d);
        }
    } else {
        const uint32_t row_base = stage_base + row * 128;
        const uint32_t x = row & 7;
#pragma unroll
        for (uint32_t j = 0; j < 8; ++j)
            ptx::st_shared_v4(row_base + ((j ^ x) << 4), v[4 * j + 0], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
}

// Shared-memory carve-up (offsets from a 1024-aligned base; identical in both
// CTAs of a pair, which cta_group::2 descriptors rely on).
struct SmemLayout {
    uint32_t d_stage_bytes, a_stage_bytes, b_stage_bytes, sf_stage_bytes;
    uint32_t off_d, off_a, off_b, off_sf, off_bar, total;
};
// Columns per output store: 128-byte rows whenever the tile allows it (full NVLink / L2 write lines).
__host__ __device__ inline uint32_t epi_cols_for(uint32_t block_n, uint32_t out_elem_bytes) {
    return (out_elem_bytes == 2 && block_n % 64 == 0) ? 64u : 32u;
}
__host__ __device__ inline SmemLayout make_smem_layout(uint32_t d_stage_bytes, uint32_t b_stage_bytes, uint32_t stages,
                                                       uint32_t sf_stage_bytes = 0) {
    SmemLayout l;
    l.d_stage_bytes = d_stage_bytes;
    l.a_stage_bytes = kBlockM * kBlockKBytes;
    l.b_stage_bytes = b_stage_bytes;
    l.off_d = 0;
    l.off_a = 2 * l.d_stage_bytes;
    l.off_b = l.off_a + stages * l.a_stage_bytes;
    l.sf_stage_bytes = sf_stage_bytes;   // per stage: 512 B of A scales + 512 B per 128 columns of B scales
    l.off_sf = l.off_b + stages * l.b_stage_bytes;
    l.off_bar = l.off_sf + stages * sf_stage_bytes;
    // full[stages] empty[stages] tmem_full[2] tmem_empty[2] + tmem ptr
    l.total = l.off_bar + (2 * stages + 4) * 8 + 16;
    return l;
}

// kMx: e4m3 operands with UE8M0 block scales applied by the tensor core (kind::mxf8f6f4.block_scale).  The
// scale words arrive pre-packed (4 consecutive 128-wide K blocks per 32-bit word, already in the 32x128-bit
// order tcgen05.cp wants) from the workspace, one TMA box per 4 K blocks, and live in TMEM behind the two
// accumulator stages.
// kMode: 0 one GEMM; 1 batched (the framework kernel's batch loop, generate_code.hpp:149-153, as one launch: the
// tile index carries the batch entry, 3-D tensor maps clip ragged M/N per entry); 2 M-grouped contiguous (MoE:
// every 128-row block of A multiplies the weight of its group; rows of one block share a group, -1 skips it).
// Modes 1 and 2 run single-CTA tiles (a CTA pair would have to share one B tile across two row blocks).
// kMode 3: multicore split-K (the reference declares it as kernel type 3, catlass_dynamic_matmul_tiling_key.h:117-130,
// and never dispatches it) for shapes whose tiles fill well under one wave: every tile's K range is cut into
// split_k slices that run as independent work items; see splitk_reduce_sm100.cuh for the second half.
constexpr int kModePlain = 0, kModeBatched = 1, kModeMGrouped = 2, kModeSplitK = 3, kModeMasked = 4;
// batched addressing (third tensor-map coordinate = tile / tiles_per_batch)
constexpr bool mode_is_batched(int mode) { return mode == kModeBatched || mode == kModeMasked; }
// Group of a tile in the M-grouped contiguous layout: the group of its first live 128-row block, -1 when every block
// is padding.  Pair tiles (256 rows) are only planned when the caller has promised that groups start on multiples of
// 256 rows, so the two blocks never belong to different groups (one of them may be padding).
__device__ __forceinline__ int32_t pair_group(const GemmParams& p, uint32_t m0, int cta) {
    int32_t g = __ldg(p.group_index + m0);
    if (cta == 2 && g < 0 && m0 + kBlockM < p.m) g = __ldg(p.group_index + m0 + kBlockM);
    return g;
}
// Turns the iterator's item into (tile, K-block range, slab = third coordinate of the 3-D maps).
template <int kMode>
__device__ __forceinline__ void resolve_item(const GemmParams& p, WorkItem& w, uint32_t& slab) {
    if constexpr (mode_is_batched(kMode)) {
        slab = w.tile / p.tiles_per_batch;
        w.tile -= slab * p.tiles_per_batch;
    } else if constexpr (kMode == kModeSplitK) {
        slab = w.tile % p.split_k;
        w.tile /= p.split_k;
        w.kb0 = static_cast<uint32_t>(static_cast<uint64_t>(slab) * p.k_blocks / p.split_k);
        w.kb1 = static_cast<uint32_t>(static_cast<uint64_t>(slab + 1) * p.k_blocks / p.split_k);
    } else {
        slab = 0;
    }
}
// The kernel body is a device function so that two entries can share it: the dynamic-shape kernel below (every
// field of GemmParams read at run time) and the shape-specialised entry a JIT build wraps around it
// (jit_entry.cuh: the shape fields are compile-time constants there and fold through the inlined body).
// The tensor maps must be the kernel's own __grid_constant__ parameters (TMA takes their address).
// kMc (north_star item 1, "TMA multicast loads"; role of the L1 staging + swizzle of CommonMatmulKernel,
// common_matmul_kernel.h:28-78): a cluster of FOUR CTAs = two cta_group::2 pairs on neighbouring N tiles of one
// 256-row block.  Both pairs need the same A rows, so each CTA fetches only 64 of its 128 A rows per stage and the
// load is multicast to the CTA of the same rank in the other pair: a third (n1 = 128) to a quarter (n1 = 256) fewer
// bytes leave L2 per flop.  A stage may be refilled once BOTH pairs have consumed it: every leader's commit goes
// to the empty barrier of all four CTAs (count 2).  Everything behind the ring (TMEM, epilogue) stays per pair.
// num_n_blocks then counts 2 * block_n wide column groups; the host only takes it for an even number of N tiles.
// kKs (DGA_KERNEL_CLUSTER_SPLITK; the reference's answer to "tiles << cores" is stream-K + ReduceAdd through the
// workspace, padding_streamk_matmul_kernel.h:33-108): a cluster of FOUR CTAs = two cta_group::2 pairs on the SAME
// 256 x n1 tile, each pair accumulating one half of K.  Afterwards CTA (pair p, rank q) owns column half p of its 128
// rows: it stages the other half of its fp32 accumulator in its (drained) ring, ships it with one bulk DSMEM copy
// into the ring of its twin (pair 1 - p, rank q), adds the half it receives to its own and stores.  No workspace, no
// second launch, no flags in global memory; one tile per cluster (the rings are busy after the main loop).  For
// shapes whose wide tiles fill under half of the machine: the wide tile keeps the operand bytes per flop that one
// SM can ingest (64 B/clk), the K split fills the SMs.
template <int kCta, typename OutT, bool kBMn, bool kMx = false, int kMode = kModePlain, bool kSk = false, bool kMc = false,
          bool kKs = false>
__device__ __forceinline__ void
gemm_sm100_body(const CUtensorMap& tm_a, const CUtensorMap& tm_b, const CUtensorMap& tm_d, const CUtensorMap& tm_sfa,
                const CUtensorMap& tm_sfb, const PeerMaps& peers, const GemmParams& p) {
    static_assert(!(kMx && kBMn), "block-scaled fp8 wants K-major operands");
    static_assert(kMode == kModePlain || kMode == kModeSplitK || !kMx, "batched / grouped modes are 16-bit");
    static_assert(kCta == 1 || kMode != kModeBatched, "the framework's batched launch uses single-CTA tiles");
    static_assert(kCta == 1 || !kBMn || kMode == kModePlain || kMode == kModeSplitK, "grouped pair tiles want [G,N,K] weights");
    static_assert(kMode != kModeSplitK || (sizeof(OutT) == 4 && !kMx), "split-K slices leave fp32 partial tiles");
    static_assert(!kSk || (kMode == kModePlain && !kMx), "stream-K is a variant of the plain 16-bit kernel");
    static_assert(!kMc || (kCta == 2 && kMode == kModePlain && !kSk && !kMx), "the multicast cluster is two whole-tile 16-bit CTA pairs");
    static_assert(!kKs || (kCta == 2 && kMode == kModePlain && !kSk && !kMx && !kMc), "the K-split cluster is two 16-bit CTA pairs on one tile");
    ptx::grid_dep_launch();   // the next grid on the stream may set itself up while this one runs
    const bool tracing = (p.debug & 16) != 0 && p.trace != nullptr;
    if (tracing && threadIdx.x == 0) {
        p.trace[blockIdx.x * 6 + 0] = ptx::globaltimer();
        p.trace[blockIdx.x * 6 + 4] = clock64();
    }
    if (p.gate != nullptr) {
        ptx::grid_dep_wait();     // the flag was written by the previous grid
        if (*p.gate != p.gate_expect) return;   // grid-uniform
    }
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (smem_base - ptx::smem_u32(smem_raw));

    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31;
    constexpr bool kQuad = kMc || kKs;                         // clusters of four CTAs = two pairs
    constexpr uint32_t kClusterCtas = kQuad ? 4 : kCta;
    const uint32_t cluster_rank = (kCta == 2) ? ptx::cluster_ctarank() : 0u;
    const uint32_t pair = kQuad ? cluster_rank >> 1 : 0u;      // which of the cluster's CTA pairs (kMc: N neighbour, kKs: K half)
    const uint32_t cta_rank = kQuad ? cluster_rank & 1u : cluster_rank;   // rank inside the pair
    const uint32_t leader_rank = pair * 2;                     // cluster rank of this pair's MMA-issuing CTA
    const bool is_leader = cta_rank == 0;
    const uint32_t num_clusters = gridDim.x / kClusterCtas;
    const uint32_t cluster_id = blockIdx.x / kClusterCtas;
    const uint32_t stages = p.num_stages;

    const uint32_t sf_stage_bytes = kMx ? 512u * (1u + p.sf_n_chunks) : 0u;
    const SmemLayout L = make_smem_layout(kBlockM * p.epi_cols * sizeof(OutT), p.b_stage_bytes, stages, sf_stage_bytes);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.off_bar);
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* tmem_full_bar = empty_bar + stages;
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
    // kKs: the exchange barriers sit where the second accumulator stage's would (one tile per cluster: stage 1 is never
    // used).  xchg_ready: the twin's ring is drained and its receive barrier armed; xchg_recv: the twin's half has landed.
    uint64_t* xchg_ready_bar = &tmem_full_bar[1];
    uint64_t* xchg_recv_bar = &tmem_empty_bar[1];

    if constexpr (kCta == 2) ptx::cluster_sync();  // peer must be resident before cta_group::2 alloc

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tm_a);
        ptx::prefetch_tensormap(&tm_b);
        ptx::prefetch_tensormap(&tm_d);
        if constexpr (kMx) {
            ptx::prefetch_tensormap(&tm_sfa);
            ptx::prefetch_tensormap(&tm_sfb);
        }
    }
    if (warp == 1 && lane == 0) {
        for (uint32_t s = 0; s < stages; ++s) {
            ptx::mbar_init(&full_bar[s], kCta);  // leader expect_tx arrive (+ peer's remote arrive)
            ptx::mbar_init(&empty_bar[s], kMc ? 2 : 1);   // one tcgen05.commit per pair that reads the stage (multicast)
        }
        for (uint32_t s = 0; s < (kKs ? 1u : 2u); ++s) {
            ptx::mbar_init(&tmem_full_bar[s], 1);
            ptx::mbar_init(&tmem_empty_bar[s], kCta * 128);  // every epilogue thread of the pair
        }
        if constexpr (kKs) {
            ptx::mbar_init(xchg_ready_bar, 1);
            ptx::mbar_init(xchg_recv_bar, 1);
        }
        ptx::fence_mbar_init();
    }
    if (warp == 2) ptx::tmem_alloc<kCta>(tmem_ptr_smem, 512);
    ptx::tc_fence_before();
    if constexpr (kCta == 2) ptx::cluster_sync(); else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    ptx::grid_dep_wait();   // first global access (TMA loads, stores, stream-K workspace) comes after this

    const uint32_t load_n = p.block_n / kCta;  // B rows this CTA loads per stage
    const uint64_t hint_a = p.hint_a != 0 ? p.hint_a : ptx::kEvictNormal;
    const uint64_t hint_b = p.hint_b != 0 ? p.hint_b : ptx::kEvictNormal;

    if (warp == 0) {
        // ------------------------------------------------ TMA producer
        // The whole warp walks the loop converged (uniform registers hold the addresses); one elected
        // lane issues.  A divergent `lane == 0` loop costs an ELECT/R2UR shuffle per operand.
        uint32_t stage = 0, phase = 0;
        WorkIter<kSk> it(p, cluster_id, num_clusters);
        WorkItem w;
        while (it.next(w)) {
            uint32_t mb, nb, slab;
            resolve_item<kMode>(p, w, slab);
            int32_t z = static_cast<int32_t>(slab);   // third tensor-map coordinate: batch entry (mode 1) or group (mode 2)
            tile_coords(p, w.tile, mb, nb);
            if constexpr (kMc) nb = nb * 2 + pair;
            if constexpr (kKs) {   // this pair's half of the K blocks
                w.kb0 = pair * (p.k_blocks / 2);
                w.kb1 = pair == 0 ? p.k_blocks / 2 : p.k_blocks;
            }
            const int32_t m_idx = static_cast<int32_t>(mb * (kBlockM * kCta) + cta_rank * kBlockM);
            const int32_t n_idx = static_cast<int32_t>(nb * p.block_n + cta_rank * load_n);
            if constexpr (kMode == kModeMGrouped) {
                z = pair_group(p, mb * (kBlockM * kCta), kCta);
                if (z < 0) continue;   // every role skips the same blocks
            }
            if constexpr (kMode == kModeMasked) {   // (pair tiles: decided by the pair's FIRST row block)
                if (static_cast<int32_t>(mb * (kBlockM * kCta)) >= __ldg(p.group_index + slab)) continue;
            }
            for (uint32_t kb = w.kb0; kb < w.kb1; ++kb) {
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                if (p.debug & 1) {
                    if (ptx::elect_one()) {
                        if (is_leader) ptx::mbar_arrive(&full_bar[stage]);
                        else ptx::mbar_arrive_remote(&full_bar[stage], leader_rank);
                    }
                } else if (ptx::elect_one()) {
                    uint8_t* sa = smem + L.off_a + stage * L.a_stage_bytes;
                    uint8_t* sb = smem + L.off_b + stage * L.b_stage_bytes;
                    const int32_t k_idx = static_cast<int32_t>(kb * p.k_block_elems);
                    uint32_t sf_bytes = 0;
                    if constexpr (kMx) {
                        if ((kb & 3) == 0) {   // one scale word covers 4 K blocks
                            uint8_t* ssf = smem + L.off_sf + stage * L.sf_stage_bytes;
                            const int32_t kg = static_cast<int32_t>(kb >> 2);
                            const int32_t row_a = kg * static_cast<int32_t>(p.sf_m_chunks) +
                                                  static_cast<int32_t>(mb * kCta + cta_rank);
                            const int32_t row_b = (kg * static_cast<int32_t>(p.num_n_blocks) + static_cast<int32_t>(nb)) *
                                                  static_cast<int32_t>(p.sf_n_chunks);
                            if constexpr (kCta == 2) {
                                ptx::tma_load_2d_pair(&tm_sfa, &full_bar[stage], ssf, 0, row_a);
                                ptx::tma_load_2d_pair(&tm_sfb, &full_bar[stage], ssf + 512, 0, row_b);
                            } else {
                                ptx::tma_load_2d(&tm_sfa, &full_bar[stage], ssf, 0, row_a);
                                ptx::tma_load_2d(&tm_sfb, &full_bar[stage], ssf + 512, 0, row_b);
                            }
                            sf_bytes = L.sf_stage_bytes;
                        }
                    }
                    if constexpr (kCta == 2) {
                        if constexpr (kMc)   // 64 of this CTA's A rows, for itself and for its twin in the other pair
                            ptx::tma_load_2d_pair_multicast(&tm_a, &full_bar[stage], sa + pair * (kBlockM / 2) * kBlockKBytes,
                                                            k_idx, m_idx + static_cast<int32_t>(pair * (kBlockM / 2)),
                                                            static_cast<uint16_t>(0x5u << cta_rank), hint_a);
                        else if constexpr (mode_is_batched(kMode))
                            ptx::tma_load_3d_pair(&tm_a, &full_bar[stage], sa, k_idx, m_idx, z, hint_a);
                        else
                            ptx::tma_load_2d_pair(&tm_a, &full_bar[stage], sa, k_idx, m_idx, hint_a);
                        if constexpr (kBMn) {
                            for (uint32_t j = 0; j < load_n / 64; ++j)
                                ptx::tma_load_2d_pair(&tm_b, &full_bar[stage], sb + j * 8192, n_idx + j * 64, k_idx, hint_b);
                        } else if constexpr (kMode == kModeMGrouped || kMode == kModeMasked) {
                            ptx::tma_load_3d_pair(&tm_b, &full_bar[stage], sb, k_idx, n_idx, z, hint_b);   // weight of group z
                        } else {
                            ptx::tma_load_2d_pair(&tm_b, &full_bar[stage], sb, k_idx, n_idx, hint_b);
                        }
                        if (is_leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * (L.a_stage_bytes + L.b_stage_bytes + sf_bytes));
                        else ptx::mbar_arrive_remote(&full_bar[stage], leader_rank);
                    } else {
                        if constexpr (mode_is_batched(kMode)) ptx::tma_load_3d(&tm_a, &full_bar[stage], sa, k_idx, m_idx, z);
                        else ptx::tma_load_2d(&tm_a, &full_bar[stage], sa, k_idx, m_idx);
                        if constexpr (kBMn) {
                            for (uint32_t j = 0; j < load_n / 64; ++j) {
                                if constexpr (kMode == kModePlain || kMode == kModeSplitK)
                                    ptx::tma_load_2d(&tm_b, &full_bar[stage], sb + j * 8192, n_idx + j * 64, k_idx);
                                else
                                    ptx::tma_load_3d(&tm_b, &full_bar[stage], sb + j * 8192, n_idx + j * 64, k_idx, z);
                            }
                        } else {
                            if constexpr (kMode == kModePlain || kMode == kModeSplitK)
                                ptx::tma_load_2d(&tm_b, &full_bar[stage], sb, k_idx, n_idx);
                            else ptx::tma_load_3d(&tm_b, &full_bar[stage], sb, k_idx, n_idx, z);
                        }
                        ptx::mbar_arrive_expect_tx(&full_bar[stage], L.a_stage_bytes + L.b_stage_bytes + sf_bytes);
                    }
                }
                __syncwarp();
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        if (is_leader) {
            // -------------------------------------------------- MMA issuer (warp-converged, one lane issues)
            // Only the low 32 bits of a descriptor change (14-bit start address field): keep the high
            // halves constant and do 32-bit arithmetic.
            const uint32_t a_hi = static_cast<uint32_t>(make_smem_desc_hi(1024, kLayoutSw128) >> 32);
            const uint32_t b_hi = a_hi;
            // K-major B: 8-row atoms 1024 B apart (SBO).  MN-major B: SBO = 1024 B
            // between 8-k-row groups, LBO = 8192 B between 64-column groups.
            const uint32_t b_lbo = kBMn ? ((8192u >> 4) << 16) : 0u;
            const uint32_t a_lo0 = ((smem_base + L.off_a) & 0x3FFFF) >> 4;
            const uint32_t b_lo0 = (((smem_base + L.off_b) & 0x3FFFF) >> 4) | b_lbo;
            const uint32_t a_step = L.a_stage_bytes >> 4, b_step = L.b_stage_bytes >> 4;
            constexpr uint32_t kBStepLo = kBMn ? (16 * 128) >> 4 : 32 >> 4;  // one UMMA_K (16 elements) along K
            const uint32_t idesc = p.idesc;
            const uint32_t tmem_sfa = tmem_base + 2 * p.block_n;   // 4 columns; B scales follow
            const uint32_t tmem_sfb = tmem_sfa + 4;
            uint32_t stage = 0, phase = 0, iter = 0;
            WorkIter<kSk> it(p, cluster_id, num_clusters);
            WorkItem w;
            for (; it.next(w); ++iter) {
                uint32_t slab_of_item;
                resolve_item<kMode>(p, w, slab_of_item);
                if constexpr (kMode == kModeMGrouped || kMode == kModeMasked) {
                    uint32_t mb, nb;
                    tile_coords(p, w.tile, mb, nb);
                    const int32_t g = kMode == kModeMasked ? __ldg(p.group_index + slab_of_item)
                                                           : pair_group(p, mb * (kBlockM * kCta), kCta);
                    if (kMode == kModeMasked ? static_cast<int32_t>(mb * (kBlockM * kCta)) >= g : g < 0) { --iter; continue; }
                }
                if constexpr (kKs) {
                    w.kb0 = pair * (p.k_blocks / 2);
                    w.kb1 = pair == 0 ? p.k_blocks / 2 : p.k_blocks;
                }
                const uint32_t as = iter & 1, aphase = (iter >> 1) & 1;
                ptx::mbar_wait(&tmem_empty_bar[as], aphase ^ 1);
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + as * p.block_n;
                const uint32_t kb0 = w.kb0, k_blocks = w.kb1;
                for (uint32_t kb = kb0; kb < k_blocks; ++kb) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    if (tracing && iter == 0 && kb == kb0 && lane == 0) p.trace[blockIdx.x * 6 + 1] = ptx::globaltimer();
                    if (ptx::elect_one()) {
                        const uint32_t a_lo = a_lo0 + stage * a_step;
                        const uint32_t b_lo = b_lo0 + stage * b_step;
                        uint32_t idesc_k = idesc;
                        if constexpr (kMx) {
                            if ((kb & 3) == 0) {
                                // scale words of the next 4 K blocks: smem -> TMEM (ordered with the MMAs in the tensor pipe)
                                const uint64_t sf_hi = make_smem_desc_hi(128, 0);
                                const uint32_t sf_lo = ((smem_base + L.off_sf + stage * L.sf_stage_bytes) & 0x3FFFF) >> 4;
                                ptx::tmem_cp_32x128b_warpx4<kCta>(tmem_sfa, sf_hi | sf_lo);
                                for (uint32_t i = 0; i < p.sf_n_chunks; ++i)
                                    ptx::tmem_cp_32x128b_warpx4<kCta>(tmem_sfb + 4 * i, sf_hi | (sf_lo + 32 * (1 + i)));
                            }
                            const uint32_t sf_id = kb & 3;
                            idesc_k = idesc | (sf_id << 29) | (sf_id << 4);
                        }
#pragma unroll
                        for (uint32_t k = 0; k < 4; ++k) {
                            const uint64_t a_desc = (static_cast<uint64_t>(a_hi) << 32) | (a_lo + k * 2);
                            const uint64_t b_desc = (static_cast<uint64_t>(b_hi) << 32) | (b_lo + k * kBStepLo);
                            if constexpr (kMx)
                                ptx::mma_mxf8<kCta>(tmem_d, a_desc, b_desc, idesc_k, tmem_sfa, tmem_sfb, ((kb - kb0) | k) != 0);
                            else if (!(p.debug & 2))
                                ptx::mma_f16<kCta>(tmem_d, a_desc, b_desc, idesc_k, ((kb - kb0) | k) != 0);
                        }
                        if constexpr (kMc) {
                            ptx::mma_commit_multicast<kCta>(&empty_bar[stage], 0xF);
                            if (kb + 1 == k_blocks)
                                ptx::mma_commit_multicast<kCta>(&tmem_full_bar[as], static_cast<uint16_t>(0x3u << leader_rank));
                        } else if constexpr (kKs) {   // this pair's two CTAs (ranks 2 pair, 2 pair + 1)
                            ptx::mma_commit_multicast<kCta>(&empty_bar[stage], static_cast<uint16_t>(0x3u << leader_rank));
                            if (kb + 1 == k_blocks)
                                ptx::mma_commit_multicast<kCta>(&tmem_full_bar[as], static_cast<uint16_t>(0x3u << leader_rank));
                        } else {
                            ptx::mma_commit<kCta>(&empty_bar[stage]);
                            if (kb + 1 == k_blocks) ptx::mma_commit<kCta>(&tmem_full_bar[as]);
                        }
                    }
                    __syncwarp();
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
            if (tracing && lane == 0) p.trace[blockIdx.x * 6 + 2] = ptx::globaltimer();
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------- epilogue
        const uint32_t epi_tid = threadIdx.x - 128;          // 0..127 == accumulator row == TMEM lane
        const uint32_t lane_base = (warp & 3) * 32;          // tcgen05.ld lane window of this warp
        uint32_t iter = 0, buf = 0;
        // one 32-column chunk of this CTA's rows: registers -> staging smem -> TMA store
        const bool wide = p.epi_cols == 64;
        int32_t zb = 0;   // batch entry of the tile being stored (mode 1)
        auto store_chunk = [&](const uint32_t (&v)[32], uint32_t m0, uint32_t col, uint32_t c, bool last) {
            const uint32_t half = wide ? (c & 1u) : 0u;
            if (half == 0) {
                // staging buffer `buf` was last read by the store issued two stores ago
                if (epi_tid == 0) ptx::tma_store_wait_read<1>();
                ptx::named_bar_sync(1, 128);
            }
            const uint32_t sbase = smem_base + L.off_d + buf * L.d_stage_bytes;
            stage_row<OutT>(sbase, epi_tid, v, half, wide);
            if (wide && half == 0 && !last) return;   // the second half of the 128-byte rows follows
            ptx::fence_async_smem();
            ptx::named_bar_sync(1, 128);
            if (epi_tid == 0) {
                const int32_t c0 = static_cast<int32_t>(col - half * kEpiCols);
                if (m0 < p.m && !(p.debug & 4)) {
                    if constexpr (mode_is_batched(kMode) || kMode == kModeSplitK)
                        ptx::tma_store_3d(&tm_d, smem + L.off_d + buf * L.d_stage_bytes, c0, static_cast<int32_t>(m0), zb);
                    else
                        ptx::tma_store_2d(&tm_d, smem + L.off_d + buf * L.d_stage_bytes, c0, static_cast<int32_t>(m0));
                    // same chunk into every peer's C (NVLink); the CTAs start their round at different peers so
                    // that at any moment the stores in flight spread over all links instead of marching in step
                    for (uint32_t i = 0, pi = blockIdx.x % max(p.num_peers, 1u); i < p.num_peers; ++i) {
                        ptx::tma_store_2d(&peers.m[pi], smem + L.off_d + buf * L.d_stage_bytes, c0,
                                          static_cast<int32_t>(m0));
                        if (++pi == p.num_peers) pi = 0;
                    }
                }
                ptx::tma_store_commit();
            }
            buf ^= 1;
        };
        WorkIter<kSk> it(p, cluster_id, num_clusters);
        WorkItem w;
        for (; it.next(w); ++iter) {
            uint32_t mb, nb, slab;
            resolve_item<kMode>(p, w, slab);
            zb = static_cast<int32_t>(slab);
            tile_coords(p, w.tile, mb, nb);
            if constexpr (kMc) nb = nb * 2 + pair;
            const uint32_t m0 = mb * (kBlockM * kCta) + cta_rank * kBlockM;
            const uint32_t n0 = nb * p.block_n;
            // pair tiles: the tile is worked on when either row block is live; a CTA whose own block is padding (or
            // starts past the group's valid rows) takes part in the TMEM handshake but stores nothing
            bool own_rows_live = true;
            if constexpr (kMode == kModeMGrouped) {
                if (pair_group(p, mb * (kBlockM * kCta), kCta) < 0) { --iter; continue; }
                own_rows_live = __ldg(p.group_index + m0) >= 0;
            }
            if constexpr (kMode == kModeMasked) {
                const int32_t valid = __ldg(p.group_index + slab);
                if (static_cast<int32_t>(mb * (kBlockM * kCta)) >= valid) { --iter; continue; }
                own_rows_live = static_cast<int32_t>(m0) < valid;
            }
            const uint32_t m0_store = own_rows_live ? m0 : p.m;   // store_chunk skips rows at or beyond p.m
            const uint32_t as = iter & 1, aphase = (iter >> 1) & 1;
            // chunks that intersect the matrix (TMA clips a partial last one)
            const uint32_t n_left = p.n - n0;
            const uint32_t chunks = (min(p.block_n, n_left) + kEpiCols - 1) / kEpiCols;
            // A tile whose width is an odd multiple of 16 (whole-tile plans only): its last chunk is moved back by 16
            // columns, so that every store stays a whole 32-column box inside the tile; the 16 columns it shares
            // with its predecessor are written twice with the same values.
            auto chunk_col = [&](uint32_t c) { return min(c * kEpiCols, p.block_n - kEpiCols); };
            ptx::mbar_wait(&tmem_full_bar[as], aphase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + (lane_base << 16) + as * p.block_n;
            auto release_tmem = [&]() {
                // accumulator fully read: hand the TMEM stage back to the issuer
                ptx::tc_fence_before();
                if constexpr (kCta == 2) ptx::mbar_arrive_remote(&tmem_empty_bar[as], leader_rank);
                else ptx::mbar_arrive(&tmem_empty_bar[as]);
            };
            if constexpr (kKs) {
                // ---- two pairs, one tile: exchange accumulator halves through DSMEM, then store the own half
                const uint32_t half_cols = p.block_n / 2;
                const uint32_t half_chunks = half_cols / kEpiCols;
                const uint32_t xchg_bytes = kBlockM * half_cols * 4;
                const uint32_t twin = (1u - pair) * 2u + cta_rank;
                // staging (to send) and landing (received) areas in the drained ring, [chunk][4-column group][row] of
                // 16-byte words: a warp's 32 rows are contiguous (no bank conflicts either side)
                uint4* send = reinterpret_cast<uint4*>(smem + L.off_a) + epi_tid;
                const uint4* recv = reinterpret_cast<const uint4*>(smem + L.off_a + xchg_bytes) + epi_tid;
                if (epi_tid == 0) {
                    // this pair's MMAs are done (tmem_full), so this CTA's ring is free: arm the landing barrier and
                    // tell the twin it may write
                    ptx::mbar_arrive_expect_tx(xchg_recv_bar, xchg_bytes);
                    ptx::mbar_arrive_remote(xchg_ready_bar, twin);
                }
                for (uint32_t c = 0; c < half_chunks; ++c) {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32(taddr + (1u - pair) * half_cols + c * kEpiCols, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (uint32_t j = 0; j < 8; ++j)
                        send[(c * 8 + j) * kBlockM] = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                }
                ptx::fence_async_smem();
                ptx::named_bar_sync(1, 128);
                if (epi_tid == 0) {
                    ptx::mbar_wait(xchg_ready_bar, 0);
                    ptx::dsmem_bulk_copy(smem_base + L.off_a + xchg_bytes, smem_base + L.off_a, xchg_bytes, xchg_recv_bar, twin);
                }
                ptx::mbar_wait(xchg_recv_bar, 0);
                const uint32_t own0 = pair * half_cols;          // first column of the own half inside the tile
                const uint32_t own_left = n0 + own0 < p.n ? p.n - (n0 + own0) : 0u;
                const uint32_t own_chunks = (min(half_cols, own_left) + kEpiCols - 1) / kEpiCols;
                for (uint32_t c = 0; c < own_chunks; ++c) {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32(taddr + own0 + c * kEpiCols, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (uint32_t j = 0; j < 8; ++j) {
                        const uint4 r = recv[(c * 8 + j) * kBlockM];
                        v[4 * j + 0] = __float_as_uint(__uint_as_float(v[4 * j + 0]) + __uint_as_float(r.x));
                        v[4 * j + 1] = __float_as_uint(__uint_as_float(v[4 * j + 1]) + __uint_as_float(r.y));
                        v[4 * j + 2] = __float_as_uint(__uint_as_float(v[4 * j + 2]) + __uint_as_float(r.z));
                        v[4 * j + 3] = __float_as_uint(__uint_as_float(v[4 * j + 3]) + __uint_as_float(r.w));
                    }
                    store_chunk(v, m0, n0 + own0 + c * kEpiCols, c, c + 1 == own_chunks);
                }
                release_tmem();
            } else if (kMode == kModeSplitK || !kSk || w.kb1 - w.kb0 == p.k_blocks) {
                for (uint32_t c = 0; c < chunks; ++c) {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32(taddr + chunk_col(c), v);
                    ptx::tmem_ld_wait();
                    if (c + 1 == chunks) release_tmem();
                    store_chunk(v, m0_store, n0 + chunk_col(c), c, c + 1 == chunks);
                }
            } else {
                // ---- stream-K slice.  Slices 0 .. S-2 of a tile ("givers") leave their fp32 partial in their
                // workspace slot and bump the tile's counter; slice S-1 (the "owner", fixed, so the summation order
                // is too) waits for S-1 arrivals, then adds the slots in slice order and its own accumulator, still
                // in TMEM, last.  No deadlock: a slice only ever waits for slices with a smaller item index, every
                // cluster walks its items in increasing order, and the whole grid is resident (<= one CTA per SM).
                const uint64_t slot_floats = static_cast<uint64_t>(kBlockM) * p.block_n;
                const uint32_t split = p.split_k;
                const uint32_t first = (w.tile - p.dp_tiles) * split;      // item index of slice 0
                const uint32_t slice = w.item - first;
                auto flag_of = [&](uint32_t item) { return p.sk_sems + item * kCta + cta_rank; };
                // slot layout [chunk][4-column group][row] of 16-byte words: a warp's 32 rows are contiguous
                auto slot_of = [&](uint32_t item) {
                    return reinterpret_cast<uint4*>(p.sk_partials + (static_cast<uint64_t>(item) * kCta + cta_rank) * slot_floats) + epi_tid;
                };
                if (slice + 1 != split) {
                    uint4* mine = slot_of(w.item);
                    for (uint32_t c = 0; c < chunks; ++c) {
                        uint32_t v[32];
                        ptx::tmem_ld_32x32(taddr + c * kEpiCols, v);
                        ptx::tmem_ld_wait();
                        if (c + 1 == chunks) release_tmem();
#pragma unroll
                        for (uint32_t j = 0; j < 8; ++j)
                            __stcg(mine + (c * 8 + j) * kBlockM, make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
                    }
                    __threadfence();
                    ptx::named_bar_sync(1, 128);
                    if (epi_tid == 0 && !(p.debug & 8)) atomicExch(flag_of(w.item), p.sk_nonce);
                } else {
                    const uint32_t others = (p.debug & 8) ? 0u : split - 1;   // debug 8: skip the reduction
                    if (others != 0) {
                        if (epi_tid < others) {
                            // every giver is resident and never waits, so this ends; the deadline turns a broken
                            // invariant into a launch failure instead of a hung device
                            const volatile int32_t* flag = flag_of(first + epi_tid);
                            const unsigned long long deadline = ptx::globaltimer() + 10000000000ull;   // 10 s
                            while (*flag != p.sk_nonce) {
                                __nanosleep(64);
                                if (ptx::globaltimer() > deadline) __trap();
                            }
                        }
                        ptx::named_bar_sync(1, 128);
                        __threadfence();
                    }
                    // software pipeline over (chunk, giver slice) pairs: the loads of pair i + 1 fly while pair i
                    // is added; two register sets, every index a compile-time constant
                    const uint32_t pairs = chunks * others;
                    uint4 ld_a[8], ld_b[8];
                    float acc[32];
                    auto fetch = [&](uint4 (&dst)[8], uint32_t i) __attribute__((always_inline)) {
                        const uint32_t c = i / others, sl = i - c * others;
                        const uint4* src = slot_of(first + sl) + c * 8 * kBlockM;
#pragma unroll
                        for (uint32_t j = 0; j < 8; ++j) dst[j] = __ldcg(src + j * kBlockM);
                    };
                    auto finish_chunk = [&](uint32_t c) __attribute__((always_inline)) {
                        uint32_t own[32];
                        ptx::tmem_ld_32x32(taddr + c * kEpiCols, own);
                        ptx::tmem_ld_wait();
                        if (c + 1 == chunks) release_tmem();
#pragma unroll
                        for (uint32_t j = 0; j < 32; ++j) own[j] = __float_as_uint(acc[j] + __uint_as_float(own[j]));
                        store_chunk(own, m0, n0 + c * kEpiCols, c, c + 1 == chunks);
                    };
                    auto consume = [&](const uint4 (&src)[8], uint32_t i) __attribute__((always_inline)) {
                        const uint32_t c = i / others, sl = i - c * others;
                        if (sl == 0) {
#pragma unroll
                            for (uint32_t j = 0; j < 32; ++j) acc[j] = 0.f;
                        }
#pragma unroll
                        for (uint32_t j = 0; j < 8; ++j) {
                            acc[4 * j + 0] += __uint_as_float(src[j].x);
                            acc[4 * j + 1] += __uint_as_float(src[j].y);
                            acc[4 * j + 2] += __uint_as_float(src[j].z);
                            acc[4 * j + 3] += __uint_as_float(src[j].w);
                        }
                        if (sl + 1 == others) finish_chunk(c);
                    };
                    if (pairs == 0) {
#pragma unroll
                        for (uint32_t j = 0; j < 32; ++j) acc[j] = 0.f;
                        for (uint32_t c = 0; c < chunks; ++c) finish_chunk(c);
                    } else {
                        fetch(ld_a, 0);
#pragma unroll 1
                        for (uint32_t i = 0; i < pairs; i += 2) {
                            if (i + 1 < pairs) fetch(ld_b, i + 1);
                            consume(ld_a, i);
                            if (i + 1 < pairs) {
                                if (i + 2 < pairs) fetch(ld_a, i + 2);
                                consume(ld_b, i + 1);
                            }
                        }
                    }
                    if (others != 0) {
                        ptx::named_bar_sync(1, 128);   // every thread has read the slots
                        if (epi_tid < others) *flag_of(first + epi_tid) = 0;   // a replay with the same nonce starts clean
                    }
                }
            }
        }
        if (epi_tid == 0) ptx::tma_store_wait_all<0>();
    }

    // ---------------------------------------------------------------- teardown
    ptx::tc_fence_before();
    if constexpr (kCta == 2) ptx::cluster_sync(); else __syncthreads();
    if (warp == 2) ptx::tmem_dealloc<kCta>(tmem_base, 512);
    if (tracing && threadIdx.x == 0) {
        p.trace[blockIdx.x * 6 + 3] = ptx::globaltimer();
        p.trace[blockIdx.x * 6 + 5] = clock64();
    }
}

template <int kCta, typename OutT, bool kBMn, bool kMx = false, int kMode = kModePlain, bool kSk = false, bool kMc = false,
          bool kKs = false>
__global__ void __launch_bounds__(kNumThreads, 1)
gemm_sm100_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                  const __grid_constant__ CUtensorMap tm_d, const __grid_constant__ CUtensorMap tm_sfa,
                  const __grid_constant__ CUtensorMap tm_sfb, const __grid_constant__ PeerMaps peers,
                  const GemmParams p) {
    gemm_sm100_body<kCta, OutT, kBMn, kMx, kMode, kSk, kMc, kKs>(tm_a, tm_b, tm_d, tm_sfa, tm_sfb, peers, p);
}

}  // namespace dga
