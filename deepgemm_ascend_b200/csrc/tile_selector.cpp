// B200 tile selector.  See tile_selector.h.
//
// Decision structure kept from the reference (select_kernel.cpp:333-369):
//   1. canonicalise degenerate layouts, 2. pick a tile that balances waves over
//   the cores (BalanceWorkload, utils.cpp:28-40), 3. walk a chain of kernel
//   variants (Small -> split-K/stream-K -> PaddingCommon -> Common) and take the
//   first that applies / the cheapest under a cost model.
// What changed for B200: 148 SMs (74 CTA pairs), tcgen05 tiles of 128*cta x N
// with N in steps of 16*cta, TMA handles ragged tiles so PADDING_* survives only
// as "leading dimension not 16-byte aligned -> repack", and the bandwidth
// polynomial (GetBandwidth, select_kernel.cpp:22-80) becomes a
// max(tensor, L2 feed, HBM) x wave-quantisation estimate.
#include "tile_selector.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <iostream>

namespace dga {

PlatformInfo& platform() {
    static PlatformInfo info;
    return info;
}

static thread_local std::string g_last_error;
void set_error(const std::string& msg) {
    g_last_error = msg;
    std::cerr << "[DGA] [ERROR] " << msg << std::endl;
}
const char* last_error() { return g_last_error.c_str(); }

uint32_t common_smem_bytes(uint32_t block_n, uint32_t cta_group, uint32_t stages, uint32_t out_bytes,
                           uint32_t in_bytes) {
    (void)in_bytes;  // a stage is always one 128-byte swizzle atom along K
    const uint32_t d_stage = 128 * epi_cols_host(block_n, out_bytes) * out_bytes;
    const uint32_t a_stage = 128 * 128;
    const uint32_t b_stage = (block_n / cta_group) * 128;
    return 2 * d_stage + stages * (a_stage + b_stage) + (2 * stages + 4) * 8 + 16 + 1024;
}

// fp32 reduction slots of the split-K kernel: split x ceil(m_pad / split) rows of 128 floats (smallm_red_bytes)
static uint32_t splitk_red_bytes(uint32_t m_pad, uint32_t split) {
    return split > 1 ? split * ceil_div(m_pad, split) * 512u : 0u;
}
uint32_t splitk_smem_bytes(uint32_t m_pad, uint32_t stages, uint32_t split) {
    const uint32_t stage = 128 * 128 + m_pad * 128;
    return stages * stage + splitk_red_bytes(m_pad, split) + (2 * stages + 2) * 8 + 16 + 1024;
}

namespace {

constexpr double kClockGHz = 1.85;        // tensor peak 2.25 PF = 148 SM x 8192 flop/clk x 1.855 GHz
constexpr double kL2BytesPerClkSm = 64.0; // sustained L2 -> smem feed per SM
constexpr double kTileOverheadClk = 1800; // pipeline fill + exposed epilogue per wave
constexpr double kLaunchUs = 2.0;

struct Cand {
    uint32_t cta, bn, stages, tiles, grid;
    double us;
};

uint32_t max_stages(uint32_t per_stage, uint32_t fixed, uint32_t budget, uint32_t cap) {
    if (fixed + per_stage > budget) return 0;
    return std::min<uint32_t>(cap, (budget - fixed) / per_stage);
}

// Time of the persistent kernel for one (cta_group, block_n).
Cand eval_common(const DgaTilingPlan& p, uint32_t cta, uint32_t bn) {
    const PlatformInfo& hw = platform();
    const uint32_t in_b = elem_bytes(p.dtype), out_b = elem_bytes(p.outDtype);
    const uint32_t k_elems = 128 / in_b;
    const uint32_t kb = ceil_div(p.k, k_elems);
    Cand c{cta, bn, 0, 0, 0, 1e30};
    const uint32_t fixed = 2 * 128 * epi_cols_host(bn, out_b) * out_b + 1024 + 256;
    c.stages = max_stages(128 * 128 + (bn / cta) * 128, fixed, hw.smemBytes, 12);
    if (c.stages < 2) return c;
    const uint32_t clusters = hw.smCount / cta;
    c.tiles = ceil_div(p.m, 128 * cta) * ceil_div(p.n, bn);
    c.grid = std::min(c.tiles, clusters) * cta;
    const double waves = std::ceil(double(c.tiles) / clusters);
    // tensor-pipe cycles per 128-byte K block: 4 (fp16) MMAs of 128*cta x bn x 16
    // (bn/2 clk each, B300_MICROARCH "tcgen05 floor"); an fp8 block holds twice the K in the same 4 MMAs.
    const double mma_clk = 2.0 * bn;
    const double feed_clk = (128.0 * 128 + (bn / cta) * 128.0) / kL2BytesPerClkSm;
    const double tile_clk = kb * std::max(mma_clk, feed_clk) + kTileOverheadClk;
    const double compute_us = waves * tile_clk / (kClockGHz * 1e3);
    const double bytes = double(p.m) * p.k * in_b + double(p.n) * p.k * in_b + double(p.m) * p.n * out_b;
    const double hbm_us = bytes / (hw.hbmGBs * 1e3);
    c.us = std::max(compute_us, hbm_us) + kLaunchUs;
    return c;
}

}  // namespace

static void fill_common(DgaTilingPlan& p, const Cand& c) {
    const uint32_t in_b = elem_bytes(p.dtype);
    p.kernelSerial = (c.tiles <= platform().smCount / c.cta) ? DGA_KERNEL_SMALL : DGA_KERNEL_COMMON;
    p.ctaGroup = static_cast<uint8_t>(c.cta);
    p.m1 = static_cast<uint16_t>(128 * c.cta);
    p.n1 = static_cast<uint16_t>(c.bn);
    p.k1 = static_cast<uint16_t>(128 / in_b);
    p.stages = static_cast<uint8_t>(c.stages);
    p.splitkFactor = 1;
    p.estTimeUs = c.us;
}

namespace {

// Swap-AB cluster split-K kernel: one CTA owns 128 weight rows x m_pad
// activations x a K slice; `split` CTAs of a cluster reduce through DSMEM.
struct SplitCand {
    uint32_t split, stages, grid;
    double us;
};

SplitCand eval_splitk(const DgaTilingPlan& p, uint32_t split) {
    const PlatformInfo& hw = platform();
    const uint32_t m_pad = round_up(p.m, 16);
    const uint32_t kb = ceil_div(p.k, 64);
    SplitCand c{split, 0, 0, 1e30};
    if (kb < split) return c;
    const uint32_t stage = 128 * 128 + m_pad * 128;
    // Two CTAs share an SM when each stays under half of the shared memory.  Measured on B200 (N=4096, K=7168,
    // profiles/r01_smallm_sweep.log): 256 CTAs at two per SM with a 3-4 stage ring stream faster than 128 CTAs with
    // the deepest ring (M=16: 11.6 vs 13.4 us), so that is preferred whenever a ring of >= 3 stages still fits.
    const uint32_t fixed = 1024 + 256 + splitk_red_bytes(m_pad, split);
    uint32_t per_sm = 2;
    c.stages = max_stages(stage, fixed, hw.smemBytes / 2 - 1024, 4);
    if (c.stages < 3) {
        per_sm = 1;
        c.stages = max_stages(stage, fixed, hw.smemBytes, 12);
    }
    // the drained ring doubles as the fp32 staging tile [m][128] of the reduction
    if (c.stages < 2 || c.stages * stage < m_pad * 512) return c;
    const uint32_t tiles = ceil_div(p.n, 128);
    c.grid = tiles * split;
    // clusters larger than 2 cannot use every SM (GPC granularity)
    const uint32_t usable = (split <= 2 ? hw.smCount : split == 4 ? 132 : 128) * per_sm;
    const double waves = std::ceil(double(c.grid) / usable);
    const double kb_per = std::ceil(double(kb) / split);
    const double cta_bytes = kb_per * stage;
    const uint32_t active = std::min(c.grid, usable);
    // each CTA streams at most ~1/100 of HBM bandwidth; all active CTAs share HBM
    const double per_cta_gbs = std::min(hw.hbmGBs / active, hw.hbmGBs / 100.0);
    const double stream_us = waves * cta_bytes / (per_cta_gbs * 1e3);
    const double mma_us = waves * kb_per * 4 * std::max(m_pad, 32u) / 2.0 / (kClockGHz * 1e3);
    // tail: partial tile -> staging -> bulk DSMEM copies -> owner's sum and store (measured 1.3 us at M=16, 2.8 at M=64)
    const double reduce_us = split > 1 ? 1.0 + m_pad * 0.03 : 0.5;
    // every wave pays its own start-up and reduction tail
    c.us = std::max(stream_us, mma_us) + waves * (reduce_us + 1.5) + kLaunchUs;
    return c;
}

bool aligned16(uint64_t elems, uint32_t bytes) { return (elems * bytes) % 16 == 0; }

}  // namespace

// Walk order for problems whose operands do not fit L2 together.  The ~74 tiles in flight cover `group` blocks of
// one dimension x the rest of the wave along the other.  The panels (block rows x K) of the grouped dimension
// stay resident across the waves of a group; every panel of the other operand is shared by the `group` tiles
// that use it at the same moment and is never touched again inside the group, so it is loaded evict-first and
// cannot push the resident set out.  Measured on fp8 config 4 (profiles/r02_fp8_raster_sweep.log): the resident
// set holds while it stays near 32 MB (the 126 MB L2 is two halves, and lines read from both dies live in both) -
// groups of 8 row blocks with evict-first weights: 373 MB of DRAM traffic against 620 MB for the round-1 walk
// (algorithmic 245 MB); 16 row blocks (64 MB resident) is back at 600 MB.
WalkOrder choose_walk(uint32_t m_blocks, uint32_t n_blocks, uint64_t panel_a_bytes, uint64_t panel_b_bytes, bool m_gt_n) {
    WalkOrder w{8, m_gt_n ? 0u : 1u, 0, 0};   // tiling_params.h:63-64: direction from m > n, fixed group
    const uint64_t a_total = panel_a_bytes * m_blocks, b_total = panel_b_bytes * n_blocks;
    const uint64_t resident = 32ull << 20;
    if (a_total + b_total <= (96ull << 20)) return w;   // both operands stay in L2 whatever the order
    auto group_for = [&](uint64_t panel, uint32_t blocks) {
        return uint32_t(std::max<uint64_t>(1, std::min<uint64_t>(blocks, resident / std::max<uint64_t>(panel, 1))));
    };
    const uint32_t g0 = group_for(panel_a_bytes, m_blocks), g1 = group_for(panel_b_bytes, n_blocks);
    const uint64_t traffic0 = a_total + b_total * ceil_div(m_blocks, g0);   // groups of row blocks, weights stream
    const uint64_t traffic1 = b_total + a_total * ceil_div(n_blocks, g1);   // groups of column blocks, activations stream
    if (traffic0 <= traffic1) w = WalkOrder{g0, 0, 0, 1};
    else w = WalkOrder{g1, 1, 1, 0};
    w.group = std::min<uint32_t>(w.group, 255);
    return w;
}

uint32_t clusters_of_four() {
    // clusters of 4 CTAs that fit beside each other with the persistent kernel's footprint (one CTA per SM): 33 on a
    // B200 (132 of 148 SMs: a cluster lives inside one GPC; tools/probes/cluster_occupancy.cu)
    return platform().smCount * 33 / 148;
}

uint32_t streamk_overhead_kb(uint32_t block_n) {
    // the exposed fix-up chain (giver drains and writes, fence + flag, owner's adding pass) is ~5 us longer than a
    // plain epilogue: ~7500 clocks, in units of one K block of this tile (2 * block_n clocks)
    return std::max(4u, 7500u / (2u * std::max(32u, block_n)));
}

uint32_t streamk_rounds_cost(uint32_t sk, uint32_t split, uint32_t clusters, uint32_t kb, uint32_t overhead_kb) {
    // K blocks of tensor time the slice phase takes: rounds x (slice length + per-slice overhead)
    const uint32_t rounds = ceil_div(sk * split, clusters);
    return rounds * (ceil_div(kb, split) + overhead_kb);
}

void streamk_split(uint32_t tiles, uint32_t clusters, uint32_t kb, uint32_t block_n, uint32_t want_split, uint32_t& dp,
                   uint32_t& sk, uint32_t& split) {
    const uint32_t rem = tiles % clusters;
    dp = tiles;
    sk = 0;
    split = 1;
    if (rem == 0 || kb < 4) return;
    uint32_t best = 0, best_cost = kb;   // cost of leaving the ragged wave whole
    if (want_split >= 2) {
        best = std::min(want_split, std::min(kb, 16u));
    } else {
        for (uint32_t s = 2; s <= 8 && s <= kb; ++s) {
            const uint32_t cost = streamk_rounds_cost(rem, s, clusters, kb, streamk_overhead_kb(block_n));
            if (cost < best_cost) { best_cost = cost; best = s; }
        }
    }
    if (best < 2) return;
    sk = rem;
    dp = tiles - rem;
    split = best;
}

uint64_t streamk_partial_bytes(uint32_t sk, uint32_t split, uint32_t cta_group, uint32_t block_n) {
    return uint64_t(sk) * split * cta_group * 128 * block_n * 4;
}

PadLayout pad_layout(const DgaTilingPlan& p) {
    const uint64_t in_b = elem_bytes(p.dtype), out_b = elem_bytes(p.outDtype);
    auto pitch = [](uint64_t cols) { return (cols + 63) / 64 * 64; };
    auto region = [](uint64_t b) { return (b + 1023) & ~1023ull; };
    PadLayout l{};
    l.ldA = pitch(p.k);
    l.ldB = pitch(p.layoutTagB ? p.k : p.n);
    l.ldC = pitch(p.n);
    uint64_t off = 0;
    l.offA = off;
    if (p.paddingTagA) off += region(uint64_t(p.m) * l.ldA * in_b);
    l.offB = off;
    if (p.paddingTagB) off += region(uint64_t(p.layoutTagB ? p.n : p.k) * l.ldB * in_b);
    l.offC = off;
    if (p.paddingTagC) off += region(uint64_t(p.m) * l.ldC * out_b);
    l.bytes = off;
    return l;
}

int finalize_plan(DgaTilingPlan& p) {
    const PlatformInfo& hw = platform();
    const uint32_t in_b = elem_bytes(p.dtype), out_b = elem_bytes(p.outDtype);
    // PaddingCommon is Common behind a repack: the tags say what is repacked, the tile fields mean the same
    if (p.kernelSerial == DGA_KERNEL_PADDING_COMMON) p.kernelSerial = DGA_KERNEL_COMMON;
    if (p.kernelSerial == DGA_KERNEL_DECODE_WS) {
        const uint32_t m_pad = round_up(p.m, 16);
        const uint32_t kb = ceil_div(p.k, 64);
        if (m_pad > 256 || p.layoutTagB != 1 || in_b != 2) {
            set_error("the decode kernel needs m <= 256, 16-bit inputs and a [N,K] weight");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        if (p.splitkFactor == 0) p.splitkFactor = 1;
        if (p.splitkFactor > 16 || p.splitkFactor > kb) {
            set_error("decode split must be <= min(16, K blocks)");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        const uint32_t stage = 128 * 128 + m_pad * 128;
        // two CTAs per SM whenever a ring of >= 3 stages fits half of the shared memory (no reduction slots here)
        uint32_t cap = max_stages(stage, 1280, hw.smemBytes / 2 - 1024, 6);
        if (cap < 3) cap = max_stages(stage, 1280, hw.smemBytes, 12);
        if (p.stages == 0 || p.stages > cap) p.stages = static_cast<uint8_t>(cap);
        if (p.stages < 2) {
            set_error("decode plan does not fit shared memory");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        p.m1 = static_cast<uint16_t>(m_pad);
        p.n1 = 128;
        p.k1 = 64;
        p.ctaGroup = 1;
        p.clusterShape = 1;
        p.blockDim = ceil_div(p.n, 128) * p.splitkFactor;
        p.smemBytes = splitk_smem_bytes(m_pad, p.stages, 1);
        p.workspaceBytes = splitk_ws_bytes(p.m, p.n, p.splitkFactor);
        p.paddingTagC = 0;   // the reduce kernel stores C with plain st.global: any pitch will do
    } else if (p.kernelSerial == DGA_KERNEL_SPLITK) {
        const uint32_t m_pad = round_up(p.m, 16);
        if (m_pad > 256 || p.layoutTagB != 1 || in_b != 2) {
            set_error("split-K small-M kernel needs m <= 256, 16-bit inputs and a [N,K] weight");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        if (p.splitkFactor == 0) p.splitkFactor = 1;
        const uint32_t fixed = 1280 + splitk_red_bytes(m_pad, p.splitkFactor);
        const uint32_t cap = max_stages(128 * 128 + m_pad * 128, fixed, hw.smemBytes, 12);
        if (p.stages == 0 || p.stages > cap) p.stages = static_cast<uint8_t>(cap);
        if (p.stages < 2 || p.stages * (128 * 128 + m_pad * 128) < m_pad * 512) {
            set_error("split-K plan does not fit shared memory");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        p.m1 = static_cast<uint16_t>(m_pad);
        p.n1 = 128;
        p.k1 = 64;
        p.ctaGroup = 1;
        p.blockDim = ceil_div(p.n, 128) * p.splitkFactor;
        p.smemBytes = splitk_smem_bytes(m_pad, p.stages, p.splitkFactor);
        p.workspaceBytes = 0;
        p.paddingTagC = 0;   // this kernel stores C with plain st.global: any pitch will do
        p.clusterShape = static_cast<uint8_t>(p.splitkFactor);   // the K slices of one weight tile form the cluster
    } else {
        if (p.ctaGroup != 1 && p.ctaGroup != 2) p.ctaGroup = 2;
        const uint32_t cta = p.ctaGroup;
        // widths that are an odd multiple of 16 (240, 208, ...: UMMA N is legal in steps of 16 for both CTA groups, each
        // CTA of a pair then holds n1 / 2 = whole 8-row swizzle atoms of B): whole-tile plans over a K-major weight
        // only; the partial-tile slots of stream-K and the wide-row epilogue assume 32-column chunks
        const bool odd16 = p.n1 % 32 == 16;
        const bool odd16_ok = p.n1 > 32 && p.layoutTagB == 1 &&
                              (p.kernelSerial == DGA_KERNEL_COMMON || p.kernelSerial == DGA_KERNEL_SMALL ||
                               p.kernelSerial == DGA_KERNEL_PADDING_COMMON);
        if (p.n1 == 0 || p.n1 > 256 || p.n1 % 16 != 0 || (odd16 && !odd16_ok) || (p.layoutTagB == 0 && (p.n1 / cta) % 64 != 0)) {
            set_error("invalid n1 for the persistent kernel: " + std::to_string(p.n1));
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        p.m1 = static_cast<uint16_t>(128 * cta);
        p.k1 = static_cast<uint16_t>(128 / in_b);
        const bool split_reduce = p.kernelSerial == DGA_KERNEL_SPLITK_REDUCE;
        const uint32_t epi_b = split_reduce ? 4u : out_b;   // split-K slices leave fp32 partial tiles
        const uint32_t fixed = 2 * 128 * epi_cols_host(p.n1, epi_b) * epi_b + 1024 + 256;
        const uint32_t cap = max_stages(128 * 128 + (p.n1 / cta) * 128, fixed, hw.smemBytes, 12);
        if (p.stages == 0 || p.stages > cap) p.stages = static_cast<uint8_t>(cap);
        if (p.stages < 2) {
            set_error("tile does not fit shared memory");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        const uint32_t tiles = ceil_div(p.m, 128 * cta) * ceil_div(p.n, p.n1);
        p.blockDim = std::min(tiles, hw.smCount / cta) * cta;
        p.smemBytes = common_smem_bytes(p.n1, cta, p.stages, epi_b, in_b);
        p.workspaceBytes = 0;
        if (p.kernelSerial == DGA_KERNEL_CLUSTER_SPLITK) {
            // two CTA pairs per cluster on ONE tile, each one half of K; halves exchanged through DSMEM (gemm_sm100.cuh, kKs)
            const uint32_t kb = ceil_div(p.k, 128 / in_b);
            const uint32_t ring = uint32_t(p.stages) * (128 * 128 + (p.n1 / cta) * 128);
            if (cta != 2 || in_b != 2 || (p.n1 != 128 && p.n1 != 256) || tiles > clusters_of_four() || kb < 2 ||
                ring < 2u * 128 * (p.n1 / 2) * 4) {
                set_error("the cluster split-K variant needs ctaGroup 2, n1 of 128 or 256, at most 33 tiles and K >= 128");
                return ACLNN_ERR_INNER_TILING_ERROR;
            }
            p.blockDim = tiles * 4;
            p.clusterShape = 4;
            p.splitkFactor = 2;
        } else if (p.clusterShape == 4 && p.kernelSerial != DGA_KERNEL_CLUSTER_SPLITK) {
            // two CTA pairs per cluster on neighbouring N tiles, A multicast between them (gemm_sm100.cuh, kMc)
            const bool whole = p.kernelSerial == DGA_KERNEL_COMMON || p.kernelSerial == DGA_KERNEL_SMALL ||
                               p.kernelSerial == DGA_KERNEL_PADDING_COMMON;
            if (cta != 2 || !whole || in_b != 2 || ceil_div(p.n, p.n1) % 2 != 0) {
                set_error("clusterShape 4 needs a whole-tile 16-bit plan with ctaGroup 2 and an even number of N tiles");
                return ACLNN_ERR_INNER_TILING_ERROR;
            }
            p.blockDim = std::min(tiles / 2, clusters_of_four()) * 4;
        } else {
            p.clusterShape = static_cast<uint8_t>(cta);
        }
        if (split_reduce) {
            const uint32_t kb = ceil_div(p.k, 128 / in_b);
            if (p.splitkFactor < 2 || p.splitkFactor > 16 || p.splitkFactor > kb) {
                set_error("split-K needs 2 <= splitkFactor <= min(16, K blocks)");
                return ACLNN_ERR_INNER_TILING_ERROR;
            }
            p.blockDim = uint32_t(std::min<uint64_t>(uint64_t(tiles) * p.splitkFactor, hw.smCount / cta)) * cta;
            p.workspaceBytes = splitk_ws_bytes(p.m, p.n, p.splitkFactor);
            p.paddingTagC = 0;   // the reduce kernel stores C with plain st.global: any pitch will do
        } else if (p.kernelSerial == DGA_KERNEL_STREAMK) {
            const uint32_t clusters = hw.smCount / cta;
            uint32_t dp, sk, split;
            streamk_split(tiles, clusters, ceil_div(p.k, 128 / in_b), p.n1, p.splitkFactor, dp, sk, split);
            if (sk == 0) {
                p.kernelSerial = tiles <= clusters ? DGA_KERNEL_SMALL : DGA_KERNEL_COMMON;
                p.splitkFactor = 1;
            } else {
                p.splitkFactor = uint16_t(split);
                p.blockDim = uint32_t(std::min<uint64_t>(uint64_t(dp) + uint64_t(sk) * split, clusters)) * cta;
                p.workspaceBytes = streamk_partial_bytes(sk, split, cta, p.n1) + ((uint64_t(sk) * split * cta * 4 + 255) & ~255ull);
            }
        } else if (p.kernelSerial == DGA_KERNEL_CLUSTER_SPLITK) {
            // grid, cluster shape and split were set above
        } else if (p.kernelSerial != DGA_KERNEL_SMALL && p.kernelSerial != DGA_KERNEL_COMMON) {
            set_error("kernel variant " + std::to_string(int(p.kernelSerial)) + " is not built");
            return ACLNN_ERR_INNER_TILING_ERROR;
        } else {
            p.kernelSerial = tiles <= (p.clusterShape == 4 ? 2 * clusters_of_four() : hw.smCount / cta)
                                 ? DGA_KERNEL_SMALL : DGA_KERNEL_COMMON;
        }
        if (p.splitkFactor == 0) p.splitkFactor = 1;
    }
    if (p.swizzleOffset == 0) p.swizzleOffset = 8;
    if (p.paddingTagA || p.paddingTagB || p.paddingTagC) {
        // repack region first, the variant's own workspace behind it (padding_common_matmul_kernel.h:77-82)
        p.workspaceBytes += pad_layout(p).bytes;
        if (p.kernelSerial == DGA_KERNEL_COMMON || p.kernelSerial == DGA_KERNEL_SMALL)
            p.kernelSerial = DGA_KERNEL_PADDING_COMMON;
    }
    return ACLNN_SUCCESS;
}

int select_kernel(DgaTilingPlan& p) {
    if (p.m == 0 || p.n == 0 || p.k == 0) {
        set_error("empty problem: m, n, k must be positive");
        return ACLNN_ERR_PARAM_INVALID;
    }
    const uint32_t in_b = elem_bytes(p.dtype), out_b = elem_bytes(p.outDtype);
    p.layoutTagA = 0;
    p.layoutTagC = 0;
    // m == 1 / n == 1 canonicalisation (select_kernel.cpp:340-347): a [K,1] matrix is both row- and
    // column-major; prefer the K-major reading, which every kernel here supports.
    if (p.n == 1) p.layoutTagB = 1;
    p.strideA = p.k;
    p.strideB = p.layoutTagB ? p.k : p.n;
    p.strideC = p.n;
    p.swizzleOffset = 8;
    p.swizzleDirection = (p.m > p.n) ? 0 : 1;  // tiling_params.h:63-64
    p.splitkFactor = 1;
    p.stages = 0;
    p.clusterShape = 0;   // = ctaGroup; the 4-CTA multicast cluster is taken by rule further down
    p.paddingTagA = aligned16(p.strideA, in_b) ? 0 : 1;
    p.paddingTagB = aligned16(p.strideB, in_b) ? 0 : 1;
    p.paddingTagC = aligned16(p.strideC, out_b) ? 0 : 1;

    // --- persistent kernel candidates (wave balancing over 74 pairs / 148 CTAs)
    Cand best{0, 0, 0, 0, 0, 1e30};
    for (uint32_t cta = 2; cta >= 1; --cta) {
        // up to 128 rows a pair tile spends half of its tensor work on rows that do not exist (audit: M=128,
        // N=32768, K=512: 10.0 us single-CTA vs 11.8 us pair; N=24576, K=1536: 17.5 vs 19.8 us)
        if (cta == 2 && p.m <= 128) continue;
        const uint32_t step = (p.layoutTagB == 0) ? 64 * cta : 16;
        for (uint32_t bn = 256; bn >= step && bn >= 32; bn -= step) {
            // Odd multiples of 16 (240, 208, ...) only where they cut N into as many column tiles as the next multiple
            // of 32: same tiles, same waves, every tile cheaper and fewer bytes through the crossbar (2048 x 2112 x
            // 7168: 240 vs 256, 44.4 vs 45.5 us; 4096 x 2112 x 7168: 85.5 vs 88.2 us).  Where they add tiles the wave
            // model over-rates them (2048 x 4096 x 7168: promised 6 %, measured 0.3 %; 1024 x 4096 x 7168: -0.6 %;
            // profiles/r01d_odd16_tiles.log), so those are left to the tuned table.
            if (bn % 32 != 0 && (bn < 48 || ceil_div(p.n, bn) != ceil_div(p.n, bn + 16))) continue;
            Cand c = eval_common(p, cta, bn);
            // prefer (on ties within 1 %) the pair mode and the larger tile
            if (c.us < best.us * 0.99) best = c;
        }
    }
    if (best.cta == 0) {
        set_error("no feasible tile");
        return ACLNN_ERR_INNER_TILING_ERROR;
    }
    // --- small-M swap-AB split-K candidates
    SplitCand sbest{0, 0, 0, 1e30};
    if (p.m <= 256 && p.layoutTagB == 1 && in_b == 2) {
        // ties go to the larger split: more CTAs in flight hide each other's start-up and reduction tails
        for (uint32_t s : {8u, 4u, 2u, 1u}) {
            SplitCand c = eval_splitk(p, s);
            if (c.us < sbest.us) sbest = c;
        }
    }
    // M <= 64 is HBM-bound by the weight: a 256-row pair tile would waste >= 3/4 of the tensor work and, worse,
    // most of the SMs' load slots, so the streaming kernel always wins there; above that the estimates decide.
    // measured on B200 (N=4096, K=7168): the swap-AB kernel also wins at M = 128 (17.7 vs 28.9 us); the estimates
    // are too coarse there, so the rule is explicit.  Beyond 128 rows the fp32 reduction slots (m x 512 B per CTA)
    // crowd out the ring and the estimates decide.
    // Multicore split-K + reduce kernel (the reference's declared kernel type 3): when even 256 x 256 pair tiles fill
    // at most a quarter of the CTA pairs and K is long, cutting K gives every pair a slice of a wide tile instead of
    // leaving most SMs idle behind narrow ones.  Measured on B200 (profiles/r01_splitk_reduce_sweep.log): M=256,
    // N=4096, K=7168: 25.2 vs 27.9 us; M=256, N=2048: 16.4 vs 27.2 us; with more tiles than that it loses
    // (M=512, N=4096: 38.5 vs 29.3 us), hence the explicit rule.
    const uint32_t pairs = platform().smCount / 2;
    const uint32_t tiles256 = ceil_div(p.m, 256) * ceil_div(p.n, 256);
    const uint32_t kb_all = ceil_div(p.k, 128 / in_b);
    const uint32_t tiles128 = ceil_div(p.n, 128);
    // slices of >= 8 K blocks once K >= 3584 (audit: 256 x 1024 x 4096 with 8 slices 12.1 us vs 13.6 with 4), >= 14 below
    const uint32_t split_r = std::min<uint32_t>(std::min<uint32_t>(8, pairs / std::max(1u, tiles256)),
                                                kb_all >= 56 ? kb_all / 8 : kb_all / 14);
    // ... except where the weight is wide and K short (N >= 6144, K <= 3072 at 16 bit): the decode kernel's CTAs then
    // run a handful of K blocks each and the persistent kernel has enough tiles to fill the machine (selector audit,
    // profiles/r01c_selector_audit.log: M=128, N=7168, K=2048: 12.0 vs 13.4 us; N=24576, K=1536: 17.5 vs 20.5 us;
    // N=32768, K=512: 10.0 vs 13.0 us).
    const bool wide_short = p.m > 64 && tiles128 >= 48 && kb_all <= 48;
    // Stream-K window (role of PaddingStreamkMatmulKernel): a pair tile 256 x n1 whose tiles, cut into two K slices,
    // fill 75-100 % of the pairs in ONE round - half the K loop per pair, short fix-up, no second round.  Of the tile
    // widths that qualify the one filling most pairs wins (ties: the wider).  Measured (profiles/
    // r01_streamk_aligned_sweep.log, r01c_selector_audit.log): M=256, N=4096, K=7168: 21.9 vs 24.8 us for split-K +
    // reduce; K=2048: 13.2 vs 18.9 us; 512 x 2112 x 7168: 21.5 vs 23.2; 768 x 1536 x 7168: 21.6 vs 23.6; 1024 x 1024 x
    // 4096: 16.4 vs 19.3; 1024 x 1536 x 7168 (n1 = 192): 24.9 vs 27.5 plain; 2048 x 576 x 7168 (n1 = 160): 23.0 vs 27.4.
    uint32_t sk_n1 = 0, sk_items = 0;
    // ... and only where the plain walk is in trouble: its best tile is narrower than 128 columns (poor L2 intensity)
    // or leaves more than 40 % of the pairs idle.  With 56-64 plain tiles of 256 x 128 the single pass wins
    // (256 x 7168 x 2048: 12.1 vs 18.2 us; 512 x 4096 x 7168: 28.0 vs 30.8 us).
    // Row-major B (the framework layout) is loaded in 64 x 64 boxes and a 256 x 128 tile is much slower there than
    // with a [N, K] weight (512 x 4096 x 7168: 35.2 us plain vs 30.6 us stream-K on 256 x 256 tiles), so anything
    // narrower than 256 counts as weak once K >= 4096 (at K = 2048 the single pass still wins: 256 x 7168 x 2048: 13.3
    // vs 18.4 us).
    const bool plain_weak = best.bn < ((p.layoutTagB == 0 && kb_all >= 64) ? 256u : 128u) ||
                            5 * best.tiles < 3 * (platform().smCount / best.cta);
    if (p.m > 128 && kb_all >= 32 && plain_weak) {
        const uint32_t step = p.layoutTagB == 0 ? 128 : 32;
        for (uint32_t bn = 256; bn >= 128; bn -= step) {
            const uint32_t items = 2 * ceil_div(p.m, 256) * ceil_div(p.n, bn);
            if (items <= pairs && 4 * items >= 3 * pairs && items > sk_items) { sk_items = items; sk_n1 = bn; }
        }
    }
    const bool rule_sk = sk_n1 != 0;
    // (with only 2-3 slices the second launch costs more than the split saves: 256 x 2048 x 2048: 15.6 vs 11.5 us)
    const bool rule_reduce = p.m > 128 && tiles256 * 4 <= pairs && split_r >= 4;
    // Row-major mat2 (the framework functions' layout) has no swap-AB decode kernel; up to 128 rows and long K the
    // weight is streamed by K slices of 128 x 128 single-CTA tiles + the reduce kernel instead, as many slices as
    // keep the grid within the SM count (audit, profiles/r01c_selector_audit.log: 16 x 4096 x 7168: 15.6 vs 31.6 us
    // plain; 16 x 2112 x 7168: 11.4 vs 31.2; 128 x 7168 x 16384: 49.6 vs 69.2).
    // Cluster split-K (kernelSerial 6): two CTA pairs per tile, one half of K each, fp32 halves exchanged through DSMEM.
    // Where 256-row tiles of width 128 or 256 fill 24..33 clusters of four (most of the machine) and K is long: the wide
    // tile keeps the bytes per flop one SM can ingest, the K split fills the SMs, and the fix-up never leaves the chip.
    // Measured (CUDA graph of 20, profiles/r02_cluster_splitk_ab.log): 512 x 4096 x 7168 26.0 vs 28.5 us plain; 256 x 4096 x
    // 7168 18.7 vs 23.3 stream-K; 512 x 2048 x 7168 18.3 vs 22.8; 1024 x 2048 x 7168 25.3 vs 27.7; 768 x 2048 x 7168 24.3 vs 26.8;
    // 256 x 7168 x 7168 26.2 vs 27.8.  Not for K <= 4096 (512 x 4096 x 4096 17.9 = 17.9; K = 2048: 13.2 vs 11.6) and not with
    // fewer tiles (256 x 2048 x 7168, 16 tiles: 17.0 vs 16.6 for split-K + reduce).
    uint32_t ks_n1 = 0;
    if (p.m > 128 && p.layoutTagB == 1 && in_b == 2 && kb_all >= 96) {
        for (uint32_t n1 : {128u, 256u}) {
            const uint32_t t = ceil_div(p.m, 256) * ceil_div(p.n, n1);
            if (t >= 24 && t <= clusters_of_four()) { ks_n1 = n1; break; }
        }
    }
    const uint32_t sms = platform().smCount;
    const bool rowmajor_small = p.layoutTagB == 0 && p.m <= 128 && in_b == 2 && kb_all >= 56 && 2 * tiles128 <= sms;
    if (ks_n1 != 0) {
        p.kernelSerial = DGA_KERNEL_CLUSTER_SPLITK;
        p.ctaGroup = 2;
        p.n1 = static_cast<uint16_t>(ks_n1);
        p.stages = 0;
        p.estTimeUs = best.us;
    } else if (rowmajor_small) {
        p.kernelSerial = DGA_KERNEL_SPLITK_REDUCE;
        p.ctaGroup = 1;
        p.n1 = 128;
        p.stages = 0;
        p.splitkFactor = static_cast<uint16_t>(8 * tiles128 <= sms ? 8 : (4 * tiles128 <= sms ? 4 : 2));
        p.estTimeUs = best.us;
    } else if (sbest.split != 0 && ((p.m <= 128 && !wide_short) || (p.m > 128 && !rule_sk && !rule_reduce && sbest.us < best.us))) {
        p.kernelSerial = DGA_KERNEL_SPLITK;
        p.splitkFactor = static_cast<uint16_t>(sbest.split);
        p.stages = static_cast<uint8_t>(sbest.stages);
        p.estTimeUs = sbest.us;
    } else if (rule_sk) {
        p.kernelSerial = DGA_KERNEL_STREAMK;
        p.ctaGroup = 2;
        p.n1 = static_cast<uint16_t>(sk_n1);
        p.stages = 0;
        p.splitkFactor = 2;
        p.estTimeUs = best.us;
    } else if (rule_reduce) {
        p.kernelSerial = DGA_KERNEL_SPLITK_REDUCE;
        p.ctaGroup = 2;
        p.n1 = 256;
        p.stages = 0;
        p.splitkFactor = static_cast<uint16_t>(split_r);
        p.estTimeUs = best.us;
    } else {
        fill_common(p, best);
        // Stream-K on the chosen tile when the last wave is ragged and K is long enough to pay for the fix-up
        // (audit: 2048 x 7168 x 16384: 279 vs 326 us; 4096 x 7168 x 16384: 575 vs 601 us; 4096 x 4096 x 7168: 143 vs 147 us).
        const uint32_t clusters = platform().smCount / best.cta;
        const uint32_t full = best.tiles / clusters, rem = best.tiles % clusters;
        if (best.cta == 2 && best.bn % 32 == 0 && full >= 1 && rem != 0 && 2 * rem <= clusters && kb_all >= 64) {
            // two slices per ragged tile, one round: the only configuration the audit shows winning (more slices or
            // a second round and the owners' adding passes eat the gain: 1024 x 7168 x 16384 with 5 slices: 234 vs 148 us)
            const uint32_t cost = streamk_rounds_cost(rem, 2, clusters, kb_all, streamk_overhead_kb(best.bn));
            if (cost < kb_all && double(kb_all - cost) >= 0.03 * double(full + 1) * kb_all) {
                p.kernelSerial = DGA_KERNEL_STREAMK;
                p.splitkFactor = 2;
                p.stages = 0;
            }
        }
    }
    // PaddingCommon survives only as a leading-dimension repack (padding_common_matmul_kernel.h:64-82): the tags
    // above describe contiguous operands; the launch path re-derives them from the caller's actual strides.
    return finalize_plan(p);
}

}  // namespace dga
