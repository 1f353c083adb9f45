// Host tile selector for B200: picks kernel variant, CTA tile, CTA-pair mode,
// K split, smem ring depth and rasterisation for a runtime (M, N, K).
// Role of CatlassTiling::SelectKernel + DoTilingLayout* + GetPaddingTag in the
// reference (op_host/op_tiling/select_kernel.cpp:333-369, do_tiling.cpp:20-204,
// select_kernel.cpp:82-268).  Pure host C++; no CUDA calls.
#pragma once
#include <cstdint>
#include <string>

#include "../../include/dga_b200.h"

namespace dga {

struct PlatformInfo {
    uint32_t smCount = 148;
    uint32_t smemBytes = 232448;  // 227 KB opt-in dynamic shared memory per CTA
    uint32_t tmemCols = 512;
    uint64_t l2Bytes = 126ull << 20;
    double hbmGBs = 6500.0;         // measured copy bandwidth class
    double denseTflops16 = 1600.0;  // measured cuBLAS burst class
    double denseTflops8 = 3200.0;
};

PlatformInfo& platform();

inline uint32_t ceil_div(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
inline uint32_t round_up(uint32_t a, uint32_t b) { return ceil_div(a, b) * b; }
inline uint32_t elem_bytes(int32_t dtype) {
    return dtype == ACL_FLOAT ? 4u : dtype == ACL_FLOAT8_E4M3FN ? 1u : 2u;
}

// Columns per output store of the persistent kernel (mirrors epi_cols_for() in gemm_sm100.cuh).
inline uint32_t epi_cols_host(uint32_t block_n, uint32_t out_bytes) { return (out_bytes == 2 && block_n % 64 == 0) ? 64u : 32u; }
// Smem bytes the persistent kernel needs for a (block_n, stages) choice;
// mirrors make_smem_layout() in gemm_sm100.cuh (+1 KB alignment slack).
uint32_t common_smem_bytes(uint32_t block_n, uint32_t cta_group, uint32_t stages, uint32_t out_bytes,
                           uint32_t in_bytes);
// Smem bytes of the swap-AB split-K kernel (gemm_smallm_sm100.cuh).
uint32_t splitk_smem_bytes(uint32_t m_pad, uint32_t stages, uint32_t split);

// Fills every field of `plan` from (m, n, k, dtype, outDtype, layoutTagB). Returns ACLNN_SUCCESS or an error.
int select_kernel(DgaTilingPlan& plan);
// Re-derives grid / smem / workspace for a plan whose tile fields were set by hand (tuning sweeps, CSV rows).
int finalize_plan(DgaTilingPlan& plan);
int select_kernel_with_cache(DgaTilingPlan& plan);

// Rasterisation of the persistent kernels (role of swizzleOffset / swizzleDirection, tiling_params.h:63-64) chosen
// from the bytes of one block row / block column panel (rows x K) and the L2 size; hints: 0 normal, 1 evict-first.
struct WalkOrder {
    uint32_t group, along_m, hint_a, hint_b;
};
WalkOrder choose_walk(uint32_t m_blocks, uint32_t n_blocks, uint64_t panel_a_bytes, uint64_t panel_b_bytes, bool m_gt_n);

// Clusters of four CTAs (clusterShape 4) the device places side by side at one CTA per SM.
uint32_t clusters_of_four();

// Stream-K decomposition of `tiles` output tiles of `kb` K blocks over `clusters` clusters: the first `dp` tiles are
// walked whole (one per cluster and wave); each of the last `sk` tiles (the ragged wave) is cut into `split` K slices
// with aligned boundaries, dealt out round-robin.  want_split >= 2 forces the slice count (hand-set plans), else the
// count that minimises rounds x slice length is chosen; sk == 0 when the tiles already fill whole waves or no
// slice count beats leaving the last wave whole.
void streamk_split(uint32_t tiles, uint32_t clusters, uint32_t kb, uint32_t block_n, uint32_t want_split, uint32_t& dp,
                   uint32_t& sk, uint32_t& split);
// Workspace layout of the stream-K variant: one fp32 partial slot per slice and CTA rank, then the per-tile counters.
uint64_t streamk_partial_bytes(uint32_t sk, uint32_t split, uint32_t cta_group, uint32_t block_n);

// Multicore split-K (DGA_KERNEL_SPLITK_REDUCE): row pitch (floats) and bytes of the [split][m][pitch] fp32 partial slabs.
inline uint32_t splitk_ws_pitch(uint32_t n) { return round_up(n, 32); }
inline uint64_t splitk_ws_bytes(uint32_t m, uint32_t n, uint32_t split) {
    return uint64_t(split) * m * splitk_ws_pitch(n) * 4;
}

// Stride repack (role of PaddingWorkspaceSize, op_kernel/kernel/kernel_utils.h:63-75, and of the workspace carve-up
// padA | padB | padC in padding_common_matmul_kernel.h:77-82): operands whose paddingTag is set are copied to the
// front of the workspace with a 128-byte-multiple row pitch; whatever else the variant needs (stream-K partials)
// follows at `bytes`.  Pure function of (m, n, k, dtypes, layoutTagB, padding tags).
struct PadLayout {
    uint64_t ldA, ldB, ldC;      // padded leading dimensions, elements
    uint64_t offA, offB, offC;   // byte offsets in the workspace
    uint64_t bytes;              // total bytes of the repack region (0 when no tag is set), 1 KB granular
};
PadLayout pad_layout(const DgaTilingPlan& plan);

// plan_cache.cpp
int cache_get(DgaTilingPlan& plan);  // keyed by (m, n, k, dtype, outDtype, layoutTagB) already set in `plan`
int cache_set(const DgaTilingPlan& plan);
void cache_clear();
uint64_t cache_size();
int cache_load_csv(const char* path);
int cache_save_csv(const char* path);

void set_error(const std::string& msg);
const char* last_error();

}  // namespace dga
