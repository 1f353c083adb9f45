// Shape-specialised entry of the persistent GEMM kernel, compiled at run time (jit.cpp) from a generated
// translation unit that defines the DGA_JIT_* constants and then includes this file.
//
// Role of the reference's generated kernel source (framework/csrc/jit/generate_code.hpp:123-369: a fixed kernel
// text behind a block of `constexpr uint32_t` shape constants formatted in per call, gemm.hpp:44-61): here the
// kernel text is the same device function the dynamic-shape library kernel wraps (gemm_sm100_body); the
// generated unit pins every shape-derived field of GemmParams to a literal, so the compiler folds the tile
// counts, the K-loop trip count, the instruction descriptor, the ring depth and the rasterisation arithmetic.
// Pointers and debug fields stay run-time values.
#pragma once
#include "gemm_sm100.cuh"

#ifndef DGA_JIT_KEEP_RT
#define DGA_JIT_KEEP_RT 0
#endif
#ifndef DGA_JIT_SPLIT_K
#define DGA_JIT_SPLIT_K 1u
#endif
#ifndef DGA_JIT_M
#error "generated unit must define the DGA_JIT_* constants before including jit_entry.cuh"
#endif

extern "C" __global__ void __launch_bounds__(dga::kNumThreads, 1)
dga_jit_gemm(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
             const __grid_constant__ CUtensorMap tm_d, const __grid_constant__ CUtensorMap tm_sfa,
             const __grid_constant__ CUtensorMap tm_sfb, const __grid_constant__ dga::PeerMaps peers,
             const dga::GemmParams rt) {
    dga::GemmParams p = rt;   // run-time part: gate, trace, debug, workspace pointers
    // DGA_JIT_KEEP_RT (developer knob, $DGA_JIT_KEEP_RT at generation time): bit set = that group of fields stays a
    // run-time value, to measure what each group of constants is worth.
    if (!(DGA_JIT_KEEP_RT & 1)) {
        p.sk_tiles = 0;
        p.sk_partials = nullptr;
        p.sk_sems = nullptr;
    }
    if (!(DGA_JIT_KEEP_RT & 2)) p.epi_cols = DGA_JIT_EPI_COLS;
    if (!(DGA_JIT_KEEP_RT & 4)) {
        p.block_n = DGA_JIT_BLOCK_N;
        p.num_stages = DGA_JIT_STAGES;
        p.b_stage_bytes = DGA_JIT_B_STAGE_BYTES;
    }
    if (!(DGA_JIT_KEEP_RT & 8)) {
        p.m = DGA_JIT_M;
        p.n = DGA_JIT_N;
        p.k = DGA_JIT_K;
        p.k_blocks = DGA_JIT_K_BLOCKS;
        p.k_block_elems = DGA_JIT_K_BLOCK_ELEMS;
    }
    if (!(DGA_JIT_KEEP_RT & 16)) {
        p.num_m_blocks = DGA_JIT_M_BLOCKS;
        p.num_n_blocks = DGA_JIT_N_BLOCKS;
        p.raster_group = DGA_JIT_RASTER_GROUP;
        p.raster_along_m = DGA_JIT_RASTER_ALONG_M;
        p.dp_tiles = DGA_JIT_DP_TILES;
        p.tiles_per_batch = DGA_JIT_TILES_PER_BATCH;
    }
    if (!(DGA_JIT_KEEP_RT & 32)) p.num_peers = 0;
    if (!(DGA_JIT_KEEP_RT & 64)) p.idesc = DGA_JIT_IDESC;
    p.sf_n_chunks = 0;
    p.sf_m_chunks = 0;
    p.split_k = DGA_JIT_SPLIT_K;
    p.group_index = nullptr;
    dga::gemm_sm100_body<DGA_JIT_CTA, DGA_JIT_OUT_T, DGA_JIT_B_MN, false, DGA_JIT_MODE>(tm_a, tm_b, tm_d, tm_sfa, tm_sfb,
                                                                                       peers, p);
}
