// oracle/_ref/libref_host.so: the REFERENCE's own host code for the path, compiled where it lies under
// /root/reference (nothing is copied), behind a tiny C ABI so tests and tests/golden/make_golden.py can call it.
// TEST INFRASTRUCTURE ONLY. Only tests/, __graft_entry__ and bench.py's reference leg may load this.
//
// Wrapped reference functions:
//   CatlassTiling::SelectKernel (aclnn_catlass_dynamic_matmul/op_host/op_tiling/select_kernel.cpp:333-369)
//     on a TilingParams built the way the aclnn tiling function builds it
//     (op_host/catlass_dynamic_matmul_tiling.cpp:102-106) or with caller-given layout tags;
//   deep_gemm_ascend::get_bench_config (deep_gemm_ascend/framework/csrc/jit/get_best_config.hpp:93-153),
//     emitted in the order mmad_bench writes params[6:28] (jit_kernels/impls/gemm_bench.hpp:66-69);
//   CeilDiv / RoundUp / GetMaxK1 / JudgeSpace / BalanceWorkload (op_tiling/utils.cpp).
#include <cstdint>

#include "select_kernel.h"
#include "tiling_params.h"
#include "utils.h"
#include "get_best_config.hpp"

extern "C" {

// out[17]: m1 n1 k1 kernelSerial paddingTagA paddingTagB paddingTagC blockDim swizzleOffset swizzleDirection
//          splitkFactor strideA strideB strideC layoutTagA layoutTagB layoutTagC (after canonicalisation)
void ref_select_kernel(uint32_t m, uint32_t n, uint32_t k, int layoutA, int layoutB, uint64_t* out) {
    TilingParams t(m, n, k, static_cast<LayoutTag>(layoutA), static_cast<LayoutTag>(layoutB), LayoutTag::TagRowMajor);
    CatlassTiling::SelectKernel(t);
    out[0] = t.m1; out[1] = t.n1; out[2] = t.k1; out[3] = t.kernelSerial;
    out[4] = t.paddingTagA; out[5] = t.paddingTagB; out[6] = t.paddingTagC; out[7] = t.blockDim;
    out[8] = t.swizzleOffset; out[9] = t.swizzleDirection; out[10] = t.splitkFactor;
    out[11] = t.strideA; out[12] = t.strideB; out[13] = t.strideC;
    out[14] = t.layoutTagA; out[15] = t.layoutTagB; out[16] = t.layoutTagC;
}

// params[0..5] = knobs in; params[6..27] out.
void ref_get_bench_config(uint32_t m, uint32_t n, uint32_t k, int32_t* params) {
    const auto c = deep_gemm_ascend::get_bench_config(m, n, k, params[0], params[1], params[2], params[3], params[4],
                                                      params[5]);
    const uint32_t v[22] = {c.m, c.n, c.k, c.batch, c.k_iters, c.m_blocks, c.n_blocks, c.k_blocks, c.m_sc_blocks,
                            c.n_sc_blocks, c.m_o_fix, c.n_o_fix, c.k_o_fix, c.db_o_num, c.m_parts, c.n_parts,
                            c.r_m_parts, c.r_n_parts, c.r_m_blocks, c.r_n_blocks, c.r_k_blocks, c.r_db_num};
    for (int i = 0; i < 22; ++i) params[6 + i] = static_cast<int32_t>(v[i]);
}

uint32_t ref_ceil_div(uint32_t a, uint32_t b) { return CatlassTiling::CeilDiv(a, b); }
uint32_t ref_round_up(uint32_t a, uint32_t b) { return CatlassTiling::RoundUp(a, b); }

}  // extern "C"
