// Second half of the multicore split-K variant: out[M,N] = sum over slices of the fp32 partial slabs the GEMM kernel
// left in the workspace ([split_k][M][ld] fp32), in slice order (deterministic), converted to the output type.
//
// Role of the reference's StreamkReduceAdd epilogue (op_kernel/kernel/padding_streamk_matmul_kernel.h:97-101), which
// also runs as a separate pass over the fp32 workspace on the vector cores.  HBM/L2-bound byte work: the slabs were
// written a few microseconds earlier and are still in L2; one thread sums four columns (16-byte loads, one per
// slice, all issued before the first add), plain stores, so C needs no particular alignment.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <type_traits>

#include "ptx.cuh"

namespace dga {

constexpr uint32_t kReduceThreads = 256;
constexpr uint32_t kReduceMaxSplit = 16;

template <typename OutT>
__device__ __forceinline__ OutT reduce_cvt(float v) {
    if constexpr (sizeof(OutT) == 4) return v;
    else if constexpr (std::is_same<OutT, __half>::value) return __float2half_rn(v);
    else return __float2bfloat16_rn(v);
}

template <typename OutT>
__global__ void __launch_bounds__(kReduceThreads)
splitk_reduce_kernel(const float* __restrict__ ws, uint64_t slab_floats, uint32_t ld_ws, OutT* __restrict__ out,
                     uint64_t ld_out, uint32_t m, uint32_t n, uint32_t split_k) {
    ptx::grid_dep_launch();
    ptx::grid_dep_wait();   // the partial slabs come from the preceding grid
    const uint32_t groups_per_row = (n + 3) / 4;
    const uint64_t total = static_cast<uint64_t>(m) * groups_per_row;
    const bool vec_out = (reinterpret_cast<uintptr_t>(out) % (4 * sizeof(OutT)) == 0) && (ld_out % 4 == 0);
    for (uint64_t i = blockIdx.x * static_cast<uint64_t>(kReduceThreads) + threadIdx.x; i < total;
         i += static_cast<uint64_t>(gridDim.x) * kReduceThreads) {
        const uint32_t row = static_cast<uint32_t>(i / groups_per_row);
        const uint32_t col = static_cast<uint32_t>(i - static_cast<uint64_t>(row) * groups_per_row) * 4;
        const float* src = ws + static_cast<uint64_t>(row) * ld_ws + col;
        float4 part[kReduceMaxSplit];
#pragma unroll
        for (uint32_t s = 0; s < kReduceMaxSplit; ++s)
            if (s < split_k) part[s] = __ldcg(reinterpret_cast<const float4*>(src + s * slab_floats));
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (uint32_t s = 0; s < kReduceMaxSplit; ++s)
            if (s < split_k) { acc.x += part[s].x; acc.y += part[s].y; acc.z += part[s].z; acc.w += part[s].w; }
        OutT* dst = out + static_cast<uint64_t>(row) * ld_out + col;
        if (vec_out && col + 4 <= n) {
            if constexpr (sizeof(OutT) == 4) {
                *reinterpret_cast<float4*>(dst) = acc;
            } else {
                OutT v4[4] = {reduce_cvt<OutT>(acc.x), reduce_cvt<OutT>(acc.y), reduce_cvt<OutT>(acc.z),
                              reduce_cvt<OutT>(acc.w)};
                *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(v4);
            }
        } else {
            const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j)
                if (col + j < n) dst[j] = reduce_cvt<OutT>(a4[j]);
        }
    }
}

}  // namespace dga
