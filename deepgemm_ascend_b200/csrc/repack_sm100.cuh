// Stride repack for operands whose leading dimension (or base pointer) is not 16-byte aligned.
//
// Role of the reference's padding kernels (PaddingCommonMatmulKernel: AIV repack of A / B into the workspace
// at offsets [0, sizeWA, sizeWA + sizeWB], C through a padded buffer that is copied back afterwards;
// op_kernel/kernel/padding_common_matmul_kernel.h:33-107, kernel_utils.h:63-75).  On Ascend the repack exists
// because GM->L1 bandwidth collapses on unaligned rows; on B200 the only thing TMA cannot do is address a
// tensor whose row pitch is not a multiple of 16 bytes, so that is the one case that is repacked:
//   copy-in : src[rows, cols] with pitch src_ld (any 2- or 4-byte alignment) -> dst with a 128-byte-multiple pitch
//   copy-out: padded C in the workspace -> the caller's C with its own pitch
// HBM-bound byte work: every thread moves one element of 16 rows (16 independent loads in flight), a warp
// touches 64 or 128 contiguous bytes per instruction, so misaligned rows cost at most one extra sector per row.
#pragma once
#include <cstdint>

namespace dga {

constexpr uint32_t kRepackRows = 16;     // rows per thread (loads issued before the first store)
constexpr uint32_t kRepackThreads = 256; // columns per block

template <typename T>
__global__ void __launch_bounds__(kRepackThreads)
repack_rows_kernel(const T* __restrict__ src, uint64_t src_ld, T* __restrict__ dst, uint64_t dst_ld, uint32_t rows,
                   uint32_t cols) {
    const uint32_t c = blockIdx.x * kRepackThreads + threadIdx.x;
    const uint32_t r0 = blockIdx.y * kRepackRows;
    if (c >= cols) return;
    T v[kRepackRows];
#pragma unroll
    for (uint32_t i = 0; i < kRepackRows; ++i)
        if (r0 + i < rows) v[i] = src[static_cast<uint64_t>(r0 + i) * src_ld + c];
#pragma unroll
    for (uint32_t i = 0; i < kRepackRows; ++i)
        if (r0 + i < rows) dst[static_cast<uint64_t>(r0 + i) * dst_ld + c] = v[i];
}

}  // namespace dga
