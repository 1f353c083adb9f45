// Run-time specialisation of the persistent GEMM kernel: code generation, on-disk + in-memory kernel cache,
// nvcc invocation and loading.  Host-only; see jit.cpp for the mapping onto the reference's JIT layer.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

namespace dga {
namespace jit {

// Everything the generated translation unit pins to a literal (the shape-derived part of GemmParams plus the
// template arguments of the kernel body).
struct Request {
    uint32_t batch = 1;
    uint32_t m = 0, n = 0, k = 0;
    uint32_t cta = 1;            // 1 or 2 (CTA pair)
    int32_t in_dtype = 0;        // ACL_FLOAT16 / ACL_BF16 (part of the key; the arithmetic is in idesc)
    int32_t out_dtype = 0;       // ACL_FLOAT / ACL_FLOAT16 / ACL_BF16
    bool b_mn = false;           // B stored [K, N] row-major
    int mode = 0;                // kModePlain, kModeBatched or kModeSplitK
    uint32_t block_n = 0, stages = 0, k_blocks = 0, m_blocks = 0, n_blocks = 0;
    uint32_t raster_group = 0, raster_along_m = 0, idesc = 0, b_stage_bytes = 0, k_block_elems = 64;
    uint32_t dp_tiles = 0, epi_cols = 32, tiles_per_batch = 0;
    uint32_t split_k = 1;        // kModeSplitK: K slices per tile
};

struct Stats {
    uint64_t compiled = 0;    // nvcc runs
    uint64_t disk_hits = 0;   // kernel directories found valid on disk and loaded
    uint64_t memory_hits = 0; // requests served from the in-process map
};

// "b{batch}m{m}n{n}k{k}_type{t}" (reference: compiler.hpp:31-32); t packs the variant knobs so that two plans for
// one shape never share a directory.
std::string signature(const Request& r);
std::string kernel_dir(const Request& r);          // $DGA_ROOT_DIR/deep_gemm_ascend/cache/kernel_<signature>/
std::string generate_code(const Request& r);       // the text written to <kernel_dir>/kernel.cu

// Makes sure <kernel_dir>/kernel.cubin exists and matches generate_code(r); compiles when it does not.  Needs a
// compiler but no GPU.
bool ensure_compiled(const Request& r, std::string* dir_out, std::string* err);

// Returns the loaded kernel for `r`, compiling it first if neither cache has it.  On failure returns false and
// leaves a one-line reason (compiler output tail included) in *err.  Thread-safe.
bool get_kernel(const Request& r, cudaKernel_t* kernel, std::string* err);

Stats stats();
void drop_memory_cache();   // forget loaded kernels (tests: forces the disk-hit path)

}  // namespace jit
}  // namespace dga
