// C ABI of the B200 dynamic-shape matmul (see include/dga_b200.h) and the
// launch glue: TMA tensor-map encoding, plan -> kernel-instance dispatch.
// Role of op_kernel/catlass_dynamic_matmul.h:32-60 (variant dispatch) and of
// op_host/catlass_dynamic_matmul_tiling.cpp:40-122 (tiling -> launch record).
// There is deliberately no CPU fallback: without a CUDA driver every compute
// entry point fails with ACLNN_ERR_RUNTIME_ERROR.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>

#include "gemm_smallm_sm100.cuh"
#include "gemm_fp8_sm100.cuh"
#include "repack_sm100.cuh"
#include "splitk_reduce_sm100.cuh"
#include "tile_selector.h"
#include "jit.h"

namespace dga {

static std::atomic<uint64_t> g_launches{0};
// tuning knobs (dgaSetOption); -1 = built-in default
static std::atomic<int64_t> g_opt_splitk_prefetch{-1};
static std::atomic<int64_t> g_opt_pdl{-1};
static std::atomic<int64_t> g_opt_splitk_batch{-1};
static std::atomic<int64_t> g_opt_raster{-1};   // raster_group override (sweeps)
static std::atomic<int64_t> g_opt_sk_coop{1};       // stream-K launches are cooperative (co-residency guaranteed)
static std::atomic<int64_t> g_opt_raster_dir{-1};   // raster_along_m override (sweeps)
static std::atomic<int64_t> g_opt_hint_a{-1}, g_opt_hint_b{-1};   // L2 eviction hints of the operand loads: 0 normal, 1 first, 2 last (-1: the walk rule's)
static uint64_t hint_code(int64_t v) { return v == 1 ? ptx::kEvictFirst : v == 2 ? ptx::kEvictLast : 0ull; }
static std::atomic<int64_t> g_opt_jit{-1};      // shape-specialised kernels compiled at run time (default: $DGA_JIT)
static bool jit_enabled() {
    static const bool env = [] { const char* e = std::getenv("DGA_JIT"); return e != nullptr && std::atoi(e) != 0; }();
    const int64_t opt = g_opt_jit.load(std::memory_order_relaxed);
    return opt >= 0 ? opt != 0 : env;
}

#define DGA_CUDA(expr)                                                                       \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            set_error(std::string(#expr) + " failed: " + cudaGetErrorString(e__));           \
            return ACLNN_ERR_RUNTIME_ERROR;                                                  \
        }                                                                                    \
    } while (0)

// ---- TMA descriptors --------------------------------------------------------------------------------
static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    });
    return fn;
}

constexpr int32_t kAclInt32 = 3;  // ACL_INT32: packed UE8M0 scale words
static CUtensorMapDataType tmap_dtype(int32_t dtype) {
    switch (dtype) {
        case ACL_FLOAT: return CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
        case kAclInt32: return CU_TENSOR_MAP_DATA_TYPE_INT32;
        case ACL_FLOAT16: return CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
        case ACL_BF16: return CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
        default: return CU_TENSOR_MAP_DATA_TYPE_UINT8;
    }
}

// 2-D row-major tensor [outer, inner] with leading dimension ld (elements); box = [box_outer, box_inner].
static int make_tmap_2d(CUtensorMap* map, const void* ptr, int32_t dtype, uint64_t inner, uint64_t outer, uint64_t ld,
                        uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle swizzle,
                        CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_L2_256B) {
    auto fn = encode_fn();
    if (fn == nullptr) {
        set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver)");
        return ACLNN_ERR_RUNTIME_ERROR;
    }
    const uint32_t eb = dtype == kAclInt32 ? 4u : elem_bytes(dtype);
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {ld * eb};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (strides[0] & 15) != 0) {
        set_error("TMA needs 16-byte aligned base pointers and leading dimensions");
        return ACLNN_ERR_PARAM_INVALID;
    }
    CUresult r = fn(map, tmap_dtype(dtype), 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string(int(r)));
        return ACLNN_ERR_RUNTIME_ERROR;
    }
    return ACLNN_SUCCESS;
}

// 3-D row-major tensor [outer2, outer, inner]: pitch `ld` between rows, `ld2` between slabs (elements); box depth 1.
static int make_tmap_3d(CUtensorMap* map, const void* ptr, int32_t dtype, uint64_t inner, uint64_t outer, uint64_t outer2,
                        uint64_t ld, uint64_t ld2, uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle swizzle) {
    auto fn = encode_fn();
    if (fn == nullptr) {
        set_error("cuTensorMapEncodeTiled unavailable (no CUDA driver)");
        return ACLNN_ERR_RUNTIME_ERROR;
    }
    const uint32_t eb = elem_bytes(dtype);
    cuuint64_t dims[3] = {inner, outer, outer2};
    cuuint64_t strides[2] = {ld * eb, ld2 * eb};
    cuuint32_t box[3] = {box_inner, box_outer, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (strides[0] & 15) != 0 || (strides[1] & 15) != 0) {
        set_error("TMA needs 16-byte aligned base pointers and pitches");
        return ACLNN_ERR_PARAM_INVALID;
    }
    CUresult r = fn(map, tmap_dtype(dtype), 3, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (3-D) failed with CUresult " + std::to_string(int(r)));
        return ACLNN_ERR_RUNTIME_ERROR;
    }
    return ACLNN_SUCCESS;
}

// ---- launch records ---------------------------------------------------------------------------------
struct DenseLaunch {
    DgaTilingPlan plan;
    const void* a;
    const void* b;
    void* d;
    uint64_t lda, ldb, ldd;  // elements
    // What the GEMM kernel itself reads and writes: the caller's operands, or their repacked copies in the
    // workspace when a padding tag is set (set by launch_dense on every call: the workspace may move).
    struct Operands {
        const void* a; const void* b; void* d;
        uint64_t lda, ldb, ldd;
    } eff{}, enc{};          // enc: what the tensor maps below were encoded for
    CUtensorMap tm_a, tm_b, tm_d;
    bool encoded = false;
    bool maps_fresh() const {
        return encoded && std::memcmp(&eff, &enc, sizeof(Operands)) == 0 && split_ws == split_ws_enc;
    }
    // kModeBatched: `batch` independent GEMMs of one shape, slabs `slab_*` elements apart.
    // kModeMGrouped: plan.m = M_total rows in 128-row blocks, weight [groups, N, K], group_index[M_total] on the device.
    int mode = kModePlain;
    void* split_ws = nullptr, *split_ws_enc = nullptr;   // fp32 partial slabs of the multicore split-K variant
    uint32_t batch = 1, groups = 0;
    uint64_t slab_a = 0, slab_b = 0, slab_d = 0;
    const int32_t* group_index = nullptr;
    // N-sharded execution: the column slice of every peer's C (device pointers mapped into this process)
    uint32_t num_peers = 0;
    void* peer_d[kMaxPeers] = {};
    PeerMaps peers{};
};

// DGA_DEBUG_FLAGS & 16: every CTA of the persistent kernel writes six timestamps (see GemmParams::trace); the last
// launch's records are read back with dgaDebugReadTrace.  Developer aid for launch-ramp / tail / clock questions.
constexpr uint32_t kTraceWords = 6 * 1024;
static int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return dev;
}
// one buffer per device (a process may drive several GPUs)
static unsigned long long* trace_buffer() {
    static std::mutex mu;
    static std::map<int, unsigned long long*> bufs;
    std::lock_guard<std::mutex> lock(mu);
    const int dev = current_device();
    auto it = bufs.find(dev);
    if (it != bufs.end()) return it->second;
    void* p = nullptr;
    if (cudaMalloc(&p, kTraceWords * 8) != cudaSuccess) return nullptr;
    cudaMemset(p, 0, kTraceWords * 8);
    return bufs[dev] = static_cast<unsigned long long*>(p);
}

// Launch attributes shared by every tensor-core kernel: the cluster shape and programmatic dependent launch (the
// kernels call griddepcontrol.wait before their first global access, so the prologue of launch i+1 overlaps the
// tail of launch i on the same stream).  DGA_PDL=0 turns the second one off (A/B measurements).
// `co_resident`: the grid's CTAs wait on one another (stream-K owners spin on their givers' flags), so every CTA must
// be on the machine at once whatever else runs there (another stream's GEMM, an NCCL kernel, a second stream-K
// launch whose owners wait for THEIR givers): a cooperative launch starts only when the whole grid fits.  Such a
// launch is not also a programmatic dependent one (it could not start under its predecessor anyway).
static void fill_launch_attrs(cudaLaunchConfig_t& cfg, cudaLaunchAttribute (&attr)[2], uint32_t cluster,
                              bool co_resident = false) {
    static const bool pdl = [] { const char* e = std::getenv("DGA_PDL"); return e == nullptr || std::atoi(e) != 0; }();
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    if (co_resident && g_opt_sk_coop.load(std::memory_order_relaxed) != 0) {
        attr[1].id = cudaLaunchAttributeCooperative;
        attr[1].val.cooperative = 1;
        cfg.numAttrs = 2;
        return;
    }
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    const int64_t opt = g_opt_pdl.load(std::memory_order_relaxed);
    cfg.numAttrs = (opt >= 0 ? opt != 0 : pdl) ? 2 : 1;
}

// cudaFuncSetAttribute costs microseconds of host time; remember the largest opt-in per kernel instance and
// device (the attribute applies to the current device only).
template <typename K>
static int set_smem_attr(K kernel, uint32_t bytes) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, uint32_t> granted;
    std::lock_guard<std::mutex> lock(mu);
    uint32_t& have = granted[std::make_pair(reinterpret_cast<const void*>(kernel), current_device())];
    if (bytes <= have) return ACLNN_SUCCESS;
    DGA_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes)));
    have = bytes;
    return ACLNN_SUCCESS;
}

// The shape-derived constants of one launch as the JIT layer wants them (jit.h).
static jit::Request make_jit_request(const DgaTilingPlan& p, int mode, uint32_t batch, const GemmParams& gp) {
    jit::Request r;
    r.batch = mode == kModeBatched ? batch : 1;
    r.m = gp.m; r.n = gp.n; r.k = gp.k;
    r.cta = p.ctaGroup;
    r.in_dtype = p.dtype;
    r.out_dtype = p.outDtype;
    r.b_mn = p.layoutTagB != 1;
    r.mode = mode;
    r.block_n = gp.block_n; r.stages = gp.num_stages; r.k_blocks = gp.k_blocks;
    r.m_blocks = gp.num_m_blocks; r.n_blocks = gp.num_n_blocks;
    r.raster_group = gp.raster_group; r.raster_along_m = gp.raster_along_m;
    r.idesc = gp.idesc; r.b_stage_bytes = gp.b_stage_bytes; r.k_block_elems = gp.k_block_elems;
    r.dp_tiles = gp.dp_tiles; r.epi_cols = gp.epi_cols; r.tiles_per_batch = gp.tiles_per_batch;
    r.split_k = mode == kModeSplitK ? gp.split_k : 1;
    if (mode == kModeSplitK) r.out_dtype = ACL_FLOAT;   // the slices leave fp32 partial tiles whatever C's type is
    return r;
}

// Launch through a kernel specialised for this exact shape and plan, compiled on first use (jit.cpp).  A compile
// or load failure is an error of the call: there is no quiet return to the dynamic-shape kernel.
static int launch_jit(const DenseLaunch& L, const GemmParams& gp, cudaStream_t stream, int mode) {
    const jit::Request req = make_jit_request(L.plan, mode, L.batch, gp);
    cudaKernel_t kernel = nullptr;
    std::string err;
    if (!jit::get_kernel(req, &kernel, &err)) {
        set_error(err);
        return ACLNN_ERR_RUNTIME_ERROR;
    }
    {
        static std::mutex mu;
        static std::map<std::pair<cudaKernel_t, int>, uint32_t> granted;
        std::lock_guard<std::mutex> lock(mu);
        uint32_t& have = granted[std::make_pair(kernel, current_device())];
        if (L.plan.smemBytes > have) {
            DGA_CUDA(cudaFuncSetAttribute(reinterpret_cast<const void*>(kernel),
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, int(L.plan.smemBytes)));
            have = L.plan.smemBytes;
        }
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(L.plan.blockDim);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = L.plan.smemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    fill_launch_attrs(cfg, attr, L.plan.ctaGroup);
    void* args[] = {const_cast<CUtensorMap*>(&L.tm_a), const_cast<CUtensorMap*>(&L.tm_b),
                    const_cast<CUtensorMap*>(&L.tm_d), const_cast<CUtensorMap*>(&L.tm_a),
                    const_cast<CUtensorMap*>(&L.tm_b), const_cast<PeerMaps*>(&L.peers),
                    const_cast<GemmParams*>(&gp)};
    DGA_CUDA(cudaLaunchKernelExC(&cfg, reinterpret_cast<const void*>(kernel), args));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

template <int kCta, typename OutT, bool kBMn, int kMode = kModePlain, bool kSk = false, bool kMc = false, bool kKs = false>
static int launch_common_inst(const DenseLaunch& L, const GemmParams& gp, cudaStream_t stream) {
    if constexpr (kMode == kModePlain && !kSk && !kMc && !kKs) {
        // a stream-K plan whose remainder came out empty is the plain walk
        if (gp.sk_tiles != 0) return launch_common_inst<kCta, OutT, kBMn, kMode, true>(L, gp, stream);
        // two CTA pairs per cluster, A multicast between them (16-bit outputs over either weight layout)
        if constexpr (kCta == 2 && sizeof(OutT) == 2) {
            if (L.plan.kernelSerial == DGA_KERNEL_CLUSTER_SPLITK)
                return launch_common_inst<kCta, OutT, kBMn, kMode, false, false, true>(L, gp, stream);
            if (L.plan.clusterShape == 4) return launch_common_inst<kCta, OutT, kBMn, kMode, false, true>(L, gp, stream);
        }
    }
    if constexpr ((kMode == kModePlain || kMode == kModeBatched || kMode == kModeSplitK) && !kSk && !kMc && !kKs) {
        if (jit_enabled() && L.num_peers == 0) return launch_jit(L, gp, stream, kMode);
    }
    auto kernel = gemm_sm100_kernel<kCta, OutT, kBMn, false, kMode, kSk, kMc, kKs>;
    int rc = set_smem_attr(kernel, L.plan.smemBytes);
    if (rc != ACLNN_SUCCESS) return rc;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(L.plan.blockDim);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = L.plan.smemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    fill_launch_attrs(cfg, attr, (kMc || kKs) ? 4 : kCta, /*co_resident=*/kSk);
    DGA_CUDA(cudaLaunchKernelEx(&cfg, kernel, L.tm_a, L.tm_b, L.tm_d, L.tm_a, L.tm_b, L.peers, gp));  // no scale maps here
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

template <int kCta, bool kBMn>
static int launch_common_out(const DenseLaunch& L, const GemmParams& gp, cudaStream_t stream) {
    switch (L.plan.outDtype) {
        case ACL_BF16: return launch_common_inst<kCta, __nv_bfloat16, kBMn>(L, gp, stream);
        case ACL_FLOAT16: return launch_common_inst<kCta, __half, kBMn>(L, gp, stream);
        case ACL_FLOAT: return launch_common_inst<kCta, float, kBMn>(L, gp, stream);
    }
    set_error("unsupported output dtype");
    return ACLNN_ERR_PARAM_INVALID;
}

// Columns per output store of the plain (whole-tile, 2-D output) launches.  $DGA_EPI_COLS=32 forces 64-byte store rows
// where 128-byte rows would be used: a measurement aid for the peer stores of the N-sharded step, whose NVLink rate
// follows the size of the write (config 5 on 8 GPUs: 3.98 ms with 64-byte rows, 2.61 ms with 128-byte rows; 256-byte
// rows - two adjacent swizzle lines per row in one store box - change nothing, the store engine writes 128-byte
// lines: profiles/r02_nshard_store_rows_g8.log).
static uint32_t plain_epi_cols(uint32_t block_n, uint32_t out_elem_bytes) {
    static const uint32_t forced = [] { const char* e = std::getenv("DGA_EPI_COLS"); return e ? uint32_t(std::atoi(e)) : 0u; }();
    const uint32_t cols = epi_cols_for(block_n, out_elem_bytes);
    return forced == 32 ? 32u : cols;
}

static int encode_common(DenseLaunch& L) {
    const DgaTilingPlan& p = L.plan;
    const uint32_t cta = p.ctaGroup, load_n = p.n1 / cta;
    if (L.mode == kModeSplitK) {
        int rcs = make_tmap_2d(&L.tm_a, L.eff.a, p.dtype, p.k, p.m, L.eff.lda, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rcs != ACLNN_SUCCESS) return rcs;
        if (p.layoutTagB == 1)
            rcs = make_tmap_2d(&L.tm_b, L.eff.b, p.dtype, p.k, p.n, L.eff.ldb, 64, load_n, CU_TENSOR_MAP_SWIZZLE_128B);
        else
            rcs = make_tmap_2d(&L.tm_b, L.eff.b, p.dtype, p.n, p.k, L.eff.ldb, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rcs != ACLNN_SUCCESS) return rcs;
        const uint64_t pitch = splitk_ws_pitch(p.n);
        return make_tmap_3d(&L.tm_d, L.split_ws, ACL_FLOAT, p.n, p.m, p.splitkFactor, pitch, uint64_t(p.m) * pitch, 32, 128,
                            CU_TENSOR_MAP_SWIZZLE_128B);
    }
    if (L.mode != kModePlain) {
        const uint32_t epi3 = epi_cols_for(p.n1, elem_bytes(p.outDtype));
        const CUtensorMapSwizzle d_swz3 = epi3 * elem_bytes(p.outDtype) == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                                                                : CU_TENSOR_MAP_SWIZZLE_64B;
        int rc3;
        if (mode_is_batched(L.mode)) {
            rc3 = make_tmap_3d(&L.tm_a, L.eff.a, p.dtype, p.k, p.m, L.batch, L.eff.lda, L.slab_a, 64, 128,
                               CU_TENSOR_MAP_SWIZZLE_128B);
            if (rc3 != ACLNN_SUCCESS) return rc3;
            rc3 = make_tmap_3d(&L.tm_d, L.eff.d, p.outDtype, p.n, p.m, L.batch, L.eff.ldd, L.slab_d, epi3, 128, d_swz3);
        } else {
            rc3 = make_tmap_2d(&L.tm_a, L.eff.a, p.dtype, p.k, p.m, L.eff.lda, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
            if (rc3 != ACLNN_SUCCESS) return rc3;
            rc3 = make_tmap_2d(&L.tm_d, L.eff.d, p.outDtype, p.n, p.m, L.eff.ldd, epi3, 128, d_swz3);
        }
        if (rc3 != ACLNN_SUCCESS) return rc3;
        const uint64_t slabs = mode_is_batched(L.mode) ? L.batch : L.groups;
        if (p.layoutTagB == 1)
            return make_tmap_3d(&L.tm_b, L.eff.b, p.dtype, p.k, p.n, slabs, L.eff.ldb, L.slab_b, 64, load_n,
                                CU_TENSOR_MAP_SWIZZLE_128B);
        return make_tmap_3d(&L.tm_b, L.eff.b, p.dtype, p.n, p.k, slabs, L.eff.ldb, L.slab_b, 64, 64,
                            CU_TENSOR_MAP_SWIZZLE_128B);
    }
    // clusterShape 4: every CTA fetches half of its 128 A rows and receives the other half from its twin
    const bool multicast_a = p.clusterShape == 4 && p.kernelSerial != DGA_KERNEL_CLUSTER_SPLITK;
    int rc = make_tmap_2d(&L.tm_a, L.eff.a, p.dtype, p.k, p.m, L.eff.lda, 64, multicast_a ? 64 : 128,
                          CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != ACLNN_SUCCESS) return rc;
    if (p.layoutTagB == 1)  // weight stored [N, K]
        rc = make_tmap_2d(&L.tm_b, L.eff.b, p.dtype, p.k, p.n, L.eff.ldb, 64, load_n, CU_TENSOR_MAP_SWIZZLE_128B);
    else  // row-major [K, N]: boxes of 64 columns x 64 k-rows
        rc = make_tmap_2d(&L.tm_b, L.eff.b, p.dtype, p.n, p.k, L.eff.ldb, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != ACLNN_SUCCESS) return rc;
    const uint32_t epi = plain_epi_cols(p.n1, elem_bytes(p.outDtype));
    const bool wide = epi * elem_bytes(p.outDtype) == 128;
    const CUtensorMapSwizzle d_swz = wide ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    for (uint32_t i = 0; i < L.num_peers; ++i) {
        rc = make_tmap_2d(&L.peers.m[i], L.peer_d[i], p.outDtype, p.n, p.m, L.ldd, epi, 128, d_swz);
        if (rc != ACLNN_SUCCESS) return rc;
    }
    return make_tmap_2d(&L.tm_d, L.eff.d, p.outDtype, p.n, p.m, L.eff.ldd, epi, 128, d_swz);
}

template <typename OutT>
static int launch_splitk_reduce(const DenseLaunch& L, cudaStream_t stream) {
    const DgaTilingPlan& p = L.plan;
    const uint32_t pitch = splitk_ws_pitch(p.n);
    const uint64_t groups = uint64_t(p.m) * ((p.n + 3) / 4);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(uint32_t(std::min<uint64_t>((groups + kReduceThreads - 1) / kReduceThreads, 148 * 8)));
    cfg.blockDim = dim3(kReduceThreads);
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    fill_launch_attrs(cfg, attr, 1);
    DGA_CUDA(cudaLaunchKernelEx(&cfg, splitk_reduce_kernel<OutT>, static_cast<const float*>(L.split_ws),
                                uint64_t(p.m) * pitch, pitch, static_cast<OutT*>(L.eff.d), L.eff.ldd, p.m, p.n,
                                uint32_t(p.splitkFactor)));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

// The part of GemmParams that follows from the plan alone (whole tiles, one per work item).
static void common_params(const DgaTilingPlan& p, GemmParams& gp) {
    const uint32_t cta = p.ctaGroup;
    gp.m = p.m; gp.n = p.n; gp.k = p.k;
    gp.block_n = p.n1;
    gp.num_stages = p.stages;
    gp.k_blocks = ceil_div(p.k, 64);
    gp.num_m_blocks = ceil_div(p.m, 128 * cta);
    gp.num_n_blocks = ceil_div(p.n, p.n1);
    if (p.clusterShape == 4 && p.kernelSerial != DGA_KERNEL_CLUSTER_SPLITK)
        gp.num_n_blocks /= 2;   // the cluster walks pairs of N tiles (finalize_plan: even count)
    gp.raster_group = p.swizzleOffset;
    if (g_opt_raster.load(std::memory_order_relaxed) > 0) gp.raster_group = uint32_t(g_opt_raster.load());
    gp.raster_along_m = p.swizzleDirection;
    if (g_opt_raster_dir.load(std::memory_order_relaxed) >= 0) gp.raster_along_m = uint32_t(g_opt_raster_dir.load());
    if (g_opt_hint_a.load(std::memory_order_relaxed) >= 0) gp.hint_a = hint_code(g_opt_hint_a.load());
    if (g_opt_hint_b.load(std::memory_order_relaxed) >= 0) gp.hint_b = hint_code(g_opt_hint_b.load());
    gp.idesc = make_idesc(128 * cta, p.n1, p.dtype == ACL_BF16 ? 1u : 0u, 0, p.layoutTagB == 1 ? 0u : 1u);
    gp.b_stage_bytes = (p.n1 / cta) * 128;
    gp.k_block_elems = 64;
    gp.epi_cols = plain_epi_cols(p.n1, elem_bytes(p.outDtype));
    gp.dp_tiles = gp.num_m_blocks * gp.num_n_blocks;
    gp.sk_tiles = 0;
}

static int launch_common(DenseLaunch& L, cudaStream_t stream, void* workspace) {
    const DgaTilingPlan& p = L.plan;
    if (p.kernelSerial == DGA_KERNEL_SPLITK_REDUCE) {
        if (L.mode != kModePlain && L.mode != kModeSplitK) {
            set_error("split-K is not combined with batched / grouped launches");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        L.mode = kModeSplitK;
        L.split_ws = workspace;
    }
    if (!L.maps_fresh()) {
        int rc = encode_common(L);
        if (rc != ACLNN_SUCCESS) return rc;
        L.encoded = true;
        L.enc = L.eff;
        L.split_ws_enc = L.split_ws;
    }
    const uint32_t cta = p.ctaGroup;
    GemmParams gp{};
    common_params(p, gp);
    gp.num_peers = L.num_peers;
    if (p.kernelSerial == DGA_KERNEL_STREAMK) {
        const uint32_t clusters = platform().smCount / cta;
        uint32_t split = 1;
        streamk_split(gp.num_m_blocks * gp.num_n_blocks, clusters, gp.k_blocks, p.n1, p.splitkFactor, gp.dp_tiles,
                      gp.sk_tiles, split);
        gp.split_k = split;
        if (gp.sk_tiles != 0) {
            const uint64_t part = streamk_partial_bytes(gp.sk_tiles, split, cta, p.n1);
            gp.sk_partials = static_cast<float*>(workspace);
            gp.sk_sems = reinterpret_cast<int32_t*>(static_cast<char*>(workspace) + part);
            // "partial is ready" flags carry a per-launch nonce (GemmParams::sk_sems): no clearing launch, whatever
            // bytes the caller's workspace holds
            static std::atomic<uint32_t> launch_seq{0};
            const uint32_t seq = launch_seq.fetch_add(1, std::memory_order_relaxed) + 1;
            gp.sk_nonce = int32_t((seq * 2654435761u) | 1u);
        }
    }
    static const uint32_t dbg = [] { const char* e = std::getenv("DGA_DEBUG_FLAGS"); return e ? uint32_t(std::atoi(e)) : 0u; }();
    gp.debug = dbg;
    if (dbg & 16) gp.trace = trace_buffer();
    if (L.mode == kModeSplitK) {
        gp.split_k = p.splitkFactor;
        gp.dp_tiles *= p.splitkFactor;
        gp.epi_cols = 32;
        int rcs;
        if (cta == 2)
            rcs = p.layoutTagB == 1 ? launch_common_inst<2, float, false, kModeSplitK>(L, gp, stream)
                                    : launch_common_inst<2, float, true, kModeSplitK>(L, gp, stream);
        else
            rcs = p.layoutTagB == 1 ? launch_common_inst<1, float, false, kModeSplitK>(L, gp, stream)
                                    : launch_common_inst<1, float, true, kModeSplitK>(L, gp, stream);
        if (rcs != ACLNN_SUCCESS) return rcs;
        switch (p.outDtype) {
            case ACL_BF16: return launch_splitk_reduce<__nv_bfloat16>(L, stream);
            case ACL_FLOAT16: return launch_splitk_reduce<__half>(L, stream);
            case ACL_FLOAT: return launch_splitk_reduce<float>(L, stream);
        }
        set_error("unsupported output dtype");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (L.mode != kModePlain) {
        // single-CTA tiles only; the instances the two callers need: the framework's batched fp32-output GEMM over
        // row-major B, and the MoE grouped GEMM over [G,N,K] weights with a 16-bit output
        gp.tiles_per_batch = gp.dp_tiles;
        gp.group_index = L.group_index;
        if (mode_is_batched(L.mode)) gp.dp_tiles *= L.batch;
        if ((cta != 1 && L.mode == kModeBatched) || p.kernelSerial == DGA_KERNEL_STREAMK) {
            set_error("batched launches use single-CTA whole tiles; grouped launches whole tiles");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        if (cta == 2 && p.layoutTagB == 1) {   // pair tiles: groups aligned to 256 rows (dgaSetExecutorGroupAlignment)
            if (L.mode == kModeMGrouped && p.outDtype == ACL_BF16)
                return launch_common_inst<2, __nv_bfloat16, false, kModeMGrouped>(L, gp, stream);
            if (L.mode == kModeMGrouped && p.outDtype == ACL_FLOAT16)
                return launch_common_inst<2, __half, false, kModeMGrouped>(L, gp, stream);
            if (L.mode == kModeMasked && p.outDtype == ACL_BF16)
                return launch_common_inst<2, __nv_bfloat16, false, kModeMasked>(L, gp, stream);
            if (L.mode == kModeMasked && p.outDtype == ACL_FLOAT16)
                return launch_common_inst<2, __half, false, kModeMasked>(L, gp, stream);
        }
        if (L.mode == kModeBatched && p.layoutTagB == 0 && p.outDtype == ACL_FLOAT)
            return launch_common_inst<1, float, true, kModeBatched>(L, gp, stream);
        if (L.mode == kModeMGrouped && p.layoutTagB == 1 && p.outDtype == ACL_BF16)
            return launch_common_inst<1, __nv_bfloat16, false, kModeMGrouped>(L, gp, stream);
        if (L.mode == kModeMGrouped && p.layoutTagB == 1 && p.outDtype == ACL_FLOAT16)
            return launch_common_inst<1, __half, false, kModeMGrouped>(L, gp, stream);
        if (L.mode == kModeMasked && p.layoutTagB == 1 && p.outDtype == ACL_BF16)
            return launch_common_inst<1, __nv_bfloat16, false, kModeMasked>(L, gp, stream);
        if (L.mode == kModeMasked && p.layoutTagB == 1 && p.outDtype == ACL_FLOAT16)
            return launch_common_inst<1, __half, false, kModeMasked>(L, gp, stream);
        set_error("this dtype / layout combination is not built for batched / grouped launches");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (cta == 2)
        return p.layoutTagB == 1 ? launch_common_out<2, false>(L, gp, stream) : launch_common_out<2, true>(L, gp, stream);
    return p.layoutTagB == 1 ? launch_common_out<1, false>(L, gp, stream) : launch_common_out<1, true>(L, gp, stream);
}

// ---- small-M swap-AB split-K ---------------------------------------------------------------------------
template <typename OutT>
static int launch_splitk_inst(const DenseLaunch& L, const SmallMParams& sp, cudaStream_t stream) {
    auto kernel = gemm_smallm_sm100_kernel<OutT>;
    int rc = set_smem_attr(kernel, L.plan.smemBytes);
    if (rc != ACLNN_SUCCESS) return rc;
    const bool ws_mode = sp.partials != nullptr;   // no cluster: the K slices meet in the workspace
    if (sp.split > 8 && !ws_mode) DGA_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(L.plan.blockDim);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = L.plan.smemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    fill_launch_attrs(cfg, attr, ws_mode ? 1u : sp.split);
    DGA_CUDA(cudaLaunchKernelEx(&cfg, kernel, L.tm_a, L.tm_b, sp));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

// The part of SmallMParams that follows from the plan alone.
static void smallm_params(const DgaTilingPlan& p, SmallMParams& sp) {
    const uint32_t m_pad = p.m1;
    sp.m = p.m; sp.n = p.n; sp.k = p.k;
    sp.m_pad = m_pad;
    sp.split = p.splitkFactor;
    sp.num_stages = p.stages;
    sp.k_blocks = ceil_div(p.k, 64);
    sp.idesc = make_idesc(128, m_pad, p.dtype == ACL_BF16 ? 1u : 0u, 0, 0);
    uint32_t cols = 32;
    while (cols < m_pad) cols <<= 1;
    sp.tmem_cols = cols;
    const int64_t pf = g_opt_splitk_prefetch.load(std::memory_order_relaxed);
    sp.prefetch_kb = pf >= 0 ? uint32_t(pf) : 4u;
    const int64_t ib = g_opt_splitk_batch.load(std::memory_order_relaxed);
    sp.issue_batch = std::max<uint32_t>(1u, std::min<uint32_t>(ib > 0 ? uint32_t(ib) : 1u, p.stages));
    static const uint32_t dbg = [] { const char* e = std::getenv("DGA_DEBUG_FLAGS"); return e ? uint32_t(std::atoi(e)) : 0u; }();
    sp.trace = (dbg & 16) && p.blockDim * 12 <= kTraceWords ? trace_buffer() : nullptr;
    sp.tiles_first = UINT32_MAX;
}

static int launch_splitk(DenseLaunch& L, cudaStream_t stream, void* workspace = nullptr) {
    const DgaTilingPlan& p = L.plan;
    const bool ws_mode = p.kernelSerial == DGA_KERNEL_DECODE_WS;
    const uint32_t m_pad = p.m1;
    if (!L.maps_fresh()) {
        // tm_a slot holds the weight map (UMMA A operand), tm_b the activation map (UMMA B operand)
        int rc = make_tmap_2d(&L.tm_a, L.eff.b, p.dtype, p.k, p.n, L.eff.ldb, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc != ACLNN_SUCCESS) return rc;
        rc = make_tmap_2d(&L.tm_b, L.eff.a, p.dtype, p.k, p.m, L.eff.lda, 64, m_pad, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc != ACLNN_SUCCESS) return rc;
        L.encoded = true;
        L.enc = L.eff;
    }
    SmallMParams sp{};
    smallm_params(p, sp);
    sp.ldc = static_cast<uint32_t>(L.eff.ldd);
    sp.out = L.eff.d;
    if (ws_mode) {
        // [split][m][pitch] fp32 slabs, the layout splitk_reduce_kernel adds up (the same reduce kernel as variant 5)
        if (sp.k_blocks < sp.split) {
            set_error("decode plan infeasible for this shape");
            return ACLNN_ERR_INNER_TILING_ERROR;
        }
        sp.partials = static_cast<float*>(workspace);
        sp.ws_pitch = splitk_ws_pitch(p.n);
        L.split_ws = L.split_ws_enc = workspace;   // (no tensor map depends on it here)
        int rcw = launch_splitk_inst<float>(L, sp, stream);   // the output type only matters to the reduce kernel
        if (rcw != ACLNN_SUCCESS) return rcw;
        switch (p.outDtype) {
            case ACL_BF16: return launch_splitk_reduce<__nv_bfloat16>(L, stream);
            case ACL_FLOAT16: return launch_splitk_reduce<__half>(L, stream);
            case ACL_FLOAT: return launch_splitk_reduce<float>(L, stream);
        }
        set_error("unsupported output dtype");
        return ACLNN_ERR_PARAM_INVALID;
    }
    // the drained ring doubles as the fp32 staging tile [m][128] of the reduction
    if (sp.k_blocks < sp.split || uint64_t(p.stages) * (16384 + m_pad * 128) < uint64_t(m_pad) * 512) {
        set_error("split-K plan infeasible for this shape");
        return ACLNN_ERR_INNER_TILING_ERROR;
    }
    switch (p.outDtype) {
        case ACL_BF16: return launch_splitk_inst<__nv_bfloat16>(L, sp, stream);
        case ACL_FLOAT16: return launch_splitk_inst<__half>(L, sp, stream);
        case ACL_FLOAT: return launch_splitk_inst<float>(L, sp, stream);
    }
    set_error("unsupported output dtype");
    return ACLNN_ERR_PARAM_INVALID;
}

static int launch_variant(DenseLaunch& L, cudaStream_t stream, void* workspace) {
    switch (L.plan.kernelSerial) {
        case DGA_KERNEL_COMMON:
        case DGA_KERNEL_SMALL:
        case DGA_KERNEL_CLUSTER_SPLITK:
        case DGA_KERNEL_PADDING_COMMON: return launch_common(L, stream, nullptr);
        case DGA_KERNEL_DECODE_WS:
            if (workspace == nullptr || (reinterpret_cast<uintptr_t>(workspace) & 15) != 0) {
                set_error("the decode variant needs its (16-byte aligned) workspace");
                return ACLNN_ERR_PARAM_INVALID;
            }
            return launch_splitk(L, stream, workspace);
        case DGA_KERNEL_STREAMK:
        case DGA_KERNEL_SPLITK_REDUCE:
            if (workspace == nullptr) {
                set_error("the stream-K / split-K variants need their workspace");
                return ACLNN_ERR_PARAM_INVALID;
            }
            if ((reinterpret_cast<uintptr_t>(workspace) & 15) != 0) {   // 16-byte vector stores / TMA slabs
                set_error("the stream-K / split-K workspace must be 16-byte aligned");
                return ACLNN_ERR_PARAM_INVALID;
            }
            return launch_common(L, stream, workspace);
        case DGA_KERNEL_SPLITK: return launch_splitk(L, stream);
    }
    set_error("kernel variant " + std::to_string(int(L.plan.kernelSerial)) + " is not built");
    return ACLNN_ERR_INNER_TILING_ERROR;
}

// ---- two decode GEMMs over the same activations in one launch (dgaDecodePairMatmul) ---------------------------
template <typename OutT>
static int launch_decode_pair_inst(const DgaTilingPlan& p, const CUtensorMap& w_first, const CUtensorMap& w_second,
                                   const CUtensorMap& x, const SmallMParams& sp, cudaStream_t stream) {
    auto kernel = gemm_smallm_pair_sm100_kernel<OutT>;
    int rc = set_smem_attr(kernel, p.smemBytes);
    if (rc != ACLNN_SUCCESS) return rc;
    if (sp.split > 8) DGA_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(p.blockDim);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = p.smemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    fill_launch_attrs(cfg, attr, sp.split);
    DGA_CUDA(cudaLaunchKernelEx(&cfg, kernel, w_first, w_second, x, sp));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

// `plan`: a finalised decode plan (kernelSerial 3) for the virtual shape m x (tiles_first + tiles_second) * 128 x k.
static int launch_decode_pair(const DenseLaunch& first, const DenseLaunch& second, const DgaTilingPlan& plan,
                              cudaStream_t stream) {
    const DgaTilingPlan& p = plan;
    const uint32_t m_pad = p.m1, n_first = first.plan.n, n_second = second.plan.n;
    CUtensorMap w_first, w_second, x;
    int rc = make_tmap_2d(&w_first, first.b, p.dtype, p.k, n_first, first.ldb, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != ACLNN_SUCCESS) return rc;
    rc = make_tmap_2d(&w_second, second.b, p.dtype, p.k, n_second, second.ldb, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != ACLNN_SUCCESS) return rc;
    rc = make_tmap_2d(&x, first.a, p.dtype, p.k, p.m, first.lda, 64, m_pad, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != ACLNN_SUCCESS) return rc;
    SmallMParams sp{};
    smallm_params(p, sp);
    sp.n = n_first;
    sp.ldc = static_cast<uint32_t>(first.ldd);
    sp.out = first.d;
    sp.tiles_first = ceil_div(n_first, 128);
    sp.n2 = n_second;
    sp.ldc2 = static_cast<uint32_t>(second.ldd);
    sp.out2 = second.d;
    if (sp.k_blocks < sp.split || uint64_t(p.stages) * (16384 + m_pad * 128) < uint64_t(m_pad) * 512 ||
        p.blockDim != (sp.tiles_first + ceil_div(n_second, 128)) * sp.split) {
        set_error("decode pair plan infeasible for this shape");
        return ACLNN_ERR_INNER_TILING_ERROR;
    }
    switch (p.outDtype) {
        case ACL_BF16: return launch_decode_pair_inst<__nv_bfloat16>(p, w_first, w_second, x, sp, stream);
        case ACL_FLOAT16: return launch_decode_pair_inst<__half>(p, w_first, w_second, x, sp, stream);
        case ACL_FLOAT: return launch_decode_pair_inst<float>(p, w_first, w_second, x, sp, stream);
    }
    set_error("unsupported output dtype");
    return ACLNN_ERR_PARAM_INVALID;
}

// rows x cols elements of `eb` bytes from a pitch of src_ld to a pitch of dst_ld (elements)
static int launch_repack(const void* src, uint64_t src_ld, void* dst, uint64_t dst_ld, uint32_t rows, uint32_t cols,
                         uint32_t eb, cudaStream_t stream) {
    const dim3 grid(ceil_div(cols, kRepackThreads), ceil_div(rows, kRepackRows));
    if (eb == 4)
        repack_rows_kernel<uint32_t><<<grid, kRepackThreads, 0, stream>>>(
            static_cast<const uint32_t*>(src), src_ld, static_cast<uint32_t*>(dst), dst_ld, rows, cols);
    else
        repack_rows_kernel<uint16_t><<<grid, kRepackThreads, 0, stream>>>(
            static_cast<const uint16_t*>(src), src_ld, static_cast<uint16_t*>(dst), dst_ld, rows, cols);
    DGA_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

// Padding tags from the operands the caller actually passed (role of GetPaddingTag, select_kernel.cpp:82-268, cut
// down to the one thing that matters on B200): TMA cannot address a tensor whose base or row pitch is not a
// multiple of 16 bytes, so exactly those operands go through the workspace.  Re-derives workspace bytes.
static int apply_padding(DenseLaunch& L, bool allow_repack = true) {
    DgaTilingPlan& p = L.plan;
    const uint32_t ib = elem_bytes(p.dtype), ob = elem_bytes(p.outDtype);
    auto unaligned = [](const void* ptr, uint64_t ld, uint32_t eb) {
        return (reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (ld * eb) % 16 != 0;
    };
    p.strideA = L.lda; p.strideB = L.ldb; p.strideC = L.ldd;
    p.paddingTagA = unaligned(L.a, L.lda, ib) ? 1 : 0;
    p.paddingTagB = unaligned(L.b, L.ldb, ib) ? 1 : 0;
    p.paddingTagC = unaligned(L.d, L.ldd, ob) ? 1 : 0;
    if (!allow_repack && (p.paddingTagA || p.paddingTagB || p.paddingTagC)) {
        set_error("base pointers and leading dimensions must be 16-byte aligned on this entry point");
        return ACLNN_ERR_INNER_TILING_ERROR;
    }
    return finalize_plan(p);
}

static int launch_dense(DenseLaunch& L, cudaStream_t stream, void* workspace = nullptr) {
    const DgaTilingPlan& p = L.plan;
    L.eff = {L.a, L.b, L.d, L.lda, L.ldb, L.ldd};
    if (!(p.paddingTagA || p.paddingTagB || p.paddingTagC)) return launch_variant(L, stream, workspace);
    if (workspace == nullptr) {
        set_error("unaligned operands are repacked through the workspace, which is missing");
        return ACLNN_ERR_PARAM_INVALID;
    }
    // PaddingCommon (padding_common_matmul_kernel.h:33-107): repack -> GEMM on the aligned copies -> copy C back
    const PadLayout pl = pad_layout(p);
    char* ws = static_cast<char*>(workspace);
    const uint32_t ib = elem_bytes(p.dtype), ob = elem_bytes(p.outDtype);
    int rc = ACLNN_SUCCESS;
    if (p.paddingTagA) {
        rc = launch_repack(L.a, L.lda, ws + pl.offA, pl.ldA, p.m, p.k, ib, stream);
        if (rc != ACLNN_SUCCESS) return rc;
        L.eff.a = ws + pl.offA; L.eff.lda = pl.ldA;
    }
    if (p.paddingTagB) {
        const uint32_t rows = p.layoutTagB ? p.n : p.k, cols = p.layoutTagB ? p.k : p.n;
        rc = launch_repack(L.b, L.ldb, ws + pl.offB, pl.ldB, rows, cols, ib, stream);
        if (rc != ACLNN_SUCCESS) return rc;
        L.eff.b = ws + pl.offB; L.eff.ldb = pl.ldB;
    }
    if (p.paddingTagC) {
        L.eff.d = ws + pl.offC; L.eff.ldd = pl.ldC;
    }
    rc = launch_variant(L, stream, ws + pl.bytes);
    if (rc != ACLNN_SUCCESS) return rc;
    if (p.paddingTagC) rc = launch_repack(ws + pl.offC, pl.ldC, L.d, L.ldd, p.m, p.n, ob, stream);
    return rc;
}

// ---- fp8 e4m3 block-scaled -------------------------------------------------------------------------------
struct Fp8Launch {
    uint32_t m = 0, n = 0, k = 0, cta = 2, bn = 256, stages = 0, grid = 0, smem = 0;
    const void* a = nullptr; const void* b = nullptr; void* d = nullptr;
    const float* sfa = nullptr; const float* sfb = nullptr;
    uint64_t lda = 0, ldb = 0, ldd = 0, sfa_ld = 0, sfb_ld = 0;
    CUtensorMap tm_a, tm_b, tm_d;
    bool encoded = false;
    // hardware block-scaled path (scales that are exact powers of two)
    uint32_t mx_cta = 2, mx_bn = 224, mx_stages = 0, mx_grid = 0, mx_smem = 0, mx_chunks = 2;
    uint64_t ws_sfa_off = 0, ws_sfb_off = 0, ws_flag_off = 0, ws_bytes = 0;
    CUtensorMap tm_b_mx, tm_d_mx, tm_sfa, tm_sfb;
    const void* encoded_ws = nullptr;
};

static const uint32_t kFp8TilesPair[] = {256, 224, 192, 160, 128};
static std::atomic<int64_t> g_opt_fp8_bn{-1};   // tile width of the fp32-scale fp8 kernel (sweeps; -1: the cost rule)
static const uint32_t kMxTilesPair[] = {224, 192, 128};   // 2 accumulator stages + scale columns must fit 512 TMEM columns

// Wave-balanced tile for the fp8 kernel: fewest (waves x block_n) column-units, ties to the wider tile.
static void plan_fp8(Fp8Launch& L) {
    const PlatformInfo& hw = platform();
    if (L.m <= 128) {
        L.cta = 1; L.bn = 128;
    } else {
        L.cta = 2;
        uint64_t best = ~0ull;
        for (uint32_t bn : kFp8TilesPair) {
            const uint64_t tiles = uint64_t(ceil_div(L.m, 256)) * ceil_div(L.n, bn);
            const uint64_t waves = (tiles + hw.smCount / 2 - 1) / (hw.smCount / 2);
            // measured time of one K block of a tile, ns (profiles/r02_fp8_tile_sweep.log: it is dominated by the
            // per-K-block cost of the MMA <-> promotion hand-over, not by the tile width, so wide tiles win unless
            // they cost a whole extra wave)
            const uint64_t step_ns = bn == 256 ? 539 : bn == 224 ? 555 : bn == 192 ? 505 : bn == 160 ? 499 : 450;
            const uint64_t cost = waves * step_ns;
            if (cost < best) { best = cost; L.bn = bn; }
        }
        const int64_t forced = g_opt_fp8_bn.load(std::memory_order_relaxed);
        for (uint32_t bn : kFp8TilesPair)
            if (forced == int64_t(bn)) L.bn = bn;
    }
    const uint32_t per_stage = 128 * 128 + (L.bn / L.cta) * 128;
    const uint32_t kb_count = ceil_div(L.k, 128);
    const uint32_t fixed = 4 * 128 * 32 * 2 + 1024 + 256 + 2 * 3 * round_up(kb_count, 4) * 4;   // + weight scales of 2 tiles
    L.stages = std::min<uint32_t>(10, (hw.smemBytes - fixed) / per_stage);
    const uint32_t tiles = ceil_div(L.m, 128 * L.cta) * ceil_div(L.n, L.bn);
    L.grid = std::min(tiles, hw.smCount / L.cta) * L.cta;
    L.smem = make_fp8_smem(L.bn, L.cta, L.stages, kb_count).total + 1024;

    // block-scaled plan
    if (L.m <= 128) {
        L.mx_cta = 1; L.mx_bn = 128;
    } else {
        L.mx_cta = 2;
        uint64_t best = ~0ull;
        for (uint32_t bn : kMxTilesPair) {
            const uint64_t tiles_bn = uint64_t(ceil_div(L.m, 256)) * ceil_div(L.n, bn);
            const uint64_t waves = (tiles_bn + hw.smCount / 2 - 1) / (hw.smCount / 2);
            const uint64_t cost = waves * bn;
            if (cost < best) { best = cost; L.mx_bn = bn; }
        }
    }
    L.mx_chunks = ceil_div(L.mx_bn, 128);
    const uint32_t sf_stage = 512 * (1 + L.mx_chunks);
    const uint32_t mx_per_stage = 128 * 128 + (L.mx_bn / L.mx_cta) * 128 + sf_stage;
    const uint32_t mx_d_stage = 128 * epi_cols_for(L.mx_bn, 2) * 2;
    const uint32_t mx_fixed = 2 * mx_d_stage + 1024 + 256;
    L.mx_stages = std::min<uint32_t>(10, (hw.smemBytes - mx_fixed) / mx_per_stage);
    const uint32_t mx_tiles = ceil_div(L.m, 128 * L.mx_cta) * ceil_div(L.n, L.mx_bn);
    L.mx_grid = std::min(mx_tiles, hw.smCount / L.mx_cta) * L.mx_cta;
    L.mx_smem = make_smem_layout(mx_d_stage, (L.mx_bn / L.mx_cta) * 128, L.mx_stages, sf_stage).total + 1024;
    // workspace: packed A scales [K/512][ceil(M/128)][128] words, packed B scales [K/512][n tiles][chunks][128], flag
    const uint64_t kg = ceil_div(L.k, 512);
    L.ws_sfa_off = 0;
    L.ws_sfb_off = kg * ceil_div(L.m, 128) * 512;
    L.ws_flag_off = L.ws_sfb_off + kg * ceil_div(L.n, L.mx_bn) * L.mx_chunks * 512;
    L.ws_bytes = L.ws_flag_off + 256;
}

// Packs fp32 scales into the UE8M0 words the block-scaled MMA reads: word = exponent bytes of 4 consecutive K
// blocks; inside every 128-entry chunk entry (r*32 + l) is stored at l*4 + r, the 32-row x 128-bit order
// tcgen05.cp.32x128b.warpx4 expects.  Sets *flag when any scale is not an exact power of two (then the
// promotion kernel, not the block-scaled one, must produce the result).
__global__ void fp8_pack_scales_kernel(const float* __restrict__ sfa, const float* __restrict__ sfb, uint32_t m,
                                       uint32_t n, uint32_t kb_total, uint32_t sfa_ld, uint32_t sfb_ld,
                                       uint32_t bn, uint32_t chunks, uint32_t* __restrict__ out_a,
                                       uint32_t* __restrict__ out_b, int32_t* __restrict__ flag) {
    const uint32_t m_chunks = (m + 127) / 128, n_tiles = (n + bn - 1) / bn, kgs = (kb_total + 3) / 4;
    const uint64_t total_a = uint64_t(kgs) * m_chunks * 128, total_b = uint64_t(kgs) * n_tiles * chunks * 128;
    bool bad = false;
    for (uint64_t i = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; i < total_a + total_b;
         i += uint64_t(gridDim.x) * blockDim.x) {
        const bool is_a = i < total_a;
        const uint64_t j = is_a ? i : i - total_a;
        const uint32_t pos = uint32_t(j & 127), l = pos >> 2, r = pos & 3, e = r * 32 + l;   // e: entry inside the chunk
        const uint64_t chunk = j >> 7;
        const float* src = nullptr;
        uint32_t kg;
        if (is_a) {
            kg = uint32_t(chunk / m_chunks);
            const uint32_t row = uint32_t(chunk % m_chunks) * 128 + e;
            if (row < m) src = sfa + uint64_t(row) * sfa_ld;
        } else {
            const uint32_t c = uint32_t(chunk % chunks);
            const uint64_t t = chunk / chunks;
            kg = uint32_t(t / n_tiles);
            const uint32_t in_tile = c * 128 + e;
            const uint32_t col = uint32_t(t % n_tiles) * bn + in_tile;
            if (in_tile < bn && col < n) src = sfb + uint64_t(col >> 7) * sfb_ld;
        }
        uint32_t word = 0x7F7F7F7Fu;   // 2^0 everywhere
        if (src != nullptr) {
            word = 0;
            for (uint32_t b = 0; b < 4; ++b) {
                const uint32_t kb = kg * 4 + b;
                uint32_t bits = 0x3F800000u;
                if (kb < kb_total) bits = __float_as_uint(__ldg(src + kb));
                const uint32_t ex = (bits >> 23) & 0xFF;
                if ((bits & 0x807FFFFFu) != 0 || ex == 0 || ex == 255) bad = true;
                word |= ex << (8 * b);
            }
        }
        (is_a ? out_a : out_b)[j] = word;
    }
    if (bad) atomicOr(flag, 1);
}

template <int kCta, uint32_t kBN>
static int launch_fp8_inst(const Fp8Launch& L, const Fp8Params& fp, cudaStream_t stream) {
    auto kernel = gemm_fp8_sm100_kernel<kCta, kBN>;
    int rc = set_smem_attr(kernel, L.smem);
    if (rc != ACLNN_SUCCESS) return rc;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(L.grid);
    cfg.blockDim = dim3(kFp8Threads);
    cfg.dynamicSmemBytes = L.smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    fill_launch_attrs(cfg, attr, kCta);
    DGA_CUDA(cudaLaunchKernelEx(&cfg, kernel, L.tm_a, L.tm_b, L.tm_d, fp));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

template <int kCta>
static int launch_mx_inst(const Fp8Launch& L, const GemmParams& gp, cudaStream_t stream) {
    auto kernel = gemm_sm100_kernel<kCta, __nv_bfloat16, false, true>;
    int rc = set_smem_attr(kernel, L.mx_smem);
    if (rc != ACLNN_SUCCESS) return rc;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(L.mx_grid);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = L.mx_smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    fill_launch_attrs(cfg, attr, kCta);
    static const PeerMaps no_peers{};
    DGA_CUDA(cudaLaunchKernelEx(&cfg, kernel, L.tm_a, L.tm_b_mx, L.tm_d_mx, L.tm_sfa, L.tm_sfb, no_peers, gp));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

// mode: 0 = decide on the device from the scale values (both kernels are enqueued, the gate flag lets exactly one
// run), 1 = promotion kernel only, 2 = block-scaled kernel only (caller guarantees power-of-two scales).
static int launch_fp8(Fp8Launch& L, void* workspace, cudaStream_t stream) {
    static const int mode = [] { const char* e = std::getenv("DGA_FP8_MODE"); return e ? std::atoi(e) : 0; }();
    if (!L.encoded) {
        int rc = make_tmap_2d(&L.tm_a, L.a, ACL_FLOAT8_E4M3FN, L.k, L.m, L.lda, 128, 128, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc != ACLNN_SUCCESS) return rc;
        // promotion kernel: two boxes of bn / (2 cta) weight rows per CTA and stage, one per column half
        rc = make_tmap_2d(&L.tm_b, L.b, ACL_FLOAT8_E4M3FN, L.k, L.n, L.ldb, 128, L.bn / (2 * L.cta), CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc != ACLNN_SUCCESS) return rc;
        rc = make_tmap_2d(&L.tm_b_mx, L.b, ACL_FLOAT8_E4M3FN, L.k, L.n, L.ldb, 128, L.mx_bn / L.mx_cta,
                          CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc != ACLNN_SUCCESS) return rc;
        rc = make_tmap_2d(&L.tm_d, L.d, ACL_BF16, L.n, L.m, L.ldd, 32, 128, CU_TENSOR_MAP_SWIZZLE_64B);
        if (rc != ACLNN_SUCCESS) return rc;
        const uint32_t epi = epi_cols_for(L.mx_bn, 2);
        rc = make_tmap_2d(&L.tm_d_mx, L.d, ACL_BF16, L.n, L.m, L.ldd, epi, 128,
                          epi == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B);
        if (rc != ACLNN_SUCCESS) return rc;
        L.encoded = true;
    }
    char* ws = static_cast<char*>(workspace);
    const uint32_t kb_total = ceil_div(L.k, 128), kgs = ceil_div(L.k, 512);
    const uint32_t m_chunks = ceil_div(L.m, 128), n_tiles = ceil_div(L.n, L.mx_bn);
    if (mode != 1 && L.encoded_ws != workspace) {
        int rc = make_tmap_2d(&L.tm_sfa, ws + L.ws_sfa_off, kAclInt32, 128, uint64_t(kgs) * m_chunks, 128, 128, 1,
                              CU_TENSOR_MAP_SWIZZLE_NONE);
        if (rc != ACLNN_SUCCESS) return rc;
        rc = make_tmap_2d(&L.tm_sfb, ws + L.ws_sfb_off, kAclInt32, 128, uint64_t(kgs) * n_tiles * L.mx_chunks, 128, 128,
                          L.mx_chunks, CU_TENSOR_MAP_SWIZZLE_NONE);
        if (rc != ACLNN_SUCCESS) return rc;
        L.encoded_ws = workspace;
    }
    int32_t* flag = reinterpret_cast<int32_t*>(ws + L.ws_flag_off);
    if (mode != 1) {
        DGA_CUDA(cudaMemsetAsync(flag, 0, sizeof(int32_t), stream));
        const uint64_t words = uint64_t(kgs) * (uint64_t(m_chunks) + uint64_t(n_tiles) * L.mx_chunks) * 128;
        const uint32_t blocks = uint32_t(std::min<uint64_t>((words + 255) / 256, 148 * 8));
        fp8_pack_scales_kernel<<<blocks, 256, 0, stream>>>(L.sfa, L.sfb, L.m, L.n, kb_total, uint32_t(L.sfa_ld),
                                                           uint32_t(L.sfb_ld), L.mx_bn, L.mx_chunks,
                                                           reinterpret_cast<uint32_t*>(ws + L.ws_sfa_off),
                                                           reinterpret_cast<uint32_t*>(ws + L.ws_sfb_off), flag);
        DGA_CUDA(cudaGetLastError());
        g_launches.fetch_add(1, std::memory_order_relaxed);
        GemmParams gp{};
        gp.m = L.m; gp.n = L.n; gp.k = L.k;
        gp.block_n = L.mx_bn;
        gp.num_stages = L.mx_stages;
        gp.k_blocks = kb_total;
        gp.num_m_blocks = ceil_div(L.m, 128 * L.mx_cta);
        gp.num_n_blocks = n_tiles;
        const WalkOrder walk = choose_walk(gp.num_m_blocks, gp.num_n_blocks, uint64_t(128) * L.mx_cta * L.k,
                                           uint64_t(L.mx_bn) * L.k, L.m > L.n);
        gp.raster_group = g_opt_raster.load(std::memory_order_relaxed) > 0 ? uint32_t(g_opt_raster.load()) : walk.group;
        gp.raster_along_m = walk.along_m;
        if (g_opt_raster_dir.load(std::memory_order_relaxed) >= 0) gp.raster_along_m = uint32_t(g_opt_raster_dir.load());
        gp.hint_a = hint_code(g_opt_hint_a.load(std::memory_order_relaxed) >= 0 ? g_opt_hint_a.load() : walk.hint_a);
        gp.hint_b = hint_code(g_opt_hint_b.load(std::memory_order_relaxed) >= 0 ? g_opt_hint_b.load() : walk.hint_b);
        // block-scaled instruction descriptor: e4m3 x e4m3, UE8M0 scales (bit 23), N >> 3 at [17,23), M >> 4 at [24,29)
        gp.idesc = (1u << 23) | ((L.mx_bn >> 3) << 17) | (((128u * L.mx_cta) >> 4) << 24);
        gp.b_stage_bytes = (L.mx_bn / L.mx_cta) * 128;
        gp.k_block_elems = 128;
        gp.sf_n_chunks = L.mx_chunks;
        gp.epi_cols = epi_cols_for(L.mx_bn, 2);
        gp.dp_tiles = gp.num_m_blocks * gp.num_n_blocks;   // whole tiles only: the scale words come in groups of 4 K blocks
        gp.sk_tiles = 0;
        gp.sf_m_chunks = m_chunks;
        gp.gate = mode == 0 ? flag : nullptr;
        gp.gate_expect = 0;
        int rc = L.mx_cta == 2 ? launch_mx_inst<2>(L, gp, stream) : launch_mx_inst<1>(L, gp, stream);
        if (rc != ACLNN_SUCCESS) return rc;
        if (mode == 2) return ACLNN_SUCCESS;
    }
    Fp8Params fp{};
    fp.m = L.m; fp.n = L.n; fp.k = L.k;
    fp.block_n = L.bn;
    fp.num_stages = L.stages;
    fp.k_blocks = kb_total;
    fp.num_m_blocks = ceil_div(L.m, 128 * L.cta);
    fp.num_n_blocks = ceil_div(L.n, L.bn);
    const WalkOrder walk_p = choose_walk(fp.num_m_blocks, fp.num_n_blocks, uint64_t(128) * L.cta * L.k, uint64_t(L.bn) * L.k,
                                         L.m > L.n);
    fp.raster_group = g_opt_raster.load(std::memory_order_relaxed) > 0 ? uint32_t(g_opt_raster.load()) : walk_p.group;
    fp.raster_along_m = walk_p.along_m;
    if (g_opt_raster_dir.load(std::memory_order_relaxed) >= 0) fp.raster_along_m = uint32_t(g_opt_raster_dir.load());
    fp.hint_a = hint_code(g_opt_hint_a.load(std::memory_order_relaxed) >= 0 ? g_opt_hint_a.load() : walk_p.hint_a);
    fp.hint_b = hint_code(g_opt_hint_b.load(std::memory_order_relaxed) >= 0 ? g_opt_hint_b.load() : walk_p.hint_b);
    fp.idesc = make_idesc(128 * L.cta, L.bn / 2, 0u, 0, 0);   // one column half per instruction
    fp.sfa_ld = uint32_t(L.sfa_ld); fp.sfb_ld = uint32_t(L.sfb_ld);
    fp.sfb_rows = ceil_div(L.n, 128);
    fp.sfa = L.sfa; fp.sfb = L.sfb;
    fp.gate = mode == 0 ? flag : nullptr;
    fp.gate_expect = 1;
    if (L.cta == 1) return launch_fp8_inst<1, 128>(L, fp, stream);
    switch (L.bn) {
        case 256: return launch_fp8_inst<2, 256>(L, fp, stream);
        case 224: return launch_fp8_inst<2, 224>(L, fp, stream);
        case 192: return launch_fp8_inst<2, 192>(L, fp, stream);
        case 160: return launch_fp8_inst<2, 160>(L, fp, stream);
        case 128: return launch_fp8_inst<2, 128>(L, fp, stream);
    }
    set_error("fp8 tile not instantiated");
    return ACLNN_ERR_INNER_TILING_ERROR;
}

__global__ void mmad_custom_stub(const void*, const void*, void*) {}

}  // namespace dga

// =========================================================================================== C ABI ====
using namespace dga;

struct aclOpExecutor {
    DenseLaunch dense;
    Fp8Launch fp8;
    bool is_fp8 = false;
    bool repeatable = false;
};

extern "C" {

aclTensor* aclCreateTensor(const int64_t* viewDims, uint64_t viewDimsNum, aclDataType dataType, const int64_t* stride,
                           int64_t offset, aclFormat, const int64_t*, uint64_t, void* tensorData) {
    if (viewDims == nullptr || viewDimsNum > 8) return nullptr;
    aclTensor* t = new aclTensor{};
    t->dtype = dataType;
    t->ndim = static_cast<int32_t>(viewDimsNum);
    int64_t run = 1;
    for (int64_t i = int64_t(viewDimsNum) - 1; i >= 0; --i) {
        t->shape[i] = viewDims[i];
        t->stride[i] = stride ? stride[i] : run;
        run *= viewDims[i];
    }
    t->data = static_cast<char*>(tensorData) + offset * int64_t(elem_bytes(dataType));
    return t;
}

aclnnStatus aclDestroyTensor(const aclTensor* tensor) {
    delete tensor;
    return ACLNN_SUCCESS;
}

static bool dtype_ok_in(int32_t d) { return d == ACL_FLOAT16 || d == ACL_BF16; }

// Shared by phase 1 and dgaMatmulWithPlan: InferShape / InferDataType / TilingFunc checks + layout detection.
static aclnnStatus validate_dense(const aclTensor* self, const aclTensor* mat2, const aclTensor* out, DenseLaunch& L) {
    if (self == nullptr || mat2 == nullptr || out == nullptr) {
        set_error("null tensor");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    if (self->ndim != 2 || mat2->ndim != 2 || out->ndim != 2) {  // catlass_dynamic_matmul.cpp:22-26
        set_error("Invalid input dimensions. self, mat2 and out must be 2-D");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (self->dtype != mat2->dtype || !dtype_ok_in(self->dtype)) {  // catlass_dynamic_matmul.cpp:37-46
        set_error("self and mat2 must share a dtype in {float16, bfloat16}");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (out->dtype != self->dtype && out->dtype != ACL_FLOAT) {
        set_error("out dtype must equal the input dtype (or float32 for the framework path)");
        return ACLNN_ERR_PARAM_INVALID;
    }
    const int64_t m = self->shape[0], ka = self->shape[1], kb = mat2->shape[0], n = mat2->shape[1];
    if (ka != kb) {  // catlass_dynamic_matmul_tiling.cpp:96-100
        set_error("K dimension mismatch: self[1]=" + std::to_string(ka) + " mat2[0]=" + std::to_string(kb));
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (out->shape[0] != m || out->shape[1] != n) {
        set_error("out shape must be [self.dim0, mat2.dim1]");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (m <= 0 || n <= 0 || ka <= 0 || m > UINT32_MAX || n > UINT32_MAX || ka > UINT32_MAX) {
        set_error("m, n, k must be positive 32-bit extents");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (self->stride[1] != 1 && !(ka == 1)) {
        set_error("self must be row-major (unit stride along K)");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (out->stride[1] != 1 && !(n == 1)) {
        set_error("out must be row-major");
        return ACLNN_ERR_PARAM_INVALID;
    }
    DgaTilingPlan& p = L.plan;
    p = DgaTilingPlan{};
    p.m = uint32_t(m); p.n = uint32_t(n); p.k = uint32_t(ka);
    p.dtype = uint8_t(self->dtype);
    p.outDtype = uint8_t(out->dtype);
    // the reference's "todo use mat stride to check transpose" (catlass_dynamic_matmul_tiling.cpp:82-84)
    if (mat2->stride[0] == 1 && (mat2->stride[1] >= ka || n == 1)) {
        p.layoutTagB = 1;
        L.ldb = n == 1 ? uint64_t(ka) : uint64_t(mat2->stride[1]);
    } else if (mat2->stride[1] == 1 && mat2->stride[0] >= n) {
        p.layoutTagB = 0;
        L.ldb = uint64_t(mat2->stride[0]);
    } else {
        set_error("mat2 must be column-major (weight stored [N,K]) or row-major [K,N]");
        return ACLNN_ERR_PARAM_INVALID;
    }
    L.lda = m == 1 ? uint64_t(ka) : uint64_t(self->stride[0]);
    L.ldd = m == 1 ? uint64_t(n) : uint64_t(out->stride[0]);
    if (L.lda < uint64_t(ka) || L.ldd < uint64_t(n)) {
        set_error("leading dimension smaller than the row length");
        return ACLNN_ERR_PARAM_INVALID;
    }
    L.a = self->data; L.b = mat2->data; L.d = out->data;
    return ACLNN_SUCCESS;
}

aclnnStatus aclnnCatlassDynamicMatmulGetWorkspaceSize(const aclTensor* self, const aclTensor* mat2,
                                                      const aclTensor* out, uint64_t* workspaceSize,
                                                      aclOpExecutor** executor) {
    if (workspaceSize == nullptr || executor == nullptr) {
        set_error("null output argument");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    aclOpExecutor* ex = new aclOpExecutor{};
    aclnnStatus rc = validate_dense(self, mat2, out, ex->dense);
    if (rc == ACLNN_SUCCESS) rc = select_kernel_with_cache(ex->dense.plan);
    if (rc == ACLNN_SUCCESS) rc = apply_padding(ex->dense);
    if (rc != ACLNN_SUCCESS) {
        delete ex;
        return rc;
    }
    *workspaceSize = ex->dense.plan.workspaceBytes;
    *executor = ex;
    return ACLNN_SUCCESS;
}

aclnnStatus aclnnCatlassDynamicMatmul(void* workspace, uint64_t workspaceSize, aclOpExecutor* executor,
                                      aclrtStream stream) {
    if (executor == nullptr || executor->is_fp8) {
        set_error("null or mismatched executor");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    aclnnStatus rc = ACLNN_SUCCESS;
    if (workspaceSize < executor->dense.plan.workspaceBytes ||
        (executor->dense.plan.workspaceBytes > 0 && workspace == nullptr)) {
        set_error("workspace too small");
        rc = ACLNN_ERR_PARAM_INVALID;
    } else {
        rc = launch_dense(executor->dense, static_cast<cudaStream_t>(stream), workspace);
    }
    if (!executor->repeatable) delete executor;
    return rc;
}

aclnnStatus dgaSetExecutorRepeatable(aclOpExecutor* executor) {
    if (executor == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    executor->repeatable = true;
    return ACLNN_SUCCESS;
}
aclnnStatus dgaDestroyExecutor(aclOpExecutor* executor) {
    delete executor;
    return ACLNN_SUCCESS;
}

aclnnStatus dgaMatmulWithPlan(const aclTensor* self, const aclTensor* mat2, const aclTensor* out,
                              const DgaTilingPlan* plan, void* workspace, uint64_t workspaceSize, aclrtStream stream) {
    if (plan == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    DenseLaunch L;
    aclnnStatus rc = validate_dense(self, mat2, out, L);
    if (rc != ACLNN_SUCCESS) return rc;
    DgaTilingPlan p = *plan;
    // shape, dtype and layout always come from the tensors
    p.m = L.plan.m; p.n = L.plan.n; p.k = L.plan.k;
    p.dtype = L.plan.dtype; p.outDtype = L.plan.outDtype; p.layoutTagB = L.plan.layoutTagB;
    L.plan = p;
    rc = apply_padding(L);   // also finalises the plan
    if (rc != ACLNN_SUCCESS) return rc;
    p = L.plan;
    if (workspaceSize < p.workspaceBytes) {
        set_error("workspace too small: the plan needs " + std::to_string(p.workspaceBytes) + " bytes");
        return ACLNN_ERR_PARAM_INVALID;
    }
    return launch_dense(L, static_cast<cudaStream_t>(stream), workspace);
}

// ---- two decode GEMMs that share their activations ---------------------------------------------------------------
static void seed_plan(DgaTilingPlan& p, uint32_t m, uint32_t n, uint32_t k, int32_t dtype, int32_t outDtype,
                      int32_t layoutTagB);
aclnnStatus dgaDecodePairMatmul(const aclTensor* self, const aclTensor* mat2First, const aclTensor* outFirst,
                                const aclTensor* mat2Second, const aclTensor* outSecond, aclrtStream stream) {
    DenseLaunch first, second;
    aclnnStatus rc = validate_dense(self, mat2First, outFirst, first);
    if (rc == ACLNN_SUCCESS) rc = validate_dense(self, mat2Second, outSecond, second);
    if (rc != ACLNN_SUCCESS) return rc;
    const DgaTilingPlan& pf = first.plan;
    if (pf.layoutTagB != 1 || second.plan.layoutTagB != 1 || pf.outDtype != second.plan.outDtype || pf.m > 128) {
        set_error("dgaDecodePairMatmul: m <= 128, both weights stored [N,K], one output dtype (larger or row-major "
                  "problems: two calls of the ordinary entry)");
        return ACLNN_ERR_PARAM_INVALID;
    }
    auto unaligned = [](const void* ptr, uint64_t ld) { return (reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (ld * 2) % 16 != 0; };
    if (unaligned(first.a, first.lda) || unaligned(first.b, first.ldb) || unaligned(second.b, second.ldb)) {
        set_error("dgaDecodePairMatmul: base pointers and row pitches of self and the weights must be 16-byte aligned");
        return ACLNN_ERR_PARAM_INVALID;
    }
    // one decode plan for the tiles of both weights: the selector's, or - where it would leave the decode kernel -
    // the widest K split that keeps two CTAs per SM busy
    const uint32_t tiles = ceil_div(pf.n, 128) + ceil_div(second.plan.n, 128);
    DgaTilingPlan p;
    seed_plan(p, pf.m, tiles * 128, pf.k, pf.dtype, pf.outDtype, 1);
    rc = select_kernel(p);
    if (rc != ACLNN_SUCCESS || p.kernelSerial != DGA_KERNEL_SPLITK) {
        seed_plan(p, pf.m, tiles * 128, pf.k, pf.dtype, pf.outDtype, 1);
        p.kernelSerial = DGA_KERNEL_SPLITK;
        const uint32_t kb = ceil_div(pf.k, 64), slots = 2 * platform().smCount;
        uint32_t split = 8;
        while (split > 1 && (tiles * split > slots || split > kb)) split >>= 1;
        p.splitkFactor = uint16_t(split);
        rc = finalize_plan(p);
        if (rc != ACLNN_SUCCESS) return rc;
    }
    return launch_decode_pair(first, second, p, static_cast<cudaStream_t>(stream));
}

// ---- N-sharded matmul with the gather of C fused into the epilogue ----------------------------------------------
aclnnStatus dgaNShardMatmulGetWorkspaceSize(const aclTensor* self, const aclTensor* mat2Shard, void* const* outPeers,
                                            uint32_t worldSize, uint32_t rank, int64_t nTotal, int64_t ldOut,
                                            uint64_t* workspaceSize, aclOpExecutor** executor) {
    if (outPeers == nullptr || workspaceSize == nullptr || executor == nullptr || self == nullptr || mat2Shard == nullptr) {
        set_error("null argument");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    if (worldSize == 0 || worldSize > kMaxPeers + 1 || rank >= worldSize) {
        set_error("N-sharded matmul supports 1.." + std::to_string(kMaxPeers + 1) + " ranks");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (mat2Shard->ndim != 2 || nTotal <= 0 || nTotal > UINT32_MAX || ldOut < nTotal) {
        set_error("bad N-shard geometry");
        return ACLNN_ERR_PARAM_INVALID;
    }
    uint32_t n_begin = 0, n_count = 0;
    aclnnStatus rc = dgaNShardRange(uint32_t(nTotal), worldSize, rank, &n_begin, &n_count);
    if (rc != ACLNN_SUCCESS) return rc;
    if (mat2Shard->shape[1] != int64_t(n_count)) {
        set_error("mat2 shard must hold columns [" + std::to_string(n_begin) + ", " + std::to_string(n_begin + n_count) +
                  ") of the weight");
        return ACLNN_ERR_PARAM_INVALID;
    }
    // out = this rank's column slice of its own full C; peers get the same slice of theirs
    const uint32_t eb = elem_bytes(self->dtype);
    aclTensor out{};
    out.dtype = self->dtype;
    out.ndim = 2;
    out.shape[0] = self->shape[0]; out.shape[1] = n_count;
    out.stride[0] = ldOut; out.stride[1] = 1;
    out.data = static_cast<char*>(outPeers[rank]) + uint64_t(n_begin) * eb;
    aclOpExecutor* ex = new aclOpExecutor{};
    rc = validate_dense(self, mat2Shard, &out, ex->dense);
    if (rc == ACLNN_SUCCESS) {
        // the persistent kernel carries the peer stores; small-M and stream-K variants are not used here
        DgaTilingPlan& p = ex->dense.plan;
        rc = select_kernel_with_cache(p);
        if (rc == ACLNN_SUCCESS && p.kernelSerial != DGA_KERNEL_COMMON && p.kernelSerial != DGA_KERNEL_SMALL) {
            DgaTilingPlan q{};
            q.m = p.m; q.n = p.n; q.k = p.k; q.dtype = p.dtype; q.outDtype = p.outDtype; q.layoutTagB = p.layoutTagB;
            q.kernelSerial = DGA_KERNEL_COMMON; q.ctaGroup = 2; q.n1 = 128;
            rc = finalize_plan(q);
            if (rc == ACLNN_SUCCESS) { q.strideA = p.strideA; q.strideB = p.strideB; q.strideC = p.strideC; p = q; }
        }
    }
    if (rc == ACLNN_SUCCESS) rc = apply_padding(ex->dense, /*allow_repack=*/false);
    if (rc != ACLNN_SUCCESS) {
        delete ex;
        return rc;
    }
    for (uint32_t r = 0, i = 0; r < worldSize; ++r) {
        if (r == rank) continue;
        ex->dense.peer_d[i++] = static_cast<char*>(outPeers[r]) + uint64_t(n_begin) * eb;
    }
    ex->dense.num_peers = worldSize - 1;
    *workspaceSize = ex->dense.plan.workspaceBytes;
    *executor = ex;
    return ACLNN_SUCCESS;
}
aclnnStatus dgaNShardMatmul(void* workspace, uint64_t workspaceSize, aclOpExecutor* executor, aclrtStream stream) {
    return aclnnCatlassDynamicMatmul(workspace, workspaceSize, executor, stream);
}

// ---- fp8 block-scaled ------------------------------------------------------------------------------------
aclnnStatus dgaFp8BlockScaledMatmulGetWorkspaceSize(const aclTensor* self, const aclTensor* selfScale,
                                                    const aclTensor* mat2, const aclTensor* mat2Scale,
                                                    const aclTensor* out, uint64_t* workspaceSize,
                                                    aclOpExecutor** executor) {
    if (!self || !selfScale || !mat2 || !mat2Scale || !out || !workspaceSize || !executor) {
        set_error("null argument");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    if (self->ndim != 2 || mat2->ndim != 2 || out->ndim != 2 || selfScale->ndim != 2 || mat2Scale->ndim != 2) {
        set_error("Invalid input dimensions. all tensors must be 2-D");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (self->dtype != ACL_FLOAT8_E4M3FN || mat2->dtype != ACL_FLOAT8_E4M3FN || selfScale->dtype != ACL_FLOAT ||
        mat2Scale->dtype != ACL_FLOAT || out->dtype != ACL_BF16) {
        set_error("fp8 matmul wants e4m3 operands, float32 scales and a bfloat16 output");
        return ACLNN_ERR_PARAM_INVALID;
    }
    const int64_t m = self->shape[0], k = self->shape[1], n = mat2->shape[1];
    if (mat2->shape[0] != k) {
        set_error("K dimension mismatch: self[1]=" + std::to_string(k) + " mat2[0]=" + std::to_string(mat2->shape[0]));
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (m <= 0 || n <= 0 || k <= 0 || m > UINT32_MAX || n > UINT32_MAX || k > UINT32_MAX || out->shape[0] != m ||
        out->shape[1] != n) {
        set_error("bad extents");
        return ACLNN_ERR_PARAM_INVALID;
    }
    const int64_t kb = (k + 127) / 128, nb = (n + 127) / 128;
    if (selfScale->shape[0] != m || selfScale->shape[1] != kb || mat2Scale->shape[0] != nb || mat2Scale->shape[1] != kb) {
        set_error("scale shapes must be [M, ceil(K/128)] and [ceil(N/128), ceil(K/128)]");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if ((self->stride[1] != 1 && k != 1) || mat2->stride[0] != 1 || (out->stride[1] != 1 && n != 1) ||
        (selfScale->stride[1] != 1 && kb != 1) || (mat2Scale->stride[1] != 1 && kb != 1)) {
        set_error("fp8 matmul layout: self row-major, mat2 column-major (weight stored [N,K]), scales and out row-major");
        return ACLNN_ERR_PARAM_INVALID;
    }
    aclOpExecutor* ex = new aclOpExecutor{};
    ex->is_fp8 = true;
    Fp8Launch& L = ex->fp8;
    L.m = uint32_t(m); L.n = uint32_t(n); L.k = uint32_t(k);
    L.a = self->data; L.b = mat2->data; L.d = out->data;
    L.sfa = static_cast<const float*>(selfScale->data);
    L.sfb = static_cast<const float*>(mat2Scale->data);
    L.lda = m == 1 ? uint64_t(k) : uint64_t(self->stride[0]);
    L.ldb = n == 1 ? uint64_t(k) : uint64_t(mat2->stride[1]);
    L.ldd = m == 1 ? uint64_t(n) : uint64_t(out->stride[0]);
    L.sfa_ld = m == 1 ? uint64_t(kb) : uint64_t(selfScale->stride[0]);
    L.sfb_ld = nb == 1 ? uint64_t(kb) : uint64_t(mat2Scale->stride[0]);
    if (L.lda % 16 || L.ldb % 16 || (L.ldd * 2) % 16) {
        set_error("leading dimensions must be 16-byte aligned");
        delete ex;
        return ACLNN_ERR_INNER_TILING_ERROR;
    }
    plan_fp8(L);
    *workspaceSize = L.ws_bytes;
    *executor = ex;
    return ACLNN_SUCCESS;
}
aclnnStatus dgaFp8BlockScaledMatmul(void* workspace, uint64_t workspaceSize, aclOpExecutor* executor,
                                    aclrtStream stream) {
    if (executor == nullptr || !executor->is_fp8) {
        set_error("not an fp8 executor");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    aclnnStatus rc;
    if (workspace == nullptr || workspaceSize < executor->fp8.ws_bytes ||
        (reinterpret_cast<uintptr_t>(workspace) & 127) != 0) {
        set_error("fp8 matmul needs the workspace phase 1 asked for (128-byte aligned)");
        rc = ACLNN_ERR_PARAM_INVALID;
    } else {
        rc = launch_fp8(executor->fp8, workspace, static_cast<cudaStream_t>(stream));
    }
    if (!executor->repeatable) delete executor;
    return rc;
}

aclnnStatus dgaFp8QueryPath(void* workspace, aclOpExecutor* executor, aclrtStream stream, int32_t* path) {
    if (workspace == nullptr || executor == nullptr || path == nullptr || !executor->is_fp8) {
        set_error("dgaFp8QueryPath wants an fp8 executor, its workspace and an output");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    DGA_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    int32_t flag = -1;
    DGA_CUDA(cudaMemcpy(&flag, static_cast<char*>(workspace) + executor->fp8.ws_flag_off, sizeof(flag), cudaMemcpyDeviceToHost));
    *path = flag != 0 ? 1 : 0;
    return ACLNN_SUCCESS;
}

// ---- M-grouped contiguous matmul (MoE) -------------------------------------------------------------------------
static void seed_plan(DgaTilingPlan& p, uint32_t m, uint32_t n, uint32_t k, int32_t dtype, int32_t outDtype,
                      int32_t layoutTagB);
static int plan_single_cta(DgaTilingPlan& p, uint32_t slabs_of_tiles);
aclnnStatus dgaMGroupedMatmulGetWorkspaceSize(const aclTensor* self, const aclTensor* weight, const aclTensor* groupIndex,
                                              const aclTensor* out, uint64_t* workspaceSize, aclOpExecutor** executor) {
    if (!self || !weight || !groupIndex || !out || !workspaceSize || !executor) {
        set_error("null argument");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    if (self->ndim != 2 || weight->ndim != 3 || groupIndex->ndim != 1 || out->ndim != 2) {
        set_error("Invalid input dimensions. self [M,K], weight [G,N,K], groupIndex [M], out [M,N]");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (self->dtype != weight->dtype || !dtype_ok_in(self->dtype) || out->dtype != self->dtype ||
        groupIndex->dtype != kAclInt32) {
        set_error("self, weight and out must share a dtype in {float16, bfloat16}; groupIndex must be int32");
        return ACLNN_ERR_PARAM_INVALID;
    }
    const int64_t m = self->shape[0], k = self->shape[1], g = weight->shape[0], n = weight->shape[1];
    if (weight->shape[2] != k) {
        set_error("K dimension mismatch: self[1]=" + std::to_string(k) + " weight[2]=" + std::to_string(weight->shape[2]));
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (out->shape[0] != m || out->shape[1] != n || groupIndex->shape[0] != m) {
        set_error("out must be [M, N] and groupIndex [M]");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (m <= 0 || n <= 0 || k <= 0 || g <= 0 || m > UINT32_MAX || n > UINT32_MAX || k > UINT32_MAX || m % 128 != 0) {
        set_error("m, n, k, groups must be positive and M_total a multiple of 128 (every 128-row block has one group)");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (self->stride[1] != 1 || weight->stride[2] != 1 || out->stride[1] != 1 || groupIndex->stride[0] != 1 ||
        self->stride[0] < k || weight->stride[1] < k || out->stride[0] < n ||
        (g > 1 && weight->stride[0] < weight->stride[1] * n)) {
        set_error("self, weight (K contiguous) and out must be row-major");
        return ACLNN_ERR_PARAM_INVALID;
    }
    aclOpExecutor* ex = new aclOpExecutor{};
    DenseLaunch& L = ex->dense;
    seed_plan(L.plan, uint32_t(m), uint32_t(n), uint32_t(k), self->dtype, out->dtype, /*layoutTagB=*/1);
    L.a = self->data; L.b = weight->data; L.d = out->data;
    L.lda = uint64_t(self->stride[0]); L.ldb = uint64_t(weight->stride[1]); L.ldd = uint64_t(out->stride[0]);
    L.plan.strideA = L.lda; L.plan.strideB = L.ldb; L.plan.strideC = L.ldd;
    L.mode = kModeMGrouped;
    L.groups = uint32_t(g);
    L.slab_b = g > 1 ? uint64_t(weight->stride[0]) : uint64_t(n) * L.ldb;
    L.group_index = static_cast<const int32_t*>(groupIndex->data);
    int rc = plan_single_cta(L.plan, 1);
    const uint32_t eb = 2;
    if (rc == ACLNN_SUCCESS &&
        (((L.lda | L.ldb | L.ldd | L.slab_b) * eb) % 16 != 0 ||
         ((reinterpret_cast<uintptr_t>(L.a) | reinterpret_cast<uintptr_t>(L.b) | reinterpret_cast<uintptr_t>(L.d)) & 15))) {
        set_error("grouped matmul wants 16-byte aligned base pointers and pitches");
        rc = ACLNN_ERR_INNER_TILING_ERROR;
    }
    if (rc != ACLNN_SUCCESS) {
        delete ex;
        return rc;
    }
    *workspaceSize = 0;
    *executor = ex;
    return ACLNN_SUCCESS;
}
// Masked layout: self [G, Mmax, K], weight [G, N, K], maskedM [G] int32, out [G, Mmax, N].
aclnnStatus dgaMGroupedMaskedMatmulGetWorkspaceSize(const aclTensor* self, const aclTensor* weight,
                                                    const aclTensor* maskedM, const aclTensor* out,
                                                    uint64_t* workspaceSize, aclOpExecutor** executor) {
    if (!self || !weight || !maskedM || !out || !workspaceSize || !executor) {
        set_error("null argument");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    if (self->ndim != 3 || weight->ndim != 3 || maskedM->ndim != 1 || out->ndim != 3) {
        set_error("Invalid input dimensions. self [G,Mmax,K], weight [G,N,K], maskedM [G], out [G,Mmax,N]");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (self->dtype != weight->dtype || !dtype_ok_in(self->dtype) || out->dtype != self->dtype ||
        maskedM->dtype != kAclInt32) {
        set_error("self, weight and out must share a dtype in {float16, bfloat16}; maskedM must be int32");
        return ACLNN_ERR_PARAM_INVALID;
    }
    const int64_t g = self->shape[0], m = self->shape[1], k = self->shape[2], n = weight->shape[1];
    if (weight->shape[2] != k) {
        set_error("K dimension mismatch: self[2]=" + std::to_string(k) + " weight[2]=" + std::to_string(weight->shape[2]));
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (weight->shape[0] != g || maskedM->shape[0] != g || out->shape[0] != g || out->shape[1] != m || out->shape[2] != n) {
        set_error("weight must be [G, N, K], maskedM [G] and out [G, Mmax, N]");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (m <= 0 || n <= 0 || k <= 0 || g <= 0 || m > UINT32_MAX || n > UINT32_MAX || k > UINT32_MAX || g > UINT32_MAX) {
        set_error("g, m, n, k must be positive 32-bit values");
        return ACLNN_ERR_PARAM_INVALID;
    }
    if (self->stride[2] != 1 || weight->stride[2] != 1 || out->stride[2] != 1 || maskedM->stride[0] != 1 ||
        self->stride[1] < k || weight->stride[1] < k || out->stride[1] < n ||
        (g > 1 && (self->stride[0] < self->stride[1] * m || weight->stride[0] < weight->stride[1] * n ||
                   out->stride[0] < out->stride[1] * m))) {
        set_error("self, weight (K contiguous) and out must be row-major per group");
        return ACLNN_ERR_PARAM_INVALID;
    }
    aclOpExecutor* ex = new aclOpExecutor{};
    DenseLaunch& L = ex->dense;
    seed_plan(L.plan, uint32_t(m), uint32_t(n), uint32_t(k), self->dtype, out->dtype, /*layoutTagB=*/1);
    L.a = self->data; L.b = weight->data; L.d = out->data;
    L.lda = uint64_t(self->stride[1]); L.ldb = uint64_t(weight->stride[1]); L.ldd = uint64_t(out->stride[1]);
    L.plan.strideA = L.lda; L.plan.strideB = L.ldb; L.plan.strideC = L.ldd;
    L.mode = kModeMasked;
    L.batch = uint32_t(g);
    L.slab_a = g > 1 ? uint64_t(self->stride[0]) : uint64_t(m) * L.lda;
    L.slab_b = g > 1 ? uint64_t(weight->stride[0]) : uint64_t(n) * L.ldb;
    L.slab_d = g > 1 ? uint64_t(out->stride[0]) : uint64_t(m) * L.ldd;
    L.group_index = static_cast<const int32_t*>(maskedM->data);
    int rc = plan_single_cta(L.plan, uint32_t(g));
    if (rc == ACLNN_SUCCESS &&
        (((L.lda | L.ldb | L.ldd | L.slab_a | L.slab_b | L.slab_d) * 2) % 16 != 0 ||
         ((reinterpret_cast<uintptr_t>(L.a) | reinterpret_cast<uintptr_t>(L.b) | reinterpret_cast<uintptr_t>(L.d)) & 15))) {
        set_error("grouped matmul wants 16-byte aligned base pointers and pitches");
        rc = ACLNN_ERR_INNER_TILING_ERROR;
    }
    if (rc != ACLNN_SUCCESS) {
        delete ex;
        return rc;
    }
    *workspaceSize = 0;
    *executor = ex;
    return ACLNN_SUCCESS;
}
aclnnStatus dgaMGroupedMatmul(void* workspace, uint64_t workspaceSize, aclOpExecutor* executor, aclrtStream stream) {
    (void)workspace; (void)workspaceSize;
    if (executor == nullptr || executor->is_fp8 ||
        (executor->dense.mode != kModeMGrouped && executor->dense.mode != kModeMasked)) {
        set_error("not a grouped-matmul executor");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    DenseLaunch& L = executor->dense;
    L.eff = {L.a, L.b, L.d, L.lda, L.ldb, L.ldd};
    aclnnStatus rc = launch_common(L, static_cast<cudaStream_t>(stream), nullptr);
    if (!executor->repeatable) delete executor;
    return rc;
}

static int plan_slab_tiles(DgaTilingPlan& p, uint32_t slabs_of_tiles, uint32_t cta);
// The caller's promise about a grouped executor's row layout, made between phase 1 and phase 2.  rows = 256: every
// group of the contiguous layout starts on a multiple of 256 rows (padding blocks, index -1, fill each group up) resp.
// Mmax of the masked layout is a multiple of 256: the launch uses 256-row CTA-pair tiles (cta_group::2, the weight
// tile shared by the pair) instead of 128-row single-CTA tiles.  rows = 128 returns to the default.  A contiguous
// layout that breaks the promise gives wrong rows in the blocks that share a pair with another group.
aclnnStatus dgaSetExecutorGroupAlignment(aclOpExecutor* executor, uint32_t rows) {
    if (executor == nullptr || executor->is_fp8 ||
        (executor->dense.mode != kModeMGrouped && executor->dense.mode != kModeMasked)) {
        set_error("not a grouped-matmul executor");
        return ACLNN_ERR_PARAM_NULLPTR;
    }
    DenseLaunch& L = executor->dense;
    if ((rows != 128 && rows != 256) || L.plan.m % rows != 0) {
        set_error("group alignment is 128 or 256 rows and must divide the row count (" + std::to_string(L.plan.m) + ")");
        return ACLNN_ERR_PARAM_INVALID;
    }
    const DgaTilingPlan keep = L.plan;
    seed_plan(L.plan, keep.m, keep.n, keep.k, keep.dtype, keep.outDtype, keep.layoutTagB);
    L.plan.strideA = keep.strideA; L.plan.strideB = keep.strideB; L.plan.strideC = keep.strideC;
    const int rc = plan_slab_tiles(L.plan, L.mode == kModeMasked ? L.batch : 1u, rows / 128);
    if (rc != ACLNN_SUCCESS) L.plan = keep;
    L.encoded = false;   // the weight box depends on the CTA group
    return rc;
}

// ---- host tiling -------------------------------------------------------------------------------------
static void seed_plan(DgaTilingPlan& p, uint32_t m, uint32_t n, uint32_t k, int32_t dtype, int32_t outDtype,
                      int32_t layoutTagB) {
    p = DgaTilingPlan{};
    p.m = m; p.n = n; p.k = k;
    p.dtype = uint8_t(dtype); p.outDtype = uint8_t(outDtype); p.layoutTagB = uint8_t(layoutTagB);
}
aclnnStatus dgaSelectKernel(uint32_t m, uint32_t n, uint32_t k, int32_t dtype, int32_t outDtype, int32_t layoutTagB,
                            DgaTilingPlan* plan) {
    if (plan == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    seed_plan(*plan, m, n, k, dtype, outDtype, layoutTagB);
    return select_kernel(*plan);
}
aclnnStatus dgaSelectKernelWithCache(uint32_t m, uint32_t n, uint32_t k, int32_t dtype, int32_t outDtype,
                                     int32_t layoutTagB, DgaTilingPlan* plan) {
    if (plan == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    seed_plan(*plan, m, n, k, dtype, outDtype, layoutTagB);
    return select_kernel_with_cache(*plan);
}
aclnnStatus dgaTilingCacheSet(const DgaTilingPlan* plan) {
    if (plan == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    DgaTilingPlan p = *plan;
    aclnnStatus rc = finalize_plan(p);
    return rc != ACLNN_SUCCESS ? rc : cache_set(p);
}
aclnnStatus dgaTilingCacheGet(uint32_t m, uint32_t n, uint32_t k, int32_t dtype, int32_t outDtype, int32_t layoutTagB,
                              DgaTilingPlan* plan) {
    if (plan == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    seed_plan(*plan, m, n, k, dtype, outDtype, layoutTagB);
    return cache_get(*plan);
}
aclnnStatus dgaTilingCacheClear(void) { cache_clear(); return ACLNN_SUCCESS; }
uint64_t dgaTilingCacheSize(void) { return cache_size(); }
aclnnStatus dgaTilingCacheLoadCsv(const char* path) { return cache_load_csv(path); }
aclnnStatus dgaTilingCacheSaveCsv(const char* path) { return cache_save_csv(path); }

aclnnStatus dgaGetPlatformInfo(DgaPlatformInfo* info) {
    if (info == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    const PlatformInfo& hw = platform();
    *info = DgaPlatformInfo{hw.smCount, hw.smemBytes, hw.tmemCols, hw.l2Bytes, hw.hbmGBs, hw.denseTflops16, hw.denseTflops8};
    return ACLNN_SUCCESS;
}
aclnnStatus dgaSetPlatformInfo(const DgaPlatformInfo* info) {
    if (info == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    if (info->smCount < 2 || info->smemBytes < 65536) {
        set_error("implausible platform info");
        return ACLNN_ERR_PARAM_INVALID;
    }
    PlatformInfo& hw = platform();
    hw.smCount = info->smCount; hw.smemBytes = info->smemBytes; hw.tmemCols = info->tmemCols;
    hw.l2Bytes = info->l2Bytes; hw.hbmGBs = info->hbmGBs; hw.denseTflops16 = info->denseTflops16;
    hw.denseTflops8 = info->denseTflops8;
    return ACLNN_SUCCESS;
}

// ---- framework entry points ------------------------------------------------------------------------------
aclnnStatus dgaRunMmadCustom(const void* x, const void* y, void* z, aclrtStream stream) {
    mmad_custom_stub<<<8, 32, 0, static_cast<cudaStream_t>(stream)>>>(x, y, z);  // blockDim 8: gemm.hpp:18
    DGA_CUDA(cudaGetLastError());
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

// z[m,n] fp32 = x[m,k] * y[k,n], both row-major contiguous 16-bit.
static aclnnStatus framework_gemm(const void* x, const void* y, void* z, uint32_t m, uint32_t n, uint32_t k,
                                  int32_t dtype, cudaStream_t stream) {
    DenseLaunch L;
    seed_plan(L.plan, m, n, k, dtype, ACL_FLOAT, /*layoutTagB=*/0);
    aclnnStatus rc = select_kernel_with_cache(L.plan);
    if (rc != ACLNN_SUCCESS) return rc;
    L.a = x; L.b = y; L.d = z;
    L.lda = k; L.ldb = n; L.ldd = n;
    rc = apply_padding(L);   // the reference's kernel takes any (m, n, k): odd n or k go through a repack
    if (rc != ACLNN_SUCCESS) return rc;
    // the framework entry points have no workspace argument (python_api.cpp:13-35): stream-ordered scratch
    void* ws = nullptr;
    if (L.plan.workspaceBytes != 0) DGA_CUDA(cudaMallocAsync(&ws, L.plan.workspaceBytes, stream));
    rc = launch_dense(L, stream, ws);
    if (ws != nullptr) cudaFreeAsync(ws, stream);
    return rc;
}

// Single-CTA tile of a batched / grouped launch: widest N tile whose wave count over all SMs is minimal.
// cta = 2: pair tiles of 256 rows (grouped launches whose groups are aligned to 256 rows).
static int plan_slab_tiles(DgaTilingPlan& p, uint32_t slabs_of_tiles, uint32_t cta) {
    const uint32_t step = p.layoutTagB == 0 ? 64 : 32;
    const uint32_t units = platform().smCount / cta;   // CTAs or CTA pairs
    uint32_t best_n1 = 0;
    uint64_t best_cost = ~0ull;
    for (uint32_t n1 = 256; n1 >= 64; n1 -= step) {
        const uint64_t tiles = uint64_t(ceil_div(p.m, 128 * cta)) * ceil_div(p.n, n1) * slabs_of_tiles;
        const uint64_t waves = (tiles + units - 1) / units;
        const uint64_t cost = waves * (n1 + 24);   // + per-tile overhead in column units
        if (cost < best_cost) { best_cost = cost; best_n1 = n1; }
    }
    p.kernelSerial = DGA_KERNEL_COMMON;
    p.ctaGroup = uint8_t(cta);
    p.clusterShape = 0;
    p.n1 = uint16_t(best_n1);
    p.stages = 0;
    p.splitkFactor = 1;
    p.paddingTagA = p.paddingTagB = p.paddingTagC = 0;
    int rc = finalize_plan(p);
    if (rc != ACLNN_SUCCESS) return rc;
    const uint64_t tiles = uint64_t(ceil_div(p.m, 128 * cta)) * ceil_div(p.n, p.n1) * slabs_of_tiles;
    p.blockDim = uint32_t(std::min<uint64_t>(tiles, units)) * cta;
    return ACLNN_SUCCESS;
}
static int plan_single_cta(DgaTilingPlan& p, uint32_t slabs_of_tiles) { return plan_slab_tiles(p, slabs_of_tiles, 1); }

static int plan_batched(DgaTilingPlan& p, uint32_t batch, uint32_t m, uint32_t n, uint32_t k) {
    seed_plan(p, m, n, k, ACL_BF16, ACL_FLOAT, /*layoutTagB=*/0);
    p.strideA = k; p.strideB = n; p.strideC = n;
    return plan_single_cta(p, batch);
}

// The framework kernel's batch loop (generate_code.hpp:149-153) as ONE launch: z[b] = x[b] * y[b], bf16 -> fp32,
// row-major, 16-byte aligned pitches (k % 8 == 0, n % 8 == 0).
static aclnnStatus framework_gemm_batched(const void* x, const void* y, void* z, uint32_t batch, uint32_t m, uint32_t n,
                                          uint32_t k, cudaStream_t stream) {
    DenseLaunch L;
    int rc = plan_batched(L.plan, batch, m, n, k);
    if (rc != ACLNN_SUCCESS) return rc;
    L.mode = kModeBatched;
    L.batch = batch;
    L.a = x; L.b = y; L.d = z;
    L.lda = k; L.ldb = n; L.ldd = n;
    L.slab_a = uint64_t(m) * k; L.slab_b = uint64_t(k) * n; L.slab_d = uint64_t(m) * n;
    L.eff = {L.a, L.b, L.d, L.lda, L.ldb, L.ldd};
    return launch_common(L, stream, nullptr);
}

aclnnStatus dgaRunMmadRtc(const void* x, const void* y, void* z, uint32_t batch, uint32_t m, uint32_t n, uint32_t k,
                          aclrtStream stream) {
    if (x == nullptr || y == nullptr || z == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool aligned = k % 8 == 0 && n % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 &&
                         (reinterpret_cast<uintptr_t>(y) & 15) == 0 && (reinterpret_cast<uintptr_t>(z) & 15) == 0;
    if (batch > 1 && aligned) {
        aclnnStatus rc = framework_gemm_batched(x, y, z, batch, m, n, k, s);
        if (rc != ACLNN_SUCCESS) return rc;
        DGA_CUDA(cudaStreamSynchronize(s));  // the reference blocks here (gemm.hpp:110)
        return ACLNN_SUCCESS;
    }
    for (uint32_t b = 0; b < batch; ++b) {  // batch loop of the reference kernel: generate_code.hpp:149-153
        const char* xb = static_cast<const char*>(x) + uint64_t(b) * m * k * 2;
        const char* yb = static_cast<const char*>(y) + uint64_t(b) * k * n * 2;
        char* zb = static_cast<char*>(z) + uint64_t(b) * m * n * 4;
        aclnnStatus rc = framework_gemm(xb, yb, zb, m, n, k, ACL_BF16, s);
        if (rc != ACLNN_SUCCESS) return rc;
    }
    DGA_CUDA(cudaStreamSynchronize(s));  // the reference blocks here (gemm.hpp:110)
    return ACLNN_SUCCESS;
}

static uint32_t align16(uint32_t x) { return (x + 15u) & ~15u; }

aclnnStatus dgaGetBenchConfig(uint32_t m, uint32_t n, uint32_t k, int32_t* params) {
    if (params == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    for (int i = 0; i < 6; ++i) {
        if (params[i] <= 0) {
            set_error("bench knobs params[0..5] must be positive");
            return ACLNN_ERR_PARAM_INVALID;
        }
    }
    // Loop bounds of the reference's 16x16-block MMAD kernel for the caller's six knobs
    // (jit/get_best_config.hpp:93-153); order of the 22 outputs as in gemm_bench.hpp:66-69.
    const uint32_t m_sections = params[0], n_sections = params[1], m_sec = params[2], n_sec = params[3];
    const uint32_t k_iter = params[4], db = params[5];
    const uint32_t mb = align16(m) / 16, nb = align16(n) / 16, kb = align16(k) / 16;
    const uint32_t db_num = k_iter / db;
    uint32_t r_mb = mb % m_sec, r_nb = nb % n_sec;
    if (r_mb == 0) r_mb = m_sec;
    if (r_nb == 0) r_nb = n_sec;
    const uint32_t k_iters = (kb + k_iter - 1) / k_iter;
    const uint32_t k_tail = kb % k_iter;
    uint32_t r_db_num, r_kb;
    if (k_tail == 0) {
        r_db_num = db_num;
        r_kb = db;
    } else {
        r_db_num = (k_tail + db - 1) / db;
        r_kb = k_tail - (r_db_num - 1) * db;
    }
    const uint32_t m_iters = (mb + m_sec - 1) / m_sec, n_iters = (nb + n_sec - 1) / n_sec;
    const uint32_t m_parts = m_iters / m_sections, n_parts = n_iters / n_sections;
    const uint32_t out[22] = {m, n, k, 1u, k_iters, mb, nb, kb, m_parts * m_sec, n_parts * n_sec,
                              align16(m) - m, align16(n) - n, align16(k) - k, db_num, m_parts, n_parts,
                              m_iters - (m_sections - 1) * m_parts, n_iters - (n_sections - 1) * n_parts,
                              r_mb, r_nb, r_kb, r_db_num};
    for (int i = 0; i < 22; ++i) params[6 + i] = static_cast<int32_t>(out[i]);
    return ACLNN_SUCCESS;
}

aclnnStatus dgaRunMmadBench(const void* x, const void* y, void* z, uint32_t m, uint32_t n, uint32_t k, int32_t* params,
                            aclrtStream stream) {
    if (x == nullptr || y == nullptr || z == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    aclnnStatus rc = dgaGetBenchConfig(m, n, k, params);
    if (rc != ACLNN_SUCCESS) return rc;
    // The six knobs tile an Ascend cube core (16x16 blocks in L1/L0); they are advisory on B200, where the
    // tile comes from the selector / plan cache.
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    rc = framework_gemm(x, y, z, m, n, k, ACL_FLOAT16, s);
    if (rc != ACLNN_SUCCESS) return rc;
    DGA_CUDA(cudaStreamSynchronize(s));  // gemm_bench.hpp:112
    return ACLNN_SUCCESS;
}

// ---- N-sharding ------------------------------------------------------------------------------------------
aclnnStatus dgaNShardRange(uint32_t n, uint32_t worldSize, uint32_t rank, uint32_t* nBegin, uint32_t* nCount) {
    if (nBegin == nullptr || nCount == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    if (worldSize == 0 || rank >= worldSize) {
        set_error("rank out of range");
        return ACLNN_ERR_PARAM_INVALID;
    }
    // contiguous slices in units of 256 columns (one CTA-pair tile) so no rank gets a ragged interior tile
    const uint32_t units = ceil_div(n, 256);
    const uint32_t base = units / worldSize, extra = units % worldSize;
    const uint32_t u0 = rank * base + std::min(rank, extra);
    const uint32_t u1 = u0 + base + (rank < extra ? 1u : 0u);
    const uint32_t b = std::min(n, u0 * 256), e = std::min(n, u1 * 256);
    *nBegin = b;
    *nCount = e - b;
    return ACLNN_SUCCESS;
}

// ---- misc --------------------------------------------------------------------------------------------------
aclnnStatus dgaSetOption(const char* name, int64_t value) {
    if (name == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    if (std::strcmp(name, "splitk_prefetch_kb") == 0) g_opt_splitk_prefetch.store(value);
    else if (std::strcmp(name, "pdl") == 0) g_opt_pdl.store(value);
    else if (std::strcmp(name, "splitk_issue_batch") == 0) g_opt_splitk_batch.store(value);
    else if (std::strcmp(name, "raster_group") == 0) g_opt_raster.store(value);
    else if (std::strcmp(name, "raster_along_m") == 0) g_opt_raster_dir.store(value);
    else if (std::strcmp(name, "fp8_bn") == 0) g_opt_fp8_bn.store(value);
    else if (std::strcmp(name, "streamk_cooperative") == 0) g_opt_sk_coop.store(value < 0 ? 1 : value);
    else if (std::strcmp(name, "hint_a") == 0) g_opt_hint_a.store(value);
    else if (std::strcmp(name, "hint_b") == 0) g_opt_hint_b.store(value);
    else if (std::strcmp(name, "jit") == 0) g_opt_jit.store(value);
    else {
        set_error(std::string("unknown option ") + name);
        return ACLNN_ERR_PARAM_INVALID;
    }
    return ACLNN_SUCCESS;
}
// ---- run-time specialisation (role of the framework's JIT layer, see jit.cpp) ---------------------------
aclnnStatus dgaJitPrepare(uint32_t batch, uint32_t m, uint32_t n, uint32_t k, int32_t dtype, int32_t outDtype,
                          int32_t layoutTagB, char* kernelDir, uint64_t kernelDirCapacity) {
    DgaTilingPlan plan;
    int mode = kModePlain;
    aclnnStatus rc;
    if (batch > 1) {
        if (dtype != ACL_BF16 || outDtype != ACL_FLOAT || layoutTagB != 0) {
            set_error("batched launches are built for bf16 x bf16 -> fp32 over row-major B (run_mmad_rtc)");
            return ACLNN_ERR_PARAM_INVALID;
        }
        rc = plan_batched(plan, batch, m, n, k);
        mode = kModeBatched;
    } else {
        seed_plan(plan, m, n, k, dtype, outDtype, layoutTagB);
        rc = select_kernel_with_cache(plan);
        if (rc == ACLNN_SUCCESS && plan.kernelSerial == DGA_KERNEL_STREAMK) plan.kernelSerial = DGA_KERNEL_COMMON;
    }
    if (rc != ACLNN_SUCCESS) return rc;
    if (plan.kernelSerial == DGA_KERNEL_SPLITK_REDUCE && mode == kModePlain) mode = kModeSplitK;
    if (plan.kernelSerial != DGA_KERNEL_COMMON && plan.kernelSerial != DGA_KERNEL_PADDING_COMMON &&
        plan.kernelSerial != DGA_KERNEL_SMALL && plan.kernelSerial != DGA_KERNEL_SPLITK_REDUCE) {
        set_error("the plan for this shape uses a variant that is not specialised at run time (kernelSerial " +
                  std::to_string(int(plan.kernelSerial)) + ")");
        return ACLNN_ERR_PARAM_INVALID;
    }
    GemmParams gp{};
    common_params(plan, gp);
    if (mode == kModeBatched) {
        gp.tiles_per_batch = gp.dp_tiles;
        gp.dp_tiles *= batch;
    }
    if (mode == kModeSplitK) {   // as launch_common sets it up
        gp.split_k = plan.splitkFactor;
        gp.dp_tiles *= plan.splitkFactor;
        gp.epi_cols = 32;
    }
    const jit::Request req = make_jit_request(plan, mode, batch, gp);
    std::string dir, err;
    if (!jit::ensure_compiled(req, &dir, &err)) {
        set_error(err);
        return ACLNN_ERR_RUNTIME_ERROR;
    }
    if (kernelDir != nullptr && kernelDirCapacity != 0) {
        std::strncpy(kernelDir, dir.c_str(), kernelDirCapacity - 1);
        kernelDir[kernelDirCapacity - 1] = '\0';
    }
    return ACLNN_SUCCESS;
}
aclnnStatus dgaJitGetStats(uint64_t* compiled, uint64_t* diskHits, uint64_t* memoryHits) {
    const jit::Stats st = jit::stats();
    if (compiled != nullptr) *compiled = st.compiled;
    if (diskHits != nullptr) *diskHits = st.disk_hits;
    if (memoryHits != nullptr) *memoryHits = st.memory_hits;
    return ACLNN_SUCCESS;
}
aclnnStatus dgaJitDropMemoryCache(void) {
    jit::drop_memory_cache();
    return ACLNN_SUCCESS;
}
aclnnStatus dgaDebugReadTrace(uint64_t* out, uint32_t words) {
    if (out == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    unsigned long long* buf = trace_buffer();
    if (buf == nullptr) return ACLNN_ERR_RUNTIME_ERROR;
    DGA_CUDA(cudaMemcpy(out, buf, size_t(std::min(words, kTraceWords)) * 8, cudaMemcpyDeviceToHost));
    return ACLNN_SUCCESS;
}
const char* dgaGetLastError(void) { return last_error(); }
uint64_t dgaKernelLaunchCount(void) { return g_launches.load(); }
const char* dgaVersion(void) { return "dga-b200 0.1 (sm_100a)"; }

}  // extern "C"

// ======================================================== matmul without any device allocation ====
// dgaSplitkMatmulNoAlloc (include/dga_b200.h).  No counterpart in the reference; added
// for callers that cannot allocate (graph capture, serving loops).  Nothing above this line changes.
//
// The fp32 partial slabs of the split-K + reduce variant, which every other entry point keeps in the caller's
// workspace.  One array per GPU that loads this module, held until the process exits, shared by every call on that
// GPU: overlapping calls race.  In HOST code the name below is only a stand-in for the device array - handing it to a
// kernel or a tensor map would pass a host address - so its device address is looked up with cudaGetSymbolAddress,
// once per device with that device current (no_alloc_scratch).
static __device__ __align__(256) unsigned char g_no_alloc_scratch[DGA_NO_ALLOC_SCRATCH_BYTES];

namespace dga {

constexpr int kMaxDevices = 64;

static int no_alloc_scratch(void** out) {
    static std::mutex mu;
    static void* addr[kMaxDevices] = {};
    const int dev = current_device();
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("device ordinal " + std::to_string(dev) + " is beyond the table of the no-allocation entry point");
        return ACLNN_ERR_RUNTIME_ERROR;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (addr[dev] == nullptr) {
        void* p = nullptr;
        DGA_CUDA(cudaGetSymbolAddress(&p, g_no_alloc_scratch));
        if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 255) != 0) {
            set_error("static scratch array has no usable device address");
            return ACLNN_ERR_RUNTIME_ERROR;
        }
        addr[dev] = p;
    }
    *out = addr[dev];
    return ACLNN_SUCCESS;
}

// The split-K GEMM launch of launch_common_inst without its run-time-compilation branch (which loads modules).
template <bool kBMn>
static int launch_splitk_gemm_static(const DenseLaunch& L, const GemmParams& gp, cudaStream_t stream) {
    auto kernel = gemm_sm100_kernel<2, float, kBMn, false, kModeSplitK, false, false, false>;
    int rc = set_smem_attr(kernel, L.plan.smemBytes);
    if (rc != ACLNN_SUCCESS) return rc;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(L.plan.blockDim);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = L.plan.smemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    fill_launch_attrs(cfg, attr, 2);
    DGA_CUDA(cudaLaunchKernelEx(&cfg, kernel, L.tm_a, L.tm_b, L.tm_d, L.tm_a, L.tm_b, L.peers, gp));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return ACLNN_SUCCESS;
}

}  // namespace dga

extern "C" aclnnStatus dgaSplitkMatmulNoAlloc(const void* self, const void* mat2, void* out, uint32_t m, uint32_t n,
                                              uint32_t k, int32_t dtype, int32_t outDtype, int32_t layoutTagB,
                                              uint32_t splitkFactor, aclrtStream stream) {
    if (self == nullptr || mat2 == nullptr || out == nullptr) return ACLNN_ERR_PARAM_NULLPTR;
    if (m == 0 || n == 0 || k == 0 || (dtype != ACL_BF16 && dtype != ACL_FLOAT16) ||
        (outDtype != dtype && outDtype != ACL_FLOAT) || (layoutTagB != 0 && layoutTagB != 1)) {
        set_error("dgaSplitkMatmulNoAlloc: m, n, k > 0, 16-bit inputs, output of the input type or fp32, layoutTagB 0 or 1");
        return ACLNN_ERR_PARAM_INVALID;
    }
    const uint32_t kb = ceil_div(k, 64);
    if (kb < 2 || (splitkFactor != 0 && (splitkFactor < 2 || splitkFactor > 16 || splitkFactor > kb))) {
        set_error("dgaSplitkMatmulNoAlloc: splitkFactor is 0 (automatic) or 2..min(16, ceil(k / 64)), and k > 64");
        return ACLNN_ERR_PARAM_INVALID;
    }
    uint32_t split = splitkFactor;
    if (split == 0) {
        // as many K slices as fill the CTA pairs once; fewer when the slabs would not fit the static array
        const uint32_t tiles = ceil_div(m, 256) * ceil_div(n, 256), clusters = platform().smCount / 2;
        split = std::min(std::min(std::max(clusters / std::max(tiles, 1u), 2u), 16u), kb);
        while (split > 2 && splitk_ws_bytes(m, n, split) > DGA_NO_ALLOC_SCRATCH_BYTES) --split;
    }
    const uint64_t need = splitk_ws_bytes(m, n, split);
    if (need > DGA_NO_ALLOC_SCRATCH_BYTES) {
        set_error("dgaSplitkMatmulNoAlloc: " + std::to_string(m) + " x " + std::to_string(n) + " with " +
                  std::to_string(split) + " K slices needs " + std::to_string(need) +
                  " bytes of scratch, the static array holds " + std::to_string(DGA_NO_ALLOC_SCRATCH_BYTES) +
                  " (DGA_NO_ALLOC_SCRATCH_BYTES); use the two-phase entry with a caller workspace");
        return ACLNN_ERR_PARAM_INVALID;
    }
    DenseLaunch L;
    seed_plan(L.plan, m, n, k, dtype, outDtype, layoutTagB);
    L.plan.kernelSerial = DGA_KERNEL_SPLITK_REDUCE;
    L.plan.ctaGroup = 2;
    L.plan.n1 = 256;
    L.plan.splitkFactor = uint16_t(split);
    L.a = self; L.b = mat2; L.d = out;
    L.lda = k; L.ldb = layoutTagB == 1 ? k : n; L.ldd = n;
    int rc = apply_padding(L, /*allow_repack=*/false);   // finalize_plan: stages, grid, shared memory, workspaceBytes
    if (rc != ACLNN_SUCCESS) return rc;
    if (L.plan.kernelSerial != DGA_KERNEL_SPLITK_REDUCE || L.plan.workspaceBytes != need) {
        set_error("dgaSplitkMatmulNoAlloc: the plan came back as another variant");
        return ACLNN_ERR_INNER_TILING_ERROR;
    }
    void* scratch = nullptr;
    rc = no_alloc_scratch(&scratch);
    if (rc != ACLNN_SUCCESS) return rc;
    L.mode = kModeSplitK;
    L.split_ws = scratch;
    L.eff = {L.a, L.b, L.d, L.lda, L.ldb, L.ldd};
    rc = encode_common(L);   // host-only: three tensor maps, the output map over the static array
    if (rc != ACLNN_SUCCESS) return rc;
    GemmParams gp{};
    common_params(L.plan, gp);
    gp.split_k = L.plan.splitkFactor;
    gp.dp_tiles *= L.plan.splitkFactor;
    gp.epi_cols = 32;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    rc = layoutTagB == 1 ? launch_splitk_gemm_static<false>(L, gp, s) : launch_splitk_gemm_static<true>(L, gp, s);
    if (rc != ACLNN_SUCCESS) return rc;
    switch (outDtype) {
        case ACL_BF16: return launch_splitk_reduce<__nv_bfloat16>(L, s);
        case ACL_FLOAT16: return launch_splitk_reduce<__half>(L, s);
        default: return launch_splitk_reduce<float>(L, s);
    }
}
