"""ctypes binding of libdga_b200.so (the C ABI declared in include/dga_b200.h).

The library is the product; this module only loads it and declares signatures. There is no Python or
CPU fallback: if the library cannot be loaded, importing this module raises.
"""
import ctypes
import os

from . import build as _build

ACL_FLOAT, ACL_FLOAT16, ACL_INT32, ACL_BF16, ACL_FLOAT8_E4M3FN = 0, 1, 3, 27, 36
ACL_FORMAT_ND = 2
ACLNN_SUCCESS = 0

KERNEL_COMMON, KERNEL_SMALL, KERNEL_PADDING_COMMON, KERNEL_SPLITK, KERNEL_STREAMK = 0, 1, 2, 3, 4
KERNEL_SPLITK_REDUCE, KERNEL_CLUSTER_SPLITK = 5, 6
NO_ALLOC_SCRATCH_BYTES = 32 << 20   # DGA_NO_ALLOC_SCRATCH_BYTES (include/dga_b200.h)
KERNEL_NAMES = {0: "Common", 1: "Small", 2: "PaddingCommon", 3: "SplitK", 4: "StreamK", 5: "SplitKReduce", 6: "ClusterSplitK"}


class AclTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("dtype", ctypes.c_int32), ("ndim", ctypes.c_int32),
                ("shape", ctypes.c_int64 * 8), ("stride", ctypes.c_int64 * 8)]


class TilingPlan(ctypes.Structure):
    """DgaTilingPlan (role of TilingParams, op_host/op_tiling/tiling_params.h:19-66)."""
    _fields_ = [("strideA", ctypes.c_uint64), ("strideB", ctypes.c_uint64), ("strideC", ctypes.c_uint64),
                ("m", ctypes.c_uint32), ("n", ctypes.c_uint32), ("k", ctypes.c_uint32),
                ("m1", ctypes.c_uint16), ("n1", ctypes.c_uint16), ("k1", ctypes.c_uint16),
                ("swizzleOffset", ctypes.c_uint8), ("swizzleDirection", ctypes.c_uint8),
                ("splitkFactor", ctypes.c_uint16),
                ("layoutTagA", ctypes.c_uint8), ("layoutTagB", ctypes.c_uint8), ("layoutTagC", ctypes.c_uint8),
                ("paddingTagA", ctypes.c_uint8), ("paddingTagB", ctypes.c_uint8), ("paddingTagC", ctypes.c_uint8),
                ("kernelSerial", ctypes.c_uint8), ("ctaGroup", ctypes.c_uint8), ("stages", ctypes.c_uint8),
                ("dtype", ctypes.c_uint8), ("outDtype", ctypes.c_uint8), ("clusterShape", ctypes.c_uint8),
                ("blockDim", ctypes.c_uint32), ("smemBytes", ctypes.c_uint32),
                ("workspaceBytes", ctypes.c_uint64), ("estTimeUs", ctypes.c_double)]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


class PlatformInfo(ctypes.Structure):
    _fields_ = [("smCount", ctypes.c_uint32), ("smemBytes", ctypes.c_uint32), ("tmemCols", ctypes.c_uint32),
                ("l2Bytes", ctypes.c_uint64), ("hbmGBs", ctypes.c_double), ("denseTflops16", ctypes.c_double),
                ("denseTflops8", ctypes.c_double)]


_P = ctypes.POINTER
_vp, _u32, _u64, _i32 = ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_int32

# name -> (restype, argtypes); every symbol include/dga_b200.h declares
SIGNATURES = {
    "aclCreateTensor": (_P(AclTensor), [_P(ctypes.c_int64), _u64, _i32, _P(ctypes.c_int64), ctypes.c_int64, _i32,
                                        _P(ctypes.c_int64), _u64, _vp]),
    "aclDestroyTensor": (_i32, [_P(AclTensor)]),
    "aclnnCatlassDynamicMatmulGetWorkspaceSize": (_i32, [_P(AclTensor), _P(AclTensor), _P(AclTensor), _P(_u64),
                                                         _P(_vp)]),
    "aclnnCatlassDynamicMatmul": (_i32, [_vp, _u64, _vp, _vp]),
    "dgaSetExecutorRepeatable": (_i32, [_vp]),
    "dgaDestroyExecutor": (_i32, [_vp]),
    "dgaFp8BlockScaledMatmulGetWorkspaceSize": (_i32, [_P(AclTensor)] * 5 + [_P(_u64), _P(_vp)]),
    "dgaFp8BlockScaledMatmul": (_i32, [_vp, _u64, _vp, _vp]),
    "dgaFp8QueryPath": (_i32, [_vp, _vp, _vp, _P(_i32)]),
    "dgaSelectKernel": (_i32, [_u32, _u32, _u32, _i32, _i32, _i32, _P(TilingPlan)]),
    "dgaSelectKernelWithCache": (_i32, [_u32, _u32, _u32, _i32, _i32, _i32, _P(TilingPlan)]),
    "dgaTilingCacheSet": (_i32, [_P(TilingPlan)]),
    "dgaTilingCacheGet": (_i32, [_u32, _u32, _u32, _i32, _i32, _i32, _P(TilingPlan)]),
    "dgaTilingCacheClear": (_i32, []),
    "dgaTilingCacheSize": (_u64, []),
    "dgaTilingCacheLoadCsv": (_i32, [ctypes.c_char_p]),
    "dgaTilingCacheSaveCsv": (_i32, [ctypes.c_char_p]),
    "dgaGetPlatformInfo": (_i32, [_P(PlatformInfo)]),
    "dgaSetPlatformInfo": (_i32, [_P(PlatformInfo)]),
    "dgaMatmulWithPlan": (_i32, [_P(AclTensor), _P(AclTensor), _P(AclTensor), _P(TilingPlan), _vp, _u64, _vp]),
    "dgaSetExecutorGroupAlignment": (_i32, [_vp, _u32]),
    "dgaDecodePairMatmul": (_i32, [_P(AclTensor)] * 5 + [_vp]),
    "dgaSplitkMatmulNoAlloc": (_i32, [_vp, _vp, _vp, _u32, _u32, _u32, _i32, _i32, _i32, _u32, _vp]),
    "dgaRunMmadCustom": (_i32, [_vp, _vp, _vp, _vp]),
    "dgaRunMmadRtc": (_i32, [_vp, _vp, _vp, _u32, _u32, _u32, _u32, _vp]),
    "dgaRunMmadBench": (_i32, [_vp, _vp, _vp, _u32, _u32, _u32, _P(_i32), _vp]),
    "dgaGetBenchConfig": (_i32, [_u32, _u32, _u32, _P(_i32)]),
    "dgaNShardRange": (_i32, [_u32, _u32, _u32, _P(_u32), _P(_u32)]),
    "dgaNShardMatmulGetWorkspaceSize": (_i32, [_P(AclTensor), _P(AclTensor), _P(_vp), _u32, _u32, ctypes.c_int64,
                                            ctypes.c_int64, _P(_u64), _P(_vp)]),
    "dgaNShardMatmul": (_i32, [_vp, _u64, _vp, _vp]),
    "dgaMGroupedMatmulGetWorkspaceSize": (_i32, [_P(AclTensor)] * 4 + [_P(_u64), _P(_vp)]),
    "dgaMGroupedMaskedMatmulGetWorkspaceSize": (_i32, [_P(AclTensor)] * 4 + [_P(_u64), _P(_vp)]),
    "dgaMGroupedMatmul": (_i32, [_vp, _u64, _vp, _vp]),
    "dgaSetOption": (_i32, [ctypes.c_char_p, ctypes.c_int64]),
    "dgaDebugReadTrace": (_i32, [_P(_u64), _u32]),
    "dgaJitPrepare": (_i32, [_u32, _u32, _u32, _u32, _i32, _i32, _i32, ctypes.c_char_p, _u64]),
    "dgaJitGetStats": (_i32, [_P(_u64), _P(_u64), _P(_u64)]),
    "dgaJitDropMemoryCache": (_i32, []),
    "dgaGetLastError": (ctypes.c_char_p, []),
    "dgaKernelLaunchCount": (_u64, []),
    "dgaVersion": (ctypes.c_char_p, []),
}


def _load():
    # Rebuild when sources changed (a no-op hash check otherwise). No silent fallback: either the real
    # library loads or the import fails.
    path = _build.build()
    # developer aid (tools/fp8_isolate.sh): a side build of the same sources with other -D flags, by NAME inside this
    # package's lib directory (never an arbitrary path); never a fallback - a variant that does not load is an error
    variant = os.environ.get("DGA_B200_LIB_VARIANT", "")
    if variant:
        if not variant.isalnum():
            raise ImportError("DGA_B200_LIB_VARIANT must be alphanumeric, got %r" % variant)
        path = os.path.join(os.path.dirname(path), "libdga_b200_%s.so" % variant)
    lib = ctypes.CDLL(path)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = restype
        fn.argtypes = argtypes
    return lib


lib = _load()


class DgaError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        msg = lib.dgaGetLastError()
        super().__init__("%s failed with status %d: %s" % (where, status, msg.decode() if msg else ""))


def check(status, where):
    if status != ACLNN_SUCCESS:
        raise DgaError(status, where)
