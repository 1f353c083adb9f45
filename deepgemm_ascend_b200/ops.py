"""Torch-facing shim over the C ABI: torch supplies device memory and streams, nothing else.

Mirrors the two call surfaces of the reference:
  * the aclnn operator CatlassDynamicMatmul(self, mat2) -> out
    (aclnn_catlass_dynamic_matmul/op_host/catlass_dynamic_matmul.cpp:50-78), two-phase underneath;
  * the host tiling entry (SelectKernel, op_host/op_tiling/select_kernel.cpp:333-378).
"""
import ctypes

import torch

from . import _lib
from ._lib import lib, check

_DTYPE_TO_ACL = {torch.float32: _lib.ACL_FLOAT, torch.float16: _lib.ACL_FLOAT16, torch.bfloat16: _lib.ACL_BF16,
                 torch.int32: _lib.ACL_INT32}
if hasattr(torch, "float8_e4m3fn"):
    _DTYPE_TO_ACL[torch.float8_e4m3fn] = _lib.ACL_FLOAT8_E4M3FN


def acl_dtype(dtype):
    try:
        return _DTYPE_TO_ACL[dtype]
    except KeyError:
        raise TypeError("unsupported dtype %s" % dtype)


def acl_tensor(t):
    """Describe a torch tensor (any strides; nothing is copied, like the reference's IgnoreContiguous inputs)."""
    d = _lib.AclTensor()
    d.data = t.data_ptr()
    d.dtype = acl_dtype(t.dtype)
    d.ndim = t.dim()
    for i in range(t.dim()):
        d.shape[i] = t.shape[i]
        d.stride[i] = t.stride(i)
    return d


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return ctypes.c_void_p(s.cuda_stream)


def _scratch(nbytes, device, stream):
    """Workspace for one asynchronous launch on `stream`. torch's caching allocator hands a freed block to the next
    allocation on the stream it was allocated on (the current one); when the launch runs on another stream the
    block must stay reserved until that stream has passed this point, or a later allocation could overwrite
    repack buffers, split-K slabs, stream-K partials and flags, or packed fp8 scales that the kernels still use."""
    ws = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=device)
    if stream is not None and stream != torch.cuda.current_stream(device):
        ws.record_stream(stream)
    return ws


def _require_cuda(*tensors):
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError("dga_b200 has no CPU path: tensors must live on a CUDA device")


def get_workspace_size(self, mat2, out):
    """Phase 1: returns (workspace_bytes, executor handle). The handle must go to `execute` or `destroy_executor`."""
    ws = ctypes.c_uint64(0)
    ex = ctypes.c_void_p(None)
    a, b, c = acl_tensor(self), acl_tensor(mat2), acl_tensor(out)
    check(lib.aclnnCatlassDynamicMatmulGetWorkspaceSize(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c),
                                                        ctypes.byref(ws), ctypes.byref(ex)),
          "aclnnCatlassDynamicMatmulGetWorkspaceSize")
    return ws.value, ex


def execute(workspace, workspace_size, executor, stream=None):
    """Phase 2: asynchronous launch on `stream` (default: torch's current stream)."""
    ptr = ctypes.c_void_p(workspace.data_ptr()) if workspace is not None else ctypes.c_void_p(None)
    check(lib.aclnnCatlassDynamicMatmul(ptr, workspace_size, executor, _stream_ptr(stream)),
          "aclnnCatlassDynamicMatmul")


def destroy_executor(executor):
    lib.dgaDestroyExecutor(executor)


def catlass_dynamic_matmul(self, mat2, out=None, stream=None):
    """out[M,N] = self[M,K] @ mat2[K,N]; fp16/bf16 in, same dtype out, fp32 accumulate.

    mat2 may be the transposed view of a weight stored [N,K] (the reference's layout) or a row-major [K,N].
    """
    _require_cuda(self, mat2)
    if self.dim() != 2 or mat2.dim() != 2:
        # same rule as InferShape (catlass_dynamic_matmul.cpp:22-26); phase 1 would report it too
        raise _lib.DgaError(161002, "catlass_dynamic_matmul (inputs must be 2-D)")
    if out is None:
        out = torch.empty((self.shape[0], mat2.shape[1]), dtype=self.dtype, device=self.device)
    ws_bytes, ex = get_workspace_size(self, mat2, out)
    ws = _scratch(ws_bytes, self.device, stream) if ws_bytes else None
    execute(ws, ws_bytes, ex, stream)
    return out


class Executor:
    """A repeatable plan: phase 1 once, phase 2 many times (CUDA-graph friendly; role of a cached aclOpExecutor)."""

    def __init__(self, self_t, mat2, out):
        _require_cuda(self_t, mat2, out)
        self._keep = (self_t, mat2, out)
        self.workspace_bytes, self._ex = get_workspace_size(self_t, mat2, out)
        check(lib.dgaSetExecutorRepeatable(self._ex), "dgaSetExecutorRepeatable")
        self._ws = torch.empty(self.workspace_bytes, dtype=torch.uint8, device=out.device) \
            if self.workspace_bytes else None

    def __call__(self, stream=None):
        execute(self._ws, self.workspace_bytes, self._ex, stream)
        return self._keep[2]

    def close(self):
        if self._ex is not None:
            destroy_executor(self._ex)
            self._ex = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def select_kernel(m, n, k, dtype=torch.bfloat16, out_dtype=None, layout_b=1, use_cache=False):
    """Host tiling decision for (m, n, k) as a dict (pure host code, works without a GPU)."""
    plan = _lib.TilingPlan()
    fn = lib.dgaSelectKernelWithCache if use_cache else lib.dgaSelectKernel
    out_dtype = dtype if out_dtype is None else out_dtype
    check(fn(m, n, k, acl_dtype(dtype), acl_dtype(out_dtype), layout_b, ctypes.byref(plan)), "dgaSelectKernel")
    return plan.as_dict()


def matmul_with_plan(self, mat2, out, stream=None, **fields):
    """Launch with hand-set plan fields (kernelSerial, ctaGroup, n1, stages, splitkFactor, swizzle*): tuning sweeps."""
    _require_cuda(self, mat2, out)
    plan = _lib.TilingPlan()
    for key, val in fields.items():
        setattr(plan, key, val)
    a, b, c = acl_tensor(self), acl_tensor(mat2), acl_tensor(out)
    global _PLAN_WS
    if _PLAN_WS is None or _PLAN_WS.device != out.device:
        _PLAN_WS = torch.empty(256 << 20, dtype=torch.uint8, device=out.device)   # covers every stream-K / split-K plan
    check(lib.dgaMatmulWithPlan(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c), ctypes.byref(plan),
                                ctypes.c_void_p(_PLAN_WS.data_ptr()), _PLAN_WS.numel(), _stream_ptr(stream)),
          "dgaMatmulWithPlan")
    return out


_PLAN_WS = None


class DecodePair:
    """Two decode matmuls over the same activations in one launch (dgaDecodePairMatmul): out_first = self @ mat2_first,
    out_second = self @ mat2_second, m <= 128, both weights stored [N, K] and passed as `w.t()`. For gate / up or
    q / kv projections kept as separate weights. The tensor descriptions are built once; calling the object enqueues
    the launch (CUDA-graph capturable, no workspace)."""

    def __init__(self, self_t, mat2_first, out_first, mat2_second, out_second):
        _require_cuda(self_t, mat2_first, out_first, mat2_second, out_second)
        self._keep = (self_t, mat2_first, out_first, mat2_second, out_second)
        self._desc = [acl_tensor(t) for t in self._keep]

    def __call__(self, stream=None):
        d = self._desc
        check(lib.dgaDecodePairMatmul(ctypes.byref(d[0]), ctypes.byref(d[1]), ctypes.byref(d[2]), ctypes.byref(d[3]),
                                      ctypes.byref(d[4]), _stream_ptr(stream)), "dgaDecodePairMatmul")
        return self._keep[2], self._keep[4]


def decode_pair_matmul(self, mat2_first, mat2_second, out_first=None, out_second=None, stream=None):
    """One-shot form of DecodePair; allocates the outputs when they are not given."""
    if out_first is None:
        out_first = torch.empty((self.shape[0], mat2_first.shape[1]), dtype=self.dtype, device=self.device)
    if out_second is None:
        out_second = torch.empty((self.shape[0], mat2_second.shape[1]), dtype=self.dtype, device=self.device)
    return DecodePair(self, mat2_first, out_first, mat2_second, out_second)(stream)


def splitk_matmul_no_alloc(self, mat2, out, splitk_factor=0, stream=None):
    """out = self @ mat2 through dgaSplitkMatmulNoAlloc: the split-K + reduce variant with its fp32 slabs in the
    library's static device array (at most _lib.NO_ALLOC_SCRATCH_BYTES), no device allocation inside the call.
    Contiguous operands; mat2 is a [N, K] tensor viewed as [K, N] (`w.t()`) or a contiguous [K, N] tensor. No
    counterpart in the reference (see include/dga_b200.h). Calls that overlap on the GPU race on the array."""
    _require_cuda(self, mat2, out)
    m, k = self.shape
    n = out.shape[1]
    if mat2.shape != (k, n) or out.shape[0] != m or not (self.is_contiguous() and out.is_contiguous()):
        raise ValueError("splitk_matmul_no_alloc: contiguous self [m,k], out [m,n]; mat2 [k,n]")
    if mat2.is_contiguous():
        layout_b = 0
    elif mat2.t().is_contiguous():
        layout_b = 1
    else:
        raise ValueError("splitk_matmul_no_alloc: mat2 is a contiguous [k,n] tensor or the .t() of a contiguous [n,k] one")
    check(lib.dgaSplitkMatmulNoAlloc(ctypes.c_void_p(self.data_ptr()), ctypes.c_void_p(mat2.data_ptr()),
                                     ctypes.c_void_p(out.data_ptr()), m, n, k, acl_dtype(self.dtype),
                                     acl_dtype(out.dtype), layout_b, int(splitk_factor), _stream_ptr(stream)),
          "dgaSplitkMatmulNoAlloc")
    return out


def set_option(name, value):
    """Process-wide tuning knob (see dgaSetOption in include/dga_b200.h); value < 0 restores the default."""
    check(lib.dgaSetOption(name.encode(), int(value)), "dgaSetOption")


def jit_prepare(m, n, k, dtype=_lib.ACL_BF16, out_dtype=None, layout_b=1, batch=1):
    """Compile (or find in $DGA_ROOT_DIR) the kernel specialised for this shape ahead of its first use; needs nvcc
    but no GPU. Returns the kernel directory (role of Compiler::build, framework/csrc/jit/compiler.hpp:26-52)."""
    import ctypes
    buf = ctypes.create_string_buffer(1024)
    out_dtype = dtype if out_dtype is None else out_dtype
    check(lib.dgaJitPrepare(batch, m, n, k, dtype, out_dtype, layout_b, buf, len(buf)), "dgaJitPrepare")
    return buf.value.decode()


def jit_stats():
    """{'compiled': nvcc runs, 'disk_hits': kernel directories reused, 'memory_hits': in-process map hits}."""
    import ctypes
    vals = [ctypes.c_uint64() for _ in range(3)]
    check(lib.dgaJitGetStats(*[ctypes.byref(v) for v in vals]), "dgaJitGetStats")
    return dict(zip(("compiled", "disk_hits", "memory_hits"), (int(v.value) for v in vals)))


def jit_drop_memory_cache():
    check(lib.dgaJitDropMemoryCache(), "dgaJitDropMemoryCache")


def kernel_launch_count():
    return int(lib.dgaKernelLaunchCount())


def nshard_range(n, world_size, rank):
    b, c = ctypes.c_uint32(0), ctypes.c_uint32(0)
    check(lib.dgaNShardRange(n, world_size, rank, ctypes.byref(b), ctypes.byref(c)), "dgaNShardRange")
    return b.value, c.value


# ---- M-grouped contiguous matmul (MoE; DeepGEMM m_grouped_gemm_nt_contiguous semantics) -------------------------
def _set_group_alignment(ex, m_alignment):
    if m_alignment != 128:
        rc = lib.dgaSetExecutorGroupAlignment(ex, int(m_alignment))
        if rc != 0:
            destroy_executor(ex)
            check(rc, "dgaSetExecutorGroupAlignment")


def m_grouped_matmul(self, weight, group_index, out=None, stream=None, m_alignment=128):
    """out[i,:] = self[i,:] @ weight[group_index[i]].T; self [M_total,K], weight [G,N,K], group_index int32 [M_total]
    (constant over every 128-row block, -1 = padding block left unwritten), M_total % 128 == 0.
    m_alignment=256: the caller's promise that every group starts on a multiple of 256 rows; the launch then uses
    256-row CTA-pair tiles (dgaSetExecutorGroupAlignment)."""
    _require_cuda(self, weight, group_index)
    if out is None:
        out = torch.empty((self.shape[0], weight.shape[1]), dtype=self.dtype, device=self.device)
    ws = ctypes.c_uint64(0)
    ex = ctypes.c_void_p(None)
    a, w, g, c = acl_tensor(self), acl_tensor(weight), acl_tensor(group_index), acl_tensor(out)
    check(lib.dgaMGroupedMatmulGetWorkspaceSize(ctypes.byref(a), ctypes.byref(w), ctypes.byref(g), ctypes.byref(c),
                                                ctypes.byref(ws), ctypes.byref(ex)),
          "dgaMGroupedMatmulGetWorkspaceSize")
    _set_group_alignment(ex, m_alignment)
    check(lib.dgaMGroupedMatmul(ctypes.c_void_p(None), 0, ex, _stream_ptr(stream)), "dgaMGroupedMatmul")
    return out


def m_grouped_matmul_masked(self, weight, masked_m, out=None, stream=None, m_alignment=128):
    """out[g,i,:] = self[g,i,:] @ weight[g].T for i < masked_m[g]; self [G,Mmax,K], weight [G,N,K], masked_m int32 [G]
    on the device (no host sync). Whole 128-row blocks past masked_m[g] are skipped and left unwritten.
    m_alignment=256 (Mmax % 256 == 0): 256-row CTA-pair tiles, worth it when most groups hold more than 128 rows."""
    _require_cuda(self, weight, masked_m)
    if out is None:
        out = torch.empty((self.shape[0], self.shape[1], weight.shape[1]), dtype=self.dtype, device=self.device)
    ws = ctypes.c_uint64(0)
    ex = ctypes.c_void_p(None)
    a, w, g, c = acl_tensor(self), acl_tensor(weight), acl_tensor(masked_m), acl_tensor(out)
    check(lib.dgaMGroupedMaskedMatmulGetWorkspaceSize(ctypes.byref(a), ctypes.byref(w), ctypes.byref(g),
                                                      ctypes.byref(c), ctypes.byref(ws), ctypes.byref(ex)),
          "dgaMGroupedMaskedMatmulGetWorkspaceSize")
    _set_group_alignment(ex, m_alignment)
    check(lib.dgaMGroupedMatmul(ctypes.c_void_p(None), 0, ex, _stream_ptr(stream)), "dgaMGroupedMatmul")
    return out


class MGroupedExecutor:
    """Repeatable grouped-matmul plan: phase 1 once, phase 2 per call."""

    def __init__(self, self_t, weight, group_index, out, m_alignment=128):
        _require_cuda(self_t, weight, group_index, out)
        self._keep = (self_t, weight, group_index, out)
        ws = ctypes.c_uint64(0)
        ex = ctypes.c_void_p(None)
        a, w, g, c = acl_tensor(self_t), acl_tensor(weight), acl_tensor(group_index), acl_tensor(out)
        check(lib.dgaMGroupedMatmulGetWorkspaceSize(ctypes.byref(a), ctypes.byref(w), ctypes.byref(g), ctypes.byref(c),
                                                    ctypes.byref(ws), ctypes.byref(ex)),
              "dgaMGroupedMatmulGetWorkspaceSize")
        _set_group_alignment(ex, m_alignment)
        check(lib.dgaSetExecutorRepeatable(ex), "dgaSetExecutorRepeatable")
        self._ex = ex

    def __call__(self, stream=None):
        check(lib.dgaMGroupedMatmul(ctypes.c_void_p(None), 0, self._ex, _stream_ptr(stream)), "dgaMGroupedMatmul")
        return self._keep[3]

    def close(self):
        if self._ex is not None:
            destroy_executor(self._ex)
            self._ex = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- fp8 e4m3 block-scaled (DeepGEMM semantics; not in the reference) -------------------------------------------
def fp8_get_workspace_size(a, a_scale, mat2, w_scale, out):
    ws = ctypes.c_uint64(0)
    ex = ctypes.c_void_p(None)
    d = [acl_tensor(t) for t in (a, a_scale, mat2, w_scale, out)]
    check(lib.dgaFp8BlockScaledMatmulGetWorkspaceSize(*[ctypes.byref(x) for x in d], ctypes.byref(ws), ctypes.byref(ex)),
          "dgaFp8BlockScaledMatmulGetWorkspaceSize")
    return ws.value, ex


def fp8_block_scaled_matmul(a, a_scale, weight, w_scale, out=None, stream=None):
    """out[M,N] (bf16) = (a * a_scale) @ (weight * w_scale)^T.

    a [M,K] float8_e4m3fn with a_scale [M, ceil(K/128)] fp32 (1x128 strips); weight [N,K] float8_e4m3fn with
    w_scale [ceil(N/128), ceil(K/128)] fp32 (128x128 blocks)."""
    _require_cuda(a, a_scale, weight, w_scale)
    if out is None:
        out = torch.empty((a.shape[0], weight.shape[0]), dtype=torch.bfloat16, device=a.device)
    ws_bytes, ex = fp8_get_workspace_size(a, a_scale, weight.t(), w_scale, out)
    ws = _scratch(ws_bytes, a.device, stream)
    check(lib.dgaFp8BlockScaledMatmul(ctypes.c_void_p(ws.data_ptr()), ws_bytes, ex, _stream_ptr(stream)),
          "dgaFp8BlockScaledMatmul")
    return out


class Fp8Executor:
    """Repeatable fp8 plan: phase 1 once, phase 2 per call."""

    def __init__(self, a, a_scale, weight, w_scale, out):
        _require_cuda(a, a_scale, weight, w_scale, out)
        self._keep = (a, a_scale, weight, w_scale, out)
        self.workspace_bytes, self._ex = fp8_get_workspace_size(a, a_scale, weight.t(), w_scale, out)
        check(lib.dgaSetExecutorRepeatable(self._ex), "dgaSetExecutorRepeatable")
        self._ws = torch.empty(max(self.workspace_bytes, 1), dtype=torch.uint8, device=out.device)

    def __call__(self, stream=None):
        check(lib.dgaFp8BlockScaledMatmul(ctypes.c_void_p(self._ws.data_ptr()), self.workspace_bytes, self._ex,
                                          _stream_ptr(stream)), "dgaFp8BlockScaledMatmul")
        return self._keep[4]

    def path(self, stream=None):
        """Which kernel produced the last result: 'block_scaled' (power-of-two scales, applied by the tensor core)
        or 'promotion' (arbitrary fp32 scales). Synchronises the stream."""
        which = ctypes.c_int32(-1)
        check(lib.dgaFp8QueryPath(ctypes.c_void_p(self._ws.data_ptr()), self._ex, _stream_ptr(stream),
                                  ctypes.byref(which)), "dgaFp8QueryPath")
        return "promotion" if which.value else "block_scaled"

    def close(self):
        if self._ex is not None:
            destroy_executor(self._ex)
            self._ex = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
