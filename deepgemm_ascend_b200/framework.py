"""deep_gemm_ascend framework entry points over CUDA tensors.

Same three names and argument meaning as the reference's pybind module deep_gemm_cpp
(deep_gemm_ascend/framework/csrc/python_api.cpp:13-35, re-exported by
deep_gemm_ascend/framework/deep_gemm_ascend/__init__.py:1-5):
  run_mmad_custom(x, y, z)          stub launch (the reference kernel returns at once, impls/mmad.cpp:79)
  run_mmad_rtc(x, y, z)             z[b,m,n] fp32 = x[b,m,k] bf16 @ y[b,k,n] bf16, blocks until done
  run_mmad_bench(x, y, z, params)   z[m,n] fp32 = x[m,k] fp16 @ y[k,n] fp16; params int32[28], [6:28] written back
Like the reference (handle.hpp:36-38) the tensors' storage is taken as contiguous row-major.
"""
import ctypes

import torch

from ._lib import lib, check
from .ops import _require_cuda, _stream_ptr


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _need(t, dtype, dim, name):
    if t.dtype != dtype or t.dim() != dim or not t.is_contiguous():
        raise TypeError("%s must be a contiguous %d-D %s tensor" % (name, dim, dtype))


def run_mmad_custom(x, y, z):
    _require_cuda(x, y, z)
    check(lib.dgaRunMmadCustom(_ptr(x), _ptr(y), _ptr(z), _stream_ptr()), "run_mmad_custom")


def run_mmad_rtc(x, y, z):
    _require_cuda(x, y, z)
    _need(x, torch.bfloat16, 3, "x")
    _need(y, torch.bfloat16, 3, "y")
    _need(z, torch.float32, 3, "z")
    b, m, k = x.shape
    if y.shape[0] != b or y.shape[1] != k or tuple(z.shape) != (b, m, y.shape[2]):
        raise ValueError("shape mismatch: x[b,m,k] y[b,k,n] z[b,m,n]")
    check(lib.dgaRunMmadRtc(_ptr(x), _ptr(y), _ptr(z), b, m, y.shape[2], k, _stream_ptr()), "run_mmad_rtc")


def run_mmad_bench(x, y, z, params):
    _require_cuda(x, y, z)
    _need(x, torch.float16, 2, "x")
    _need(y, torch.float16, 2, "y")
    _need(z, torch.float32, 2, "z")
    if params.dtype != torch.int32 or params.numel() < 28:
        raise TypeError("params must be an int32 tensor with 28 entries")
    m, k = x.shape
    if y.shape[0] != k or tuple(z.shape) != (m, y.shape[1]):
        raise ValueError("shape mismatch: x[m,k] y[k,n] z[m,n]")
    # the reference reads the six knobs with .item() (a device sync, gemm_bench.hpp:52-57) and writes the 22
    # derived ints back into params[6:28] (gemm_bench.hpp:79-81)
    host = params.detach().to("cpu", copy=True).contiguous()
    arr = (ctypes.c_int32 * 28)(*host[:28].tolist())
    check(lib.dgaRunMmadBench(_ptr(x), _ptr(y), _ptr(z), m, y.shape[1], k, arr, _stream_ptr()), "run_mmad_bench")
    params[6:28].copy_(torch.tensor(list(arr)[6:28], dtype=torch.int32))


def get_bench_config(m, n, k, knobs):
    """Host-only: the 28-int parameter vector for six knobs (works without a GPU)."""
    arr = (ctypes.c_int32 * 28)(*(list(knobs) + [0] * 22))
    check(lib.dgaGetBenchConfig(m, n, k, arr), "dgaGetBenchConfig")
    return list(arr)
