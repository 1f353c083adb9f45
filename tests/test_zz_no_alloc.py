"""dgaSplitkMatmulNoAlloc (include/dga_b200.h): the split-K + reduce matmul with its fp32 slabs in the library's
static device array instead of a caller workspace. No counterpart in the reference. Collected last on purpose (the
file name sorts after every other test file).

Checked through the C ABI (ctypes on libdga_b200.so) against the existing entry points on two calls in a row:
  * the two-phase operator call (aclnnCatlassDynamicMatmul) on the same inputs, at the tolerance the parity tests use
    for 16-bit outputs (max relative difference <= 1e-2, positive inputs) resp. 1e-5 for fp32 outputs;
  * dgaMatmulWithPlan with the same split-K plan and a caller workspace: the same two kernels, so bit-identical;
  * the fp64 oracle.
The second call has another shape, dtype and weight layout than the first and runs over the first call's leftovers
in the shared array; then the first problem again."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import matmul_oracle as orc

pytestmark = pytest.mark.gpu

DEV = "cuda:0"

# (m, n, k, dtype, out dtype, layout_b, K slices): 16 MiB, 7 MiB and 14 MiB of the 32 MiB array
CALLS = [(512, 2048, 4096, torch.bfloat16, torch.bfloat16, 1, 4),
         (300, 1000, 7168, torch.float16, torch.float32, 0, 6),
         (512, 2048, 4096, torch.bfloat16, torch.bfloat16, 1, 4),
         (129, 4104, 2048, torch.bfloat16, torch.bfloat16, 1, 0)]   # automatic slice count, ragged tiles


def _operands(m, n, k, dtype, layout_b, seed):
    rng = np.random.RandomState(seed)
    a = rng.uniform(1, 10, [m, k]).astype(np.float32)
    b = rng.uniform(1, 10, [k, n]).astype(np.float32)
    if dtype == torch.float16:
        a, b = a.astype(np.float16).astype(np.float32), b.astype(np.float16).astype(np.float32)
    else:
        a, b = orc.round_to_bf16(a), orc.round_to_bf16(b)
    ta = torch.from_numpy(a).to(DEV).to(dtype)
    if layout_b == 1:
        mat2 = torch.from_numpy(np.ascontiguousarray(b.T)).to(DEV).to(dtype).t()   # weight stored [N, K]
    else:
        mat2 = torch.from_numpy(b).to(DEV).to(dtype)
    return a, b, ta, mat2


def _no_alloc_call(dga, ta, mat2, out, layout_b, split):
    """Straight through the C ABI: plain pointers and sizes."""
    lib = dga._lib.lib
    m, k = ta.shape
    n = out.shape[1]
    rc = lib.dgaSplitkMatmulNoAlloc(ctypes.c_void_p(ta.data_ptr()), ctypes.c_void_p(mat2.data_ptr()),
                                    ctypes.c_void_p(out.data_ptr()), m, n, k, dga.ops.acl_dtype(ta.dtype),
                                    dga.ops.acl_dtype(out.dtype), layout_b, split,
                                    ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.dgaGetLastError().decode()


def test_no_alloc_entry_matches_the_existing_entry_points(dga):
    torch.cuda.set_device(0)
    prepared = []
    for i, (m, n, k, dtype, out_dtype, layout_b, split) in enumerate(CALLS):
        a, b, ta, mat2 = _operands(m, n, k, dtype, layout_b, seed=70 + i)
        # the existing entry points first, each with its own workspace
        want = torch.empty((m, n), dtype=out_dtype, device=DEV)
        dga.catlass_dynamic_matmul(ta, mat2, out=want)
        same_plan = None
        if split != 0:
            same_plan = torch.empty((m, n), dtype=out_dtype, device=DEV)
            dga.matmul_with_plan(ta, mat2, same_plan, kernelSerial=5, ctaGroup=2, n1=256, splitkFactor=split)
        got = torch.full((m, n), float("nan"), dtype=out_dtype, device=DEV)
        prepared.append((a, b, ta, mat2, want, same_plan, got, layout_b, split))
    torch.cuda.synchronize()

    # the calls under test, back to back on one stream; from the second call on nothing on the device may be
    # allocated or freed (the first call loads the kernels and looks the array up)
    launches0 = dga.kernel_launch_count()
    free_after_first = None
    for a, b, ta, mat2, want, same_plan, got, layout_b, split in prepared:
        _no_alloc_call(dga, ta, mat2, got, layout_b, split)
        if free_after_first is None:
            torch.cuda.synchronize()
            free_after_first = torch.cuda.mem_get_info(0)[0]
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free_after_first, "device memory moved during the no-allocation calls"
    assert dga.kernel_launch_count() - launches0 == 2 * len(prepared)   # GEMM + reduce per call, nothing else

    for a, b, ta, mat2, want, same_plan, got, layout_b, split in prepared:
        g, w = got.float().cpu().numpy(), want.float().cpu().numpy()
        assert np.isfinite(g).all()
        tol = 1e-5 if got.dtype == torch.float32 else 1e-2
        assert orc.max_rel_err(g, w.astype(np.float64)) <= tol          # vs the existing operator call
        assert orc.max_rel_err(g, orc.golden_fp64(a, b)) <= tol         # vs the oracle
        if same_plan is not None:
            assert np.array_equal(g, same_plan.float().cpu().numpy())   # same kernels, caller workspace: bit-identical


def test_no_alloc_entry_rejects_what_does_not_fit(dga):
    lib = dga._lib.lib
    limit = dga._lib.NO_ALLOC_SCRATCH_BYTES
    m, n, k = 2048, 4096, 1024   # 2 slices x 2048 x 4096 x 4 B = 64 MiB
    assert 2 * m * n * 4 > limit
    ta = torch.ones((m, k), dtype=torch.bfloat16, device=DEV)
    w = torch.ones((n, k), dtype=torch.bfloat16, device=DEV)
    out = torch.zeros((m, n), dtype=torch.bfloat16, device=DEV)
    torch.cuda.synchronize()
    before = dga.kernel_launch_count()
    for split in (0, 2):
        rc = lib.dgaSplitkMatmulNoAlloc(ctypes.c_void_p(ta.data_ptr()), ctypes.c_void_p(w.data_ptr()),
                                        ctypes.c_void_p(out.data_ptr()), m, n, k, dga.ops.acl_dtype(ta.dtype),
                                        dga.ops.acl_dtype(out.dtype), 1, split, None)
        assert rc == 161002   # ACLNN_ERR_PARAM_INVALID
        msg = lib.dgaGetLastError().decode()
        assert str(limit) in msg and str(2 * m * n * 4) in msg
    # unaligned rows (k = 1001 bf16 elements) would need the repack workspace: rejected as well
    rc = lib.dgaSplitkMatmulNoAlloc(ctypes.c_void_p(ta.data_ptr()), ctypes.c_void_p(w.data_ptr()),
                                    ctypes.c_void_p(out.data_ptr()), 256, 512, 1001, dga.ops.acl_dtype(ta.dtype),
                                    dga.ops.acl_dtype(out.dtype), 1, 2, None)
    assert rc != 0
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() == before
    assert float(out.float().abs().max()) == 0.0
