"""GPU parity of the run-time specialised kernels (option "jit"): same inputs, same oracle and tolerances as
test_gpu_parity.py, plus the cache behaviour the reference's JIT layer has (compile once, in-process map, directory
found again: framework/csrc/jit/compiler.hpp:26-52, cache.hpp:17-30)."""
import numpy as np
import pytest
import torch

from oracle import matmul_oracle as orc
from test_gpu_parity import _inputs, _run_aclnn, DEV

pytestmark = pytest.mark.gpu


@pytest.fixture()
def jit_on(dga, tmp_path, monkeypatch):
    monkeypatch.setenv("DGA_ROOT_DIR", str(tmp_path))
    dga.set_option("jit", 1)
    dga.jit_drop_memory_cache()
    yield dga
    dga.set_option("jit", -1)
    dga.jit_drop_memory_cache()


@pytest.mark.parametrize("m,n,k,dtype,layout_b", [
    (512, 1024, 512, torch.bfloat16, 1),     # CTA pair, weight [N, K]
    (300, 520, 264, torch.bfloat16, 1),      # ragged tiles
    (1024, 1024, 1024, torch.float16, 1),    # BASELINE config 1
    (256, 384, 320, torch.float16, 0),       # row-major B
    (1100, 777, 203, torch.bfloat16, 1),     # unaligned pitches: repack + specialised kernel
    (256, 2048, 7168, torch.bfloat16, 1),    # split-K + reduce: the GEMM half is specialised (8 slices)
    (16, 2112, 7168, torch.bfloat16, 0),     # row-major mat2, few rows: K slices of single-CTA tiles
])
def test_specialised_kernel_matches_oracle(jit_on, m, n, k, dtype, layout_b):
    dga = jit_on
    a, b = _inputs(m, n, k, dtype, seed=m + n + k)
    before = dga.jit_stats()
    out = _run_aclnn(dga, a, b, dtype, layout_b)
    after = dga.jit_stats()
    assert after["compiled"] == before["compiled"] + 1, "the launch did not go through a freshly compiled kernel"
    assert orc.max_rel_err(out, orc.golden_fp64(a, b)) <= 1e-2
    # bit-identical to the dynamic-shape kernel (same tiles, same accumulation order)
    dga.set_option("jit", 0)
    assert np.array_equal(out, _run_aclnn(dga, a, b, dtype, layout_b))
    dga.set_option("jit", 1)


def test_memory_then_disk_cache(jit_on):
    dga = jit_on
    a, b = _inputs(384, 512, 256, torch.bfloat16, seed=5)
    s0 = dga.jit_stats()
    first = _run_aclnn(dga, a, b, torch.bfloat16, 1)
    s1 = dga.jit_stats()
    second = _run_aclnn(dga, a, b, torch.bfloat16, 1)
    s2 = dga.jit_stats()
    dga.jit_drop_memory_cache()
    third = _run_aclnn(dga, a, b, torch.bfloat16, 1)
    s3 = dga.jit_stats()
    assert s1["compiled"] == s0["compiled"] + 1
    assert (s2["compiled"], s2["memory_hits"]) == (s1["compiled"], s1["memory_hits"] + 1)
    assert (s3["compiled"], s3["disk_hits"]) == (s2["compiled"], s2["disk_hits"] + 1)
    assert np.array_equal(first, second) and np.array_equal(first, third)


def test_rtc_goes_through_the_specialised_kernel(jit_on):
    """run_mmad_rtc is the reference's runtime-compiled entry (gemm.hpp:63-111): batched, bf16 -> fp32."""
    dga = jit_on
    rng = np.random.default_rng(11)
    bsz, m, n, k = 3, 200, 328, 264
    x = torch.from_numpy(orc.heavy_tail(rng, [bsz, m, k])).to(DEV).to(torch.bfloat16)
    y = torch.from_numpy(orc.heavy_tail(rng, [bsz, k, n])).to(DEV).to(torch.bfloat16)
    out = torch.zeros((bsz, m, n), dtype=torch.float32, device=DEV)
    before = dga.jit_stats()["compiled"]
    dga.run_mmad_rtc(x, y, out)
    assert dga.jit_stats()["compiled"] == before + 1
    for i in range(bsz):
        ref = orc.golden_fp64(x[i].float().cpu().numpy(), y[i].float().cpu().numpy())
        assert orc.max_rel_err(out[i].cpu().numpy(), ref) < 1e-5


def test_compile_failure_fails_the_call(jit_on, monkeypatch):
    dga = jit_on
    monkeypatch.setenv("NVCC", "/bin/false")
    a, b = _inputs(640, 256, 128, torch.bfloat16, seed=1)
    with pytest.raises(dga.DgaError, match="kernel compilation failed"):
        _run_aclnn(dga, a, b, torch.bfloat16, 1)
