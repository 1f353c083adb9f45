"""Run-time specialisation cache, host side (no GPU): directory scheme, generated source, reuse, stale-directory
rebuild and loud failure. Mirrors what the reference's JIT layer promises (framework/csrc/jit/compiler.hpp:26-52,
cache.hpp:11-33, kernel_runtime.hpp:30-33); nvcc cross-compiles sm_100a without a device."""
import os
import re

import pytest

import deepgemm_ascend_b200 as dga
from deepgemm_ascend_b200 import _lib


@pytest.fixture()
def jit_root(tmp_path, monkeypatch):
    monkeypatch.setenv("DGA_ROOT_DIR", str(tmp_path))
    return tmp_path


def test_prepare_writes_reference_directory_scheme_and_reuses_it(jit_root):
    before = dga.jit_stats()["compiled"]
    d = dga.jit_prepare(512, 1024, 256)
    # $DGA_ROOT_DIR/deep_gemm_ascend/cache/kernel_b{}m{}n{}k{}_type{}/ (compiler.hpp:31-33)
    assert re.fullmatch(re.escape(str(jit_root)) + r"/deep_gemm_ascend/cache/kernel_b1m512n1024k256_type\d+/", d)
    assert sorted(os.listdir(d)) == ["kernel.cu", "kernel.cubin"]
    assert os.path.getsize(d + "kernel.cubin") > 10000
    code = open(d + "kernel.cu").read()
    for token in ("#define DGA_JIT_M 512u", "#define DGA_JIT_N 1024u", "#define DGA_JIT_K 256u",
                  "#define DGA_JIT_K_BLOCKS 4u", '#include "jit_entry.cuh"'):
        assert token in code
    assert dga.jit_stats()["compiled"] == before + 1
    assert dga.jit_prepare(512, 1024, 256) == d
    assert dga.jit_stats()["compiled"] == before + 1   # valid directory: no second compile


def test_variants_of_one_shape_get_their_own_directory(jit_root):
    d_bf16 = dga.jit_prepare(256, 512, 256, _lib.ACL_BF16)
    d_fp16 = dga.jit_prepare(256, 512, 256, _lib.ACL_FLOAT16)
    d_rowmajor = dga.jit_prepare(256, 512, 256, _lib.ACL_BF16, _lib.ACL_FLOAT, 0)
    assert len({d_bf16, d_fp16, d_rowmajor}) == 3
    assert "__half" in open(d_fp16 + "kernel.cu").read()
    assert "#define DGA_JIT_B_MN true" in open(d_rowmajor + "kernel.cu").read()


def test_batched_signature(jit_root):
    d = dga.jit_prepare(48, 128, 320, _lib.ACL_BF16, _lib.ACL_FLOAT, 0, batch=3)
    assert "/kernel_b3m48n128k320_type" in d
    assert "#define DGA_JIT_MODE 1" in open(d + "kernel.cu").read()


def test_stale_directory_is_rebuilt(jit_root):
    d = dga.jit_prepare(640, 256, 128)
    with open(d + "kernel.cu", "a") as f:
        f.write("// left by another version\n")
    before = dga.jit_stats()["compiled"]
    dga.jit_prepare(640, 256, 128)
    assert dga.jit_stats()["compiled"] == before + 1
    assert "another version" not in open(d + "kernel.cu").read()


def test_compiler_failure_is_an_error(jit_root, monkeypatch):
    monkeypatch.setenv("NVCC", "/bin/false")
    with pytest.raises(dga.DgaError, match="kernel compilation failed"):
        dga.jit_prepare(384, 256, 128)


def test_unspecialised_variant_is_refused(jit_root):
    with pytest.raises(dga.DgaError, match="not specialised"):
        dga.jit_prepare(16, 4096, 7168)   # decode kernel


def _prepare_in_child(root, queue):
    os.environ["DGA_ROOT_DIR"] = root
    import deepgemm_ascend_b200 as child_dga
    try:
        queue.put(("ok", child_dga.jit_prepare(896, 512, 192)))
    except Exception as e:   # noqa: BLE001 - reported to the parent
        queue.put(("error", repr(e)))


def test_two_processes_racing_for_one_directory(jit_root):
    """The binary appears under its final name only by rename, so concurrent builders of one kernel both succeed and
    leave a loadable directory (the reference's cache has no such guard)."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    queue = ctx.Queue()
    procs = [ctx.Process(target=_prepare_in_child, args=(str(jit_root), queue)) for _ in range(2)]
    for p in procs:
        p.start()
    results = [queue.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert [r[0] for r in results] == ["ok", "ok"], results
    d = results[0][1]
    assert results[1][1] == d
    assert sorted(os.listdir(d)) == ["kernel.cu", "kernel.cubin"]   # no temporary files left behind
    before = dga.jit_stats()["compiled"]
    assert dga.jit_prepare(896, 512, 192) == d
    assert dga.jit_stats()["compiled"] == before                     # the directory is valid for a third process
