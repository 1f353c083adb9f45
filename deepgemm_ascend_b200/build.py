"""Builds libdga_b200.so (CUDA kernels + C ABI) in-tree with nvcc for sm_100a.

Replaces the reference's Ascend build scripts (aclnn_catlass_dynamic_matmul/build.sh:151-163,
deep_gemm_ascend/build_dga.sh:16-17): one shared library, no generated code, no JIT.
nvcc cross-compiles without a GPU, so this runs in the CPU container too.
"""
import hashlib
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_DIR = os.path.join(PKG_DIR, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libdga_b200.so")
STAMP = os.path.join(LIB_DIR, "libdga_b200.stamp")

SOURCES = ["api.cu", "tile_selector.cpp", "plan_cache.cpp", "jit.cpp"]
HEADERS = ["ptx.cuh", "gemm_sm100.cuh", "gemm_smallm_sm100.cuh", "gemm_fp8_sm100.cuh", "repack_sm100.cuh", "splitk_reduce_sm100.cuh", "tile_selector.h", "jit.h", "jit_entry.cuh",
           os.path.join("..", "..", "include", "dga_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
    "-Xcompiler", "-fPIC", "-shared", "--expt-relaxed-constexpr",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest():
    h = hashlib.sha256()
    for name in SOURCES + HEADERS:
        path = os.path.join(CSRC, name)
        if os.path.exists(path):
            with open(path, "rb") as f:
                h.update(name.encode())
                h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_fresh():
    if not (os.path.exists(LIB_PATH) and os.path.exists(STAMP)):
        return False
    with open(STAMP) as f:
        return f.read().strip() == _digest()


def build(force=False, verbose=False):
    """Compile the library if sources changed. Returns the path of the .so."""
    os.makedirs(LIB_DIR, exist_ok=True)
    if not force and is_fresh():
        return LIB_PATH
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + \
          [os.path.join(CSRC, s) for s in SOURCES]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout + proc.stderr)
        raise RuntimeError("nvcc failed building libdga_b200.so")
    if verbose:
        sys.stderr.write(proc.stderr)
    with open(STAMP, "w") as f:
        f.write(_digest())
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
