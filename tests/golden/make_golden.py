#!/usr/bin/env python3
"""Generates the golden fixtures in this directory by RUNNING THE REFERENCE ITSELF in the build container.

    python tests/golden/make_golden.py            (needs /root/reference and oracle/_ref/libref_host.so)

Nothing here is needed at test time: the fixtures are committed and /root/reference does not exist on the GPU box.

Fixtures
  numeric_uniform.npz   deep_gemm_ascend/scripts/gen_golden.py:10-23 gen_golden_data(M,N,K) run with
                        np.random.seed(seed): fp16 U[1,10] inputs + the fp32 numpy golden it writes to output/golden.bin
  numeric_heavy.npz     deep_gemm_ascend/framework/tests/test.py:23-38 gen_golden_data() (M=1,N=512,K=128, lognormal
                        clipped to [1,10]); the function body is lifted out of the file with ast (the module imports
                        torch_npu/acl at the top) and run with np.random.default_rng patched to a seeded generator
  comparator.json       deep_gemm_ascend/scripts/verify.py:15-35 verify_result() verdicts on perturbed outputs
  ref_tiling.json       CatlassTiling::SelectKernel (op_tiling/select_kernel.cpp:333-369) through oracle/_ref, for the
                        BASELINE shapes, the reference benchmark shape list (benchmark.py:24-44) and the unit-test
                        shapes, 4 layout combinations, 24 and 20 cores
  ref_tiling_py.json    get_best_config.MatmulTilingCalculator.calculate (tiling_calculator.py:601-650) on the same
                        shapes: the Python mirror, recorded to document where it diverges from the C++ selector
  bench_config.json     deep_gemm_ascend::get_bench_config (jit/get_best_config.hpp:93-153) through oracle/_ref
"""
import ast
import ctypes
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("DGA_REFERENCE", "/root/reference")

SHAPES = [
    # BASELINE.json configs
    (1024, 1024, 1024), (4096, 7168, 2048), (1, 4096, 7168), (16, 4096, 7168), (64, 4096, 7168), (128, 4096, 7168),
    (512, 4096, 7168), (2048, 4096, 7168), (8192, 4096, 7168), (4096, 7168, 16384), (16384, 57344, 7168),
    (16384, 7168, 7168),
    # deep_gemm_ascend/framework/benchmark/benchmark.py:24-44
    (4096, 4096, 4096), (8, 7168, 18432), (8, 18432, 7168), (64, 7168, 18432), (64, 18432, 7168), (64, 24576, 1536),
    (64, 32768, 512), (64, 7168, 16384), (128, 7168, 18432), (128, 18432, 7168), (1024, 4096, 7168),
    (1024, 18432, 7168), (1279, 5003, 7681), (3511, 6151, 8191), (5119, 6997, 9901),
    # unit-test shapes (aclnn tests/select_kernel_test.cpp:29-144, test_tiling_calculator.py:210-220) and edge cases
    (128, 128, 128), (16, 16, 16), (8192, 8192, 8192), (1, 1024, 1024), (1024, 1, 1024), (16, 578, 2014), (1, 512, 128),
    (1, 1, 1), (300, 520, 136),
]


def load_script(rel):
    path = os.path.join(REF, rel)
    with open(path, encoding="utf-8") as f:
        return f.read(), path


def numeric_uniform():
    src, path = load_script("deep_gemm_ascend/scripts/gen_golden.py")
    ns = {"__name__": "ref_gen_golden"}
    exec(compile(src, path, "exec"), ns)
    out = {}
    cases = [(0, 64, 128, 32), (1, 33, 50, 70), (2, 1, 512, 128), (3, 128, 96, 256)]
    cwd = os.getcwd()
    for seed, m, n, k in cases:
        with tempfile.TemporaryDirectory() as tmp:
            os.chdir(tmp)
            try:
                np.random.seed(seed)
                ns["gen_golden_data"](m, n, k)
                x1 = np.fromfile("input/x1_gm.bin", dtype=np.float16).reshape(m, k)
                x2 = np.fromfile("input/x2_gm.bin", dtype=np.float16).reshape(k, n)
                g = np.fromfile("output/golden.bin", dtype=np.float32).reshape(m, n)
            finally:
                os.chdir(cwd)
        tag = "s%d_%dx%dx%d" % (seed, m, n, k)
        out[tag + "_x1"], out[tag + "_x2"], out[tag + "_golden"] = x1, x2, g
    np.savez_compressed(os.path.join(HERE, "numeric_uniform.npz"), **out)
    return sorted(out)


def lift_function(src, name):
    tree = ast.parse(src)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            return ast.get_source_segment(src, node)
    raise KeyError(name)


def numeric_heavy():
    src, path = load_script("deep_gemm_ascend/framework/tests/test.py")
    body = lift_function(src, "gen_golden_data")
    out = {}
    real_default_rng = np.random.default_rng
    for seed in (0, 1):
        np.random.default_rng = lambda *a, _s=seed, **k: real_default_rng(_s)
        try:
            ns = {"np": np}
            exec(compile(body, path, "exec"), ns)
            x1, x2, g = ns["gen_golden_data"]()
        finally:
            np.random.default_rng = real_default_rng
        out["s%d_x1" % seed], out["s%d_x2" % seed], out["s%d_golden" % seed] = x1, x2, g
    np.savez_compressed(os.path.join(HERE, "numeric_heavy.npz"), **out)
    return sorted(out)


def comparator():
    """verify.py verdicts: the perturbations mirror deep_gemm_ascend/scripts/tests/test_all_scripts.py:39-65,153-166."""
    src, path = load_script("deep_gemm_ascend/scripts/verify.py")
    ns = {"__name__": "ref_verify"}
    exec(compile(src, path, "exec"), ns)
    rng = np.random.RandomState(7)
    golden = rng.uniform(100, 1000, 20000).astype(np.float32)
    cases = {}

    def run(tag, output):
        with tempfile.TemporaryDirectory() as tmp:
            o, g = os.path.join(tmp, "o.bin"), os.path.join(tmp, "g.bin")
            output.astype(np.float32).tofile(o)
            golden.tofile(g)
            devnull = open(os.devnull, "w")
            old = sys.stdout
            sys.stdout = devnull
            try:
                ok = bool(ns["verify_result"](o, g))
            finally:
                sys.stdout = old
                devnull.close()
        cases[tag] = ok

    run("exact", golden.copy())
    run("rel_1e-7", golden * np.float32(1 + 1e-7))
    run("rel_5e-6", (golden.astype(np.float64) * (1 + 5e-6)).astype(np.float32))
    run("times_2", golden * 2)
    bad = golden.copy(); bad[:200] *= 1.5          # 1 % corrupted
    run("corrupt_1pct", bad)
    bad = golden.copy(); bad[:2] *= 1.5            # exactly 1e-4 corrupted: still passes (<=)
    run("corrupt_1e-4", bad)
    bad = golden.copy(); bad[:3] *= 1.5            # 1.5e-4 corrupted
    run("corrupt_1.5e-4", bad)
    run("short", golden[:-1].copy())
    with open(os.path.join(HERE, "comparator.json"), "w") as f:
        json.dump({"seed": 7, "n": 20000, "low": 100, "high": 1000, "verdicts": cases}, f, indent=1)
    return cases


def ref_lib():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])
    return os.path.join(ROOT, "oracle", "_ref", "libref_host.so")


def ref_tiling():
    """SelectKernel holds a function-local static PlatformInfo (select_kernel.cpp:349), so each core count runs in
    its own process."""
    lib_path = ref_lib()
    probe = (
        "import ctypes, json, sys\n"
        "l = ctypes.CDLL(%r)\n"
        "shapes = json.loads(sys.argv[1])\n"
        "out = (ctypes.c_uint64 * 17)()\n"
        "rows = []\n"
        "for m, n, k in shapes:\n"
        "    for la in (0, 1):\n"
        "        for lb in (0, 1):\n"
        "            l.ref_select_kernel(m, n, k, la, lb, out)\n"
        "            rows.append([m, n, k, la, lb] + [int(v) for v in out])\n"
        "print(json.dumps(rows))\n" % lib_path)
    result = {"columns": ["m", "n", "k", "layoutA_in", "layoutB_in", "m1", "n1", "k1", "kernelSerial", "paddingTagA",
                          "paddingTagB", "paddingTagC", "blockDim", "swizzleOffset", "swizzleDirection",
                          "splitkFactor", "strideA", "strideB", "strideC", "layoutTagA", "layoutTagB", "layoutTagC"]}
    for cores in (24, 20):
        env = dict(os.environ, REF_CORE_NUM=str(cores))
        txt = subprocess.check_output([sys.executable, "-c", probe, json.dumps(SHAPES)], env=env, text=True)
        result["cores%d" % cores] = json.loads(txt.strip().splitlines()[-1])
    with open(os.path.join(HERE, "ref_tiling.json"), "w") as f:
        json.dump(result, f, separators=(",", ":"))
    return len(result["cores24"])


def ref_tiling_py():
    sys.path.insert(0, os.path.join(REF, "get_best_config"))
    from tiling_calculator import MatmulTilingCalculator  # noqa: E402  (the reference's Python mirror)
    rows = {}
    for cores in (24, 20):
        from tiling_calculator import PlatformInfo
        info = PlatformInfo()
        info.coreNum = cores
        calc = MatmulTilingCalculator(info)
        lst = []
        for m, n, k in SHAPES:
            for la in (0, 1):
                for lb in (0, 1):
                    r = calc.calculate(m, n, k, la, lb)
                    lst.append([m, n, k, la, lb, int(r["tiling"][0]), int(r["tiling"][1]), int(r["tiling"][2]),
                                int(r["kernelSerial"]), int(r["paddingTagA"]), int(r["paddingTagB"]),
                                int(r["paddingTagC"]), int(r["blockDim"])])
        rows["cores%d" % cores] = lst
    rows["columns"] = ["m", "n", "k", "layoutA", "layoutB", "m1", "n1", "k1", "kernelSerial", "paddingTagA",
                       "paddingTagB", "paddingTagC", "blockDim"]
    with open(os.path.join(HERE, "ref_tiling_py.json"), "w") as f:
        json.dump(rows, f, separators=(",", ":"))
    return len(rows["cores24"])


def bench_config():
    lib = ctypes.CDLL(ref_lib())
    rng = np.random.RandomState(11)
    rows = []
    fixed = [((1024, 4096, 7168), (2, 4, 8, 16, 20, 10)), ((1, 512, 128), (1, 1, 3, 8, 20, 10)),
             ((4096, 4096, 4096), (4, 5, 16, 16, 8, 4)), ((1279, 5003, 7681), (3, 7, 5, 9, 12, 5)),
             ((16, 16, 16), (1, 1, 1, 1, 1, 1))]
    for _ in range(40):
        shape = tuple(int(v) for v in rng.randint(1, 9000, 3))
        kit = int(rng.randint(1, 33))
        knobs = (int(rng.randint(1, 6)), int(rng.randint(1, 6)), int(rng.randint(1, 17)), int(rng.randint(1, 17)), kit,
                 int(rng.randint(1, kit + 1)))
        fixed.append((shape, knobs))
    for shape, knobs in fixed:
        p = (ctypes.c_int32 * 28)(*knobs)
        lib.ref_get_bench_config(shape[0], shape[1], shape[2], p)
        rows.append({"shape": list(shape), "params": list(p)})
    with open(os.path.join(HERE, "bench_config.json"), "w") as f:
        json.dump(rows, f, separators=(",", ":"))
    return len(rows)


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("reference tree not found at %s" % REF)
    print("numeric_uniform:", numeric_uniform())
    print("numeric_heavy:", numeric_heavy())
    print("comparator:", comparator())
    print("ref_tiling rows:", ref_tiling())
    print("ref_tiling_py rows:", ref_tiling_py())
    print("bench_config rows:", bench_config())
