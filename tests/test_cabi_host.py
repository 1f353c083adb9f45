"""Host side of the C ABI without a GPU: the library loads, exports every symbol include/dga_b200.h declares,
and the pure-host entry points (tile selector, plan cache, CSV, bench config, N-shard ranges, validation) behave.
Selector/cache tests restate the invariants of the reference's gtest suite
(aclnn_catlass_dynamic_matmul/tests/{select_kernel,do_tiling,cache,csv,utils}_test.cpp)."""
import ctypes
import json
import os
import re

import pytest
import torch

from oracle import matmul_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(dga):
    header = open(os.path.join(ROOT, "include", "dga_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b((?:acl|dga)[A-Za-z0-9]+)\s*\(", header))
    assert {"aclnnCatlassDynamicMatmulGetWorkspaceSize", "aclnnCatlassDynamicMatmul", "dgaRunMmadRtc"} <= declared
    lib = dga._lib.lib
    for name in sorted(declared):
        assert hasattr(lib, name), "library does not export %s" % name
    assert declared == set(dga._lib.SIGNATURES), "ctypes table and header disagree"
    assert b"sm_100a" in lib.dgaVersion()


def test_no_cpu_fallback(dga):
    a = torch.ones(16, 16, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        dga.catlass_dynamic_matmul(a, a)
    if not torch.cuda.is_available():
        # phase 2 on a box without a CUDA driver must fail loudly, not compute on the host
        plan = dga.select_kernel(128, 128, 128)
        assert plan["blockDim"] > 0


def _desc(dga, shape, stride, dtype):
    d = dga._lib.AclTensor()
    d.data = 0x10000
    d.dtype = dtype
    d.ndim = len(shape)
    for i, (s, st) in enumerate(zip(shape, stride)):
        d.shape[i], d.stride[i] = s, st
    return d


def _phase1(dga, a, b, c):
    ws, ex = ctypes.c_uint64(0), ctypes.c_void_p(None)
    rc = dga._lib.lib.aclnnCatlassDynamicMatmulGetWorkspaceSize(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c),
                                                               ctypes.byref(ws), ctypes.byref(ex))
    if rc == 0:
        dga._lib.lib.dgaDestroyExecutor(ex)
    return rc, ws.value


def test_phase1_validation_matches_reference_rules(dga, capfd):
    BF16, F16 = 27, 1
    ok = (_desc(dga, (64, 128), (128, 1), BF16), _desc(dga, (128, 256), (1, 128), BF16),
          _desc(dga, (64, 256), (256, 1), BF16))
    rc, ws = _phase1(dga, *ok)
    assert rc == 0
    # K mismatch: catlass_dynamic_matmul_tiling.cpp:96-100
    rc, _ = _phase1(dga, ok[0], _desc(dga, (96, 256), (1, 96), BF16), ok[2])
    assert rc != 0
    assert "[DGA] [ERROR]" in capfd.readouterr().err
    # not 2-D: catlass_dynamic_matmul.cpp:22-26
    rc, _ = _phase1(dga, _desc(dga, (2, 64, 128), (8192, 128, 1), BF16), ok[1], ok[2])
    assert rc != 0
    # dtype mismatch / unsupported dtype: catlass_dynamic_matmul.cpp:37-46
    rc, _ = _phase1(dga, _desc(dga, (64, 128), (128, 1), F16), ok[1], ok[2])
    assert rc != 0
    rc, _ = _phase1(dga, _desc(dga, (64, 128), (128, 1), 0), _desc(dga, (128, 256), (1, 128), 0),
                    _desc(dga, (64, 256), (256, 1), 0))
    assert rc != 0
    # null arguments
    assert dga._lib.lib.aclnnCatlassDynamicMatmulGetWorkspaceSize(None, None, None, None, None) != 0
    # empty operands (a zero-sized M, N or K): refused with a message, never a zero-sized launch
    for shp in [((0, 128), (128, 256), (0, 256)), ((64, 128), (128, 0), (64, 0)), ((64, 0), (0, 256), (64, 256))]:
        (am, ak), (bk, bn), (cm, cn) = shp
        capfd.readouterr()
        rc, _ = _phase1(dga, _desc(dga, (am, ak), (max(ak, 1), 1), BF16), _desc(dga, (bk, bn), (1, max(bk, 1)), BF16),
                        _desc(dga, (cm, cn), (max(cn, 1), 1), BF16))
        assert rc != 0
        assert "[DGA] [ERROR]" in capfd.readouterr().err
    # row-major mat2 (framework layout) is accepted too
    rc, _ = _phase1(dga, ok[0], _desc(dga, (128, 256), (256, 1), BF16), ok[2])
    assert rc == 0


def test_grouped_phase1_validation(dga, capfd):
    """Shape / dtype rules of the two MoE layouts (contiguous: self [M,K], weight [G,N,K], index [M]; masked: self
    [G,Mmax,K], weight [G,N,K], maskedM [G]); phase 1 needs no GPU."""
    BF16, I32 = 27, 3
    L = dga._lib.lib

    def phase1(fn, *tensors):
        ws, ex = ctypes.c_uint64(0), ctypes.c_void_p(None)
        rc = fn(*[ctypes.byref(t) for t in tensors], ctypes.byref(ws), ctypes.byref(ex))
        if rc == 0:
            assert ws.value == 0
            L.dgaDestroyExecutor(ex)
        return rc

    a = _desc(dga, (256, 128), (128, 1), BF16)
    w = _desc(dga, (4, 512, 128), (512 * 128, 128, 1), BF16)
    idx = _desc(dga, (256,), (1,), I32)
    c = _desc(dga, (256, 512), (512, 1), BF16)
    assert phase1(L.dgaMGroupedMatmulGetWorkspaceSize, a, w, idx, c) == 0
    assert phase1(L.dgaMGroupedMatmulGetWorkspaceSize, _desc(dga, (200, 128), (128, 1), BF16), w,
                  _desc(dga, (200,), (1,), I32), _desc(dga, (200, 512), (512, 1), BF16)) != 0   # M_total % 128
    assert phase1(L.dgaMGroupedMatmulGetWorkspaceSize, a, w, _desc(dga, (256,), (1,), BF16), c) != 0  # index dtype

    am = _desc(dga, (4, 200, 128), (200 * 128, 128, 1), BF16)
    mm = _desc(dga, (4,), (1,), I32)
    cm = _desc(dga, (4, 200, 512), (200 * 512, 512, 1), BF16)
    assert phase1(L.dgaMGroupedMaskedMatmulGetWorkspaceSize, am, w, mm, cm) == 0
    assert phase1(L.dgaMGroupedMaskedMatmulGetWorkspaceSize, am, w, _desc(dga, (3,), (1,), I32), cm) != 0   # G mismatch
    assert phase1(L.dgaMGroupedMaskedMatmulGetWorkspaceSize, am, _desc(dga, (4, 512, 96), (512 * 96, 96, 1), BF16),
                  mm, cm) != 0                                                                      # K mismatch
    assert phase1(L.dgaMGroupedMaskedMatmulGetWorkspaceSize, a, w, mm, cm) != 0                     # self not 3-D
    assert "[DGA] [ERROR]" in capfd.readouterr().err


def test_round2_entry_points_validate_on_the_host(dga, capfd):
    """The entries added in round 2 refuse what they do not serve BEFORE anything touches a device: the group
    alignment promise, the two-GEMM decode launch, the matmul that allocates nothing."""
    BF16, F16, I32 = 27, 1, 3
    L = dga._lib.lib
    INVALID, NULLPTR = 161002, 161001
    # --- dgaSetExecutorGroupAlignment: 128 or 256, must divide the row count; only grouped executors
    w = _desc(dga, (4, 512, 128), (512 * 128, 128, 1), BF16)

    def grouped(rows):
        ws, ex = ctypes.c_uint64(0), ctypes.c_void_p(None)
        rc = L.dgaMGroupedMatmulGetWorkspaceSize(ctypes.byref(_desc(dga, (rows, 128), (128, 1), BF16)), ctypes.byref(w),
                                                 ctypes.byref(_desc(dga, (rows,), (1,), I32)),
                                                 ctypes.byref(_desc(dga, (rows, 512), (512, 1), BF16)),
                                                 ctypes.byref(ws), ctypes.byref(ex))
        assert rc == 0
        return ex
    ex = grouped(512)
    assert L.dgaSetExecutorGroupAlignment(ex, 256) == 0
    assert L.dgaSetExecutorGroupAlignment(ex, 128) == 0
    assert L.dgaSetExecutorGroupAlignment(ex, 64) == INVALID
    L.dgaDestroyExecutor(ex)
    ex = grouped(384)
    assert L.dgaSetExecutorGroupAlignment(ex, 256) == INVALID      # 384 rows are not whole pairs
    L.dgaDestroyExecutor(ex)
    assert L.dgaSetExecutorGroupAlignment(None, 256) == NULLPTR
    ok = (_desc(dga, (64, 128), (128, 1), BF16), _desc(dga, (128, 256), (1, 128), BF16),
          _desc(dga, (64, 256), (256, 1), BF16))
    ws, ex = ctypes.c_uint64(0), ctypes.c_void_p(None)
    assert L.aclnnCatlassDynamicMatmulGetWorkspaceSize(*[ctypes.byref(t) for t in ok], ctypes.byref(ws), ctypes.byref(ex)) == 0
    assert L.dgaSetExecutorGroupAlignment(ex, 256) == NULLPTR      # a dense executor
    L.dgaDestroyExecutor(ex)

    # --- dgaDecodePairMatmul
    def pair(self_t, w1, o1, w2, o2):
        return L.dgaDecodePairMatmul(*[None if t is None else ctypes.byref(t) for t in (self_t, w1, o1, w2, o2)], None)
    x = _desc(dga, (256, 128), (128, 1), BF16)
    wa, oa = _desc(dga, (128, 256), (1, 128), BF16), _desc(dga, (256, 256), (256, 1), BF16)
    assert pair(x, wa, oa, wa, oa) == INVALID                                      # m > 128
    x64, o64 = _desc(dga, (64, 128), (128, 1), BF16), _desc(dga, (64, 256), (256, 1), BF16)
    assert pair(x64, wa, o64, _desc(dga, (128, 256), (256, 1), BF16), o64) == INVALID   # row-major second weight
    assert pair(x64, wa, o64, _desc(dga, (96, 256), (1, 96), BF16), o64) == INVALID     # K mismatch
    assert pair(x64, wa, o64, wa, _desc(dga, (64, 256), (256, 1), 0)) == INVALID        # two output dtypes
    x_odd = _desc(dga, (64, 100), (100, 1), BF16)                                       # 200-byte rows: no repack here
    assert pair(x_odd, _desc(dga, (100, 256), (1, 100), BF16), o64, _desc(dga, (100, 256), (1, 100), BF16), o64) == INVALID
    assert pair(None, wa, o64, wa, o64) == NULLPTR

    # --- dgaSplitkMatmulNoAlloc: the size limit and the argument rules come before the first CUDA call
    p = ctypes.c_void_p(0x10000)
    limit = dga._lib.NO_ALLOC_SCRATCH_BYTES
    header = open(os.path.join(ROOT, "include", "dga_b200.h")).read()
    assert re.search(r"#define\s+DGA_NO_ALLOC_SCRATCH_BYTES\s+\(32u\s*<<\s*20\)", header) and limit == 32 << 20
    capfd.readouterr()
    assert L.dgaSplitkMatmulNoAlloc(p, p, p, 4096, 4096, 4096, BF16, BF16, 1, 4, None) == INVALID
    err = capfd.readouterr().err
    assert str(4 * 4096 * 4096 * 4) in err and str(limit) in err
    assert L.dgaSplitkMatmulNoAlloc(p, p, p, 2048, 4096, 1024, BF16, BF16, 1, 0, None) == INVALID   # automatic split: 64 MiB
    assert L.dgaSplitkMatmulNoAlloc(None, p, p, 256, 256, 256, BF16, BF16, 1, 2, None) == NULLPTR
    assert L.dgaSplitkMatmulNoAlloc(p, p, p, 256, 256, 256, 0, 0, 1, 2, None) == INVALID            # fp32 inputs
    assert L.dgaSplitkMatmulNoAlloc(p, p, p, 256, 256, 256, BF16, F16, 1, 2, None) == INVALID       # bf16 in, fp16 out
    assert L.dgaSplitkMatmulNoAlloc(p, p, p, 256, 256, 256, BF16, BF16, 1, 17, None) == INVALID     # > 16 slices
    assert L.dgaSplitkMatmulNoAlloc(p, p, p, 256, 256, 64, BF16, BF16, 1, 2, None) == INVALID       # one K block
    assert L.dgaSplitkMatmulNoAlloc(p, p, p, 256, 256, 1001, BF16, BF16, 1, 2, None) != 0           # unaligned rows
    assert L.dgaSplitkMatmulNoAlloc(p, p, p, 0, 256, 256, BF16, BF16, 1, 2, None) == INVALID        # empty


SWEEP = [1, 16, 64, 128, 512, 2048, 8192]


def test_selector_invariants_on_reference_shapes(dga, golden_dir):
    with open(os.path.join(golden_dir, "ref_tiling.json")) as f:
        ref = json.load(f)
    shapes = sorted({tuple(r[:3]) for r in ref["cores24"]})
    assert len(shapes) >= 30
    for m, n, k in shapes:
        p = dga.select_kernel(m, n, k)
        if (k * 2) % 16 or (n * 2) % 16:
            # unaligned leading dimensions: repack through the workspace (PADDING_ND, tiling_params.h:24-29)
            assert p["paddingTagA"] == p["paddingTagB"] == (1 if (k * 2) % 16 else 0)
            assert p["paddingTagC"] == (1 if (n * 2) % 16 and p["kernelSerial"] != 3 else 0)
            if p["paddingTagA"] or p["paddingTagB"] or p["paddingTagC"]:
                assert p["kernelSerial"] in (2, 3, 4) and p["workspaceBytes"] > 0
        else:
            assert p["paddingTagA"] == p["paddingTagB"] == p["paddingTagC"] == 0
        # the invariants select_kernel_test.cpp:29-144 and do_tiling_test.cpp:29-153 assert
        assert p["m1"] > 0 and p["n1"] > 0 and p["k1"] > 0 and p["blockDim"] > 0
        assert p["blockDim"] <= 148 or p["kernelSerial"] == 3
        assert p["smemBytes"] <= 232448
        assert p["stages"] >= 2
        assert p["strideA"] == k and p["strideB"] == k and p["strideC"] == n   # tiling_params.h:45-65
        assert p["swizzleDirection"] == (0 if m > n else 1)                     # tiling_params.h:64


def test_padding_workspace_layout(dga):
    """padA | padB | padC at the front of the workspace (padding_common_matmul_kernel.h:77-82), 128-byte row pitch."""
    m, n, k = 300, 1001, 515
    p = dga.select_kernel(m, n, k)
    assert (p["paddingTagA"], p["paddingTagB"], p["paddingTagC"]) == (1, 1, 1) and p["kernelSerial"] == 2
    kp, np_ = -(-k // 64) * 64, -(-n // 64) * 64
    up = lambda b: (b + 1023) // 1024 * 1024  # noqa: E731
    assert p["workspaceBytes"] == up(m * kp * 2) + up(n * kp * 2) + up(m * np_ * 2)
    # aligned K, odd N: only C goes through the workspace (weight stored [N,K]) ...
    q = dga.select_kernel(300, 1001, 512)
    assert (q["paddingTagA"], q["paddingTagB"], q["paddingTagC"]) == (0, 0, 1)
    # ... and with a row-major [K,N] weight its pitch is N as well
    q = dga.select_kernel(300, 1001, 512, layout_b=0)
    assert (q["paddingTagA"], q["paddingTagB"], q["paddingTagC"]) == (0, 1, 1)
    # the decode kernel stores C directly: no C repack
    q = dga.select_kernel(16, 1001, 512)
    assert q["kernelSerial"] == 3 and q["paddingTagC"] == 0 and q["workspaceBytes"] == 0


def test_selector_decision_structure(dga):
    # decode shapes stream the weight with the swap-AB split-K kernel (role of the reference's small-M clamping,
    # do_tiling.cpp:20-204, and of its declared multicore split-K type 3)
    for m in (1, 16, 64):
        p = dga.select_kernel(m, 4096, 7168)
        assert p["kernelSerial"] == 3 and p["splitkFactor"] >= 2 and p["m1"] == max(16, (m + 15) // 16 * 16)
    # prefill shapes use the persistent CTA-pair kernel with a wave-balanced N tile
    p = dga.select_kernel(4096, 7168, 2048)
    assert p["kernelSerial"] == 0 and p["ctaGroup"] == 2 and p["m1"] == 256
    tiles = -(-4096 // 256) * -(-7168 // p["n1"])
    waves = -(-tiles // 74)
    assert tiles / (waves * 74) > 0.95, "wave quantisation: BalanceWorkload's job (utils.cpp:28-40)"
    # under-filled and long K -> multicore split-K + reduce (the reference's declared kernel type 3 + StreamkReduceAdd)
    p = dga.select_kernel(256, 2048, 7168)
    assert p["kernelSerial"] == 5 and p["ctaGroup"] == 2 and p["n1"] == 256 and p["splitkFactor"] == 8
    assert p["workspaceBytes"] == 8 * 256 * 2048 * 4 and p["blockDim"] == 128
    # one 256-row block whose 128-wide tiles, cut in two K slices, fill the pairs in one round -> stream-K (the
    # reference's PaddingStreamk pick at its own ragged-wave shapes, select_kernel.cpp:290-301)
    for m, k in ((192, 4096), (256, 2048)):
        p = dga.select_kernel(m, 4096, k)
        assert p["kernelSerial"] == 4 and p["ctaGroup"] == 2 and p["n1"] == 128 and p["splitkFactor"] == 2
        assert p["blockDim"] == 128 and p["workspaceBytes"] >= 64 * 2 * 128 * 128 * 4
    # the same block with K >= 6144: cluster split-K on 256 x 128 tiles (32 clusters of four), no workspace
    p = dga.select_kernel(256, 4096, 7168)
    assert p["kernelSerial"] == 6 and p["n1"] == 128 and p["blockDim"] == 128 and p["workspaceBytes"] == 0
    # long K, 256 x 256 tiles fill 32 of the 33 clusters of four: two CTA pairs per tile, halves exchanged through DSMEM
    ks = dga.select_kernel(512, 4096, 7168)
    assert ks["kernelSerial"] == 6 and ks["clusterShape"] == 4 and ks["blockDim"] == 128 and ks["workspaceBytes"] == 0
    assert dga.select_kernel(512, 4096, 4096)["kernelSerial"] in (0, 1)      # shorter K, enough tiles: single pass
    assert dga.select_kernel(256, 4096, 512)["kernelSerial"] in (0, 1)       # short K: nothing to split
    # at most one tile per cluster -> Small (select_kernel.cpp:285-289)
    assert dga.select_kernel(128, 128, 128)["kernelSerial"] in (1, 3)


def test_tile_widths_that_are_odd_multiples_of_16(dga):
    """n1 = 240, 208, ... (UMMA N is legal in steps of 16): accepted for whole-tile plans over a K-major weight, refused
    for stream-K (its partial-tile slots are cut in 32-column chunks), for row-major mat2 and below 48."""
    lib = dga._lib.lib

    def try_plan(layout_b=1, **fields):
        lib.dgaTilingCacheClear()
        plan = dga._lib.TilingPlan()
        for key, val in dga.select_kernel(2048, 4096, 7168, layout_b=layout_b).items():
            setattr(plan, key, val)
        plan.stages = 0
        for key, val in fields.items():
            setattr(plan, key, val)
        rc = lib.dgaTilingCacheSet(ctypes.byref(plan))
        got = dga.select_kernel(2048, 4096, 7168, layout_b=layout_b, use_cache=True) if rc == 0 else None
        lib.dgaTilingCacheClear()
        return rc, got

    rc, got = try_plan(kernelSerial=0, ctaGroup=2, n1=240)
    assert rc == 0 and got["n1"] == 240 and got["stages"] >= 4 and got["smemBytes"] <= 232448
    rc, got = try_plan(kernelSerial=0, ctaGroup=1, n1=208)
    assert rc == 0 and got["n1"] == 208
    assert try_plan(kernelSerial=4, ctaGroup=2, n1=240, splitkFactor=2)[0] != 0
    assert try_plan(kernelSerial=0, ctaGroup=2, n1=16)[0] != 0
    assert try_plan(kernelSerial=0, ctaGroup=2, n1=40)[0] != 0
    assert try_plan(layout_b=0, kernelSerial=0, ctaGroup=1, n1=240)[0] != 0
    # the tuned-table sweep offers them (after every older candidate: indices are checkpoint keys)
    import deepgemm_ascend_b200.autotune as at
    cands = at.candidate_plans(2048, 4096, 7168)
    odd = [i for i, p in enumerate(cands) if p.get("n1", 0) % 32 == 16]
    round2 = [i for i, p in enumerate(cands) if p.get("clusterShape") == 4 or p["kernelSerial"] == 6]
    older = [i for i, p in enumerate(cands) if (p.get("n1", 0) % 32 == 0 or "n1" not in p) and i not in round2]
    assert odd and odd[0] > max(older)
    assert round2 and round2[0] > odd[-1]          # the round-2 cluster candidates come behind those again


def test_cluster_shape_rules(dga, tmp_path):
    """clusterShape (north_star item 2, "chooses ... the cluster shape"): every plan reports the CTAs per cluster; 4 =
    two CTA pairs sharing their A rows by TMA multicast, accepted only for whole-tile pair plans with an even number
    of N tiles, at most 33 such clusters side by side, and it survives the CSV round trip."""
    lib = dga._lib.lib
    assert dga.select_kernel(4096, 7168, 2048)["clusterShape"] == 2
    assert dga.select_kernel(128, 32768, 512)["clusterShape"] == 1
    p64 = dga.select_kernel(64, 4096, 7168)
    assert p64["kernelSerial"] == 3 and p64["clusterShape"] == p64["splitkFactor"]

    def try_plan(m, n, k, **fields):
        lib.dgaTilingCacheClear()
        plan = dga._lib.TilingPlan()
        for key, val in dga.select_kernel(m, n, k).items():
            setattr(plan, key, val)
        plan.stages = 0
        plan.splitkFactor = 1
        for key, val in fields.items():
            setattr(plan, key, val)
        rc = lib.dgaTilingCacheSet(ctypes.byref(plan))
        return rc, (dga.select_kernel(m, n, k, use_cache=True) if rc == 0 else None)

    rc, got = try_plan(512, 4096, 7168, kernelSerial=0, ctaGroup=2, n1=128, clusterShape=4)
    assert rc == 0 and got["clusterShape"] == 4 and got["blockDim"] == 128          # 64 pair tiles = 32 clusters of 4
    path = str(tmp_path / "c4.csv").encode()
    assert lib.dgaTilingCacheSaveCsv(path) == 0
    assert open(path.decode()).read().splitlines()[0].endswith(",layoutTagB,clusterShape,origin")
    rc, got = try_plan(4096, 7168, 2048, kernelSerial=0, ctaGroup=2, n1=256, clusterShape=4)
    assert rc == 0 and got["blockDim"] == 33 * 4 and got["kernelSerial"] == 0      # 224 column groups on 33 clusters
    assert try_plan(4096, 7168, 2048, kernelSerial=0, ctaGroup=1, n1=256, clusterShape=4)[0] != 0   # needs CTA pairs
    assert try_plan(512, 384, 7168, kernelSerial=0, ctaGroup=2, n1=128, clusterShape=4)[0] != 0     # 3 N tiles: odd
    assert try_plan(2048, 4096, 7168, kernelSerial=4, ctaGroup=2, n1=256, splitkFactor=2, clusterShape=4)[0] != 0
    lib.dgaTilingCacheClear()
    assert lib.dgaTilingCacheLoadCsv(path) == 0
    assert dga.select_kernel(512, 4096, 7168, use_cache=True)["clusterShape"] == 4
    lib.dgaTilingCacheClear()


def test_cache_first_writer_wins_and_csv_round_trip(dga, tmp_path):
    lib = dga._lib.lib
    lib.dgaTilingCacheClear()
    assert lib.dgaTilingCacheSize() == 0
    p1 = dga.select_kernel(2048, 4096, 7168, use_cache=True)
    assert lib.dgaTilingCacheSize() == 1
    p2 = dga.select_kernel(2048, 4096, 7168, use_cache=True)   # select_kernel_test.cpp:94-110: a hit returns the same tile
    assert p1 == p2
    # an explicit SetTiling replaces what the analytic selector inserted by itself on a miss (a tuned table loaded
    # after the first call for a shape must not be ignored) ...
    plan = dga._lib.TilingPlan()
    for key, val in p1.items():
        setattr(plan, key, val)
    first = 128 if p1["n1"] != 128 else 64
    plan.n1 = first
    plan.stages = 0
    assert lib.dgaTilingCacheSet(ctypes.byref(plan)) == 0
    assert dga.select_kernel(2048, 4096, 7168, use_cache=True)["n1"] == first
    # ... and cache_test.cpp:63-85: a second SetTiling for the same key does not replace the first
    plan.n1 = 96
    plan.stages = 0
    assert lib.dgaTilingCacheSet(ctypes.byref(plan)) == 0
    assert dga.select_kernel(2048, 4096, 7168, use_cache=True)["n1"] == first
    p1 = dga.select_kernel(2048, 4096, 7168, use_cache=True)
    dga.select_kernel(64, 4096, 7168, use_cache=True)
    # degenerate shapes are stored under the caller's layout (the selector canonicalises n == 1 to the K-major reading)
    size = lib.dgaTilingCacheSize()
    dga.select_kernel(256, 1, 512, layout_b=0, use_cache=True)
    dga.select_kernel(256, 1, 512, layout_b=0, use_cache=True)
    assert lib.dgaTilingCacheSize() == size + 1
    got = dga._lib.TilingPlan()
    assert lib.dgaTilingCacheGet(256, 1, 512, dga._lib.ACL_BF16, dga._lib.ACL_BF16, 0, ctypes.byref(got)) == 0
    path = str(tmp_path / "tiling.csv").encode()
    assert lib.dgaTilingCacheSaveCsv(path) == 0
    text = open(path.decode()).read().splitlines()
    # the reference's 11 columns come first (csv.cpp:23-26, csv_test.cpp:31-37)
    assert text[0].startswith("m,n,k,m1,n1,k1,kernelSerial,paddingTagA,paddingTagB,paddingTagC,blockDim")
    assert text[0].endswith(",clusterShape,origin")
    assert len(text) == 4
    assert sorted(row.rsplit(",", 1)[1] for row in text[1:]) == ["0", "0", "1"]   # two analytic rows, one explicit
    lib.dgaTilingCacheClear()
    # a table loaded AFTER the analytic selector has answered for a shape still takes effect
    assert dga.select_kernel(2048, 4096, 7168, use_cache=True)["n1"] != first
    assert lib.dgaTilingCacheLoadCsv(path) == 0
    assert lib.dgaTilingCacheSize() == 3
    back = dga.select_kernel(2048, 4096, 7168, use_cache=True)
    back.pop("estTimeUs"); want = dict(p1); want.pop("estTimeUs")   # the estimate is not persisted
    assert back == want
    lib.dgaTilingCacheClear()


def test_bench_config_matches_reference_golden(dga, golden_dir):
    with open(os.path.join(golden_dir, "bench_config.json")) as f:
        rows = json.load(f)
    for row in rows:
        m, n, k = row["shape"]
        assert dga.get_bench_config(m, n, k, row["params"][:6]) == row["params"]
        assert orc.get_bench_config(m, n, k, *row["params"][:6]) == row["params"]


def test_nshard_ranges_cover_n(dga):
    for n in (57344, 7168, 4096, 1000, 255):
        for g in (1, 2, 4, 8):
            got, end = [], 0
            for r in range(g):
                b, c = dga.nshard_range(n, g, r)
                assert b == end
                end = b + c
                got.append(c)
            assert end == n
            assert all(c % 256 == 0 for c in got[:-1]) or n < 256 * g
    assert dga.nshard_range(57344, 8, 3) == (3 * 7168, 7168)
    with pytest.raises(dga.DgaError):
        dga.nshard_range(1024, 2, 2)
