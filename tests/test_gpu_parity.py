"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs and against the
committed golden vectors of the reference. Tolerances (BASELINE.json north_star):
  * 16-bit outputs: max element-wise relative error vs the fp64 oracle <= 1e-2 (positive reference distributions);
  * fp32 outputs (framework path): the reference's own comparator np.isclose(rtol, atol=1e-9) with <= 1e-4 mismatching
    elements at rtol 2e-4 (framework/tests/test.py:19-21); the tighter rtol 1e-6/1.5e-6 verdicts are reported, and
    the result must stay within 1e-5 max relative error of fp64 (tensor-core accumulation order differs from numpy's).
"""
import os

import numpy as np
import pytest
import torch

from oracle import matmul_oracle as orc

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _to_dev(x, dtype):
    return torch.from_numpy(np.ascontiguousarray(x)).to(DEV).to(dtype)


def _inputs(m, n, k, dtype, seed, dist="uniform"):
    """Reference input distributions (gen_golden.py:11-12, test.py:30-32), rounded to `dtype` so that the oracle and
    the device see exactly the same values."""
    if dist == "uniform":
        rng = np.random.RandomState(seed)
        a = rng.uniform(1, 10, [m, k]).astype(np.float32)
        b = rng.uniform(1, 10, [k, n]).astype(np.float32)
    else:
        rng = np.random.default_rng(seed)
        a, b = orc.heavy_tail(rng, [m, k]), orc.heavy_tail(rng, [k, n])
    if dtype == torch.float16:
        a, b = a.astype(np.float16).astype(np.float32), b.astype(np.float16).astype(np.float32)
    else:
        a, b = orc.round_to_bf16(a), orc.round_to_bf16(b)
    return a, b


def _run_aclnn(dga, a, b, dtype, layout_b, **plan):
    """out = self @ mat2 through the two-phase C ABI. layout_b 1: weight stored [N,K], mat2 = W.t() (reference layout)."""
    ta = _to_dev(a, dtype)
    if layout_b == 1:
        w = _to_dev(b.T, dtype)   # [N, K]
        mat2 = w.t()
    else:
        mat2 = _to_dev(b, dtype)
    out = torch.empty((a.shape[0], b.shape[1]), dtype=dtype, device=DEV)
    before = dga.kernel_launch_count()
    if plan:
        dga.matmul_with_plan(ta, mat2, out, **plan)
    else:
        dga.catlass_dynamic_matmul(ta, mat2, out=out)
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() > before, "no kernel of this library was launched"
    return out.float().cpu().numpy()


ACLNN_CASES = [
    # (m, n, k, dtype, layout_b, plan)  -- BASELINE config 1 first
    (1024, 1024, 1024, torch.float16, 1, {}),
    (1024, 1024, 1024, torch.float16, 0, {}),
    (128, 128, 64, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 1, "n1": 128}),
    (256, 256, 512, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 256}),
    (300, 520, 136, torch.bfloat16, 1, {}),                  # ragged M, N and K tiles
    (1000, 1000, 1000, torch.bfloat16, 1, {}),
    (1000, 1000, 1000, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 224}),
    (512, 1024, 512, torch.float16, 0, {"kernelSerial": 0, "ctaGroup": 1, "n1": 128}),
    # tile widths that are an odd multiple of 16: the last 32-column store is moved back by 16 columns
    (2048, 4096, 2048, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 240}),
    (1000, 1000, 1000, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 208}),
    (300, 520, 1992, torch.float16, 1, {"kernelSerial": 0, "ctaGroup": 1, "n1": 240}),
    (257, 4088, 640, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 48}),
    (127, 509, 1027, torch.bfloat16, 1, {"kernelSerial": 2, "ctaGroup": 1, "n1": 80}),   # + repack of A, B, C
    (1, 512, 128, torch.bfloat16, 1, {}),                    # the reference's own test shape (test.py:24-26)
    (1, 4096, 7168, torch.bfloat16, 1, {}),
    (16, 4096, 7168, torch.bfloat16, 1, {}),
    (64, 4096, 7168, torch.bfloat16, 1, {}),
    (128, 4096, 7168, torch.bfloat16, 1, {}),
    (100, 1000, 1000, torch.bfloat16, 1, {"kernelSerial": 3, "splitkFactor": 4}),
    (16, 256, 512, torch.bfloat16, 1, {"kernelSerial": 3, "splitkFactor": 2}),
    # cluster sizes that are not powers of two; rows not divisible by the split; ragged N and K
    (50, 1000, 1992, torch.bfloat16, 1, {"kernelSerial": 3, "splitkFactor": 3}),
    (64, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 3, "splitkFactor": 5}),
    (7, 520, 1992, torch.float16, 1, {"kernelSerial": 3, "splitkFactor": 7}),
    (16, 584, 2016, torch.float16, 1, {}),                   # custom_catlass demo shape 16x578x2014, 16 B-aligned
    (512, 4096, 7168, torch.bfloat16, 1, {}),
    (256, 4096, 7168, torch.bfloat16, 1, {}),                # the selector's stream-K window (one 256-row block)
    (192, 4096, 2048, torch.bfloat16, 1, {}),
    (200, 4000, 4096, torch.bfloat16, 0, {}),
    (16, 2112, 7168, torch.bfloat16, 0, {}),                 # row-major mat2, few rows, long K: K slices + reduce kernel
    (128, 4096, 7168, torch.bfloat16, 0, {}),
    (512, 4096, 7168, torch.bfloat16, 0, {}),                # row-major mat2: stream-K on 256 x 256 tiles
    (2048, 4096, 2048, torch.bfloat16, 1, {}),
    # stream-K variant (role of PaddingStreamkMatmulKernel): K blocks of the ragged wave dealt out to every cluster
    (512, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 2, "n1": 256}),
    (128, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 1, "n1": 128}),
    (1000, 1000, 1000, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 2, "n1": 128}),
    (300, 520, 1992, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 1, "n1": 64}),
    (2048, 4096, 2048, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 2, "n1": 256}),
    (640, 1024, 1024, torch.float16, 0, {"kernelSerial": 4, "ctaGroup": 2, "n1": 128}),
    # hand-set slice counts (splitkFactor of a stream-K plan), incl. one that does not divide the K blocks
    (512, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 2, "n1": 256, "splitkFactor": 2}),
    (2048, 4096, 2048, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 2, "n1": 256, "splitkFactor": 4}),
    (1000, 1000, 1000, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 1, "n1": 128, "splitkFactor": 3}),
    (384, 512, 4096, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 2, "n1": 64, "splitkFactor": 7}),
    # multicore split-K + reduce kernel (the reference's declared kernel type 3 + StreamkReduceAdd)
    (512, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 5, "ctaGroup": 2, "n1": 256, "splitkFactor": 2}),
    (128, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 5, "ctaGroup": 1, "n1": 256, "splitkFactor": 8}),
    (256, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 5, "ctaGroup": 2, "n1": 256, "splitkFactor": 4}),
    (300, 520, 1992, torch.bfloat16, 1, {"kernelSerial": 5, "ctaGroup": 1, "n1": 64, "splitkFactor": 3}),
    (1000, 1000, 1000, torch.bfloat16, 1, {"kernelSerial": 5, "ctaGroup": 2, "n1": 128, "splitkFactor": 5}),
    (640, 1024, 1024, torch.float16, 0, {"kernelSerial": 5, "ctaGroup": 2, "n1": 128, "splitkFactor": 2}),
    (127, 509, 1027, torch.bfloat16, 1, {"kernelSerial": 5, "ctaGroup": 1, "n1": 128, "splitkFactor": 4}),   # + repack of A, B
    # clusters of four CTAs: two pairs on neighbouring N tiles, A rows fetched once and multicast (clusterShape 4)
    (512, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 128, "clusterShape": 4}),
    (256, 512, 512, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 256, "clusterShape": 4}),
    (2048, 4096, 2048, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 256, "clusterShape": 4}),  # several tiles per cluster
    (1000, 1000, 1000, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 128, "clusterShape": 4}),  # ragged M, N, K
    (300, 900, 1992, torch.float16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 240, "clusterShape": 4}),     # odd-16 width
    (640, 1024, 1024, torch.float16, 0, {"kernelSerial": 0, "ctaGroup": 2, "n1": 128, "clusterShape": 4}),    # row-major mat2
    (127, 509, 1027, torch.bfloat16, 1, {"kernelSerial": 2, "ctaGroup": 2, "n1": 64, "clusterShape": 4}),     # + repack of A, B, C
    # cluster split-K (kernelSerial 6): two CTA pairs on one tile, one half of K each, fp32 halves exchanged through DSMEM
    (512, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 6, "ctaGroup": 2, "n1": 256}),
    (256, 2048, 2048, torch.bfloat16, 1, {"kernelSerial": 6, "ctaGroup": 2, "n1": 128}),
    (300, 1000, 1992, torch.float16, 1, {"kernelSerial": 6, "ctaGroup": 2, "n1": 256}),     # ragged M, N and K
    (256, 512, 192, torch.bfloat16, 1, {"kernelSerial": 6, "ctaGroup": 2, "n1": 256}),      # 3 K blocks: halves of 1 and 2
    (512, 1024, 1024, torch.float16, 0, {"kernelSerial": 6, "ctaGroup": 2, "n1": 128}),     # row-major mat2
    (127, 509, 1027, torch.bfloat16, 1, {"kernelSerial": 6, "ctaGroup": 2, "n1": 128}),     # + repack of A, B, C
    # decode through the workspace (kernelSerial 7): no cluster, fp32 K-slice slabs + the reduce kernel
    (64, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 7, "splitkFactor": 8}),
    (128, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 7, "splitkFactor": 9}),           # slices of 12 and 13 K blocks
    (1, 4096, 7168, torch.bfloat16, 1, {"kernelSerial": 7, "splitkFactor": 8}),
    (16, 256, 512, torch.bfloat16, 1, {"kernelSerial": 7, "splitkFactor": 2}),
    (50, 1000, 1992, torch.bfloat16, 1, {"kernelSerial": 7, "splitkFactor": 3}),            # ragged N and K, rows not a multiple of 16
    (7, 520, 1992, torch.float16, 1, {"kernelSerial": 7, "splitkFactor": 7}),
    (100, 1000, 1000, torch.bfloat16, 1, {"kernelSerial": 7, "splitkFactor": 1}),
    (200, 2048, 4096, torch.bfloat16, 1, {"kernelSerial": 7, "splitkFactor": 4}),
    (33, 509, 1027, torch.bfloat16, 1, {"kernelSerial": 7, "splitkFactor": 4}),             # + repack of A and B
]


@pytest.mark.parametrize("m,n,k,dtype,layout_b,plan", ACLNN_CASES)
def test_aclnn_matches_fp64_oracle(dga, m, n, k, dtype, layout_b, plan):
    if dtype == torch.float16 and k > 2048:
        pytest.skip("fp16 output overflows for U[1,10] inputs beyond K ~ 2100")
    a, b = _inputs(m, n, k, dtype, seed=m + n + k)
    got = _run_aclnn(dga, a, b, dtype, layout_b, **plan)
    ref = orc.golden_fp64(a, b)
    err = orc.max_rel_err(got, ref)
    assert err <= orc.MAX_REL_16BIT, "max rel err %.3e" % err
    # and it must be as good as rounding the exact result: <= 1 ulp of the output type
    ulp = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
    assert err <= ulp


def test_config2_sampled_rows(dga):
    """BASELINE config 2 at full size: bf16 4096 x 7168 x 2048, checked on sampled rows against fp64."""
    m, n, k = 4096, 7168, 2048
    a, b = _inputs(m, n, k, torch.bfloat16, seed=2)
    got = _run_aclnn(dga, a, b, torch.bfloat16, 1)
    rows = np.random.RandomState(0).choice(m, 64, replace=False)
    ref = orc.golden_fp64(a[rows], b)
    assert orc.max_rel_err(got[rows], ref) <= 2.0 ** -7
    # linearity / checksum over the whole output: sum_j C[i,j] == A[i,:] . (B 1)
    row_sum = a.astype(np.float64) @ b.astype(np.float64).sum(axis=1)
    assert np.max(np.abs(got.astype(np.float64).sum(axis=1) - row_sum) / row_sum) < 2e-4


def _sampled_full_size(dga, m, n, k, seed):
    """Full-size case generated on the device (the reference's U[1,10] distribution), checked on sampled rows against
    the fp64 oracle and over the whole output through the row-sum identity sum_j C[i,j] == A[i,:] . (B 1)."""
    gen = torch.Generator(device=DEV).manual_seed(seed)
    a = (torch.rand((m, k), device=DEV, generator=gen) * 9 + 1).to(torch.bfloat16)
    w = (torch.rand((n, k), device=DEV, generator=gen) * 9 + 1).to(torch.bfloat16)   # weight stored [N,K]
    out = torch.empty((m, n), dtype=torch.bfloat16, device=DEV)
    before = dga.kernel_launch_count()
    dga.catlass_dynamic_matmul(a, w.t(), out=out)
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() > before
    rows = np.random.RandomState(seed).choice(m, min(m, 48), replace=False)
    rows_t = torch.from_numpy(rows).to(DEV)
    w_host = w.float().cpu().numpy()
    ref = orc.golden_fp64(a[rows_t].float().cpu().numpy(), w_host.T)
    assert orc.max_rel_err(out[rows_t].float().cpu().numpy(), ref) <= 2.0 ** -7
    w_sum = w_host.astype(np.float64).sum(axis=0)                      # B 1, exact in fp64
    row_sum = a.double().cpu().numpy() @ w_sum
    got_sum = out.double().sum(dim=1).cpu().numpy()
    assert np.max(np.abs(got_sum - row_sum) / row_sum) < 2e-4


@pytest.mark.parametrize("m", [512, 2048, 8192])
def test_config3_sweep_full_size(dga, m):
    """BASELINE config 3, the prefill end of the dynamic-M sweep at N=4096, K=7168 (decode points are in ACLNN_CASES)."""
    _sampled_full_size(dga, m, 4096, 7168, seed=m)


@pytest.mark.parametrize("m,n,k,serial", [(4096, 4096, 7168, 4),      # ragged last wave, long K -> stream-K, 2 slices
                                          (2048, 7168, 16384, 4),
                                          (128, 7168, 2048, 1),       # wide weight, short K -> single-CTA tiles, not decode
                                          (128, 24576, 1536, 1)])
def test_selector_audit_shapes_full_size(dga, m, n, k, serial):
    """Shapes whose plan comes from the selector audit (profiles/r01c_selector_audit.log), through the default path."""
    assert dga.select_kernel(m, n, k)["kernelSerial"] == serial
    _sampled_full_size(dga, m, n, k, seed=m + n)


def test_config5_one_shard_full_size(dga):
    """BASELINE config 5 as one rank of eight sees it: M=16384, N=57344/8, K=7168."""
    _sampled_full_size(dga, 16384, 7168, 7168, seed=5)


def test_config5_full_width_one_gpu(dga):
    """BASELINE config 5 whole (16384 x 57344 x 7168) on one GPU: the 1-GPU anchor of the strong-scaling curve."""
    _sampled_full_size(dga, 16384, 57344, 7168, seed=55)


def test_golden_vectors_framework_path(dga, golden_dir):
    """run_mmad_bench (fp16 -> fp32) on the reference's own generated vectors (gen_golden.py run with a seed)."""
    z = np.load(os.path.join(golden_dir, "numeric_uniform.npz"))
    tags = sorted({k.rsplit("_", 1)[0] for k in z.files})
    for tag in tags:
        x1, x2, golden = z[tag + "_x1"], z[tag + "_x2"], z[tag + "_golden"]
        m, k = x1.shape
        n = x2.shape[1]
        x = torch.from_numpy(x1).to(DEV)
        y = torch.from_numpy(x2).to(DEV)
        out = torch.zeros((m, n), dtype=torch.float32, device=DEV)
        params = torch.tensor([1, 1, 4, 4, 8, 4] + [0] * 22, dtype=torch.int32, device=DEV)
        dga.run_mmad_bench(x, y, out, params)   # blocks, like the reference
        got = out.cpu().numpy()
        ok, ratio = orc.verify_result(got, golden, rtol=orc.RTOL_TEST)
        assert ok, (tag, ratio)
        assert orc.max_rel_err(got, orc.golden_fp64(x1, x2)) < 1e-5
        want = orc.get_bench_config(m, n, k, 1, 1, 4, 4, 8, 4)
        assert params.cpu().tolist() == want   # params[6:28] written back (gemm_bench.hpp:79-81)


def test_golden_vectors_rtc_path(dga, golden_dir):
    """run_mmad_rtc (bf16 -> fp32, batched) on the reference's test.py vectors (M=1, N=512, K=128)."""
    z = np.load(os.path.join(golden_dir, "numeric_heavy.npz"))
    for seed in (0, 1):
        x1, x2 = z["s%d_x1" % seed], z["s%d_x2" % seed]
        # test.py:70-75 casts the fp32 host data to bf16 on the device; the golden stays fp32-of-fp32
        x = torch.from_numpy(x1).to(DEV).to(torch.bfloat16).unsqueeze(0).contiguous()
        y = torch.from_numpy(x2).to(DEV).to(torch.bfloat16).unsqueeze(0).contiguous()
        out = torch.zeros((1, 1, 512), dtype=torch.float32, device=DEV)
        dga.run_mmad_rtc(x, y, out)
        got = out.cpu().numpy().reshape(1, 512)
        exact = orc.golden_fp64(x.float().cpu().numpy()[0], y.float().cpu().numpy()[0])
        ok, ratio = orc.verify_result(got, exact.astype(np.float32), rtol=orc.RTOL_TEST)
        assert ok, ratio
        assert orc.max_rel_err(got, exact) < 1e-5


def test_rtc_batched(dga):
    rng = np.random.default_rng(3)
    b, m, n, k = 3, 48, 128, 320
    x = torch.from_numpy(orc.heavy_tail(rng, [b, m, k])).to(DEV).to(torch.bfloat16)
    y = torch.from_numpy(orc.heavy_tail(rng, [b, k, n])).to(DEV).to(torch.bfloat16)
    out = torch.zeros((b, m, n), dtype=torch.float32, device=DEV)
    dga.run_mmad_rtc(x, y, out)
    for i in range(b):
        ref = orc.golden_fp64(x[i].float().cpu().numpy(), y[i].float().cpu().numpy())
        assert orc.max_rel_err(out[i].cpu().numpy(), ref) < 1e-5


def test_stream_k_tolerates_a_dirty_workspace_and_graph_replay(dga):
    """The selector's stream-K plan keeps "partial is ready" flags in the caller's workspace; whatever bytes the
    workspace held before (all ones, random, the leftovers of a previous run) the result is the same, and a captured
    launch replays correctly (the flags are written back to zero)."""
    m, n, k = 256, 4096, 2048
    assert dga.select_kernel(m, n, k)["kernelSerial"] == 4
    a, b = _inputs(m, n, k, torch.bfloat16, seed=21)
    ta, w = _to_dev(a, torch.bfloat16), _to_dev(b.T, torch.bfloat16)
    ref = orc.golden_fp64(a, b)
    out = torch.zeros((m, n), dtype=torch.bfloat16, device=DEV)
    ws_bytes, ex = dga.get_workspace_size(ta, w.t(), out)
    dga.destroy_executor(ex)
    assert ws_bytes > 0
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)
    gen = torch.Generator(device=DEV).manual_seed(1)
    results = []
    for fill in ("ones", "random", "leftover", "leftover"):
        if fill == "ones":
            ws.fill_(255)
        elif fill == "random":
            ws.copy_(torch.randint(0, 256, (ws_bytes,), dtype=torch.uint8, device=DEV, generator=gen))
        out.zero_()
        _, ex = dga.get_workspace_size(ta, w.t(), out)
        dga.execute(ws, ws_bytes, ex)
        torch.cuda.synchronize()
        results.append(out.float().cpu().numpy())
        assert orc.max_rel_err(results[-1], ref) <= 1e-2
    assert all(np.array_equal(results[0], r) for r in results[1:])   # fixed summation order: bit-identical
    exe = dga.Executor(ta, w.t(), out)
    stream = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(stream):
        exe(stream)
        stream.synchronize()
        with torch.cuda.graph(graph, stream=stream):
            exe(stream)
    for _ in range(3):
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert np.array_equal(out.float().cpu().numpy(), results[0])
    exe.close()


def test_side_stream_launch_keeps_its_workspace(dga):
    """The one-shot operator call allocates its workspace itself. Launched on a stream that is not torch's current
    one, the block must stay reserved until that stream is done with it: here the current stream keeps allocating
    and overwriting same-sized blocks while a queue of split-K + reduce launches (fp32 slabs in the workspace) is
    still pending on the side stream."""
    m, n, k = 256, 2048, 7168
    plan = dga.select_kernel(m, n, k)
    assert plan["kernelSerial"] == 5 and plan["workspaceBytes"] > 0
    a, b = _inputs(m, n, k, torch.bfloat16, seed=31)
    ta, w = _to_dev(a, torch.bfloat16), _to_dev(b.T, torch.bfloat16)
    ref = orc.golden_fp64(a, b)
    side = torch.cuda.Stream()
    outs = [torch.zeros((m, n), dtype=torch.bfloat16, device=DEV) for _ in range(12)]
    blocker = torch.randn((8192, 8192), device=DEV)
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        for _ in range(6):
            blocker = blocker @ blocker * 1e-4          # keeps the side stream busy: the GEMMs below queue up
    for out in outs:
        dga.catlass_dynamic_matmul(ta, w.t(), out=out, stream=side)
        junk = torch.empty(plan["workspaceBytes"], dtype=torch.uint8, device=DEV)   # current stream: same size class
        junk.fill_(0x7F)
        del junk
    torch.cuda.synchronize()
    for out in outs:
        assert orc.max_rel_err(out.float().cpu().numpy(), ref) <= 2.0 ** -7


def test_concurrent_stream_k_launches_do_not_wait_on_each_other(dga):
    """Stream-K CTAs spin on flags that other CTAs of the SAME launch set, so every CTA of a launch has to be on the
    device at once. The launch is cooperative: two stream-K launches queued on two streams, each with
    its own workspace, run one after the other instead of interleaving half-resident grids that would wait on each
    other until the in-kernel deadline traps. Both must finish and match the oracle."""
    m, n, k = 256, 4096, 2048
    assert dga.select_kernel(m, n, k)["kernelSerial"] == 4
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    jobs = []
    for i, s in enumerate(streams):
        a, b = _inputs(m, n, k, torch.bfloat16, seed=41 + i)
        ta, w = _to_dev(a, torch.bfloat16), _to_dev(b.T, torch.bfloat16)
        out = torch.zeros((m, n), dtype=torch.bfloat16, device=DEV)
        jobs.append((dga.Executor(ta, w.t(), out), out, orc.golden_fp64(a, b), s))
    torch.cuda.synchronize()
    for _ in range(50):
        for exe, _, _, s in jobs:
            exe(s)
    torch.cuda.synchronize()
    for exe, out, ref, _ in jobs:
        assert orc.max_rel_err(out.float().cpu().numpy(), ref) <= 1e-2
        exe.close()


# (m, n_first, n_second, k, dtype): decode pairs; ragged N on either side, one-tile weights, a K too short for the
# selector's decode rule (128 x 8192 x 2048 goes to the persistent kernel: the pair entry forces its own split).
# fp16 cases keep K short: an fp16 output of U(1, 10) inputs overflows beyond K ~ 2000.
PAIR_CASES = [(1, 4096, 4096, 7168, torch.bfloat16), (16, 2048, 512, 7168, torch.bfloat16),
              (64, 4096, 4096, 7168, torch.bfloat16), (64, 1536, 576, 1536, torch.float16),
              (128, 1000, 520, 1992, torch.bfloat16), (50, 136, 2050, 1024, torch.float16),
              (128, 4096, 4096, 2048, torch.bfloat16), (7, 128, 128, 128, torch.bfloat16)]


@pytest.mark.parametrize("m,n_first,n_second,k,dtype", PAIR_CASES)
def test_decode_pair_matches_fp64_oracle(dga, m, n_first, n_second, k, dtype):
    """dgaDecodePairMatmul: two weights, one launch; both outputs against the oracle, and against the ordinary entry
    at the same tolerance."""
    a, b1 = _inputs(m, n_first, k, dtype, seed=5)
    _, b2 = _inputs(m, n_second, k, dtype, seed=6)
    ta = _to_dev(a, dtype)
    w1, w2 = _to_dev(b1.T, dtype), _to_dev(b2.T, dtype)
    o1 = torch.full((m, n_first), float("nan"), dtype=dtype, device=DEV)
    o2 = torch.full((m, n_second), float("nan"), dtype=dtype, device=DEV)
    pair = dga.DecodePair(ta, w1.t(), o1, w2.t(), o2)
    before = dga.kernel_launch_count()
    pair()
    pair()   # repeatable
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() - before == 2   # one kernel per pair
    for out, b, w in ((o1, b1, w1), (o2, b2, w2)):
        got = out.float().cpu().numpy()
        assert np.isfinite(got).all()
        assert orc.max_rel_err(got, orc.golden_fp64(a, b)) <= 1e-2
        single = dga.catlass_dynamic_matmul(ta, w.t()).float().cpu().numpy()
        assert orc.max_rel_err(got, single.astype(np.float64)) <= 1e-2


def test_decode_pair_rejects_what_it_does_not_serve(dga):
    ta = torch.ones((256, 512), dtype=torch.bfloat16, device=DEV)
    w = torch.ones((256, 512), dtype=torch.bfloat16, device=DEV)
    with pytest.raises(dga.DgaError):    # m > 128
        dga.decode_pair_matmul(ta, w.t(), w.t())
    with pytest.raises(dga.DgaError):    # row-major [K, N] weight
        dga.decode_pair_matmul(ta[:64], w.t(), torch.ones((512, 256), dtype=torch.bfloat16, device=DEV))
    with pytest.raises(dga.DgaError):    # K extents differ
        dga.decode_pair_matmul(ta[:64], w.t(), torch.ones((256, 256), dtype=torch.bfloat16, device=DEV).t())


def test_executor_repeatable_and_graph_capturable(dga):
    a, b = _inputs(256, 512, 256, torch.bfloat16, seed=9)
    ta, w = _to_dev(a, torch.bfloat16), _to_dev(b.T, torch.bfloat16)
    out = torch.zeros((256, 512), dtype=torch.bfloat16, device=DEV)
    ex = dga.Executor(ta, w.t(), out)
    ex()
    torch.cuda.synchronize()
    first = out.clone()
    out.zero_()
    stream = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(stream):
        ex(stream)
        stream.synchronize()
        with torch.cuda.graph(graph, stream=stream):
            ex(stream)
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, first)
    ex.close()


def test_strided_views_are_not_copied(dga):
    """IgnoreContiguous (op_host/catlass_dynamic_matmul.cpp:59,65): leading dimensions larger than the row."""
    a, b = _inputs(200, 384, 192, torch.bfloat16, seed=4)
    big_a = torch.zeros((200, 256), dtype=torch.bfloat16, device=DEV)
    big_a[:, :192] = _to_dev(a, torch.bfloat16)
    big_w = torch.zeros((384, 320), dtype=torch.bfloat16, device=DEV)
    big_w[:, :192] = _to_dev(b.T, torch.bfloat16)
    big_c = torch.zeros((200, 512), dtype=torch.bfloat16, device=DEV)
    dga.catlass_dynamic_matmul(big_a[:, :192], big_w[:, :192].t(), out=big_c[:, :384])
    torch.cuda.synchronize()
    got = big_c[:, :384].float().cpu().numpy()
    assert orc.max_rel_err(got, orc.golden_fp64(a, b)) <= 2.0 ** -7
    assert float(big_c[:, 384:].abs().max()) == 0.0   # nothing written outside the view


# ---- stride repack (role of PaddingCommonMatmulKernel / PADDING_ND, padding_common_matmul_kernel.h:33-107) ----------
PADDING_CASES = [
    # (m, n, k, dtype, layout_b, plan): leading dimensions that are not multiples of 16 bytes
    (16, 578, 2014, torch.float16, 1, {}),                    # the custom_catlass demo shape as it is
    (127, 509, 251, torch.bfloat16, 1, {}),
    (127, 509, 251, torch.bfloat16, 0, {}),
    (33, 77, 130, torch.float16, 0, {}),
    (300, 1001, 515, torch.bfloat16, 1, {"kernelSerial": 0, "ctaGroup": 2, "n1": 128}),
    (300, 1001, 515, torch.bfloat16, 1, {"kernelSerial": 4, "ctaGroup": 1, "n1": 64}),   # PaddingStreamk
    (1, 4097, 1027, torch.bfloat16, 1, {}),
    (1279, 5003, 1537, torch.bfloat16, 1, {}),                # benchmark.py:41-43 class of odd shapes (K cut for the oracle)
]


@pytest.mark.parametrize("m,n,k,dtype,layout_b,plan", PADDING_CASES)
def test_unaligned_leading_dimensions_take_the_repack_path(dga, m, n, k, dtype, layout_b, plan):
    if not plan:
        p = dga.select_kernel(m, n, k, dtype=dtype, layout_b=layout_b)
        assert p["paddingTagA"] or p["paddingTagB"] or p["paddingTagC"]
        assert p["workspaceBytes"] > 0
    a, b = _inputs(m, n, k, dtype, seed=m + n + k)
    got = _run_aclnn(dga, a, b, dtype, layout_b, **plan)
    err = orc.max_rel_err(got, orc.golden_fp64(a, b))
    assert err <= (2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7), "max rel err %.3e" % err


def test_misaligned_base_pointers_are_repacked(dga):
    """A view that starts 2 bytes into an allocation: aligned pitch, unaligned base."""
    m, n, k = 96, 200, 136
    a, b = _inputs(m, n, k, torch.bfloat16, seed=11)
    flat_a = torch.zeros(m * k + 8, dtype=torch.bfloat16, device=DEV)
    flat_w = torch.zeros(n * k + 8, dtype=torch.bfloat16, device=DEV)
    flat_c = torch.zeros(m * n + 8, dtype=torch.bfloat16, device=DEV)
    va = flat_a[1:1 + m * k].view(m, k); va.copy_(_to_dev(a, torch.bfloat16))
    vw = flat_w[3:3 + n * k].view(n, k); vw.copy_(_to_dev(b.T, torch.bfloat16))
    vc = flat_c[5:5 + m * n].view(m, n)
    dga.catlass_dynamic_matmul(va, vw.t(), out=vc)
    torch.cuda.synchronize()
    assert orc.max_rel_err(vc.float().cpu().numpy(), orc.golden_fp64(a, b)) <= 2.0 ** -7
    assert float(flat_c[:5].abs().max()) == 0.0 and float(flat_c[5 + m * n:].abs().max()) == 0.0


def test_framework_path_odd_shapes(dga):
    """run_mmad_bench takes any (m, n, k) (benchmark.py:41-43 sweeps odd ones): fp16 -> fp32 with odd n and k."""
    for m, n, k in ((48, 130, 321), (1, 77, 33), (257, 1003, 514)):
        x1, x2, golden = orc.gen_golden_data(m, n, k, np.random.RandomState(m + n + k))
        z = torch.zeros((m, n), dtype=torch.float32, device=DEV)
        params = torch.tensor([1, 1, 4, 4, 8, 4] + [0] * 22, dtype=torch.int32, device=DEV)
        dga.run_mmad_bench(torch.from_numpy(x1).to(DEV), torch.from_numpy(x2).to(DEV), z, params)
        ok, ratio = orc.verify_result(z.cpu().numpy(), golden, rtol=orc.RTOL_TEST)
        assert ok, ((m, n, k), ratio)
        assert orc.max_rel_err(z.cpu().numpy(), orc.golden_fp64(x1, x2)) < 1e-5


def test_framework_path_long_k_variants(dga):
    """The framework layout (row-major mat2, fp32 out) through the variants the selector gives it for long K: K slices
    + reduce kernel for few rows, stream-K for one ragged round; scratch workspace comes from the stream-ordered pool.
    Same comparator as the reference's tests plus the fp64 bound."""
    for m, n, k in ((64, 512, 4096), (16, 2112, 7168), (512, 4096, 4096)):
        x1, x2, golden = orc.gen_golden_data(m, n, k, np.random.RandomState(m + n + k))
        z = torch.zeros((m, n), dtype=torch.float32, device=DEV)
        params = torch.tensor([1, 1, 4, 4, 8, 4] + [0] * 22, dtype=torch.int32, device=DEV)
        before = dga.kernel_launch_count()
        dga.run_mmad_bench(torch.from_numpy(x1).to(DEV), torch.from_numpy(x2).to(DEV), z, params)
        assert dga.kernel_launch_count() > before
        ok, ratio = orc.verify_result(z.cpu().numpy(), golden, rtol=orc.RTOL_TEST)
        assert ok, ((m, n, k), ratio)
        # fp32 accumulation over 4096-7168 products: the bound grows with K (1e-5 holds up to K ~ 1024 above)
        err = orc.max_rel_err(z.cpu().numpy(), orc.golden_fp64(x1, x2))
        assert err < 5e-5, ((m, n, k), err)


# ---- batched framework launch and the M-grouped contiguous matmul ------------------------------------
def test_rtc_batched_single_launch(dga):
    """run_mmad_rtc with batch > 1 and aligned pitches is ONE launch (3-D tensor maps), ragged m and n per entry."""
    rng = np.random.default_rng(7)
    b, m, n, k = 5, 200, 328, 264
    x = torch.from_numpy(orc.heavy_tail(rng, [b, m, k])).to(DEV).to(torch.bfloat16)
    y = torch.from_numpy(orc.heavy_tail(rng, [b, k, n])).to(DEV).to(torch.bfloat16)
    out = torch.zeros((b, m, n), dtype=torch.float32, device=DEV)
    before = dga.kernel_launch_count()
    dga.run_mmad_rtc(x, y, out)
    assert dga.kernel_launch_count() - before == 1
    for i in range(b):
        ref = orc.golden_fp64(x[i].float().cpu().numpy(), y[i].float().cpu().numpy())
        assert orc.max_rel_err(out[i].cpu().numpy(), ref) < 1e-5


@pytest.mark.parametrize("groups,n,k,dtype", [(4, 512, 256, torch.bfloat16), (8, 1000, 520, torch.bfloat16),
                                                (3, 384, 1024, torch.float16)])
def test_m_grouped_matmul_matches_oracle(dga, groups, n, k, dtype):
    rng = np.random.RandomState(groups + n + k)
    blocks = rng.randint(0, 4, size=groups)          # 128-row blocks per group; some groups are empty
    blocks[0] = max(blocks[0], 1)
    index = []
    for g, nb in enumerate(blocks):
        index += [g] * (128 * nb)
        if g % 2 == 1:
            index += [-1] * 128                       # a padding block between groups
    index = np.array(index, dtype=np.int32)
    m_total = index.size
    a, _ = _inputs(m_total, 8, k, dtype, seed=1)
    w = np.stack([_inputs(8, n, k, dtype, seed=10 + g)[1].T for g in range(groups)])   # [G, N, K]
    ta, tw = _to_dev(a, dtype), _to_dev(w, dtype)
    out = torch.full((m_total, n), 768.0, dtype=dtype, device=DEV)
    before = dga.kernel_launch_count()
    dga.m_grouped_matmul(ta, tw, torch.from_numpy(index).to(DEV), out=out)
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() - before == 1
    got = out.float().cpu().numpy()
    ref = orc.m_grouped_golden_fp64(a, w, index)
    live = index >= 0
    assert orc.max_rel_err(got[live], ref[live]) <= (2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7)
    assert np.all(got[~live] == 768.0)               # padding blocks are left untouched


def test_m_grouped_matmul_deepseek_shape(dga):
    """MoE up-projection class shape: 8 experts x 512 rows, N=4096, K=7168, sampled rows against fp64."""
    g, per, n, k = 8, 512, 4096, 7168
    gen = torch.Generator(device=DEV).manual_seed(3)
    a = (torch.rand((g * per, k), device=DEV, generator=gen) * 9 + 1).to(torch.bfloat16)
    w = (torch.rand((g, n, k), device=DEV, generator=gen) * 9 + 1).to(torch.bfloat16)
    index = torch.arange(g, device=DEV, dtype=torch.int32).repeat_interleave(per)
    out = dga.m_grouped_matmul(a, w, index)
    torch.cuda.synchronize()
    rows = np.random.RandomState(0).choice(g * per, 24, replace=False)
    for r in rows:
        ref = orc.golden_fp64(a[r:r + 1].float().cpu().numpy(), w[r // per].float().cpu().numpy().T)
        assert orc.max_rel_err(out[r:r + 1].float().cpu().numpy(), ref) <= 2.0 ** -7


@pytest.mark.parametrize("groups,m_max,n,k,dtype", [(4, 256, 512, 256, torch.bfloat16),
                                                      (6, 384, 1000, 520, torch.bfloat16),
                                                      (3, 200, 384, 1024, torch.float16)])
def test_m_grouped_masked_matches_oracle(dga, groups, m_max, n, k, dtype):
    rng = np.random.RandomState(groups + m_max + n + k)
    masked = rng.randint(0, m_max + 1, size=groups).astype(np.int32)
    masked[0], masked[-1] = 0, m_max                    # an empty group and a full one
    a = np.stack([_inputs(m_max, 8, k, dtype, seed=40 + g)[0] for g in range(groups)])       # [G, Mmax, K]
    w = np.stack([_inputs(8, n, k, dtype, seed=50 + g)[1].T for g in range(groups)])         # [G, N, K]
    ta, tw = _to_dev(a, dtype), _to_dev(w, dtype)
    out = torch.full((groups, m_max, n), 768.0, dtype=dtype, device=DEV)
    before = dga.kernel_launch_count()
    dga.m_grouped_matmul_masked(ta, tw, torch.from_numpy(masked).to(DEV), out=out)
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() - before == 1
    got = out.float().cpu().numpy()
    ref = orc.m_grouped_masked_golden_fp64(a, w, masked)
    tol = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
    for g in range(groups):
        rows = int(masked[g])
        if rows:
            assert orc.max_rel_err(got[g, :rows], ref[g, :rows]) <= tol
        first_skipped = -(-rows // 128) * 128            # whole blocks past masked_m are left untouched
        assert np.all(got[g, first_skipped:] == 768.0)


@pytest.mark.parametrize("groups,n,k,dtype", [(4, 512, 256, torch.bfloat16), (7, 1000, 520, torch.bfloat16),
                                                (3, 384, 1024, torch.float16)])
def test_m_grouped_pair_tiles_matches_oracle(dga, groups, n, k, dtype):
    """m_alignment=256: every group occupies a multiple of 256 rows (its tail filled with padding blocks, index -1),
    the launch uses 256-row CTA-pair tiles. A pair may hold one live block and one padding block, in either order
    (a group of 128 live rows; a leading padding block), or two padding blocks: padding is never written."""
    rng = np.random.RandomState(groups * 7 + n + k)
    index = [-1] * 128 + [0] * 128                      # padding FIRST inside a pair, then group 0's only block
    for g in range(1, groups):
        live = int(rng.randint(0, 5))                    # live 128-row blocks of this group (0: the group is empty)
        slot = -(-live // 2) * 2                         # padded to whole pairs
        index += [g] * (128 * live) + [-1] * (128 * (slot - live))
        if g == 2:
            index += [-1] * 256                          # a pair of padding blocks
    index = np.array(index, dtype=np.int32)
    m_total = index.size
    assert m_total % 256 == 0
    a, _ = _inputs(m_total, 8, k, dtype, seed=2)
    w = np.stack([_inputs(8, n, k, dtype, seed=20 + g)[1].T for g in range(groups)])   # [G, N, K]
    ta, tw, ti = _to_dev(a, dtype), _to_dev(w, dtype), torch.from_numpy(index).to(DEV)
    out = torch.full((m_total, n), 768.0, dtype=dtype, device=DEV)
    before = dga.kernel_launch_count()
    dga.m_grouped_matmul(ta, tw, ti, out=out, m_alignment=256)
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() - before == 1
    got = out.float().cpu().numpy()
    ref = orc.m_grouped_golden_fp64(a, w, index)
    live = index >= 0
    assert orc.max_rel_err(got[live], ref[live]) <= (2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7)
    assert np.all(got[~live] == 768.0)                   # padding blocks are left untouched
    # the same problem through the default single-CTA tiles: identical sums (same K order per row)
    out128 = torch.full((m_total, n), 768.0, dtype=dtype, device=DEV)
    dga.m_grouped_matmul(ta, tw, ti, out=out128)
    torch.cuda.synchronize()
    assert np.array_equal(got, out128.float().cpu().numpy())
    ex = dga.MGroupedExecutor(ta, tw, ti, out, m_alignment=256)   # repeatable form
    out.fill_(768.0)
    ex(); ex()
    torch.cuda.synchronize()
    assert np.array_equal(got, out.float().cpu().numpy())
    ex.close()
    with pytest.raises(dga.DgaError):                    # 384 rows are not a multiple of 256
        dga.m_grouped_matmul(ta[:384], tw, ti[:384], m_alignment=256)


@pytest.mark.parametrize("groups,m_max,n,k,dtype", [(4, 256, 512, 256, torch.bfloat16),
                                                      (6, 512, 1000, 520, torch.bfloat16),
                                                      (3, 768, 384, 1024, torch.float16)])
def test_m_grouped_masked_pair_tiles_matches_oracle(dga, groups, m_max, n, k, dtype):
    rng = np.random.RandomState(groups + m_max + n + k)
    masked = rng.randint(0, m_max + 1, size=groups).astype(np.int32)
    masked[0], masked[-1] = 0, m_max                    # an empty group and a full one
    if groups > 2:
        masked[1] = 100                                  # only the first block of the first pair is live
    a = np.stack([_inputs(m_max, 8, k, dtype, seed=60 + g)[0] for g in range(groups)])       # [G, Mmax, K]
    w = np.stack([_inputs(8, n, k, dtype, seed=70 + g)[1].T for g in range(groups)])         # [G, N, K]
    ta, tw = _to_dev(a, dtype), _to_dev(w, dtype)
    out = torch.full((groups, m_max, n), 768.0, dtype=dtype, device=DEV)
    before = dga.kernel_launch_count()
    dga.m_grouped_matmul_masked(ta, tw, torch.from_numpy(masked).to(DEV), out=out, m_alignment=256)
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() - before == 1
    got = out.float().cpu().numpy()
    ref = orc.m_grouped_masked_golden_fp64(a, w, masked)
    tol = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
    for g in range(groups):
        rows = int(masked[g])
        if rows:
            assert orc.max_rel_err(got[g, :rows], ref[g, :rows]) <= tol
        first_skipped = -(-rows // 128) * 128            # whole blocks past masked_m are left untouched
        assert np.all(got[g, first_skipped:] == 768.0)


# ---- seeded fuzz over shapes, layouts, dtypes and plans (edge cases: tiny, ragged, odd pitches) ------------------------
def _fuzz_cases(count=48, seed=20261104):
    rng = np.random.RandomState(seed)
    cases = []
    for _ in range(count):
        m = int(rng.choice([1, 2, 7, 16, 17, 63, 64, 100, 128, 129, 255, 256, 300, 513, 640]))
        n = int(rng.choice([1, 8, 24, 64, 100, 128, 136, 250, 256, 384, 520, 777, 1024]))
        k = int(rng.choice([1, 8, 16, 40, 64, 72, 128, 200, 256, 520, 1000, 1992]))
        dtype = torch.bfloat16 if rng.rand() < 0.6 else torch.float16
        layout_b = int(rng.rand() < 0.7)
        kind = rng.randint(0, 6)
        plan = {}
        if kind == 1:
            plan = {"kernelSerial": 0, "ctaGroup": int(rng.choice([1, 2])), "n1": int(rng.choice([64, 128, 192, 256]))}
        elif kind == 2:
            plan = {"kernelSerial": 4, "ctaGroup": int(rng.choice([1, 2])), "n1": int(rng.choice([64, 128, 256]))}
        elif kind == 3 and k >= 128:
            plan = {"kernelSerial": 5, "ctaGroup": int(rng.choice([1, 2])), "n1": int(rng.choice([64, 128, 256])),
                    "splitkFactor": 2}
        elif kind == 4 and m <= 256 and layout_b == 1 and k >= 256:
            plan = {"kernelSerial": 3, "splitkFactor": int(rng.choice([1, 2, 4]))}
        if (layout_b == 0 or k == 1) and plan.get("ctaGroup") == 2 and plan.get("n1") in (64, 192):
            plan["n1"] = 128   # row-major B wants 64-column boxes per CTA (a [1,N] mat2 reads as row-major too)
        cases.append((m, n, k, dtype, layout_b, plan))
    return cases


@pytest.mark.parametrize("m,n,k,dtype,layout_b,plan", _fuzz_cases())
def test_fuzz_against_fp64_oracle(dga, m, n, k, dtype, layout_b, plan):
    if dtype == torch.float16 and k > 1000:
        dtype = torch.bfloat16
    a, b = _inputs(m, n, k, dtype, seed=3 * m + 5 * n + 7 * k)
    got = _run_aclnn(dga, a, b, dtype, layout_b, **plan)
    err = orc.max_rel_err(got, orc.golden_fp64(a, b))
    assert err <= (2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7), "max rel err %.3e" % err


# ---- second fuzz: the selector's own plan on long-K shapes, where its measured rules (stream-K windows, K slices +
# reduce kernel, decode kernel, single-CTA tiles) switch from one variant to the next ---------------------------------
def _selector_fuzz_cases(count=32, seed=4096):
    rng = np.random.RandomState(seed)
    cases = []
    for _ in range(count):
        m = int(rng.choice([1, 16, 60, 64, 65, 128, 129, 192, 250, 256, 257, 384, 500, 512, 768]))
        n = int(rng.choice([512, 576, 1000, 1024, 1536, 2048, 2112, 3000, 4096, 4104]))
        k = int(rng.choice([2048, 2304, 3584, 4096, 5000, 7168, 8192]))
        layout_b = int(rng.rand() < 0.6)
        if layout_b == 0:
            n = (n + 7) // 8 * 8      # keep the row-major pitch 16-byte aligned half of the time
            if rng.rand() < 0.3:
                n += 1                # ... and exercise the repack path on the rest
        cases.append((m, n, k, layout_b))
    return cases


@pytest.mark.parametrize("m,n,k,layout_b", _selector_fuzz_cases())
def test_selector_plan_fuzz_long_k(dga, m, n, k, layout_b):
    a, b = _inputs(m, n, k, torch.bfloat16, seed=m + 3 * n + 5 * k)
    got = _run_aclnn(dga, a, b, torch.bfloat16, layout_b)
    err = orc.max_rel_err(got, orc.golden_fp64(a, b))
    assert err <= 2.0 ** -7, "max rel err %.3e with plan %s" % (
        err, {key: val for key, val in dga.select_kernel(m, n, k, layout_b=layout_b).items()
              if key in ("kernelSerial", "ctaGroup", "n1", "splitkFactor", "paddingTagB", "paddingTagC")})
