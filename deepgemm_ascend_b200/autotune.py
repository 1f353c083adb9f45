"""Tuned-table sweep with checkpoint/resume: the B200 counterpart of the reference's tiling grid search.

Role of `GEMMBenchmarkRunner` (deep_gemm_ascend/framework/benchmark/benchmark.py:227-341) and of the CSV plan cache it
feeds (aclnn_catlass_dynamic_matmul/op_host/op_tiling/cache.cpp:28-67, csv.cpp:23-26):

  * per shape, a list of candidate plans (there: the six Ascend L1/L0 knobs, benchmark.py:49-225; here: kernel
    variant x CTA-pair mode x N tile x ring depth x K split, the knobs `DgaTilingPlan` exposes);
  * the list is cut into contiguous ranges, one per process (`rank_id`, `num_processes`; benchmark.py:247-253);
  * before a candidate runs, its global index goes to `shape_<M>_<N>_<K>_rank_<r>_checkpoint.jsonl`; a restart that
    finds its own checkpoint pointing at an unfinished candidate records it as failed (time = -1) and moves on
    (benchmark.py:255-292), so one crashing candidate cannot wedge the sweep;
  * every candidate's result is appended to `shape_<M>_<N>_<K>_rank_<r>.jsonl` with the reference's field names
    (idx, M, N, K, time, diff, parameters; benchmark.py:318-329) after a correctness check against a golden
    (benchmark.py:391-405);
  * `collect()` picks the fastest correct candidate per shape and writes the plan-cache CSV (the reference's 11
    columns first), which `$CACHE_FILE_PATH` / `dgaTilingCacheLoadCsv` feed back into phase 1.

The timing and checking functions are injectable so that the bookkeeping is testable without a GPU.
"""
import argparse
import ctypes
import json
import math
import os

from . import _lib
from ._lib import lib, check

KERNEL_COMMON, KERNEL_SPLITK, KERNEL_STREAMK, KERNEL_SPLITK_REDUCE, KERNEL_CLUSTER_SPLITK = 0, 3, 4, 5, 6
ERROR_TOLERANCE = 1e-2   # max relative error of a candidate against the fp32 golden on sampled rows


def candidate_plans(m, n, k, layout_b=1):
    """All plans worth timing for one shape, in a fixed order (the index is the checkpoint key)."""
    out = []
    for cta in (2, 1):
        step = 64 * cta if layout_b == 0 else 32
        for n1 in range(256, 31, -step):
            if n1 % (16 * cta) or n1 % 32:
                continue
            if n1 > max(64, -(-n // 32) * 32) and n1 != 64:
                continue   # tile wider than the matrix
            for stages in (0, 4):          # 0 = as deep as shared memory allows
                out.append({"kernelSerial": KERNEL_COMMON, "ctaGroup": cta, "n1": n1, "stages": stages})
            if m > 128:
                for split in (0, 2, 4):    # 0 = slice count chosen by streamk_split
                    out.append({"kernelSerial": KERNEL_STREAMK, "ctaGroup": cta, "n1": n1, "stages": 0,
                                "splitkFactor": split})
    if m > 128 and -(-m // 256) * -(-n // 256) * 4 <= 74:
        for split in (2, 4, 6, 8):   # multicore split-K + reduce kernel on 256 x 256 pair tiles
            if -(-k // 64) >= 14 * split // 2:
                out.append({"kernelSerial": KERNEL_SPLITK_REDUCE, "ctaGroup": 2, "n1": 256, "stages": 0,
                            "splitkFactor": split})
    if m <= 128 and layout_b == 0 and -(-k // 64) >= 56:
        # no swap-AB decode kernel for row-major mat2: K slices of the persistent kernel instead
        for cta in (2, 1):
            for n1 in (256, 128):
                for split in (2, 4, 8):
                    out.append({"kernelSerial": KERNEL_SPLITK_REDUCE, "ctaGroup": cta, "n1": n1, "stages": 0,
                                "splitkFactor": split})
    if m <= 256 and layout_b == 1:
        for split in (1, 2, 4, 8):
            if -(-k // 64) < split:
                continue
            for stages in (0, 3, 4, 6):
                out.append({"kernelSerial": KERNEL_SPLITK, "splitkFactor": split, "stages": stages})
    if m > 128 and layout_b == 1:
        # widths that are an odd multiple of 16 (whole-tile plans over a K-major weight): appended last so that the
        # indices of the candidates above, which are checkpoint keys, stay what they were.  Measured 2-3 % ahead of
        # the best multiple of 32 where they fill the last wave better (profiles/r01d_odd16_tiles.log), so they are
        # for the tuned table to pick, not for the analytic selector.
        for cta in (2, 1):
            for n1 in range(240, 47, -32):
                if n1 > -(-n // 16) * 16:
                    continue
                out.append({"kernelSerial": KERNEL_COMMON, "ctaGroup": cta, "n1": n1, "stages": 0})
    if m > 128:
        # round 2, again appended behind everything older: clusters of four CTAs.  (1) two pairs on neighbouring N
        # tiles with their A rows multicast (clusterShape 4): +-3 % against the pair kernel depending on how the waves
        # fall on 66 instead of 74 pairs (profiles/r02_cluster_shapes_ab_interleaved.log) - a tuned-table pick, not an
        # analytic one; (2) cluster split-K (two pairs on one tile, halves exchanged through DSMEM) outside the
        # selector's window of 24-33 tiles and K >= 6144.
        for n1 in (256, 224, 192, 128):
            if -(-n // n1) % 2 == 0 and (layout_b == 1 or n1 % 128 == 0):
                out.append({"kernelSerial": KERNEL_COMMON, "ctaGroup": 2, "n1": n1, "stages": 0, "clusterShape": 4})
        for n1 in (256, 128):
            if -(-m // 256) * -(-n // n1) <= 33 and -(-k // 64) >= 2:
                out.append({"kernelSerial": KERNEL_CLUSTER_SPLITK, "ctaGroup": 2, "n1": n1, "stages": 0})
    return out


def process_range(total, rank_id, num_processes):
    """Contiguous slice of candidate indices owned by `rank_id` (benchmark.py:247-253)."""
    per = math.ceil(total / num_processes) if total else 0
    start = rank_id * per
    return start, min(start + per, total)


class AutotuneRunner:
    def __init__(self, shape_group, rank_id=0, num_processes=1, result_dir="./results", layout_b=1,
                 time_fn=None, check_fn=None):
        self.shape_group = [tuple(s) for s in shape_group]
        self.rank_id, self.num_processes = rank_id, num_processes
        self.result_dir = result_dir
        self.layout_b = layout_b
        self.time_fn = time_fn if time_fn is not None else _GpuBench().time
        self.check_fn = check_fn if check_fn is not None else _GpuBench().diff
        os.makedirs(result_dir, exist_ok=True)

    def _paths(self, shape):
        stem = "shape_%s_rank_%d" % ("_".join(map(str, shape)), self.rank_id)
        return (os.path.join(self.result_dir, stem + ".jsonl"),
                os.path.join(self.result_dir, stem + "_checkpoint.jsonl"))

    @staticmethod
    def _append(path, record):
        with open(path, "a", encoding="utf-8") as f:
            f.write(json.dumps(record) + "\n")

    def benchmark_shape(self, shape):
        """Time this rank's share of the candidates of one shape; resumable."""
        result_path, checkpoint_path = self._paths(shape)
        m, n, k = shape
        cands = candidate_plans(m, n, k, self.layout_b)
        start_idx, end_idx = process_range(len(cands), self.rank_id, self.num_processes)
        last_idx = -1
        if os.path.exists(checkpoint_path):
            with open(checkpoint_path, encoding="utf-8") as f:
                last_idx = json.load(f).get("last_process_idx", -1)
        done = set()
        if os.path.exists(result_path):
            with open(result_path, encoding="utf-8") as f:
                done = {json.loads(line)["idx"] for line in f if line.strip()}
        for idx in range(start_idx, end_idx):
            if idx in done:
                continue
            params = cands[idx]
            if idx == last_idx:
                # the previous run died inside this candidate: record it as failed and skip (benchmark.py:280-292)
                self._append(result_path, {"idx": idx, "M": m, "N": n, "K": k, "time": -1, "diff": -1,
                                           "parameters": params})
                continue
            with open(checkpoint_path, "w", encoding="utf-8") as f:
                json.dump({"last_process_idx": idx}, f)
            try:
                diff = float(self.check_fn(shape, params, self.layout_b))
                time_us = float(self.time_fn(shape, params, self.layout_b)) if diff <= ERROR_TOLERANCE \
                    else float("inf")
            except _lib.DgaError:
                diff, time_us = -1.0, float("inf")   # infeasible plan (tile does not fit, split > K blocks, ...)
            self._append(result_path, {"idx": idx, "M": m, "N": n, "K": k, "time": time_us, "diff": diff,
                                       "parameters": params})
        with open(checkpoint_path, "w", encoding="utf-8") as f:
            json.dump({"last_process_idx": -1}, f)

    def run(self):
        for shape in self.shape_group:
            self.benchmark_shape(shape)


def collect(result_dir):
    """{(M, N, K): (time_us, parameters)} of the fastest correct candidate over all ranks' result files."""
    best = {}
    for name in sorted(os.listdir(result_dir)):
        if not (name.startswith("shape_") and name.endswith(".jsonl")) or name.endswith("_checkpoint.jsonl"):
            continue
        with open(os.path.join(result_dir, name), encoding="utf-8") as f:
            for line in f:
                if not line.strip():
                    continue
                r = json.loads(line)
                t = r["time"]
                if t is None or t <= 0 or t == float("inf") or r["diff"] < 0 or r["diff"] > ERROR_TOLERANCE:
                    continue
                key = (r["M"], r["N"], r["K"])
                if key not in best or t < best[key][0]:
                    best[key] = (t, r["parameters"])
    return best


def write_csv(best, path, dtype=_lib.ACL_BF16, layout_b=1):
    """Finalise every winning plan through the library and save the plan cache as CSV (reference schema first)."""
    check(lib.dgaTilingCacheClear(), "dgaTilingCacheClear")
    for (m, n, k), (time_us, params) in sorted(best.items()):
        plan = _lib.TilingPlan()
        # defaults (rasterisation, strides) from the selector, tile fields from the sweep
        check(lib.dgaSelectKernel(m, n, k, dtype, dtype, layout_b, ctypes.byref(plan)), "dgaSelectKernel")
        plan.stages = 0
        plan.splitkFactor = 0
        for key, val in params.items():
            setattr(plan, key, val)
        plan.estTimeUs = time_us
        check(lib.dgaTilingCacheSet(ctypes.byref(plan)), "dgaTilingCacheSet")
    check(lib.dgaTilingCacheSaveCsv(path.encode()), "dgaTilingCacheSaveCsv")
    n_rows = int(lib.dgaTilingCacheSize())
    check(lib.dgaTilingCacheClear(), "dgaTilingCacheClear")
    return n_rows


class _GpuBench:
    """Default timer / checker: CUDA-graph replay of 10 launches over buffers rotating beyond L2; sampled-row check
    against an fp32 torch matmul of the same inputs (a tool-side check, not a product path)."""
    _cache = {}

    def _buffers(self, shape, layout_b=1):
        import torch
        key = (shape, layout_b)
        if key not in self._cache:
            self._cache.clear()
            m, n, k = shape
            dev = "cuda:0"
            nbytes = 2 * (m * k + n * k + m * n)
            sets = max(2, min(16, int(2 * (126 << 20) / nbytes) + 2))
            gen = torch.Generator(device=dev).manual_seed(0)
            a = [torch.rand((m, k), device=dev, generator=gen).add_(0.5).to(torch.bfloat16) for _ in range(sets)]
            # layout_b 1: weight stored [N, K], mat2 is its transposed view (the aclnn operator's layout);
            # layout_b 0: mat2 row-major [K, N] (the framework functions' layout)
            w = [torch.rand((n, k), device=dev, generator=gen).add_(0.5).to(torch.bfloat16) for _ in range(sets)]
            w = [x.t() if layout_b == 1 else x.t().contiguous() for x in w]
            c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
            self._cache[key] = (a, w, c)
        return self._cache[key]

    def diff(self, shape, params, layout_b):
        import torch
        from .ops import matmul_with_plan
        a, w, c = self._buffers(shape, layout_b)
        c[0].zero_()
        matmul_with_plan(a[0], w[0], c[0], **params)
        torch.cuda.synchronize()
        rows = torch.linspace(0, shape[0] - 1, min(shape[0], 32)).long().to(a[0].device)
        ref = a[0][rows].float() @ w[0].float()
        return float(((c[0][rows].float() - ref).abs() / ref.abs().clamp_min(1e-6)).max())

    def time(self, shape, params, layout_b, per=10, iters=10):
        import torch
        from .ops import matmul_with_plan
        a, w, c = self._buffers(shape, layout_b)
        sets = len(a)
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            for i in range(min(sets, 3)):
                matmul_with_plan(a[i], w[i], c[i], **params)
            side.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for i in range(per):
                    matmul_with_plan(a[i % sets], w[i % sets], c[i % sets], **params)
        graph.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            graph.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (iters * per)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--shapes", nargs="+", default=["4096x7168x2048"], help="MxNxK ...")
    ap.add_argument("--rank-id", type=int, default=0)
    ap.add_argument("--num-processes", type=int, default=1)
    ap.add_argument("--result-dir", default="./results")
    ap.add_argument("--csv", default=None, help="after the sweep, write the winners as a plan-cache CSV here")
    args = ap.parse_args(argv)
    shapes = [tuple(int(v) for v in s.lower().split("x")) for s in args.shapes]
    AutotuneRunner(shapes, args.rank_id, args.num_processes, args.result_dir).run()
    best = collect(args.result_dir)
    for key in sorted(best):
        print("%s: %.2f us %s" % ("x".join(map(str, key)), best[key][0], best[key][1]))
    if args.csv:
        print("wrote %d rows to %s" % (write_csv(best, args.csv), args.csv))


if __name__ == "__main__":
    main()
