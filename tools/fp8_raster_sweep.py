#!/usr/bin/env python3
"""Walk order x L2 eviction hints of the block-scaled fp8 kernel on BASELINE config 4 (4096 x 7168 x 16384).
Round 1 left this kernel reading 557 MB from DRAM for 245 MB of algorithmic traffic: the K panels of the ~74 tiles in
flight do not stay in L2 between the waves that share them. Knobs (dgaSetOption): raster_group, raster_along_m,
hint_a / hint_b (0 normal, 1 evict-first, 2 evict-last).
  fp8_raster_sweep.py time    -> CUDA-event time of every configuration (10 launches each, rotating buffers)
  fp8_raster_sweep.py once    -> every configuration launched twice, in the printed order (run under
                                 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
                                 -k regex:gemm_sm100_kernel; the second launch of each pair is the warm one)"""
import itertools
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402

GROUPS = (2, 4, 8, 16, 32)
DIRS = (1, 0)
HINTS = ((0, 0), (1, 2), (2, 1), (0, 2), (2, 0), (1, 0), (0, 1))


def configs():
    if os.environ.get("SWEEP_ONE"):   # "dir,group,hint_a,hint_b": one configuration per process (evict-last lines of an
        d, g, ha, hb = (int(v) for v in os.environ["SWEEP_ONE"].split(","))   # earlier configuration survive in L2)
        yield {"raster_along_m": d, "raster_group": g, "hint_a": ha, "hint_b": hb}
        return
    if os.environ.get("SWEEP_SHORT"):
        yield {"raster_along_m": -1, "raster_group": -1, "hint_a": -1, "hint_b": -1}     # the library's own rule
        yield {"raster_along_m": 0, "raster_group": 8, "hint_a": 0, "hint_b": 1}
        yield {"raster_along_m": 1, "raster_group": 8, "hint_a": 0, "hint_b": 0}
        yield {"raster_along_m": -1, "raster_group": -1, "hint_a": -1, "hint_b": -1}
        return
    for d, g, (ha, hb) in itertools.product(DIRS, GROUPS, HINTS):
        yield {"raster_along_m": d, "raster_group": g, "hint_a": ha, "hint_b": hb}


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "time"
    m, n, k = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (4096, 7168, 16384)
    dev = "cuda:0"
    gen = torch.Generator(device=dev).manual_seed(0)
    sets = 3
    a = [torch.randn((m, k), device=dev, generator=gen).to(torch.float8_e4m3fn) for _ in range(sets)]
    w = [torch.randn((n, k), device=dev, generator=gen).to(torch.float8_e4m3fn) for _ in range(sets)]
    sa = [torch.pow(2.0, torch.randint(-3, 4, (m, k // 128), device=dev, generator=gen).float()) for _ in range(sets)]
    sw = [torch.pow(2.0, torch.randint(-3, 4, (n // 128, k // 128), device=dev, generator=gen).float()) for _ in range(sets)]
    c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    ex = [dga.Fp8Executor(a[i], sa[i], w[i], sw[i], c[i]) for i in range(sets)]
    for i in range(sets):
        ex[i]()
    torch.cuda.synchronize()
    assert ex[0].path() == "block_scaled"
    ref = c[0].clone()
    if mode == "ab":
        # a few candidates, interleaved over several rounds: single passes drift with the power-capped clock
        cands = [(1, 8, 0, 0), (1, 2, 0, 0), (0, 8, 0, 1), (0, 8, 2, 1), (1, 8, 1, 2), (0, 8, 0, 0), (1, 11, 1, 2)]
        total = {c: 0.0 for c in cands}
        rounds = 8
        for _ in range(rounds):
            for cand in cands:
                for key, val in zip(("raster_along_m", "raster_group", "hint_a", "hint_b"), cand):
                    dga.set_option(key, val)
                for i in range(3):
                    ex[i % sets]()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(20):
                    ex[i % sets]()
                e1.record()
                torch.cuda.synchronize()
                total[cand] += e0.elapsed_time(e1) * 50.0
        for cand in cands:
            us = total[cand] / rounds
            print("dir=%d group=%2d hint_a=%d hint_b=%d: %7.1f us  %6.1f TFLOP/s (mean of %d interleaved rounds x 20)" % (
                cand + (us, 2.0 * m * n * k / us / 1e6, rounds)), flush=True)
        return
    for cfg in configs():
        for key, val in cfg.items():
            dga.set_option(key, val)
        if mode == "once":
            ex[1]()
            ex[0]()
            torch.cuda.synchronize()
            print("cfg", cfg, flush=True)
            continue
        for i in range(3):
            ex[i % sets]()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(10):
            ex[i % sets]()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 100.0
        ok = torch.equal(c[0], ref)
        print("dir=%d group=%2d hint_a=%d hint_b=%d: %7.1f us  %6.1f TFLOP/s  %s" % (
            cfg["raster_along_m"], cfg["raster_group"], cfg["hint_a"], cfg["hint_b"], us, 2.0 * m * n * k / us / 1e6,
            "same result" if ok else "RESULT DIFFERS"), flush=True)


if __name__ == "__main__":
    main()
