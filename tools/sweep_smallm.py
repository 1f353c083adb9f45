#!/usr/bin/env python3
"""Decode-kernel sweep in one process: (M) x split x ring depth x L2-prefetch depth x PDL, CUDA graph of 20 launches,
weights rotating beyond L2.  Developer tool; prints one line per point and the best per M."""
import itertools
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402


def time_point(m, n, k, fields, per=20, iters=60):
    dev = "cuda:0"
    nbytes = 2 * (m * k + n * k + m * n)
    sets = max(2, int(2 * (126 << 20) / nbytes) + 2)
    a = [torch.randn((m, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    w = [torch.randn((n, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]

    def run(i):
        dga.matmul_with_plan(a[i % sets], w[i % sets].t(), c[i % sets], **fields)

    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for i in range(sets):
            run(i)
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for i in range(per):
                run(i)
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters / per
    return us, nbytes / us / 1e3


def main():
    n, k = 4096, 7168
    ms = [int(v) for v in (sys.argv[1].split(",") if len(sys.argv) > 1 else "1,16,64,128".split(","))]
    # optional: comma lists of split factors (any cluster size up to 8, not only powers of two), ring depths, prefetch depths
    arg = lambda i, d: tuple(int(v) for v in sys.argv[i].split(",")) if len(sys.argv) > i else d
    splits, depths, pfs = arg(2, (4, 8)), arg(3, (3, 4, 6, 0)), arg(4, (0, 4, 8, 16, 64))
    for m in ms:
        best = None
        for split, st, pf, pdl in itertools.product(splits, depths, pfs, (1,)):
            dga.set_option("pdl", pdl)
            dga.set_option("splitk_prefetch_kb", pf)
            try:
                us, gbs = time_point(m, n, k, dict(kernelSerial=3, splitkFactor=split, stages=st))
            except Exception as e:  # infeasible plan
                print("m=%d split=%d stages=%d pf=%d: %s" % (m, split, st, pf, str(e)[:80]))
                continue
            print("m=%d split=%d stages=%d pf=%d pdl=%d: %.2f us %.0f GB/s" % (m, split, st, pf, pdl, us, gbs), flush=True)
            if best is None or us < best[0]:
                best = (us, gbs, split, st, pf)
        print("BEST m=%d: %.2f us %.0f GB/s split=%d stages=%d pf=%d" % ((m,) + best), flush=True)


if __name__ == "__main__":
    main()
