#!/usr/bin/env python3
"""Persistent kernel: effect of the producer's lookahead L2 prefetch distance (option common_prefetch_kb)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402


def bench(m, n, k, iters):
    dev = "cuda:0"
    nbytes = 2 * (m * k + n * k + m * n)
    sets = max(2, int(2 * (126 << 20) / nbytes) + 2)
    a = [torch.randn((m, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    w = [torch.randn((n, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    ex = [dga.Executor(a[i], w[i].t(), c[i]) for i in range(sets)]
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for i in range(sets):
            ex[i](side)
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for i in range(20):
                ex[i % sets](side)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters / 20
    return us, 2.0 * m * n * k / us / 1e6


for m, n, k, iters in ((4096, 7168, 2048, 20), (2048, 4096, 7168, 20), (8192, 4096, 7168, 6), (512, 4096, 7168, 40),
                       (4096, 4096, 4096, 20)):
    for pf in (0, 4, 8, 12, 16, 0):
        dga.set_option("common_prefetch_kb", pf)
        us, tf = bench(m, n, k, iters)
        print("%d %d %d pf=%d: %.2f us %.1f TFLOP/s" % (m, n, k, pf, us, tf), flush=True)
