#!/usr/bin/env python3
"""Times one (m, n, k) with optional hand-set plan fields, rotating buffers beyond L2. Developer tool.
usage: time_plan.py M N K [key=value ...]   e.g. time_plan.py 4096 7168 2048 kernelSerial=0 ctaGroup=2 n1=256 stages=5"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402


def main():
    m, n, k = (int(v) for v in sys.argv[1:4])
    fields = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in sys.argv[4:]}
    iters = fields.pop("iters", 200)
    use_graph = fields.pop("graph", 0)   # graph=1: replay 20 back-to-back launches as one CUDA graph per iteration
    dev = "cuda:0"
    nbytes = 2 * (m * k + n * k + m * n)
    sets = max(2, int(2 * (126 << 20) / nbytes) + 2)
    a = [torch.randn((m, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    w = [torch.randn((n, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]

    def run(i):
        if fields:
            dga.matmul_with_plan(a[i % sets], w[i % sets].t(), c[i % sets], **fields)
        else:
            dga.catlass_dynamic_matmul(a[i % sets], w[i % sets].t(), out=c[i % sets])

    per = 1
    if use_graph:
        per = 20
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            for i in range(sets):
                run(i)
            side.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for i in range(per):
                    run(i)
        inner = run
        run = lambda i: g.replay()  # noqa: E731
        iters = max(10, iters // per)
    for i in range(10):
        run(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        run(i)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters / per
    print("%d %d %d %s dbg=%s pdl=%s: %.2f us  %.1f TFLOP/s  %.1f GB/s" % (
        m, n, k, fields, os.environ.get("DGA_DEBUG_FLAGS", "0"), os.environ.get("DGA_PDL", "1"), us, 2.0 * m * n * k / us / 1e6, nbytes / us / 1e3))


if __name__ == "__main__":
    main()
