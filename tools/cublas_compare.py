#!/usr/bin/env python3
"""Comparator only (never in the product path): torch.matmul (cuBLAS) on the same shapes, same rotation scheme."""
import sys
import torch

shapes = [(4096, 7168, 2048), (2048, 4096, 7168), (8192, 4096, 7168), (512, 4096, 7168), (128, 4096, 7168), (64, 4096, 7168)]
dev = "cuda:0"
for m, n, k in shapes:
    nbytes = 2 * (m * k + n * k + m * n)
    sets = max(2, int(2 * (126 << 20) / nbytes) + 2)
    a = [torch.randn((m, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    w = [torch.randn((n, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    for i in range(10):
        torch.matmul(a[i % sets], w[i % sets].t(), out=c[i % sets])
    torch.cuda.synchronize()
    iters = 200
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        torch.matmul(a[i % sets], w[i % sets].t(), out=c[i % sets])
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    print("cublas %d %d %d: %.2f us %.1f TFLOP/s %.1f GB/s" % (m, n, k, us, 2.0 * m * n * k / us / 1e6, nbytes / us / 1e3))
