#!/usr/bin/env python3
"""Phase timeline of the decode kernel inside a chain of back-to-back launches (run with DGA_DEBUG_FLAGS=16).
usage: trace_smallm.py M N K split stages [pdl]"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402

m, n, k, split, stages = (int(v) for v in sys.argv[1:6])
pdl = int(sys.argv[6]) if len(sys.argv) > 6 else 1
dga.set_option("pdl", pdl)
dev = "cuda:0"
sets = 6
a = [torch.randn((m, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
w = [torch.randn((n, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
fields = dict(kernelSerial=3, splitkFactor=split, stages=stages)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for i in range(sets):
        dga.matmul_with_plan(a[i], w[i].t(), c[i], **fields)
    side.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for i in range(20):
            dga.matmul_with_plan(a[i % sets], w[i % sets].t(), c[i % sets], **fields)
for rep in range(3):
    g.replay()
    g.replay()
    torch.cuda.synchronize()
    grid = -(-n // 128) * split
    buf = (ctypes.c_uint64 * (12 * grid))()
    dga._lib.check(dga._lib.lib.dgaDebugReadTrace(buf, 12 * grid), "trace")
    t = np.array(buf[:], dtype=np.float64).reshape(grid, 12)
    g0 = t[:, 0].min()
    ghz = ((t[:, 10] - t[:, 2]) / np.maximum(t[:, 1] - t[:, 0], 1)).mean()
    us = lambda a, b: ((t[:, a] - t[:, b]) / ghz / 1e3)  # noqa: E731
    rd = t[:, 8] > 0
    print("rep %d pdl=%d: grid %d | last launch spans %.2f us (first CTA start -> last CTA end) | clock %.2f GHz" % (
        rep, pdl, grid, (t[:, 1].max() - g0) / 1e3, ghz))
    print("   CTA start skew %.2f us | wait returned at %.2f..%.2f us after first start" % (
        (t[:, 0].max() - g0) / 1e3, (t[:, 11].min() - g0) / 1e3, (t[:, 11].max() - g0) / 1e3))
    print("   per CTA means (us): setup %.2f | dep-wait %.2f | wait->first data %.2f | stream (first data->last MMA issued) %.2f"
          " | ->tmem_full seen %.2f | ->cluster sync 1 %.2f | reduce+store %.2f | sync 2 + exit %.2f | total %.2f" % (
              us(3, 2).mean(), us(4, 3).mean(), us(5, 4).mean(), us(6, 5).mean(), us(7, 6).mean(),
              us(8, 7)[rd].mean() if rd.any() else 0, us(9, 8)[rd].mean() if rd.any() else 0,
              us(10, 9)[rd].mean() if rd.any() else 0, us(10, 2).mean()))
    print("   max over CTAs (us): wait->first data %.2f | stream %.2f | total %.2f" % (
        us(5, 4).max(), us(6, 5).max(), us(10, 2).max()))
