#!/usr/bin/env python3
"""Per-CTA timeline of the persistent kernel (run with DGA_DEBUG_FLAGS=16): SM clock actually held during the launch,
start skew, time to first operands, tail.  usage: trace_common.py M N K [n1]"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402

m, n, k = (int(v) for v in sys.argv[1:4])
fields = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in sys.argv[4:]}
dev = "cuda:0"
sets = 6
a = [torch.randn((m, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
w = [torch.randn((n, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
plan = dga.select_kernel(m, n, k)
plan.update(fields)
for rep in range(3):
    for i in range(60):
        if fields:
            dga.matmul_with_plan(a[i % sets], w[i % sets].t(), c[i % sets], **fields)
        else:
            dga.catlass_dynamic_matmul(a[i % sets], w[i % sets].t(), out=c[i % sets])
    torch.cuda.synchronize()
    grid = plan["blockDim"] if not fields else 148
    buf = (ctypes.c_uint64 * (6 * grid))()
    dga._lib.check(dga._lib.lib.dgaDebugReadTrace(buf, 6 * grid), "trace")
    t = np.array(buf[:], dtype=np.float64).reshape(grid, 6)
    t = t[t[:, 3] > 0]
    t0 = t[:, 0].min()
    dur = t[:, 3].max() - t0
    ghz = ((t[:, 5] - t[:, 4]) / (t[:, 3] - t[:, 0])).mean()
    print("rep %d: ctas=%d kernel span %.2f us | SM clock %.3f GHz | start skew max %.2f us | first data after %.2f us "
          "(max %.2f) | last MMA issued %.2f us before end (mean) | end skew %.2f us" % (
              rep, len(t), dur / 1e3, ghz, (t[:, 0].max() - t0) / 1e3, (t[:, 1][t[:, 1] > 0] - t[:, 0][t[:, 1] > 0]).mean() / 1e3,
              (t[:, 1][t[:, 1] > 0] - t0).max() / 1e3, (t[:, 3][t[:, 2] > 0] - t[:, 2][t[:, 2] > 0]).mean() / 1e3,
              (t[:, 3].max() - t[:, 3].min()) / 1e3))
