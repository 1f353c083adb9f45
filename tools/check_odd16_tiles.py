#!/usr/bin/env python3
"""Tiles whose width is an odd multiple of 16 (n1 = 240, 208, 48, ...): parity against fp32 torch on whole-tile plans
(both CTA groups, ragged M / N / K, fp32 output of the framework path), then timings of the shapes where such a width
balances the waves better than the selector's multiple of 32.  Developer tool behind the selector's candidate list."""
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import deepgemm_ascend_b200 as dga  # noqa: E402

ok = True
for m, n, k, cta, n1 in [(2048, 4096, 7168, 2, 240), (512, 4096, 512, 2, 240), (1000, 1000, 1000, 2, 208), (300, 520, 1992, 1, 240),
                         (256, 4096, 256, 2, 48), (128, 1000, 512, 1, 80), (4096, 7168, 2048, 2, 240), (257, 4090 // 8 * 8, 640, 2, 176)]:
    a = torch.randn((m, k), device="cuda:0", dtype=torch.bfloat16)
    w = torch.randn((n, k), device="cuda:0", dtype=torch.bfloat16)
    ref = a.float() @ w.float().t()
    c = torch.full((m, n), float("nan"), device="cuda:0", dtype=torch.bfloat16)
    try:
        dga.matmul_with_plan(a, w.t(), c, kernelSerial=0, ctaGroup=cta, n1=n1)
        torch.cuda.synchronize()
        err = ((c.float() - ref).abs().max() / ref.abs().max()).item()
    except Exception as e:
        err = float("inf")
        print(str(e)[:120])
    good = err < 1e-2
    ok &= good
    print("m=%d n=%d k=%d cta=%d n1=%d: max|d|/max|ref| = %.3g %s" % (m, n, k, cta, n1, err, "ok" if good else "FAIL"), flush=True)
if not ok:
    sys.exit(1)
for m, n, k, plans in [(2048, 4096, 7168, (256, 240)), (4096, 7168, 2048, (224, 240)), (1024, 4096, 7168, (256, 240, 208)),
                       (2048, 7168, 2048, (256, 240, 224, 208)), (4096, 4096, 7168, (256, 240)), (2048, 2112, 7168, (256, 240, 192, 176))]:
    for rep in range(2):
        for n1 in plans:
            out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "time_plan.py"), str(m), str(n), str(k),
                                  "kernelSerial=0", "ctaGroup=2", "n1=%d" % n1], capture_output=True, text=True)
            print("m=%d n=%d k=%d n1=%d: %s" % (m, n, k, n1, (out.stdout.strip().splitlines() or [out.stderr[-200:]])[-1]), flush=True)
