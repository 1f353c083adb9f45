#!/usr/bin/env python3
"""Decode kernel with cluster sizes that are not powers of two (split 3, 5, 6, 7): parity against fp32 torch on a few
shapes incl. ragged N / K and M not divisible by the split.  Developer tool behind the selector's split list."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402

for m, n, k in [(1, 4096, 7168), (16, 4096, 7168), (64, 4096, 7168), (50, 1000, 1992), (128, 512, 4096)]:
    a = torch.randn((m, k), device="cuda:0", dtype=torch.bfloat16)
    w = torch.randn((n, k), device="cuda:0", dtype=torch.bfloat16)
    ref = a.float() @ w.float().t()
    for split in (3, 5, 6, 7):
        c = torch.full((m, n), float("nan"), device="cuda:0", dtype=torch.bfloat16)
        try:
            dga.matmul_with_plan(a, w.t(), c, kernelSerial=3, splitkFactor=split)
        except Exception as e:
            print("m=%d n=%d k=%d split=%d: %s" % (m, n, k, split, str(e)[:100]))
            continue
        torch.cuda.synchronize()
        err = ((c.float() - ref).abs().max() / ref.abs().max()).item()
        print("m=%d n=%d k=%d split=%d: max|d|/max|ref| = %.3g %s" % (m, n, k, split, err, "ok" if err < 1e-2 else "FAIL"),
              flush=True)
