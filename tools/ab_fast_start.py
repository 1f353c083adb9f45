#!/usr/bin/env python3
"""Decode kernel A/B: producer warp that skips the CTA-wide rendezvous (splitk_fast_start 1) against the plain prologue,
the selector's plan per M (hand-set), alternating passes, best of three; parity of the new prologue first."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402
from sweep_smallm import time_point  # noqa: E402


def check(m, n, k, fields):
    a = torch.randn((m, k), device="cuda:0", dtype=torch.bfloat16)
    w = torch.randn((n, k), device="cuda:0", dtype=torch.bfloat16)
    ref = a.float() @ w.float().t()
    for threads in (256, 192):
        dga.set_option("splitk_threads", threads)
        c = torch.full((m, n), float("nan"), device="cuda:0", dtype=torch.bfloat16)
        for _ in range(4):
            dga.matmul_with_plan(a, w.t(), c, **fields)
        torch.cuda.synchronize()
        err = ((c.float() - ref).abs().max() / ref.abs().max()).item()
        print("check m=%d n=%d k=%d %s threads=%d: %.3g" % (m, n, k, fields, threads, err), flush=True)
        assert err < 1e-2
    dga.set_option("splitk_threads", -1)


def main():
    n, k = 4096, 7168
    dga.set_option("splitk_fast_start", 1)
    sel = {1: dict(kernelSerial=3, splitkFactor=8, stages=4), 16: dict(kernelSerial=3, splitkFactor=8, stages=4),
           64: dict(kernelSerial=3, splitkFactor=8, stages=3), 128: dict(kernelSerial=3, splitkFactor=4, stages=5)}
    for m, nn, kk, f in [(1, n, k, sel[1]), (64, n, k, sel[64]), (128, n, k, sel[128]), (50, 1000, 1992, dict(kernelSerial=3, splitkFactor=3)),
                         (16, 520, 4096, dict(kernelSerial=3, splitkFactor=1)), (100, 384, 512, dict(kernelSerial=3, splitkFactor=8))]:
        check(m, nn, kk, f)
    for m in (1, 16, 64, 128):
        res = {}
        for rep in range(3):
            for fast in (0, 1):
                dga.set_option("splitk_fast_start", fast)
                us, gbs = time_point(m, n, k, sel[m], iters=40)
                if fast not in res or us < res[fast][0]:
                    res[fast] = (us, gbs)
        print("m=%d: plain %.2f us %.0f GB/s | fast start %.2f us %.0f GB/s" % ((m,) + res[0] + res[1]), flush=True)


if __name__ == "__main__":
    main()
