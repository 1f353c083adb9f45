#!/usr/bin/env python3
"""Tile width of the fp32-scale fp8 kernel (option fp8_bn: 256 / 224 / 192 with two TMEM buffers, 160 with three, 128 with
four) on one shape, interleaved over several rounds, checked against the first width's result.
usage: fp8_tile_sweep.py [M N K]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402


def main():
    m, n, k = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (4096, 7168, 16384)
    dev = "cuda:0"
    gen = torch.Generator(device=dev).manual_seed(0)
    sets = 3
    a = [torch.randn((m, k), device=dev, generator=gen).to(torch.float8_e4m3fn) for _ in range(sets)]
    w = [torch.randn((n, k), device=dev, generator=gen).to(torch.float8_e4m3fn) for _ in range(sets)]
    sa = [torch.rand((m, -(-k // 128)), device=dev, generator=gen) + 0.5 for _ in range(sets)]
    sw = [torch.rand((-(-n // 128), -(-k // 128)), device=dev, generator=gen) + 0.5 for _ in range(sets)]
    c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    widths = (224, 256, 192, 160, 128)
    ex, ref, total = {}, None, {}
    for bn in widths:
        dga.set_option("fp8_bn", bn)
        ex[bn] = [dga.Fp8Executor(a[i], sa[i], w[i], sw[i], c[i]) for i in range(sets)]   # the plan is made in phase 1
        ex[bn][0]()
        torch.cuda.synchronize()
        assert ex[bn][0].path() == "promotion"
        if ref is None:
            ref = c[0].clone()
        diff = (c[0].float() - ref.float()).abs().max().item() / max(ref.float().abs().max().item(), 1e-30)
        print("bn=%d: max |diff| / max |ref| against bn=%d: %.3e" % (bn, widths[0], diff), flush=True)
        total[bn] = 0.0
    rounds = 6
    for _ in range(rounds):
        for bn in widths:
            for i in range(3):
                ex[bn][i % sets]()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(20):
                ex[bn][i % sets]()
            e1.record()
            torch.cuda.synchronize()
            total[bn] += e0.elapsed_time(e1) * 50.0
    for bn in widths:
        us = total[bn] / rounds
        print("%d x %d x %d  bn=%d: %8.1f us  %7.1f TFLOP/s" % (m, n, k, bn, us, 2.0 * m * n * k / us / 1e6), flush=True)


if __name__ == "__main__":
    main()
