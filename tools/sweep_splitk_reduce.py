#!/usr/bin/env python3
"""Mid-M shapes: multicore split-K + reduce (kernelSerial 5) against the selector's current choice."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402
from sweep_smallm import time_point  # noqa: E402

shapes = [(m, 4096, 7168) for m in (64, 128, 192, 256, 384, 512, 768, 1024)] + [(128, 7168, 2048), (256, 7168, 2048),
                                                                              (512, 7168, 2048), (256, 2048, 7168)]
for m, n, k in shapes:
    cur = dga.select_kernel(m, n, k)
    keys = ("kernelSerial", "ctaGroup", "n1", "stages", "splitkFactor")
    us, _ = time_point(m, n, k, {key: cur[key] for key in keys}, iters=30)
    print("%d %d %d current %s: %.2f us" % (m, n, k, {key: cur[key] for key in keys}, us), flush=True)
    best = None
    for cta in (2, 1):
        if cta == 2 and m <= 128:
            continue
        for n1 in (256, 192, 128):
            tiles = -(-m // (128 * cta)) * -(-n // n1)
            for split in (2, 3, 4, 6, 8, 12, 16):
                if tiles * split > 2 * (148 // cta) or -(-k // 64) < 8 * split:
                    continue
                try:
                    us, _ = time_point(m, n, k, dict(kernelSerial=5, ctaGroup=cta, n1=n1, splitkFactor=split), iters=30)
                except Exception as e:
                    print("   cta%d n1=%d split=%d: %s" % (cta, n1, split, str(e)[:70]))
                    continue
                print("   cta%d n1=%d split=%d (items %d): %.2f us" % (cta, n1, split, tiles * split, us), flush=True)
                if best is None or us < best[0]:
                    best = (us, cta, n1, split)
    print("BEST %d %d %d: %.2f us cta%d n1=%d split=%d" % ((m, n, k) + best), flush=True)
