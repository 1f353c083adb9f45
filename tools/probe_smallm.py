#!/usr/bin/env python3
"""Where does the decode kernel's time go: fixed cost vs streaming rate (time against bytes), producer issue batching."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402
from sweep_smallm import time_point  # noqa: E402

dga.set_option("pdl", 1)
for m in (16, 64):
    for n, k in ((4096, 7168), (4096, 14336), (4096, 28672), (8192, 7168), (16384, 7168), (2048, 7168), (7168, 2048)):
        for split, st in ((8, 6), (4, 6), (4, 0), (8, 4)):
            if m == 64 and (split, st) == (8, 6):
                continue
            for batch in (1, 2, 3):
                dga.set_option("splitk_issue_batch", batch)
                dga.set_option("splitk_prefetch_kb", 4)
                try:
                    us, gbs = time_point(m, n, k, dict(kernelSerial=3, splitkFactor=split, stages=st), iters=30)
                except Exception as e:
                    print("m=%d n=%d k=%d split=%d st=%d batch=%d: %s" % (m, n, k, split, st, batch, str(e)[:60]))
                    continue
                print("m=%d n=%d k=%d split=%d st=%d batch=%d: %.2f us %.0f GB/s" % (m, n, k, split, st, batch, us, gbs),
                      flush=True)
