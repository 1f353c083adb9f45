#!/usr/bin/env python3
"""A/B of hand-set plans in one process, interleaved repetitions (clock/power drift cancels).
usage: ab_plans.py M N K "k=v,k=v" "k=v,..." ..."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402,F401
from sweep_smallm import time_point  # noqa: E402

m, n, k = (int(v) for v in sys.argv[1:4])
plans = [dict((kv.split("=")[0], int(kv.split("=")[1])) for kv in arg.split(",")) for arg in sys.argv[4:]]
res = [[] for _ in plans]
for rep in range(4):
    for i, pl in enumerate(plans):
        us, _ = time_point(m, n, k, pl, per=20, iters=20)
        res[i].append(us)
for pl, r in zip(plans, res):
    best = min(r)
    print("%d %d %d %s: min %.2f us (%.1f TFLOP/s) all %s" % (m, n, k, pl, best, 2.0 * m * n * k / best / 1e6,
                                                            " ".join("%.2f" % v for v in r)), flush=True)
