#!/usr/bin/env python3
"""Selector audit: for a grid of shapes, sweep the candidate plans (the tuned-table runner) and time the selector's own
plan the same way; prints winner, selector, and how much the selector leaves on the table.
usage: selector_audit.py RESULT_DIR MxNxK ...   (env LAYOUT_B=0 audits row-major mat2, the framework layout)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402
from deepgemm_ascend_b200 import autotune as at  # noqa: E402

KEYS = ("kernelSerial", "ctaGroup", "n1", "stages", "splitkFactor")


def main():
    result_dir = sys.argv[1]
    shapes = [tuple(int(v) for v in s.lower().split("x")) for s in sys.argv[2:]]
    layout_b = int(os.environ.get("LAYOUT_B", "1"))
    at.AutotuneRunner(shapes, 0, 1, result_dir, layout_b=layout_b).run()
    best = at.collect(result_dir)
    bench = at._GpuBench()
    worst = []
    for shape in shapes:
        plan = dga.select_kernel(*shape, layout_b=layout_b)
        fields = {k: plan[k] for k in KEYS}
        if plan["kernelSerial"] == 3:
            fields = {"kernelSerial": 3, "splitkFactor": plan["splitkFactor"], "stages": plan["stages"]}
        t_sel = min(bench.time(shape, fields, layout_b) for _ in range(2))
        t_best, p_best = best.get(shape, (float("nan"), None))
        loss = t_sel / t_best - 1.0
        worst.append((loss, shape))
        print("%-18s selector %8.2f us %-70s | best %8.2f us %s | selector is %+.1f %%" % (
            "x".join(map(str, shape)), t_sel, fields, t_best, p_best, 100 * loss), flush=True)
    worst.sort(reverse=True)
    print("largest gaps:", [("x".join(map(str, s)), "%.1f %%" % (100 * l)) for l, s in worst[:8]])


if __name__ == "__main__":
    main()
