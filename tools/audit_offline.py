#!/usr/bin/env python3
"""Offline half of the selector audit: compares, inside one sweep's own records (same warm-up state for every
candidate), the time of the candidate that equals the selector's plan with the fastest candidate.
usage: audit_offline.py RESULT_DIR"""
import glob
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402


def same(plan, cand):
    if plan["kernelSerial"] in (0, 1, 2):
        return cand["kernelSerial"] == 0 and cand.get("ctaGroup") == plan["ctaGroup"] and cand.get("n1") == plan["n1"] \
            and cand.get("stages", 0) == 0
    if plan["kernelSerial"] == 3:
        return cand["kernelSerial"] == 3 and cand.get("splitkFactor") == plan["splitkFactor"] and \
            cand.get("stages", 0) in (0, plan["stages"])
    if plan["kernelSerial"] == 4:
        return cand["kernelSerial"] == 4 and cand.get("ctaGroup") == plan["ctaGroup"] and cand.get("n1") == plan["n1"] \
            and cand.get("splitkFactor", 0) in (0, plan["splitkFactor"])
    if plan["kernelSerial"] == 5:
        return cand["kernelSerial"] == 5 and cand.get("splitkFactor") == plan["splitkFactor"] and \
            cand.get("ctaGroup") == plan["ctaGroup"] and cand.get("n1") == plan["n1"]
    return False


def main():
    rows = []
    for path in sorted(glob.glob(os.path.join(sys.argv[1], "shape_*_rank_0.jsonl"))):
        recs = [json.loads(line) for line in open(path) if line.strip()]
        recs = [r for r in recs if r["time"] and 0 < r["time"] < 1e9 and 0 <= r["diff"] <= 1e-2]
        if not recs:
            continue
        m, n, k = recs[0]["M"], recs[0]["N"], recs[0]["K"]
        plan = dga.select_kernel(m, n, k, layout_b=int(os.environ.get("LAYOUT_B", "1")))
        best = min(recs, key=lambda r: r["time"])
        mine = [r for r in recs if same(plan, r["parameters"])]
        t_mine = min(r["time"] for r in mine) if mine else float("nan")
        rows.append((t_mine / best["time"] - 1.0, (m, n, k), t_mine, best, plan))
    rows.sort(key=lambda r: (r[1][1], r[1][2], r[1][0]))
    for loss, shape, t_mine, best, plan in rows:
        print("%-18s selector ks%d cta%d n1=%-3d sk%d %8.2f us | best %8.2f us %-75s | %+6.1f %%" % (
            "x".join(map(str, shape)), plan["kernelSerial"], plan["ctaGroup"], plan["n1"], plan["splitkFactor"], t_mine,
            best["time"], best["parameters"], 100 * loss))


if __name__ == "__main__":
    main()
