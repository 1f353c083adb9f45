#!/usr/bin/env python3
"""Interleaved A/B of library options (dgaSetOption) and hand-set plan fields on one bf16 shape: the candidates
take turns over several rounds, so that the drift of the power-capped clock hits all of them alike.
usage: ab_options.py M N K [rounds=6] [iters=20] -- "name=value,name=value" "..." ...
       names that dgaSetOption knows (raster_group, raster_along_m, hint_a, hint_b, pdl, ...) are options, the rest
       (kernelSerial, ctaGroup, n1, clusterShape, stages, ...) plan fields; an empty string is the default path."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402

OPTIONS = ("raster_group", "raster_along_m", "hint_a", "hint_b", "pdl", "jit", "splitk_prefetch_kb", "splitk_issue_batch",
           "streamk_cooperative")


def main():
    sep = sys.argv.index("--")
    head, cands = sys.argv[1:sep], sys.argv[sep + 1:]
    m, n, k = (int(v) for v in head[:3])
    kw = dict(kv.split("=") for kv in head[3:])
    rounds, iters = int(kw.get("rounds", 6)), int(kw.get("iters", 20))
    dev = "cuda:0"
    nbytes = 2 * (m * k + n * k + m * n)
    sets = max(2, int(2 * (126 << 20) / nbytes) + 2)
    a = [torch.randn((m, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    w = [torch.randn((n, k), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    c = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
    parsed = []
    for cand in cands:
        fields = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in cand.split(",") if kv}
        parsed.append(({k2: v for k2, v in fields.items() if k2 in OPTIONS}, {k2: v for k2, v in fields.items() if k2 not in OPTIONS}))
    total = [0.0] * len(cands)
    for _ in range(rounds):
        for ci, (opts, plan) in enumerate(parsed):
            for name in ("raster_group", "raster_along_m", "hint_a", "hint_b"):
                dga.set_option(name, opts.get(name, -1))
            for name, val in opts.items():
                dga.set_option(name, val)

            def run(i):
                if plan:
                    dga.matmul_with_plan(a[i % sets], w[i % sets].t(), c[i % sets], **plan)
                else:
                    dga.catlass_dynamic_matmul(a[i % sets], w[i % sets].t(), out=c[i % sets])
            for i in range(3):
                run(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(iters):
                run(i)
            e1.record()
            torch.cuda.synchronize()
            total[ci] += e0.elapsed_time(e1) * 1e3 / iters
    for cand, t in zip(cands, total):
        us = t / rounds
        print("%d x %d x %d  %-60s %9.2f us  %7.1f TFLOP/s" % (m, n, k, cand or "(default)", us, 2.0 * m * n * k / us / 1e6), flush=True)


if __name__ == "__main__":
    main()
