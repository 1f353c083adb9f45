#!/usr/bin/env python3
"""Bring-up / sweep harness for the GPU box: runs each case in its own process (a device trap in one case
must not poison the rest), checks the result against the fp64 numpy oracle and times it with CUDA events.

    python tools/gpu_check.py [suite]            suites: bringup (default), sweep, smallm
Writes one JSON line per case to stdout and to gpurun_out/gpu_check_<suite>.jsonl.
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child(case):
    import numpy as np
    import torch
    import deepgemm_ascend_b200 as dga
    from oracle import matmul_oracle as orc

    m, n, k = case["m"], case["n"], case["k"]
    dtype = {"bf16": torch.bfloat16, "fp16": torch.float16}[case.get("dtype", "bf16")]
    out_dtype = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32, None: dtype}[case.get("out")]
    layout_b = case.get("layout_b", 1)
    rng = np.random.RandomState(case.get("seed", 0))
    lo, hi = case.get("range", [1, 10])
    x1 = rng.uniform(lo, hi, [m, k]).astype(np.float32)
    x2 = rng.uniform(lo, hi, [k, n]).astype(np.float32)
    a = torch.from_numpy(x1).to(dtype).cuda()
    if layout_b == 1:
        w = torch.from_numpy(np.ascontiguousarray(x2.T)).to(dtype).cuda()  # stored [N,K]
        b = w.t()
    else:
        w = torch.from_numpy(x2).to(dtype).cuda()
        b = w
    out = torch.zeros((m, n), dtype=out_dtype, device="cuda")
    plan = case.get("plan")
    sel = dga.select_kernel(m, n, k, dtype, out_dtype, layout_b)

    def run():
        if plan:
            dga.matmul_with_plan(a, b, out, **plan)
        else:
            dga.catlass_dynamic_matmul(a, b, out=out)

    run()
    torch.cuda.synchronize()
    res = {"case": case, "selected": {key: sel[key] for key in
                                      ("kernelSerial", "ctaGroup", "m1", "n1", "stages", "splitkFactor", "blockDim")}}
    # oracle: fp64 matmul of the exactly representable inputs
    a64 = a.float().cpu().numpy().astype(np.float64)
    b64 = w.float().cpu().numpy().astype(np.float64)
    rows = case.get("check_rows")
    got = out.float().cpu().numpy()
    if rows and rows < m:
        idx = np.unique(np.concatenate([np.arange(min(m, 8)), np.arange(max(0, m - 8), m),
                                        rng.randint(0, m, rows)]))
        a64, got = a64[idx], got[idx]
    ref = a64 @ (b64.T if layout_b == 1 else b64)
    res["max_rel"] = orc.max_rel_err(got, ref)
    res["calc_diff"] = orc.calc_diff(got, ref)
    tol = 1e-2 if out_dtype != torch.float32 else 1e-4
    res["ok"] = bool(res["max_rel"] <= tol) and bool(np.isfinite(got).all())
    if out_dtype == torch.float32:
        g32 = (a64.astype(np.float32) @ (b64.T if layout_b == 1 else b64).astype(np.float32))
        for name, rtol in (("1e-6", 1e-6), ("1.5e-6", 1.5e-6), ("2e-4", 2e-4)):
            res["ref_cmp_rtol_" + name] = orc.verify_result(got, g32, rtol=rtol)[1]
    iters = case.get("iters", 20)
    if iters:
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            run()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / iters
        res["us"] = us
        res["tflops"] = 2.0 * m * n * k / us / 1e6
        eb, ob = 2, out.element_size()
        res["gbs"] = (m * k * eb + n * k * eb + m * n * ob) / us / 1e3
    print("RESULT " + json.dumps(res), flush=True)


def P(ks, cta, n1, **kw):
    d = {"kernelSerial": ks, "ctaGroup": cta, "n1": n1}
    d.update(kw)
    return d


SUITES = {
    "bringup": [
        # smallest thing first: 1-CTA, one tile, one k-block ... then widen
        {"m": 128, "n": 128, "k": 64, "plan": P(0, 1, 128), "iters": 0},
        {"m": 128, "n": 256, "k": 512, "plan": P(0, 1, 256), "iters": 0},
        {"m": 256, "n": 256, "k": 512, "plan": P(0, 2, 256), "iters": 0},
        {"m": 1024, "n": 1024, "k": 1024, "dtype": "fp16", "plan": P(0, 1, 128)},
        {"m": 1024, "n": 1024, "k": 1024, "dtype": "fp16", "plan": P(0, 2, 256)},
        {"m": 1024, "n": 1024, "k": 1024, "dtype": "fp16"},
        {"m": 1000, "n": 1000, "k": 1000, "plan": P(0, 2, 224)},
        {"m": 300, "n": 520, "k": 136, "plan": P(0, 1, 96)},
        {"m": 4096, "n": 7168, "k": 2048, "plan": P(0, 1, 256)},
        {"m": 4096, "n": 7168, "k": 2048, "plan": P(0, 2, 256)},
        {"m": 4096, "n": 7168, "k": 2048, "plan": P(0, 2, 224)},
        {"m": 4096, "n": 7168, "k": 2048},
        # framework layout: row-major B, fp32 out
        {"m": 128, "n": 128, "k": 64, "layout_b": 0, "out": "fp32", "plan": P(0, 1, 128), "iters": 0},
        {"m": 512, "n": 1024, "k": 512, "layout_b": 0, "out": "fp32", "dtype": "fp16"},
        {"m": 1024, "n": 1024, "k": 1024, "layout_b": 0, "out": "fp32", "dtype": "fp16", "plan": P(0, 2, 256)},
        # small-M swap-AB split-K
        {"m": 16, "n": 128, "k": 64, "plan": P(3, 1, 128, splitkFactor=1), "iters": 0},
        {"m": 16, "n": 256, "k": 512, "plan": P(3, 1, 128, splitkFactor=2), "iters": 0},
        {"m": 64, "n": 4096, "k": 7168, "plan": P(3, 1, 128, splitkFactor=1)},
        {"m": 64, "n": 4096, "k": 7168, "plan": P(3, 1, 128, splitkFactor=2)},
        {"m": 64, "n": 4096, "k": 7168, "plan": P(3, 1, 128, splitkFactor=4)},
        {"m": 64, "n": 4096, "k": 7168, "plan": P(3, 1, 128, splitkFactor=8)},
        {"m": 1, "n": 4096, "k": 7168},
        {"m": 16, "n": 4096, "k": 7168},
        {"m": 100, "n": 1000, "k": 1000, "plan": P(3, 1, 128, splitkFactor=4)},
        {"m": 256, "n": 4096, "k": 7168, "plan": P(3, 1, 128, splitkFactor=4)},
    ],
    "sweep": [
        {"m": mm, "n": 4096, "k": 7168, "check_rows": 64} for mm in (1, 16, 64, 128, 512, 2048, 8192)
    ],
}


def main():
    suite = sys.argv[1] if len(sys.argv) > 1 else "bringup"
    if suite == "--child":
        child(json.loads(sys.argv[2]))
        return
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    out_path = os.path.join(ROOT, "gpurun_out", "gpu_check_%s.jsonl" % suite)
    n_fail = 0
    with open(out_path, "w") as log:
        for case in SUITES[suite]:
            t0 = time.time()
            try:
                proc = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", json.dumps(case)],
                                      capture_output=True, text=True, timeout=int(os.environ.get("CASE_TIMEOUT", 180)))
                line = [ln for ln in proc.stdout.splitlines() if ln.startswith("RESULT ")]
                if line:
                    rec = json.loads(line[-1][7:])
                else:
                    rec = {"case": case, "ok": False, "rc": proc.returncode,
                           "err": (proc.stdout[-600:] + proc.stderr[-1500:])}
            except subprocess.TimeoutExpired:
                rec = {"case": case, "ok": False, "err": "timeout"}
            rec["wall_s"] = round(time.time() - t0, 1)
            n_fail += 0 if rec.get("ok") else 1
            s = json.dumps(rec)
            print(s, flush=True)
            log.write(s + "\n")
            log.flush()
    print("FAILED %d of %d" % (n_fail, len(SUITES[suite])))


if __name__ == "__main__":
    main()
