#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200 dynamic-shape matmul (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--sustain-seconds S]

A step is one pass of the hot path over one batch: one GEMM launch through the reference-facing two-phase C ABI
(aclnnCatlassDynamicMatmulGetWorkspaceSize once, aclnnCatlassDynamicMatmul per step).
Workloads (BASELINE.json configs):
  auto (default)     one GPU: config2, the configuration BASELINE.json's metric is quoted on.  --gpus g > 1: config5,
                     the N-sharded shape north_star item 4 names, STRONG-scaled over the g ranks.  Either way the line
                     carries a "config5" block with the 1-GPU anchor (the whole 16384 x 57344 x 7168 GEMM on one GPU,
                     measured in the same process), so the strong-scaling curve has its N = 1 point at every N.
  config2            bf16 M=4096 N=7168 K=2048, weight stored [N,K]; with --gpus g the weight is N-sharded:
                     every rank owns a 7168-column slice of a 7168*g wide weight (weak scaling, no data-path
                     collective; the optional all-gather of C is timed separately and reported under "allgather")
  config5            bf16 M=16384 N=57344 K=7168; one GPU: the whole GEMM; g ranks: weight N-sharded, a step = GEMM
                     with the all-gather of C fused into its epilogue (dgaNShardMatmul) + cross-rank barrier
  repack             bf16 1279 x 5003 x 7681, the reference's first odd benchmark shape (benchmark.py:41-43): no operand
                     has a 16-byte aligned pitch, all three go through the stride repack (PaddingCommon)
  m<M>               bf16 M in {1,16,...} at N=4096 K=7168 (dynamic-M sweep point; CUDA-graph of 20 launches per step)
  fp8                e4m3 block-scaled M=4096 N=7168 K=16384, UE8M0 (power-of-two) scales -> tensor-core block scaling
  fp8_fp32scales     same shape, arbitrary fp32 scales -> software promotion kernel
  moe                M-grouped contiguous bf16 (8 experts x 1024 rows, N=4096, K=7168; SURVEY 8(f) f4, not a BASELINE config)
Timing: CUDA events on the launching stream, K timed steps after W >= 3 warm-ups, barrier + synchronize on both
sides, max over ranks. Buffers rotate over enough distinct (A, B, C) sets to exceed 2x the 126 MB L2.
The oracle (oracle/) is used only for the cpu_baseline leg and for --impl reference.
"""
import argparse
import json
import re
import os
import sys
import threading
import time

# torch.distributed.run exports OMP_NUM_THREADS=1 to every rank it starts; the CPU golden (the reference arm, and the
# cpu_baseline leg of ours) runs "with all the host threads it can use", so the BLAS pool must be sized before numpy
# loads it
for _var in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
    os.environ.pop(_var, None)

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "GEMM TFLOP/s on DeepSeek-V3 shapes (bf16 dense, fp32 accumulate)"
L2_BYTES = 126 << 20
NOMINAL = {"bf16_tflops": 2250.0, "fp8_tflops": 4500.0, "hbm_gbs": 8000.0}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        d["source"] = "measured"
        return d
    # fallback stated by /opt/skills/guides/B200_PROFILING.md
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def resolve_workload(name, world):
    """"auto": BASELINE.json's headline configuration on one GPU, its N-sharded configuration on several."""
    if name == "auto":
        return "config2" if world == 1 else "config5"
    return name


def workload_shape(name, world):
    name = resolve_workload(name, world)
    if name == "repack":
        return dict(m=1279, n=5003, k=7681, dtype="bf16", n_total=5003 * world, scaling="weak")
    if name == "config2":
        return dict(m=4096, n=7168, k=2048, dtype="bf16", n_total=7168 * world, scaling="weak")
    if name == "config5":
        return dict(m=16384, n=57344 // world, k=7168, dtype="bf16", n_total=57344, scaling="strong")
    if name in ("fp8", "fp8_fp32scales"):
        return dict(m=4096, n=7168, k=16384, dtype="fp8", n_total=7168 * world, scaling="weak")
    if name == "moe256":   # the same MoE problem with the groups promised to start on multiples of 256 rows: pair tiles
        return dict(m=8192, n=4096, k=7168, dtype="bf16", n_total=4096 * world, scaling="weak", groups=8, m_alignment=256)
    if name == "moe":
        return dict(m=8192, n=4096, k=7168, dtype="bf16", n_total=4096 * world, scaling="weak", groups=8)
    if re.fullmatch(r"m\d+pair", name):    # two decode GEMMs (N = 4096 each) sharing X in one launch: dgaDecodePairMatmul
        return dict(m=int(name[1:-4]), n=8192, k=7168, dtype="bf16", n_total=8192 * world, scaling="weak", pair=True)
    if re.fullmatch(r"m\d+n\d+", name):   # decode-sized probes at another N (same K), e.g. m64n8192
        mm, nn = (int(x) for x in name[1:].split("n"))
        return dict(m=mm, n=nn, k=7168, dtype="bf16", n_total=nn * world, scaling="weak")
    if name.startswith("m") and name[1:].isdigit():
        return dict(m=int(name[1:]), n=4096, k=7168, dtype="bf16", n_total=4096 * world, scaling="weak")
    raise SystemExit("unknown workload %s" % name)


def algorithmic(shape):
    """FLOPs = 2MNK; bytes(16-bit) = 2(MK+NK+MN); bytes(fp8) adds the scale tensors."""
    m, n, k = shape["m"], shape["n"], shape["k"]
    flops = 2.0 * m * n * k
    if shape["dtype"] == "fp8":
        kb, nb = -(-k // 128), -(-n // 128)
        nbytes = m * k + n * k + 2 * m * n + 4 * m * kb + 4 * nb * kb
    else:
        nbytes = 2.0 * (m * k + shape.get("groups", 1) * n * k + m * n)   # moe: every expert's weight is read once
    return flops, float(nbytes)


# ------------------------------------------------------------------------------------------- clocks sampler
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.max_mhz, self.power = [], set(), None, []
        self._stop_evt = threading.Event()
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._h = None

    def run(self):
        if self._h is None:
            return
        nv = self._nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                self.power.append(nv.nvmlDeviceGetPowerUsage(self._h) / 1000.0)
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.002)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "power_w_max": max(self.power) if self.power else None, "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------- CPU baseline
_GOLDEN_INPUTS = {}   # inputs of the last sample (generated once per run, not per step)


def cpu_golden_run(shape, budget_s, max_reps):
    """The reference's CPU golden for this path: np.matmul of the fp16 inputs widened to fp32
    (deep_gemm_ascend/scripts/gen_golden.py:14-15), restated in oracle.matmul_oracle.golden_fp32, on all host cores.
    The whole problem when one repetition is <= 1.5 TFLOP (config 2: all 4096 rows, ~0.13 s); beyond that a bounded
    sample: M is cut (to a multiple of 128 rows) so that one repetition is ~1.2 TFLOP over the full N and K (config 5:
    1408 of 16384 rows, 1-2 s; a thinner slab would time the fp16 -> fp32 widening of the 822 MB weight, not the
    GEMM). Repeated until the budget is used."""
    from oracle import matmul_oracle as orc
    m, n, k = shape["m"], shape["n_total"] if shape["scaling"] == "strong" else shape["n"], shape["k"]
    rows = m if 2.0 * m * n * k <= 1.5e12 else max(1, min(m, int(1.2e12 / (2.0 * n * k)) // 128 * 128))
    if _GOLDEN_INPUTS.get("key") != (rows, n, k):
        rng = np.random.RandomState(0)
        _GOLDEN_INPUTS.update(key=(rows, n, k), x1=rng.uniform(1, 10, [rows, k]).astype(np.float16),
                              x2=rng.random_sample([k, n]).astype(np.float16) * np.float16(9) + np.float16(1))
        orc.golden_fp32(_GOLDEN_INPUTS["x1"][: max(1, rows // 8)], _GOLDEN_INPUTS["x2"])  # warm the BLAS threads
    x1, x2 = _GOLDEN_INPUTS["x1"], _GOLDEN_INPUTS["x2"]
    t0 = time.perf_counter()
    reps = 0
    while reps < max_reps and (time.perf_counter() - t0 < budget_s or reps == 0):
        orc.golden_fp32(x1, x2)
        reps += 1
    dt = time.perf_counter() - t0
    tflops = 2.0 * rows * n * k * reps / dt / 1e12
    try:
        from threadpoolctl import threadpool_info
        threads = max([p.get("num_threads", 1) for p in threadpool_info()] or [os.cpu_count()])
    except Exception:
        threads = os.cpu_count()
    return {"value": tflops, "unit": "TFLOP/s", "cores": int(threads), "kind": "port",
            "sample": "np.matmul(fp16->fp32) golden, %d of %d rows x N=%d x K=%d, %d reps in %.1f s"
                      % (rows, m, n, k, reps, dt)}, dt / reps


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (its numpy golden), all host threads,
    bounded sample per step. Rank 0 only."""
    if rank != 0:
        return
    shape = workload_shape(args.workload, world)
    base, _ = cpu_golden_run(shape, budget_s=2.0, max_reps=2)  # warm-up, also sizes the sample
    per_step = []
    rows = None
    for i in range(args.warmup + args.steps):
        res, dt = cpu_golden_run(shape, budget_s=0.0, max_reps=1)
        if i >= args.warmup:
            per_step.append((res["value"], dt))
        rows = res["sample"]
    value = float(np.mean([v for v, _ in per_step]))
    ms = float(np.mean([d for _, d in per_step])) * 1e3
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": shape["scaling"], "vs_baseline": None, "dtype": "f32 (CPU golden of fp16/bf16 inputs)",
            "data": "synthetic", "config": config_of(args.workload, shape, world),
            "cpu_baseline": {"value": value, "unit": "TFLOP/s", "cores": base["cores"], "kind": "port",
                             "sample": rows},
            "e2e": {"value": value, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def describe(name, shape, world):
    return "%s: %s M=%d N=%d K=%d per GPU (N_total=%d over %d GPU%s), weight stored [N,K]" % (
        resolve_workload(name, world), shape["dtype"], shape["m"], shape["n"], shape["k"], shape["n_total"], world,
        "" if world == 1 else "s")


def config_of(name, shape, world):
    """The `config` object: what is measured, the same in both arms (arm-specific detail goes under "run")."""
    return {"workload": describe(name, shape, world), "m": shape["m"], "n_per_gpu": shape["n"], "k": shape["k"],
            "n_total": shape["n_total"], "scaling": shape["scaling"],
            "l2_policy": "GPU arm: operands of consecutive steps rotate over buffer sets whose total exceeds 2x the "
                         "126 MB L2, or one set exceeds it (config5); CPU arm: operands exceed the host's last-level cache"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sustain-seconds", type=float, default=2.0,
                    help="length of the back-to-back run behind the `sustained` block of the default line (0 = skip)")
    ap.add_argument("--no-config5", action="store_true", help="skip the 1-GPU config-5 anchor of the default line")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and args.gpus > 1:
        # not under torchrun: re-launch ourselves one rank per GPU
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29511"),
               os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    from bench_gpu import run_ours
    run_ours(args, rank, world)


if __name__ == "__main__":
    main()
