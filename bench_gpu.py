"""GPU arm of bench.py (kept separate so that `bench.py --impl reference` never imports torch.cuda or the library)."""
import json
import os
import time

import numpy as np
import torch

import bench as B


def _rotation(bytes_per_set):
    """Enough distinct buffer sets that consecutive steps never find their inputs in the 126 MB L2."""
    return max(2, int(np.ceil(2.0 * B.L2_BYTES / bytes_per_set)) + 1)


def _fence(dist):
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()


def run_ours(args, rank, world):
    import deepgemm_ascend_b200 as dga   # raises if libdga_b200.so cannot be built/loaded: no fallback
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    workload = B.resolve_workload(args.workload, world)
    if workload == "config5" and world > 1:
        return run_nshard(args, rank, world, dga, dev, dist)
    shape = B.workload_shape(workload, world)
    m, n, k = shape["m"], shape["n"], shape["k"]
    fp8 = shape["dtype"] == "fp8"
    flops, nbytes = B.algorithmic(shape)
    # launches shorter than ~40 us are replayed from a CUDA graph: the Python loop cannot enqueue them fast enough
    small_m = flops < 40e9 and not fp8
    launches_per_step = 20 if small_m else 1
    sets = _rotation(nbytes)
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    sampler = B.ClockSampler(local)   # from before the warm-up to after the per-step probe pass
    sampler.start()
    if fp8:
        def quant(rows):
            x = torch.randn((rows, k), device=dev, generator=gen)
            return x.to(torch.float8_e4m3fn)
        a_bufs = [quant(m) for _ in range(sets)]
        w_bufs = [quant(n) for _ in range(sets)]
        def scales(rows):
            if workload == "fp8_fp32scales":   # arbitrary fp32 scales: software promotion kernel
                return torch.rand((rows, -(-k // 128)), device=dev, generator=gen) + 0.5
            # UE8M0 (power-of-two) scales, DeepGEMM's format on Blackwell: tensor-core block scaling
            e = torch.randint(-3, 4, (rows, -(-k // 128)), device=dev, generator=gen)
            return torch.pow(2.0, e.float())
        sa = [scales(m) for _ in range(sets)]
        sw = [scales(-(-n // 128)) for _ in range(sets)]
        c_bufs = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
        execs = [dga.Fp8Executor(a_bufs[i], sa[i], w_bufs[i], sw[i], c_bufs[i]) for i in range(sets)]
    else:
        a_bufs = [torch.randn((m, k), device=dev, dtype=torch.bfloat16, generator=gen) for _ in range(sets)]
        w_rows = n // 2 if shape.get("pair") else n
        w_bufs = [torch.randn((w_rows, k), device=dev, dtype=torch.bfloat16, generator=gen) for _ in range(sets)]  # [N,K]
        c_bufs = [torch.empty((m, n), device=dev, dtype=torch.bfloat16) for _ in range(sets)]
        if shape.get("pair"):
            # the two weights are separate allocations; C_first and C_second are the column halves of one [m, n] buffer
            h = n // 2
            w_second = [torch.randn((h, k), device=dev, dtype=torch.bfloat16, generator=gen) for _ in range(sets)]
            execs = [dga.DecodePair(a_bufs[i], w_bufs[i].t(), c_bufs[i][:, :h], w_second[i].t(), c_bufs[i][:, h:])
                     for i in range(sets)]
        elif shape.get("groups"):
            g = shape["groups"]
            w_bufs = [torch.randn((g, n, k), device=dev, dtype=torch.bfloat16, generator=gen) for _ in range(sets)]
            index = torch.arange(g, device=dev, dtype=torch.int32).repeat_interleave(m // g)
            execs = [dga.MGroupedExecutor(a_bufs[i], w_bufs[i], index, c_bufs[i], m_alignment=shape.get("m_alignment", 128))
                     for i in range(sets)]
        else:
            execs = [dga.Executor(a_bufs[i], w_bufs[i].t(), c_bufs[i]) for i in range(sets)]   # phase 1, once per set
    plan = dga.select_kernel(m, n, k, use_cache=True) if not fp8 else {"kernelSerial": 0}
    if shape.get("groups"):
        pair = shape.get("m_alignment", 128) == 256
        plan = {"kernelSerial": 0, "ctaGroup": 2 if pair else 1, "m1": 256 if pair else 128}

    stream = torch.cuda.current_stream()
    graph = None
    if small_m:
        # host enqueue cost (~10-20 us per call through Python) exceeds the kernel time: replay 20 back-to-back
        # launches as one CUDA graph
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            for i in range(sets):
                execs[i](side)
            side.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for j in range(launches_per_step):
                    execs[j % sets](side)

    def step(i):
        if graph is not None:
            graph.replay()
        else:
            execs[i % sets](stream)

    for i in range(args.warmup):
        step(i)
    _fence(dist)
    launches0 = dga.kernel_launch_count()
    # the timed region: exactly K steps between two events, nothing else on the stream (an event record between two
    # launches would also break their programmatic-dependent-launch overlap)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(args.steps):
        step(i)
    e1.record(stream)
    _fence(dist)
    launches = dga.kernel_launch_count() - launches0
    if graph is not None:
        launches = args.steps * launches_per_step   # replays do not pass through the host counter
    total_ms = e0.elapsed_time(e1)
    # per-step spread (median / min), from a second pass with an event after every step; not the headline
    n_probe = min(args.steps, 50)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n_probe + 1)]
    ev[0].record(stream)
    for i in range(n_probe):
        step(i)
        ev[i + 1].record(stream)
    _fence(dist)
    clocks = sampler.finish()
    per_step = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(n_probe))
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    n_launch = args.steps * launches_per_step
    launch_us = total_ms * 1e3 / n_launch
    value = flops * n_launch * world / (total_ms_max * 1e-3) / 1e12

    # ---- sustained: >= S seconds of back-to-back launches with the clock sampled throughout (default line only)
    sustained = None
    if world == 1 and workload == "config2" and args.sustain_seconds > 0:
        sustained = sustained_run(step, stream, launch_us * launches_per_step * 1e-6, args.sustain_seconds, flops * launches_per_step,
                                  local)

    # ---- e2e: host buffers in, host buffer out, through the public operator call
    e2e = e2e_run(dga, dev, m, n, k, steps=max(3, min(args.steps, 20)), flops=flops, world=world, dist=dist, fp8=fp8,
                  groups=shape.get("groups", 0), pair=bool(shape.get("pair")), m_alignment=shape.get("m_alignment", 128))

    # ---- optional: all-gather of C over NVLink (only when the caller needs the full output)
    gather = None
    if dist is not None:
        gather = allgather_run(dist, dev, c_bufs[0], world)

    # ---- BASELINE config 5 on this one GPU: the N = 1 anchor of the strong-scaling curve the multi-GPU lines report
    config5 = None
    if world == 1 and workload == "config2" and not args.no_config5:
        del execs, a_bufs, w_bufs, c_bufs
        torch.cuda.empty_cache()
        config5 = config5_single_gpu(dga, dev, steps=10, warmup=3)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    peaks = B.measured_peaks()
    timed_s = total_ms * 1e-3
    if m <= 128 and not fp8:
        achieved = nbytes / (launch_us * 1e-6) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": achieved / peaks["hbm_gbs"], "traffic": None,
                "peak_source": "%s hbm_gbs" % peaks["source"], "frac_of_nominal_8TBs": achieved / B.NOMINAL["hbm_gbs"]}
    else:
        which = "bf16_tflops" if timed_s < 1.0 else "bf16_tflops_sustained"
        achieved = flops / (launch_us * 1e-6) / 1e12
        # no fp8 figure is measured by the driver: the fp8 tensor rate is exactly 2x the bf16 one on this part
        peak = peaks[which] * (2.0 if fp8 else 1.0)
        roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak, "traffic": None,
                "peak_source": "%s %s%s (timed region %.2f s)" % (peaks["source"], which, " x2 for fp8" if fp8 else "",
                                                                  timed_s),
                "frac_of_nominal": achieved / (B.NOMINAL["fp8_tflops"] if fp8 else B.NOMINAL["bf16_tflops"])}
    roof["kernel"] = "gemm_fp8_sm100_kernel" if workload == "fp8_fp32scales" else \
        "gemm_smallm_sm100_kernel" if plan["kernelSerial"] == 3 else "gemm_sm100_kernel"
    roof["launch_us"] = launch_us
    roof["algorithmic_bytes"] = nbytes
    roof["algorithmic_flops"] = flops
    if workload == "repack":
        # PaddingCommon: A, W and C all pass through the workspace. The copies are overhead, not algorithmic bytes:
        # every repacked operand is read and written once more (C: written padded, read back, written out)
        extra = 2.0 * 2 * (m * k + n * k + m * n)
        roof["repack_overhead_bytes"] = extra
        roof["repack_overhead_us_at_hbm_peak"] = extra / (peaks["hbm_gbs"] * 1e9) * 1e6
        roof["launches_per_gemm"] = launches / max(1, n_launch)
    traffic_path = os.path.join(B.ROOT, "profiles", "traffic_%s.json" % workload)
    if os.path.exists(traffic_path):
        with open(traffic_path) as f:
            roof["traffic"] = json.load(f).get("dram_bytes_per_launch")
    line = {"metric": B.METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms_max / args.steps, "higher_is_better": True,
            "scaling": shape["scaling"], "vs_baseline": None, "dtype": "fp8_e4m3" if fp8 else "bf16",
            "data": "synthetic", "config": B.config_of(workload, shape, world),
            "run": {"buffer_sets": "rotating %d buffer sets (%.0f MB > 2x L2)" % (sets, sets * nbytes / 1e6),
                    "launches_per_step": launches_per_step,
                    "jit": os.environ.get("DGA_JIT", "0") not in ("", "0"),   # shape-specialised kernels
                    "plan": {key: plan.get(key) for key in ("kernelSerial", "ctaGroup", "clusterShape", "m1", "n1", "k1",
                                                            "stages", "splitkFactor", "blockDim")}},
            # single launches bracketed by their own events (no overlap with the neighbouring launches' prologue and
            # tail, which the back-to-back timed region above has): an upper bound per launch, not the bench value
            "ms_per_isolated_step_median": per_step[len(per_step) // 2], "ms_per_isolated_step_min": per_step[0],
            "roofline": roof, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
    if sustained is not None:
        sustained["frac_of_measured_sustained"] = sustained["tflops"] / peaks["bf16_tflops_sustained"]
        sustained["frac_of_nominal"] = sustained["tflops"] / B.NOMINAL["bf16_tflops"]
        line["sustained"] = sustained
    if config5 is not None:
        line["config5"] = config5
    if gather is not None:
        line["allgather"] = gather
    if not args.no_cpu_baseline:
        line["cpu_baseline"], _ = B.cpu_golden_run(shape, budget_s=10.0, max_reps=400)
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def sustained_run(step, stream, step_s, seconds, flops_per_step, device_index):
    """>= `seconds` of back-to-back steps between two events, SM clock and throttle reasons sampled every 2 ms for
    the whole run: the figure to hold against MEASURED_PEAKS' bf16_tflops_sustained (the burst figure above is
    timed over milliseconds, before the power limiter settles)."""
    n = max(100, int(seconds / max(step_s, 1e-6)) + 1)
    torch.cuda.synchronize()
    sampler = B.ClockSampler(device_index)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(n):
        step(i)
    e1.record(stream)
    torch.cuda.synchronize()
    clocks = sampler.finish()
    ms = e0.elapsed_time(e1)
    return {"seconds": ms * 1e-3, "steps": n, "tflops": flops_per_step * n / (ms * 1e-3) / 1e12, "clocks": clocks}


def config5_single_gpu(dga, dev, steps, warmup):
    """The whole 16384 x 57344 x 7168 bf16 GEMM on one GPU (A 235 MB, W 822 MB, C 1.88 GB: one set exceeds L2 many
    times over, no rotation needed)."""
    m, n, k = 16384, 57344, 7168
    gen = torch.Generator(device=dev).manual_seed(99)
    a = torch.randn((m, k), device=dev, dtype=torch.bfloat16, generator=gen)
    w = torch.randn((n, k), device=dev, dtype=torch.bfloat16, generator=gen)
    c = torch.empty((m, n), device=dev, dtype=torch.bfloat16)
    ex = dga.Executor(a, w.t(), c)
    stream = torch.cuda.current_stream()
    for _ in range(warmup):
        ex(stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        ex(stream)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    plan = dga.select_kernel(m, n, k, use_cache=True)
    flops = 2.0 * m * n * k
    ex.close()
    return {"workload": "config5: bf16 M=16384 N=57344 K=7168, whole problem on ONE GPU (the N = 1 anchor of the "
                        "strong-scaling curve)", "n_gpus": 1, "ms": ms, "tflops": flops / (ms * 1e-3) / 1e12,
            "frac_of_nominal": flops / (ms * 1e-3) / 1e12 / B.NOMINAL["bf16_tflops"], "steps": steps,
            "plan": {key: plan.get(key) for key in ("kernelSerial", "ctaGroup", "clusterShape", "n1", "stages", "blockDim")}}


def _copy_rates(dev, nbytes, dist):
    """One large pinned H2D and D2H copy on their own (GB/s per rank). With several ranks a barrier right before each
    timed copy makes them run at the same moment: the figure is what the host gives every GPU when all ask at once."""
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    devb = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    out = {}
    for name, (dst, src) in (("h2d", (devb, host)), ("d2h", (host, devb))):
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            dst.copy_(src, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        out[name] = 3 * nbytes / (e0.elapsed_time(e1) * 1e-3) / 1e9
    return out


def _e2e_rates(e2e, ms_per_step, world, dev, dist):
    """Achieved host<->device rates inside the e2e pipeline (per rank and over all ranks) and, beside them, what one
    big copy alone reaches while every rank copies at once: names the limiter of the end-to-end figure by a number."""
    h2d, d2h = e2e["h2d_bytes_per_step"], e2e["d2h_bytes_per_step"]
    alone = _copy_rates(dev, 256 << 20, dist)
    e2e["h2d_GBs_per_rank"] = h2d / (ms_per_step * 1e-3) / 1e9
    e2e["d2h_GBs_per_rank"] = d2h / (ms_per_step * 1e-3) / 1e9
    e2e["h2d_GBs_all_ranks"] = e2e["h2d_GBs_per_rank"] * world
    e2e["d2h_GBs_all_ranks"] = e2e["d2h_GBs_per_rank"] * world
    e2e["copy_alone_GBs"] = {"h2d": alone["h2d"], "d2h": alone["d2h"],
                             "note": "256 MB pinned copies, one direction at a time, every rank at the same moment (barrier), nothing else running"}
    return e2e


def e2e_run(dga, dev, m, n, k, steps, flops, world, dist, fp8=False, groups=0, pair=False, m_alignment=128):
    """Same metric end to end: every step copies its A and W from pinned host memory to the device, calls the
    operator and copies C back to pinned host memory, all inside the timed region.  The three legs run on three
    streams over a ring of 2 device buffer sets, so the copies of step i+1 and i-1 overlap the GEMM of step i
    (PCIe is full duplex); every step still moves all of its own bytes."""
    in_dtype = torch.float8_e4m3fn if fp8 else torch.bfloat16
    depth = 2
    ha = torch.randn((m, k)).to(in_dtype).pin_memory()
    hw = torch.randn((groups, n, k) if groups else (n, k)).to(in_dtype).pin_memory()
    index = torch.arange(groups, device=dev, dtype=torch.int32).repeat_interleave(m // groups) if groups else None
    hc = [torch.empty((m, n), dtype=torch.bfloat16).pin_memory() for _ in range(depth)]
    da = [torch.empty_like(ha, device=dev) for _ in range(depth)]
    dw = [torch.empty_like(hw, device=dev) for _ in range(depth)]
    dc = [torch.empty((m, n), dtype=torch.bfloat16, device=dev) for _ in range(depth)]
    extra = 0
    if fp8:
        hsa = (torch.rand((m, -(-k // 128))) + 0.5).pin_memory()
        hsw = (torch.rand((-(-n // 128), -(-k // 128))) + 0.5).pin_memory()
        dsa = [torch.empty_like(hsa, device=dev) for _ in range(depth)]
        dsw = [torch.empty_like(hsw, device=dev) for _ in range(depth)]
        extra = hsa.numel() * 4 + hsw.numel() * 4
    main = torch.cuda.current_stream()
    s_in, s_mm, s_out = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
    in_done = [torch.cuda.Event() for _ in range(depth)]
    mm_done = [torch.cuda.Event() for _ in range(depth)]
    out_done = [torch.cuda.Event() for _ in range(depth)]

    def one(j):
        i = j % depth
        with torch.cuda.stream(s_in):
            s_in.wait_event(mm_done[i])          # the GEMM that last read this slot's inputs
            da[i].copy_(ha, non_blocking=True)
            dw[i].copy_(hw, non_blocking=True)
            if fp8:
                dsa[i].copy_(hsa, non_blocking=True)
                dsw[i].copy_(hsw, non_blocking=True)
            in_done[i].record(s_in)
        with torch.cuda.stream(s_mm):
            s_mm.wait_event(in_done[i])
            s_mm.wait_event(out_done[i])         # the copy-out that last read this slot's C
            if fp8:
                dga.fp8_block_scaled_matmul(da[i], dsa[i], dw[i], dsw[i], out=dc[i])
            elif groups:
                dga.m_grouped_matmul(da[i], dw[i], index, out=dc[i], m_alignment=m_alignment)
            elif pair:   # the two weights are the row halves of the host buffer, the two outputs the column halves
                h = n // 2
                dga.decode_pair_matmul(da[i], dw[i][:h].t(), dw[i][h:].t(), dc[i][:, :h], dc[i][:, h:])
            else:
                dga.catlass_dynamic_matmul(da[i], dw[i].t(), out=dc[i])     # two-phase C ABI underneath
            mm_done[i].record(s_mm)
        with torch.cuda.stream(s_out):
            s_out.wait_event(mm_done[i])
            hc[i].copy_(dc[i], non_blocking=True)
            out_done[i].record(s_out)

    def drain():
        for st in (s_in, s_mm, s_out):
            main.wait_stream(st)

    for j in range(3):
        one(j)
    drain()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    for st in (s_in, s_mm, s_out):
        st.wait_stream(main)
    for j in range(steps):
        one(j)
    drain()
    e1.record(main)
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    checksum = float(sum(h.float().abs().mean() for h in hc) / depth)   # the device->host results are really read
    out = {"value": flops * steps * world / (float(ms.item()) * 1e-3) / 1e12, "unit": "TFLOP/s",
           "h2d_bytes_per_step": int((ha.numel() + hw.numel()) * ha.element_size() + extra),
           "d2h_bytes_per_step": int(hc[0].numel() * 2), "steps": steps, "result_abs_mean": checksum,
           "pipeline": "3 streams (H2D | GEMM | D2H) over %d buffer sets" % depth}
    return _e2e_rates(out, float(ms.item()) / steps, world, dev, dist)


def allgather_run(dist, dev, c_local, world):
    """NCCL all-gather of the rank's C slice into a [g, M, N/g] staging buffer (contiguous messages)."""
    full = torch.empty((world,) + tuple(c_local.shape), dtype=c_local.dtype, device=dev)
    for _ in range(3):
        dist.all_gather_into_tensor(full, c_local)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    e0.record()
    for _ in range(reps):
        dist.all_gather_into_tensor(full, c_local)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / reps], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    recv = c_local.numel() * c_local.element_size() * (world - 1)
    return {"ms": float(ms.item()), "recv_bytes_per_rank": int(recv),
            "recv_GBs_per_rank": recv / (float(ms.item()) * 1e-3) / 1e9}


def run_nshard(args, rank, world, dga, dev, dist):
    """BASELINE config 5 on `world` > 1 ranks: M=16384, N=57344, K=7168, weight N-sharded, every rank ends up with the
    whole C. A step = fused GEMM + gather (NShardMatmul) + the cross-rank barrier that makes C complete everywhere.
    STRONG scaling: the same 13.47 TFLOP whatever the rank count; the line carries the 1-GPU anchor (the whole GEMM
    on this rank's GPU alone, measured in this process) and the efficiencies t1 / (g * t_g) against it."""
    from deepgemm_ascend_b200.nshard import NShardMatmul
    shape = B.workload_shape("config5", world)
    m, k, n_total = shape["m"], shape["k"], shape["n_total"]
    n_begin, n_count = dga.nshard_range(n_total, world, rank)
    local = int(os.environ.get("LOCAL_RANK", "0"))
    sampler = B.ClockSampler(local)
    sampler.start()
    gen = torch.Generator(device=dev).manual_seed(99)          # A is replicated: same seed on every rank
    a = torch.randn((m, k), device=dev, dtype=torch.bfloat16, generator=gen)
    gen_w = torch.Generator(device=dev).manual_seed(1000 + rank)
    w = torch.randn((n_count, k), device=dev, dtype=torch.bfloat16, generator=gen_w)
    op = NShardMatmul(a, w, n_total)
    stream = torch.cuda.current_stream()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        _fence(dist)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        _fence(dist)
        t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def fused():
        op(stream)
        op.finish()

    steps = min(args.steps, 50)
    launches0 = dga.kernel_launch_count()
    fused_ms = timed(fused, steps, args.warmup)
    launches = dga.kernel_launch_count() - launches0
    clocks = sampler.finish()

    # what it is measured against: compact local GEMM (compute only), NCCL all-gather of C, and the two in sequence
    c_local = torch.empty((m, n_count), device=dev, dtype=torch.bfloat16)
    staging = torch.empty((world, m, n_count), device=dev, dtype=torch.bfloat16)
    ex = dga.Executor(a, w.t(), c_local)
    compute_ms = timed(lambda: ex(stream), steps, 3)
    gather_ms = timed(lambda: dist.all_gather_into_tensor(staging, c_local), min(steps, 10), 3)

    def unfused():
        ex(stream)
        dist.all_gather_into_tensor(staging, c_local)
    unfused_ms = timed(unfused, min(steps, 10), 3)
    ex.close()
    del staging, c_local

    # the 1-GPU anchor: the whole problem on this rank's GPU alone (every rank does it at the same time; rank 0's
    # time is reported).  C goes into the symmetric buffer, which is free between fused steps.
    w_full = torch.randn((n_total, k), device=dev, dtype=torch.bfloat16, generator=gen_w)
    ex1 = dga.Executor(a, w_full.t(), op.c)
    t1_ms = timed(lambda: ex1(stream), 10, 3)
    ex1.close()
    del w_full

    e2e = nshard_e2e(dga, NShardMatmul, dev, dist, rank, world, m, k, n_total, steps=max(3, min(args.steps, 8)))

    if rank == 0:
        flops = 2.0 * m * n_total * k
        peaks = B.measured_peaks()
        per_rank_tflops = flops / world / (fused_ms * 1e-3) / 1e12
        sent = m * n_count * 2 * (world - 1)
        recv = m * (n_total - n_count) * 2
        link_ms = recv / 770e9 * 1e3            # measured peer-copy bandwidth per direction (B200_PROFILING.md)
        gemm_ms = flops / world / (peaks["bf16_tflops_sustained"] * 1e12) * 1e3
        target_ms = max(link_ms, gemm_ms)
        plan = dga.select_kernel(m, n_count, k)
        line = {"metric": B.METRIC, "value": flops / (fused_ms * 1e-3) / 1e12, "unit": "TFLOP/s", "n_gpus": world,
                "steps": steps, "warmup": args.warmup, "ms_per_step": fused_ms, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": B.config_of("config5", shape, world),
                "run": {"step": "GEMM with the all-gather of C fused into the epilogue (TMA stores to peer C over "
                                "NVLink) + cross-rank barrier; inputs (A 235 MB, W shard) exceed L2",
                        "plan": {key: plan[key] for key in ("kernelSerial", "ctaGroup", "clusterShape", "m1", "n1", "stages",
                                                            "blockDim")}},
                "roofline": {"bound": "nvlink" if link_ms >= gemm_ms else "tensor", "achieved": per_rank_tflops,
                             "peak": flops / world / (target_ms * 1e-3) / 1e12, "unit": "TFLOP/s per GPU",
                             "frac": target_ms / fused_ms, "traffic": None, "kernel": "gemm_sm100_kernel (peer stores)",
                             "target_ms": target_ms, "gemm_ms_at_measured_peak": gemm_ms, "nvlink_ms_at_770GBs": link_ms,
                             "launch_us": fused_ms * 1e3, "algorithmic_flops": flops / world,
                             "algorithmic_bytes_over_nvlink_in": float(recv),
                             "peak_source": "slower of FLOPs / measured sustained cuBLAS and bytes received / 770 GB/s"},
                "config5": {"workload": "config5 strong-scaled over %d GPUs, every rank ends with the whole C" % world,
                            "n_gpus": world, "fused_ms": fused_ms, "compute_only_ms": compute_ms,
                            "nccl_allgather_ms": gather_ms, "gemm_then_nccl_ms": unfused_ms,
                            "one_gpu_ms": t1_ms, "one_gpu_tflops": flops / (t1_ms * 1e-3) / 1e12,
                            "strong_efficiency_compute_only": t1_ms / (world * compute_ms),
                            "strong_efficiency_fused": t1_ms / (world * fused_ms),
                            "strong_efficiency_gemm_then_nccl": t1_ms / (world * unfused_ms),
                            "link_floor_ms": link_ms, "fused_frac_of_link_floor": link_ms / fused_ms,
                            "bytes_sent_per_rank": int(sent), "bytes_received_per_rank": int(recv),
                            "tflops_per_gpu_fused": per_rank_tflops,
                            "tflops_per_gpu_compute_only": flops / world / (compute_ms * 1e-3) / 1e12},
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
        if not args.no_cpu_baseline:
            line["cpu_baseline"], _ = B.cpu_golden_run(shape, budget_s=10.0, max_reps=400)
        print(json.dumps(line), flush=True)
    op.close()
    dist.destroy_process_group()


def nshard_e2e(dga, NShardMatmul, dev, dist, rank, world, m, k, n_total, steps):
    """Config 5 end to end on `world` ranks: every step each rank copies A (replicated, 235 MB) and its weight shard
    from pinned host memory, runs the fused GEMM + gather, and copies ITS column slice of C back to pinned host memory
    (the host then holds the whole C once, spread over the ranks' buffers).  Two buffer sets; a set's C is written by
    the peers too, so a cross-rank barrier precedes every GEMM: it orders the peers' device->host copy of that set
    (two steps earlier) before this rank's stores into it."""
    n_begin, n_count = dga.nshard_range(n_total, world, rank)
    depth = 2
    ha = torch.randn((m, k)).to(torch.bfloat16).pin_memory()
    hw = torch.randn((n_count, k)).to(torch.bfloat16).pin_memory()
    hc = [torch.empty((m, n_count), dtype=torch.bfloat16).pin_memory() for _ in range(depth)]
    da = [torch.empty((m, k), dtype=torch.bfloat16, device=dev) for _ in range(depth)]
    dw = [torch.empty((n_count, k), dtype=torch.bfloat16, device=dev) for _ in range(depth)]
    ops = [NShardMatmul(da[i], dw[i], n_total) for i in range(depth)]
    main = torch.cuda.current_stream()
    s_in, s_mm, s_out = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
    in_done = [torch.cuda.Event() for _ in range(depth)]
    mm_done = [torch.cuda.Event() for _ in range(depth)]
    out_done = [torch.cuda.Event() for _ in range(depth)]

    def one(j):
        i = j % depth
        with torch.cuda.stream(s_in):
            s_in.wait_event(mm_done[i])
            da[i].copy_(ha, non_blocking=True)
            dw[i].copy_(hw, non_blocking=True)
            in_done[i].record(s_in)
        with torch.cuda.stream(s_mm):
            s_mm.wait_event(in_done[i])
            s_mm.wait_event(out_done[i])
            ops[i].handle.barrier(channel=1)     # every rank's copy-out of this set (two steps ago) is behind us
            ops[i](s_mm)
            ops[i].finish()
            mm_done[i].record(s_mm)
        with torch.cuda.stream(s_out):
            s_out.wait_event(mm_done[i])
            hc[i].copy_(ops[i].local_slice(), non_blocking=True)
            out_done[i].record(s_out)

    def drain():
        for st in (s_in, s_mm, s_out):
            main.wait_stream(st)

    for j in range(2):
        one(j)
    drain()
    _fence(dist)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    for st in (s_in, s_mm, s_out):
        st.wait_stream(main)
    for j in range(steps):
        one(j)
    drain()
    e1.record(main)
    _fence(dist)
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    checksum = float(sum(h.float().abs().mean() for h in hc) / depth)
    flops = 2.0 * m * n_total * k
    out = {"value": flops * steps / (float(ms.item()) * 1e-3) / 1e12, "unit": "TFLOP/s",
           "h2d_bytes_per_step": int((ha.numel() + hw.numel()) * 2), "d2h_bytes_per_step": int(hc[0].numel() * 2),
           "bytes_are": "per rank; every rank copies A (replicated) and its own weight shard in, its own C slice out",
           "steps": steps, "result_abs_mean": checksum,
           "pipeline": "3 streams (H2D | fused GEMM + gather + barrier | D2H) over %d symmetric buffer sets" % depth}
    out = _e2e_rates(out, float(ms.item()) / steps, world, dev, dist)
    for o in ops:
        o.close()
    return out
