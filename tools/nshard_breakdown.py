#!/usr/bin/env python3
"""Where the fused N-shard step spends its time beyond the local GEMM (run under torchrun, 2+ GPUs):
  fused            GEMM with peer stores + cross-rank barrier          (what bench.py --gpus g times)
  kernel_only      the same launches without the barrier between them
  barrier_only     the barrier alone
  local_wide       the local GEMM into this rank's column slice of the full-width C (pitch N_total, no peer stores)
  local_compact    the local GEMM into a compact [M, N/g] C
usage: torchrun --nproc-per-node g tools/nshard_breakdown.py [M N_total K]"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepgemm_ascend_b200 as dga  # noqa: E402
from deepgemm_ascend_b200.nshard import NShardMatmul  # noqa: E402


def main():
    m, n_total, k = (int(v) for v in sys.argv[1:4]) if len(sys.argv) > 3 else (16384, 57344, 7168)
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    n_begin, n_count = dga.nshard_range(n_total, world, rank)
    gen = torch.Generator(device=dev).manual_seed(99)
    a = torch.randn((m, k), device=dev, dtype=torch.bfloat16, generator=gen)
    w = torch.randn((n_count, k), device=dev, dtype=torch.bfloat16, generator=gen)
    op = NShardMatmul(a, w, n_total)
    stream = torch.cuda.current_stream()

    def timed(fn, steps=20, warmup=3):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def fused():
        op(stream)
        op.finish()

    res = {}
    if os.environ.get("FUSED_ONLY"):
        for rnd in range(3):
            res.setdefault("fused", []).append(timed(fused))
        if rank == 0:
            print("fused %s ms (DGA_EPI_COLS=%s)" % ("  ".join("%.3f" % v for v in res["fused"]), os.environ.get("DGA_EPI_COLS", "")),
                  flush=True)
        dist.destroy_process_group()
        return
    for rnd in range(2):
        res.setdefault("fused", []).append(timed(fused))
        res.setdefault("kernel_only", []).append(timed(lambda: op(stream)))
        op.finish()
        res.setdefault("barrier_only", []).append(timed(op.finish))
        ex_wide = dga.Executor(a, w.t(), op.local_slice())
        res.setdefault("local_wide", []).append(timed(lambda: ex_wide(stream)))
        ex_wide.close()
        c_local = torch.empty((m, n_count), device=dev, dtype=torch.bfloat16)
        ex = dga.Executor(a, w.t(), c_local)
        res.setdefault("local_compact", []).append(timed(lambda: ex(stream)))
        ex.close()
        del c_local
    if rank == 0:
        for key, vals in res.items():
            print("%-14s %s ms" % (key, "  ".join("%.3f" % v for v in vals)), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
