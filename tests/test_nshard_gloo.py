"""Multi-rank host logic on CPU: world_size 2, gloo. The compute is the CPU oracle (tests only); what is checked is the
sharding arithmetic the product uses (dgaNShardRange through the C ABI) and the gather layout."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, m, k, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from deepgemm_ascend_b200.nshard import shard_ranges, gather_layout_to_full
        from oracle import matmul_oracle as orc
        ranges = shard_ranges(n, world)
        assert ranges[0][0] == 0 and sum(c for _, c in ranges) == n
        rng = np.random.RandomState(0)                       # A and W replicated by seed
        a = rng.uniform(1, 10, [m, k]).astype(np.float16)
        w = rng.uniform(1, 10, [n, k]).astype(np.float16)
        b, c = ranges[rank]
        mine = torch.from_numpy(orc.golden_fp32(a, w[b:b + c].T.copy()))    # this rank's column slice
        width = max(cnt for _, cnt in ranges)
        padded = torch.zeros((m, width))
        padded[:, :c] = mine
        parts = [torch.zeros((m, width)) for _ in range(world)]
        dist.all_gather(parts, padded)                       # what ncclAllGather does on the GPU box
        full = gather_layout_to_full(parts, n, world)
        want = torch.from_numpy(orc.golden_fp32(a, w.T.copy()))
        ok = torch.equal(full, want)
        flag = torch.tensor([1 if ok else 0])
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            out.put(int(flag.item()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [1024, 1000, 4096 + 256])
def test_nshard_gather_two_ranks(n):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, 48, 64, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert out.get(timeout=5) == 1
