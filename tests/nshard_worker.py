"""One rank of the 2-GPU N-shard parity test (tests/test_gpu_nshard.py launches it with torch.distributed.run).

Every rank runs dgaNShardMatmul (GEMM with the all-gather of C fused into the epilogue: TMA stores into the peers' C
over NVLink) and then checks ITS OWN full C - its slice and the slices its peers stored into it - against the fp64 CPU
oracle on sampled rows. Not a comparison with an NCCL gather of the library's own GEMM."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import deepgemm_ascend_b200 as dga
    from deepgemm_ascend_b200.nshard import NShardMatmul, shard_ranges
    from oracle import matmul_oracle as orc

    worst = 0.0
    # (m, n_total, k, sampled rows): a ragged shape (unequal slices, ragged tiles), then BASELINE config 5 as two of
    # eight ranks see it
    for m, n_total, k, n_rows in ((300, 1000, 520, 300), (16384, 2 * 7168, 7168, 24)):
        ranges = shard_ranges(n_total, world)
        gen = torch.Generator(device=dev).manual_seed(1000 + m)          # A replicated: same seed on every rank
        a = (torch.rand((m, k), device=dev, generator=gen) * 9 + 1).to(torch.bfloat16)
        w_all = []
        for r, (b, c) in enumerate(ranges):                               # every rank can rebuild every slice (oracle side)
            gw = torch.Generator(device=dev).manual_seed(2000 + 17 * r + m)
            w_all.append((torch.rand((c, k), device=dev, generator=gw) * 9 + 1).to(torch.bfloat16))
        op = NShardMatmul(a, w_all[rank], n_total)
        op.c.fill_(float("nan"))                                           # anything not stored shows up
        torch.cuda.synchronize()
        dist.barrier()
        before = dga.kernel_launch_count()
        op()
        op.finish()
        torch.cuda.synchronize()
        assert dga.kernel_launch_count() > before
        rows = np.random.RandomState(7 + rank).choice(m, n_rows, replace=False)   # different rows on every rank
        rows_t = torch.from_numpy(rows).to(dev)
        a_rows = a[rows_t].float().cpu().numpy()
        got = op.c[rows_t].float().cpu().numpy()
        assert np.isfinite(got).all(), "rank %d: part of C was never written" % rank
        for r, (b, c) in enumerate(ranges):
            ref = orc.golden_fp64(a_rows, w_all[r].float().cpu().numpy().T)
            err = orc.max_rel_err(got[:, b:b + c], ref)
            worst = max(worst, err)
            assert err <= 2.0 ** -7, "rank %d, slice of rank %d, shape %s: max rel err %g" % (rank, r, (m, n_total, k), err)
        op.close()
        del op
    flag = torch.tensor([1], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    print("nshard parity ok: rank %d of %d, worst max rel err vs fp64 %.3e" % (rank, world, worst), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
