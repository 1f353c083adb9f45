"""N-sharded matmul over the GPUs of one box (BASELINE config 5; north_star item 4).

Rank r owns weight rows dgaNShardRange(N, g, r) and computes C[:, slice] = A @ W_r^T; A is replicated; there is no
reduction across ranks. When the caller needs the whole C on every rank there are two ways to get it:

  * fused (the product path): C lives in symmetric memory (torch.distributed._symmetric_memory supplies the peer
    mappings, nothing else); the GEMM epilogue TMA-stores every finished chunk into the same column slice of every
    peer's C over NVLink while the remaining tiles are still being computed (dgaNShardMatmul). A cross-rank barrier
    after the launch makes the result complete everywhere.
  * nccl (the baseline it is measured against): compact local C, then ncclAllGather into a [g, M, N/g] staging buffer.

The reference has nothing comparable: its only multi-device use is sharding the tuning grid by rank_id
(deep_gemm_ascend/framework/benchmark/benchmark.py:247-253).
"""
import ctypes

import torch

from . import _lib
from ._lib import lib, check
from .ops import acl_tensor, _stream_ptr, _require_cuda, destroy_executor, nshard_range


def shard_ranges(n, world_size):
    """[(begin, count)] per rank: contiguous, in units of 256 columns (host-only)."""
    return [nshard_range(n, world_size, r) for r in range(world_size)]


def gather_layout_to_full(staging, n, world_size):
    """[g, M, N/g]-style staging (what an all-gather of compact slices produces, ragged last slice allowed) -> [M, N].
    `staging` is a list of per-rank [M, count_r] tensors. Host-side helper, works on any device."""
    ranges = shard_ranges(n, world_size)
    m = staging[0].shape[0]
    full = staging[0].new_empty((m, n))
    for (b, c), part in zip(ranges, staging):
        full[:, b:b + c] = part[:, :c]
    return full


class NShardMatmul:
    """C = A @ W^T with W sharded over the ranks of `group` along N and the gather of C fused into the kernel."""

    def __init__(self, a, weight_shard, n_total, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        _require_cuda(a, weight_shard)
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self.n_total = int(n_total)
        self.n_begin, self.n_count = nshard_range(self.n_total, self.world, self.rank)
        if weight_shard.shape[0] != self.n_count:
            raise ValueError("weight shard must hold rows [%d, %d) of the weight" % (self.n_begin,
                                                                                      self.n_begin + self.n_count))
        m = a.shape[0]
        self.a, self.w = a, weight_shard
        # the full C of this rank, reachable from every peer
        self.c = symm_mem.empty((m, self.n_total), dtype=a.dtype, device=a.device)
        self.handle = symm_mem.rendezvous(self.c, self.group)
        ptrs = [int(p) for p in self.handle.buffer_ptrs]
        arr = (ctypes.c_void_p * self.world)(*ptrs)
        ws = ctypes.c_uint64(0)
        ex = ctypes.c_void_p(None)
        ta, tb = acl_tensor(a), acl_tensor(weight_shard.t())
        check(lib.dgaNShardMatmulGetWorkspaceSize(ctypes.byref(ta), ctypes.byref(tb), arr, self.world, self.rank,
                                                  self.n_total, self.n_total, ctypes.byref(ws), ctypes.byref(ex)),
              "dgaNShardMatmulGetWorkspaceSize")
        check(lib.dgaSetExecutorRepeatable(ex), "dgaSetExecutorRepeatable")
        self._ex = ex
        self.workspace_bytes = ws.value
        self._ws = torch.empty(max(1, ws.value), dtype=torch.uint8, device=a.device)

    def __call__(self, stream=None):
        """Asynchronous launch; the whole C is valid on every rank after `finish()`."""
        check(lib.dgaNShardMatmul(ctypes.c_void_p(self._ws.data_ptr()), self.workspace_bytes, self._ex,
                                  _stream_ptr(stream)), "dgaNShardMatmul")
        return self.c

    def finish(self):
        """Device-side barrier across the ranks on the current stream: every peer's stores have landed."""
        self.handle.barrier(channel=0)

    def local_slice(self):
        return self.c[:, self.n_begin:self.n_begin + self.n_count]

    def close(self):
        if self._ex is not None:
            destroy_executor(self._ex)
            self._ex = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
