"""dga-b200: B200-native (sm_100a) dynamic-shape matmul behind DeepGEMM_Ascend's operator surface."""
from . import _lib
from ._lib import DgaError, TilingPlan, KERNEL_NAMES
from .ops import (catlass_dynamic_matmul, get_workspace_size, execute, destroy_executor, Executor, select_kernel,
                  matmul_with_plan, kernel_launch_count, set_option, nshard_range, acl_tensor, fp8_block_scaled_matmul,
                  Fp8Executor, m_grouped_matmul, m_grouped_matmul_masked, MGroupedExecutor, jit_prepare, jit_stats, jit_drop_memory_cache,
                  splitk_matmul_no_alloc, DecodePair, decode_pair_matmul)
from .framework import run_mmad_custom, run_mmad_rtc, run_mmad_bench, get_bench_config

__all__ = ["catlass_dynamic_matmul", "get_workspace_size", "execute", "destroy_executor", "Executor",
           "select_kernel", "matmul_with_plan", "kernel_launch_count", "set_option", "nshard_range", "acl_tensor",
           "fp8_block_scaled_matmul", "Fp8Executor", "m_grouped_matmul", "m_grouped_matmul_masked", "MGroupedExecutor", "jit_prepare", "jit_stats",
           "jit_drop_memory_cache", "run_mmad_custom", "run_mmad_rtc", "run_mmad_bench", "get_bench_config", "DgaError", "TilingPlan",
           "KERNEL_NAMES", "splitk_matmul_no_alloc", "DecodePair", "decode_pair_matmul"]
