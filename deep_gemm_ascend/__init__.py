"""Drop-in for the reference's Python package `deep_gemm_ascend`
(deep_gemm_ascend/framework/deep_gemm_ascend/__init__.py:1-5): the same three names, CUDA tensors."""
from deepgemm_ascend_b200.framework import run_mmad_custom, run_mmad_rtc, run_mmad_bench

__all__ = ["run_mmad_custom", "run_mmad_rtc", "run_mmad_bench"]
