"""`GetBestConfig.predict(m, n, k, layout_tag_a, layout_tag_b)` for B200: the Python entry of the reference's tuner
(get_best_config/get_best_config.py:673-748, which `Predictor::Predict` calls back into,
aclnn_catlass_dynamic_matmul/.../predictor.cpp:140-157) over this library's C++ tile selector.

The reference's three steps and what stands in their place here:
  1. `MatmulTilingCalculator.calculate` (tiling_calculator.py:452-689: operator type, tile, padding tags, blockDim) ->
     `dgaSelectKernel`, the C++ selector of csrc/tile_selector.cpp;
  2. the MLP time predictors per operator type (get_best_config.py:676-708; their weights are not in the reference
     tree) -> the measured tuned table: a plan recorded by `deepgemm_ascend_b200.autotune` in `$CACHE_FILE_PATH`
     wins over the analytic choice (`dgaSelectKernelWithCache`, first writer wins like cache.cpp:69-77);
  3. the fall-back to the analytic tile when the model has no opinion (get_best_config.py:725-726) -> the same:
     without a table entry `predict_tiling` equals `tiling`.
The result carries the reference's keys with the reference's meaning (kernelSerial 0-4 and their operator names as in
tiling_calculator.py:415-449; 5 is this library's split-K + reduce plan) plus the B200-only fields a caller needs to
reproduce the launch (`ctaGroup`, `stages`, `smemBytes`, `workspaceBytes`, `estTimeUs`).  Pure host code: works
without a GPU.
"""
import torch

from . import ops

OPERATOR_TYPES = {
    0: "CommonMatmul",
    1: "SmallMatmul",
    2: "PaddingCommonMatmul",
    3: "PaddingMultiCoreSplitkMatmul",
    4: "PaddingStreamkMatmul",
    5: "PaddingMultiCoreSplitkReduceMatmul",
    6: "ClusterSplitkMatmul",
}
DISPATCH_POLICY_TAGS = {1: "DynamicSmall", 4: "DynamicStreamk"}   # everything else: "DynamicCommon"


def _describe(plan, m, n, k, layout_tag_a, layout_tag_b):
    serial = plan["kernelSerial"]
    tile = (plan["m1"], plan["n1"], plan["k1"])
    return {
        "tiling": tile,
        "operator_type": OPERATOR_TYPES[serial],
        "shape": (m, n, k),
        "layout": (layout_tag_a, layout_tag_b),
        "original_m1": tile[0], "original_n1": tile[1], "original_k1": tile[2],
        "splitkFactor": plan["splitkFactor"],
        "paddingTagA": plan["paddingTagA"], "paddingTagB": plan["paddingTagB"], "paddingTagC": plan["paddingTagC"],
        "blockDim": plan["blockDim"],
        "kernelSerial": serial,
        "dispatchPolicyTag": DISPATCH_POLICY_TAGS.get(serial, "DynamicCommon"),
        "ctaGroup": plan["ctaGroup"], "clusterShape": plan["clusterShape"], "stages": plan["stages"],
        "smemBytes": plan["smemBytes"],
        "workspaceBytes": plan["workspaceBytes"], "estTimeUs": plan["estTimeUs"],
    }


class GetBestConfig:
    """Same call as the reference's class of this name; `dtype` is the one thing the reference leaves implicit
    (its tuner is fp16-only, benchmark.py:343-367)."""

    def __init__(self, dtype=torch.bfloat16, out_dtype=None):
        self.dtype = dtype
        self.out_dtype = out_dtype

    def predict(self, m, n, k, layout_tag_a, layout_tag_b):
        """layout tags as in the reference: 0 row-major, 1 column-major.  A must be row-major: that is the operator's
        contract (catlass_dynamic_matmul_tiling.cpp:82-84 fixes A = RowMajor), so tag 1 for A is an error here rather
        than a silently different plan."""
        if layout_tag_a != 0:
            raise ValueError("the B200 operator takes a row-major A (layout_tag_a = 0)")
        if layout_tag_b not in (0, 1):
            raise ValueError("layout_tag_b must be 0 (row-major [K,N]) or 1 (column-major, weight stored [N,K])")
        kw = dict(dtype=self.dtype, out_dtype=self.out_dtype, layout_b=layout_tag_b)
        analytic = ops.select_kernel(m, n, k, use_cache=False, **kw)
        tuned = ops.select_kernel(m, n, k, use_cache=True, **kw)
        result = _describe(tuned, m, n, k, layout_tag_a, layout_tag_b)
        # the reference keeps the calculator's tile beside the predicted one
        result["tiling"] = (analytic["m1"], analytic["n1"], analytic["k1"])
        result["original_m1"], result["original_n1"], result["original_k1"] = result["tiling"]
        result["predict_tiling"] = {"m1": tuned["m1"], "n1": tuned["n1"], "k1": tuned["k1"]}
        result["m1"], result["n1"], result["k1"] = tuned["m1"], tuned["n1"], tuned["k1"]
        return result


def main():
    """The reference's own two demo calls (get_best_config.py:751-761)."""
    cfg = GetBestConfig()
    for args in [dict(m=72, n=7392, k=8192, layout_tag_a=0, layout_tag_b=1), dict(m=1, n=4, k=4, layout_tag_a=0, layout_tag_b=0)]:
        for key, value in cfg.predict(**args).items():
            print("%s: %s" % (key, value))


if __name__ == "__main__":
    main()
