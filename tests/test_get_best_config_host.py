"""Python entry of the tuner (`GetBestConfig.predict`) over the C++ selector: the reference's result keys and meaning
(get_best_config/get_best_config.py:712-748; tiling_calculator.py:415-449), the tuned table standing in for the MLP
predictors, the fall-back to the analytic tile.  No GPU."""
import ctypes

import pytest

# keys of MatmulTilingCalculator.calculate() plus what predict() adds; listed by running the reference's
# `calculate(512, 4096, 7168, 0, 1)` (tiling_calculator.py:452-689) and reading get_best_config.py:725-748
REFERENCE_KEYS = {"tiling", "operator_type", "shape", "layout", "original_m1", "original_n1", "original_k1",
                  "splitkFactor", "paddingTagA", "paddingTagB", "paddingTagC", "blockDim", "kernelSerial",
                  "dispatchPolicyTag", "predict_tiling", "m1", "n1", "k1"}


def test_predict_returns_the_reference_keys_and_the_selector_plan(dga):
    from deepgemm_ascend_b200.get_best_config import GetBestConfig, OPERATOR_TYPES
    dga._lib.lib.dgaTilingCacheClear()
    cfg = GetBestConfig()
    for m, n, k, lb in [(72, 7392, 8192, 1), (1, 4, 4, 0), (4096, 7168, 2048, 1), (512, 4096, 7168, 1), (64, 4096, 7168, 1),
                        (1000, 1000, 1000, 0)]:
        r = cfg.predict(m, n, k, 0, lb)
        assert REFERENCE_KEYS <= set(r)
        plan = dga.select_kernel(m, n, k, layout_b=lb)
        assert r["shape"] == (m, n, k) and r["layout"] == (0, lb)
        assert (r["m1"], r["n1"], r["k1"]) == (plan["m1"], plan["n1"], plan["k1"]) == tuple(r["tiling"])
        assert r["predict_tiling"] == {"m1": plan["m1"], "n1": plan["n1"], "k1": plan["k1"]}   # no table entry: fall-back
        assert r["kernelSerial"] == plan["kernelSerial"] and r["operator_type"] == OPERATOR_TYPES[plan["kernelSerial"]]
        assert r["blockDim"] == plan["blockDim"] and r["splitkFactor"] == plan["splitkFactor"]
        assert r["dispatchPolicyTag"] == {1: "DynamicSmall", 4: "DynamicStreamk"}.get(plan["kernelSerial"], "DynamicCommon")
    dga._lib.lib.dgaTilingCacheClear()


def test_operator_names_follow_the_reference_serials():
    from deepgemm_ascend_b200.get_best_config import OPERATOR_TYPES
    # tiling_calculator.py:424-430
    assert [OPERATOR_TYPES[i] for i in range(5)] == ["CommonMatmul", "SmallMatmul", "PaddingCommonMatmul",
                                                      "PaddingMultiCoreSplitkMatmul", "PaddingStreamkMatmul"]


def test_a_tuned_table_entry_wins_over_the_analytic_tile(dga):
    from deepgemm_ascend_b200.get_best_config import GetBestConfig
    lib = dga._lib.lib
    lib.dgaTilingCacheClear()
    base = dga.select_kernel(2048, 4096, 7168)
    plan = dga._lib.TilingPlan()
    for key, val in base.items():
        setattr(plan, key, val)
    plan.n1 = 128 if base["n1"] != 128 else 64
    plan.stages = 0
    assert lib.dgaTilingCacheSet(ctypes.byref(plan)) == 0
    try:
        r = GetBestConfig().predict(2048, 4096, 7168, 0, 1)
        assert r["predict_tiling"]["n1"] == plan.n1 and r["n1"] == plan.n1
        assert tuple(r["tiling"]) == (base["m1"], base["n1"], base["k1"])   # the calculator's tile is kept beside it
    finally:
        lib.dgaTilingCacheClear()


def test_column_major_a_is_refused():
    from deepgemm_ascend_b200.get_best_config import GetBestConfig
    with pytest.raises(ValueError):
        GetBestConfig().predict(128, 128, 128, 1, 1)
    with pytest.raises(ValueError):
        GetBestConfig().predict(128, 128, 128, 0, 2)
