"""The oracle (oracle/matmul_oracle.py) against the golden vectors produced by running the reference itself
(tests/golden/make_golden.py). CPU only. Bit-exact for integers and for the fp32 numpy golden."""
import json
import os

import numpy as np
import pytest

from oracle import matmul_oracle as orc


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_uniform_golden_bit_exact(golden_dir):
    z = _load(golden_dir, "numeric_uniform.npz")
    tags = sorted({k.rsplit("_", 1)[0] for k in z.files})
    assert len(tags) == 4
    for tag in tags:
        seed = int(tag.split("_")[0][1:])
        m, n, k = (int(v) for v in tag.split("_")[1].split("x"))
        # gen_golden.py:11-15 uses the global RandomState; np.random.RandomState(seed) reproduces np.random.seed(seed)
        x1, x2, g = orc.gen_golden_data(m, n, k, np.random.RandomState(seed))
        assert x1.dtype == np.float16 and x2.dtype == np.float16 and g.dtype == np.float32
        assert np.array_equal(x1, z[tag + "_x1"])
        assert np.array_equal(x2, z[tag + "_x2"])
        assert np.array_equal(g, z[tag + "_golden"])
        assert x1.min() >= 1 and x1.max() <= 10


def test_heavy_tail_golden_bit_exact(golden_dir):
    z = _load(golden_dir, "numeric_heavy.npz")
    for seed in (0, 1):
        rng = np.random.default_rng(seed)
        x1 = orc.heavy_tail(rng, [1, 128])   # test.py:23-35: M=1, N=512, K=128, x1 drawn first
        x2 = orc.heavy_tail(rng, [128, 512])
        assert np.array_equal(x1, z["s%d_x1" % seed])
        assert np.array_equal(x2, z["s%d_x2" % seed])
        assert np.array_equal(orc.golden_fp32(x1, x2), z["s%d_golden" % seed])


def test_comparator_matches_reference_verdicts(golden_dir):
    with open(os.path.join(golden_dir, "comparator.json")) as f:
        spec = json.load(f)
    rng = np.random.RandomState(spec["seed"])
    golden = rng.uniform(spec["low"], spec["high"], spec["n"]).astype(np.float32)

    def corrupt(count):
        bad = golden.copy()
        bad[:count] *= 1.5
        return bad

    outputs = {
        "exact": golden.copy(),
        "rel_1e-7": golden * np.float32(1 + 1e-7),
        "rel_5e-6": (golden.astype(np.float64) * (1 + 5e-6)).astype(np.float32),
        "times_2": golden * 2,
        "corrupt_1pct": corrupt(200),
        "corrupt_1e-4": corrupt(2),
        "corrupt_1.5e-4": corrupt(3),
        "short": golden[:-1].copy(),
    }
    assert set(outputs) == set(spec["verdicts"])
    for tag, want in spec["verdicts"].items():
        ok, _ = orc.verify_result(outputs[tag], golden, rtol=orc.RTOL_VERIFY)
        assert ok == want, tag


def test_comparator_empty_is_pass():
    # the reference divides by zero here (verify.py:33; its own test fails); an empty result has nothing to mismatch
    ok, ratio = orc.verify_result(np.zeros(0, np.float32), np.zeros(0, np.float32))
    assert ok and ratio == 0.0


def test_bench_config_matches_reference(golden_dir):
    with open(os.path.join(golden_dir, "bench_config.json")) as f:
        rows = json.load(f)
    assert len(rows) >= 40
    for row in rows:
        m, n, k = row["shape"]
        assert orc.get_bench_config(m, n, k, *row["params"][:6]) == row["params"]


def test_fp64_vs_fp32_golden_close():
    x1, x2, g32 = orc.gen_golden_data(96, 80, 1024, np.random.RandomState(5))
    g64 = orc.golden_fp64(x1, x2)
    assert orc.max_rel_err(g32, g64) < 2e-6


def test_bf16_round_trip_and_rne():
    v = np.array([1.0, 1.00390625, 1.01171875, -2.5, 3.0e38, 0.0], dtype=np.float32)
    r = orc.round_to_bf16(v)
    assert r[0] == 1.0 and r[3] == -2.5 and r[5] == 0.0
    assert r[1] == 1.0            # 1 + 2^-8: tie -> even (1.0)
    assert r[2] == np.float32(1.015625)  # 1 + 3*2^-8: tie -> even (1 + 2^-6)
    try:
        import ml_dtypes
    except ImportError:
        return
    x = np.random.RandomState(0).randn(4096).astype(np.float32) * 100
    assert np.array_equal(orc.round_to_bf16(x), x.astype(ml_dtypes.bfloat16).astype(np.float32))


def test_e4m3_codec_against_ml_dtypes():
    ml_dtypes = pytest.importorskip("ml_dtypes")
    x = np.random.RandomState(1).randn(20000).astype(np.float32) * 60
    x = np.clip(x, -448, 448)
    want = x.astype(ml_dtypes.float8_e4m3fn)
    got = orc.to_e4m3_bits(x)
    assert np.array_equal(got, want.view(np.uint8))
    assert np.array_equal(orc.from_e4m3_bits(got), want.astype(np.float64))


def test_fp8_casts_follow_deepgemm_semantics():
    rng = np.random.RandomState(2)
    x = rng.randn(7, 300).astype(np.float32)
    codes, sf = orc.per_token_cast_to_fp8(x)
    assert codes.shape == (7, 300) and sf.shape == (7, 3)
    deq = orc.from_e4m3_bits(codes) * np.repeat(sf.astype(np.float64), 128, axis=1)[:, :300]
    assert np.max(np.abs(deq - x)) / np.max(np.abs(x)) < 0.07
    # amax of every strip maps to +-448
    strip = np.abs(orc.from_e4m3_bits(codes[:, :128])).max(axis=1)
    assert np.all(strip == 448.0)
    w = rng.randn(130, 260).astype(np.float32)
    wc, wsf = orc.per_block_cast_to_fp8(w)
    assert wc.shape == (130, 260) and wsf.shape == (2, 3)
    a_c, a_sf = orc.per_token_cast_to_fp8(rng.randn(5, 260).astype(np.float32))
    out = orc.fp8_block_scaled_golden(a_c, a_sf, wc, wsf)
    assert out.shape == (5, 130)


def test_calc_diff():
    a = np.random.RandomState(3).randn(1000)
    assert orc.calc_diff(a, a) < 1e-15
    assert orc.calc_diff(a, -a) > 1.9
