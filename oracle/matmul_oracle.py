"""CPU oracle for the dynamic-shape matmul hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import this
module, and only as the checker or the reported CPU baseline. The product (deepgemm_ascend_b200) never does.

Parity status
  * framework path (row-major x row-major -> fp32): PINNED. This restates the reference's own pure-numpy
    goldens and comparator; tests/golden/ holds vectors produced by importing the reference scripts
    (tests/golden/make_golden.py) and tests/test_oracle.py checks this module against them bit for bit.
  * aclnn path (fp16/bf16 -> same dtype): the arithmetic lives in CATLASS, a dependency cloned at build time
    with no pinned tag (aclnn_catlass_dynamic_matmul/build.sh:8,19-29) and absent from /root/reference; the
    reference ships no numeric golden for it. PARITY UNPINNED BY THE REFERENCE: anchored instead on
    north_star's rule (max relative error vs an fp64 CPU matmul <= 1e-2) with the reference's input
    distributions.
  * fp8 block-scaled: not in the reference at all; semantics restated from upstream DeepGEMM
    (deep_gemm/utils/math.py per_token_cast_to_fp8 / per_block_cast_to_fp8, testing/numeric.py calc_diff).
    PARITY UNPINNED.
"""
import numpy as np

# --- comparator tolerances the reference uses -------------------------------------------------------------
RTOL_VERIFY = 1e-6        # deep_gemm_ascend/scripts/verify.py:10
RTOL_BENCH = 1.5e-6       # framework/tests/bench_main.py:30, framework/benchmark/benchmark.py:20
RTOL_TEST = 2e-4          # framework/tests/test.py:19
ATOL = 1e-9               # verify.py:11
ERROR_TOL = 1e-4          # verify.py:12  (allowed fraction of mismatching elements)
MAX_REL_16BIT = 1e-2      # BASELINE.json north_star: max rel err vs fp64 for bf16/fp16 outputs
CALC_DIFF_FP8 = 1e-3      # upstream DeepGEMM tests: calc_diff < 1e-3


def gen_golden_data(m, n, k, rng):
    """fp16 inputs U[1,10], fp32 numpy golden (deep_gemm_ascend/scripts/gen_golden.py:10-15).

    `rng` is a numpy RandomState-like object with .uniform (the reference uses the unseeded global one,
    gen_golden.py:11; pass np.random.RandomState(seed) to reproduce a fixture).
    """
    x1 = rng.uniform(1, 10, [m, k]).astype(np.float16)
    x2 = rng.uniform(1, 10, [k, n]).astype(np.float16)
    golden = golden_fp32(x1, x2)
    return x1, x2, golden


def heavy_tail(rng, shape):
    """lognormal(1, 1.2) clipped to [1, 10] (framework/tests/test.py:30-32, benchmark.py:343-367).
    `rng` is a numpy Generator (np.random.default_rng)."""
    v = rng.lognormal(mean=1.0, sigma=1.2, size=shape)
    return np.clip(v, 1, 10).astype(np.float32)


def golden_fp32(x1, x2):
    """The reference golden: fp32 matmul of the inputs widened to fp32 (gen_golden.py:14-15, test.py:37)."""
    return np.matmul(x1.astype(np.float32), x2.astype(np.float32)).astype(np.float32)


def golden_fp64(x1, x2):
    """fp64 matmul of the exactly representable inputs: the north_star oracle."""
    return np.matmul(x1.astype(np.float64), x2.astype(np.float64))


def verify_result(output, golden, rtol=RTOL_VERIFY, atol=ATOL, error_tol=ERROR_TOL):
    """The reference comparator (verify.py:15-35; test.py:40-64): element-wise np.isclose, pass when the
    mismatching fraction is <= error_tol. Sizes must match (verify.py:18-20). Returns (ok, mismatch_fraction)."""
    output = np.asarray(output).reshape(-1)
    golden = np.asarray(golden).reshape(-1)
    if output.size != golden.size:
        return False, 1.0
    if golden.size == 0:
        return True, 0.0  # the reference divides by zero here (verify.py:33); an empty result has no mismatch
    close = np.isclose(output, golden, rtol=rtol, atol=atol, equal_nan=True)
    ratio = float(np.count_nonzero(~close)) / golden.size
    return ratio <= error_tol, ratio


def max_rel_err(output, ref64):
    """max |out - ref| / |ref| over elements (positive-data parity metric)."""
    out = np.asarray(output, dtype=np.float64)
    ref = np.asarray(ref64, dtype=np.float64)
    denom = np.maximum(np.abs(ref), np.finfo(np.float64).tiny)
    return float(np.max(np.abs(out - ref) / denom)) if ref.size else 0.0


def max_err_over_max(output, ref64):
    """max |out - ref| / max |ref|: for zero-mean (randn) data where element-wise relative error is ill-defined."""
    out = np.asarray(output, dtype=np.float64)
    ref = np.asarray(ref64, dtype=np.float64)
    return float(np.max(np.abs(out - ref)) / max(np.max(np.abs(ref)), 1e-300)) if ref.size else 0.0


def calc_diff(x, y):
    """1 - 2<x,y>/(|x|^2+|y|^2)  (upstream deep_gemm/testing/numeric.py:5-11)."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    denom = (x * x + y * y).sum()
    return float(1.0 - 2.0 * (x * y).sum() / denom) if denom != 0 else 0.0


# --- bf16 / fp8 helpers (numpy has neither) -------------------------------------------------------------------
def to_bf16_bits(x):
    """fp32 -> bf16 bit pattern (uint16), round to nearest even."""
    u = np.asarray(x, dtype=np.float32).view(np.uint32)
    rounded = u + (np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1)))
    out = (rounded >> np.uint32(16)).astype(np.uint16)
    nan = np.isnan(np.asarray(x, dtype=np.float32))
    out[nan] = 0x7FC0
    return out


def from_bf16_bits(bits):
    return (np.asarray(bits, dtype=np.uint16).astype(np.uint32) << np.uint32(16)).view(np.float32)


def round_to_bf16(x):
    return from_bf16_bits(to_bf16_bits(x))


def _e4m3_table():
    vals = np.zeros(256, dtype=np.float64)
    for b in range(256):
        s = -1.0 if b & 0x80 else 1.0
        e = (b >> 3) & 0xF
        m = b & 0x7
        if e == 0xF and m == 0x7:
            vals[b] = np.nan
        elif e == 0:
            vals[b] = s * (m / 8.0) * 2.0 ** -6
        else:
            vals[b] = s * (1 + m / 8.0) * 2.0 ** (e - 7)
    return vals


E4M3_VALUES = _e4m3_table()
_E4M3_POS = E4M3_VALUES[:0x7F]  # 0 .. 448, monotonically increasing


def to_e4m3_bits(x):
    """fp32 -> float8_e4m3fn bit pattern (uint8), round to nearest even, saturating at +-448
    (inputs here are pre-scaled into range, as in per_token_cast_to_fp8)."""
    x = np.asarray(x, dtype=np.float64)
    a = np.minimum(np.abs(x), 448.0)
    hi = np.searchsorted(_E4M3_POS, a, side="left").clip(0, 0x7E)
    lo = np.maximum(hi - 1, 0)
    d_hi = _E4M3_POS[hi] - a
    d_lo = a - _E4M3_POS[lo]
    pick_lo = (d_lo < d_hi) | ((d_lo == d_hi) & (lo % 2 == 0))
    code = np.where(pick_lo, lo, hi).astype(np.uint8)
    code = np.where(_E4M3_POS[hi] == a, hi, code).astype(np.uint8)
    return (code | np.where(np.signbit(x), 0x80, 0).astype(np.uint8)).astype(np.uint8)


def from_e4m3_bits(bits):
    return E4M3_VALUES[np.asarray(bits, dtype=np.uint8)]


def ceil_to_ue8m0(sf):
    """Round a positive fp32 scale up to the next power of two (upstream deep_gemm/utils/math.py ceil_to_ue8m0):
    the scale format Blackwell's block-scaled MMA consumes."""
    bits = np.asarray(sf, dtype=np.float32).view(np.uint32)
    exp = ((bits >> np.uint32(23)) & np.uint32(0xFF)) + ((bits & np.uint32(0x7FFFFF)) != 0).astype(np.uint32)
    return (np.clip(exp, 1, 254).astype(np.uint32) << np.uint32(23)).view(np.float32)


def per_token_cast_to_fp8(x, use_ue8m0=False):
    """1x128 activation scaling (upstream deep_gemm/utils/math.py:25-37): per row and 128-column strip,
    scale = max(|x|, 1e-4) / 448; codes = e4m3(x / scale). Returns (codes uint8 [m,k], scales fp32 [m, ceil(k/128)])."""
    x = np.asarray(x, dtype=np.float32)
    m, k = x.shape
    kp = (k + 127) // 128 * 128
    xp = np.zeros((m, kp), dtype=np.float32)
    xp[:, :k] = x
    v = xp.reshape(m, kp // 128, 128)
    amax = np.maximum(np.abs(v).max(axis=2), 1e-4).astype(np.float32)
    sf = (amax / np.float32(448.0)).astype(np.float32)
    if use_ue8m0:
        sf = ceil_to_ue8m0(sf)
    codes = to_e4m3_bits(v * (np.float32(1.0) / sf)[:, :, None]).reshape(m, kp)[:, :k]
    return np.ascontiguousarray(codes), sf


def per_block_cast_to_fp8(w, use_ue8m0=False):
    """128x128 weight scaling (upstream deep_gemm/utils/math.py:50-60). w is [n, k].
    Returns (codes uint8 [n,k], scales fp32 [ceil(n/128), ceil(k/128)])."""
    w = np.asarray(w, dtype=np.float32)
    n, k = w.shape
    npad, kpad = (n + 127) // 128 * 128, (k + 127) // 128 * 128
    wp = np.zeros((npad, kpad), dtype=np.float32)
    wp[:n, :k] = w
    v = wp.reshape(npad // 128, 128, kpad // 128, 128)
    amax = np.maximum(np.abs(v).max(axis=(1, 3)), 1e-4).astype(np.float32)
    sf = (amax / np.float32(448.0)).astype(np.float32)
    if use_ue8m0:
        sf = ceil_to_ue8m0(sf)
    codes = to_e4m3_bits(v * (np.float32(1.0) / sf)[:, None, :, None]).reshape(npad, kpad)[:n, :k]
    return np.ascontiguousarray(codes), sf


def fp8_block_scaled_golden(a_codes, a_sf, w_codes, w_sf):
    """fp64 oracle of the fp8 GEMM: dequantise (code value x scale) then matmul. a [m,k], w [n,k] -> [m,n]."""
    a = from_e4m3_bits(a_codes) * np.repeat(a_sf.astype(np.float64), 128, axis=1)[:, :a_codes.shape[1]]
    wsf = np.repeat(np.repeat(w_sf.astype(np.float64), 128, axis=0), 128, axis=1)
    w = from_e4m3_bits(w_codes) * wsf[:w_codes.shape[0], :w_codes.shape[1]]
    return a @ w.T


# --- framework host math ------------------------------------------------------------------------------------
def m_grouped_golden_fp64(a, w, group_index):
    """M-grouped contiguous matmul (MoE): out[i,:] = a[i,:] . w[group_index[i]]^T in fp64; rows whose index is
    negative are padding and come back as NaN (the device leaves them unwritten).  Not in the reference; semantics of
    upstream deepseek-ai/DeepGEMM `m_grouped_*_gemm_nt_contiguous` (a [M_total,K], b [G,N,K], m_indices [M_total]);
    parity unpinned by the reference, anchored on the fp64 rule of north_star."""
    a64, w64 = a.astype(np.float64), w.astype(np.float64)
    out = np.full((a.shape[0], w.shape[1]), np.nan, dtype=np.float64)
    for g in np.unique(group_index):
        if g < 0:
            continue
        rows = np.nonzero(group_index == g)[0]
        out[rows] = a64[rows] @ w64[g].T
    return out


def m_grouped_masked_golden_fp64(a, w, masked_m):
    """M-grouped masked matmul (MoE decode): out[g,i,:] = a[g,i,:] . w[g]^T in fp64 for i < masked_m[g]; the other
    rows come back as NaN (don't-care on the device).  Not in the reference; semantics of upstream
    deepseek-ai/DeepGEMM `m_grouped_*_gemm_nt_masked` (a [G,Mmax,K], b [G,N,K], masked_m [G]); parity unpinned by the
    reference, anchored on the fp64 rule of north_star."""
    a64, w64 = a.astype(np.float64), w.astype(np.float64)
    out = np.full((a.shape[0], a.shape[1], w.shape[1]), np.nan, dtype=np.float64)
    for g in range(a.shape[0]):
        rows = int(masked_m[g])
        out[g, :rows] = a64[g, :rows] @ w64[g].T
    return out


def align16(x):
    return (x + 15) & ~15


def get_bench_config(m, n, k, m_sections, n_sections, m_sec_o_blocks, n_sec_o_blocks, k_o_iter_blocks, db_o_blocks):
    """The 28-int parameter vector of run_mmad_bench: six knobs followed by the 22 derived loop bounds, in the
    order gemm_bench.hpp:66-69 writes them (derivation: framework/csrc/jit/get_best_config.hpp:93-153)."""
    mb, nb, kb = align16(m) // 16, align16(n) // 16, align16(k) // 16
    db_num = k_o_iter_blocks // db_o_blocks
    r_mb = mb % m_sec_o_blocks or m_sec_o_blocks
    r_nb = nb % n_sec_o_blocks or n_sec_o_blocks
    k_iters = -(-kb // k_o_iter_blocks)
    k_tail = kb % k_o_iter_blocks
    if k_tail == 0:
        r_db_num, r_kb = db_num, db_o_blocks
    else:
        r_db_num = -(-k_tail // db_o_blocks)
        r_kb = k_tail - (r_db_num - 1) * db_o_blocks
    m_iters, n_iters = -(-mb // m_sec_o_blocks), -(-nb // n_sec_o_blocks)
    m_parts, n_parts = m_iters // m_sections, n_iters // n_sections
    return [m_sections, n_sections, m_sec_o_blocks, n_sec_o_blocks, k_o_iter_blocks, db_o_blocks,
            m, n, k, 1, k_iters, mb, nb, kb, m_parts * m_sec_o_blocks, n_parts * n_sec_o_blocks,
            align16(m) - m, align16(n) - n, align16(k) - k, db_num, m_parts, n_parts,
            m_iters - (m_sections - 1) * m_parts, n_iters - (n_sections - 1) * n_parts,
            r_mb, r_nb, r_kb, r_db_num]
