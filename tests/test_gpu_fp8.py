"""fp8 e4m3 block-scaled GEMM (DeepGEMM semantics) against the fp64 oracle of the dequantised operands.
Tolerance: calc_diff < 1e-3 (upstream DeepGEMM's test threshold); the kernel differs from the oracle only by fp32
accumulation order, so we also require max |err| / max |ref| < 1e-2 on the bf16 output (bf16 rounding is 2^-8)."""
import numpy as np
import pytest
import torch

from oracle import matmul_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _case(m, n, k, seed, ue8m0=False):
    rng = np.random.RandomState(seed)
    # a wide dynamic range per strip so that the scales really differ between K blocks
    x = (rng.randn(m, k) * np.exp(rng.randn(m, 1) + rng.randn(1, k // 128 + 1).repeat(128, axis=1)[:, :k])).astype(np.float32)
    w = (rng.randn(n, k) * np.exp(rng.randn(n // 128 + 1, 1).repeat(128, axis=0)[:n])).astype(np.float32)
    a_codes, a_sf = orc.per_token_cast_to_fp8(x, ue8m0)
    w_codes, w_sf = orc.per_block_cast_to_fp8(w, ue8m0)
    return a_codes, a_sf, w_codes, w_sf


def _run(dga, a_codes, a_sf, w_codes, w_sf):
    a = torch.from_numpy(a_codes).to(DEV).view(torch.float8_e4m3fn)
    w = torch.from_numpy(w_codes).to(DEV).view(torch.float8_e4m3fn)
    before = dga.kernel_launch_count()
    out = dga.fp8_block_scaled_matmul(a, torch.from_numpy(a_sf).to(DEV), w, torch.from_numpy(w_sf).to(DEV))
    torch.cuda.synchronize()
    assert dga.kernel_launch_count() > before
    return out.float().cpu().numpy()


@pytest.mark.parametrize("m,n,k", [
    (128, 128, 128), (256, 256, 256), (256, 512, 1024), (300, 520, 400), (64, 384, 512), (1, 256, 384),
    (512, 448, 2048), (1024, 1536, 1024), (257, 7168, 512),
])
def test_fp8_matches_oracle(dga, m, n, k):
    a_codes, a_sf, w_codes, w_sf = _case(m, n, k, seed=m + n + k)
    got = _run(dga, a_codes, a_sf, w_codes, w_sf)
    ref = orc.fp8_block_scaled_golden(a_codes, a_sf, w_codes, w_sf)
    assert got.shape == ref.shape
    diff = orc.calc_diff(got, ref)
    assert diff < orc.CALC_DIFF_FP8, "calc_diff %.3e" % diff
    assert orc.max_err_over_max(got, ref) < 1e-2


@pytest.mark.parametrize("m,n,k", [
    (128, 128, 128), (256, 256, 512), (256, 512, 1024), (300, 520, 400), (64, 384, 640), (1, 256, 384),
    (512, 448, 2048), (1024, 1536, 1152), (257, 7168, 512), (2048, 2048, 4096),
])
def test_fp8_ue8m0_block_scaled_path(dga, m, n, k):
    """Power-of-two scales (DeepGEMM's use_ue8m0 on Blackwell): the tensor core applies them (kind::mxf8f6f4.block_scale)."""
    a_codes, a_sf, w_codes, w_sf = _case(m, n, k, seed=7 * m + n + k, ue8m0=True)
    got = _run(dga, a_codes, a_sf, w_codes, w_sf)
    ref = orc.fp8_block_scaled_golden(a_codes, a_sf, w_codes, w_sf)
    diff = orc.calc_diff(got, ref)
    assert diff < orc.CALC_DIFF_FP8, "calc_diff %.3e" % diff
    assert orc.max_err_over_max(got, ref) < 1e-2


def test_fp8_scales_are_applied_per_block(dga):
    """Unit codes with hand-made scales: out[m,n] = sum_kb sfa[m,kb] * sfb[n/128,kb] * 128 exactly."""
    m, n, k = 256, 384, 512
    one = np.full((1,), 0x38, dtype=np.uint8)  # e4m3 1.0
    a_codes = np.broadcast_to(one, (m, k)).copy()
    w_codes = np.broadcast_to(one, (n, k)).copy()
    rng = np.random.RandomState(0)
    a_sf = (2.0 ** rng.randint(-3, 3, size=(m, 4))).astype(np.float32)
    w_sf = (2.0 ** rng.randint(-3, 3, size=(3, 4))).astype(np.float32)
    got = _run(dga, a_codes, a_sf, w_codes, w_sf)
    ref = orc.fp8_block_scaled_golden(a_codes, a_sf, w_codes, w_sf)
    assert np.array_equal(got, orc.round_to_bf16(ref.astype(np.float32)))


def test_fp8_config4_sampled(dga):
    """BASELINE config 4 at full size (4096 x 7168 x 16384), sampled rows against the oracle."""
    m, n, k = 4096, 7168, 16384
    gen = torch.Generator(device=DEV).manual_seed(4)
    a = torch.randn((m, k), device=DEV, generator=gen)
    w = torch.randn((n, k), device=DEV, generator=gen)
    # quantise on the device with the same rule as the oracle (amax.clamp(1e-4) / 448; x * (1 / scale))
    av = a.view(m, k // 128, 128)
    a_sf = av.abs().amax(dim=2).clamp(min=1e-4) / 448.0
    a8 = (av * (1.0 / a_sf).unsqueeze(2)).view(m, k).to(torch.float8_e4m3fn)
    wv = w.view(n // 128, 128, k // 128, 128)
    w_sf = wv.abs().amax(dim=(1, 3)).clamp(min=1e-4) / 448.0
    w8 = (wv * (1.0 / w_sf)[:, None, :, None]).view(n, k).to(torch.float8_e4m3fn)
    del a, w, av, wv
    out = dga.fp8_block_scaled_matmul(a8, a_sf.contiguous(), w8, w_sf.contiguous())
    torch.cuda.synchronize()
    rows = np.random.RandomState(1).choice(m, 16, replace=False)
    rows_t = torch.from_numpy(rows).to(DEV)
    c0, c1 = 1024, 1792   # a 768-column window (six 128-blocks, straddling 224-wide tiles) keeps the fp64 oracle small
    ref = orc.fp8_block_scaled_golden(a8[rows_t].view(torch.uint8).cpu().numpy(), a_sf[rows_t].cpu().numpy(),
                                      w8[c0:c1].view(torch.uint8).cpu().numpy(), w_sf[c0 // 128:c1 // 128].cpu().numpy())
    got = out[rows_t][:, c0:c1].float().cpu().numpy()
    assert orc.calc_diff(got, ref) < orc.CALC_DIFF_FP8
    assert orc.max_err_over_max(got, ref) < 1e-2


def _ceil_pow2(sf):
    """ceil_to_ue8m0 on the device: the next power of two at or above a positive fp32 scale (oracle: ceil_to_ue8m0)."""
    bits = sf.contiguous().view(torch.int32)
    exp = ((bits >> 23) & 0xFF) + ((bits & 0x7FFFFF) != 0).to(torch.int32)
    return (exp.clamp(1, 254) << 23).view(torch.float32)


def test_fp8_config4_ue8m0_sampled(dga):
    """BASELINE config 4 at full size with UE8M0 (power-of-two) scales, i.e. per_token_cast_to_fp8(..., use_ue8m0=True):
    the shape and the kernel (tcgen05.mma kind::mxf8f6f4.block_scale) behind the bench's `fp8` line. Sampled rows x a
    column window against the fp64 oracle, and the library must report that the tensor-core kernel, not the software
    promotion kernel, produced the result."""
    m, n, k = 4096, 7168, 16384
    gen = torch.Generator(device=DEV).manual_seed(44)
    a = torch.randn((m, k), device=DEV, generator=gen)
    w = torch.randn((n, k), device=DEV, generator=gen)
    av = a.view(m, k // 128, 128)
    a_sf = _ceil_pow2(av.abs().amax(dim=2).clamp(min=1e-4) / 448.0)
    a8 = (av * (1.0 / a_sf).unsqueeze(2)).view(m, k).to(torch.float8_e4m3fn)
    wv = w.view(n // 128, 128, k // 128, 128)
    w_sf = _ceil_pow2(wv.abs().amax(dim=(1, 3)).clamp(min=1e-4) / 448.0)
    w8 = (wv * (1.0 / w_sf)[:, None, :, None]).view(n, k).to(torch.float8_e4m3fn)
    del a, w, av, wv
    out = torch.empty((m, n), device=DEV, dtype=torch.bfloat16)
    ex = dga.Fp8Executor(a8, a_sf, w8, w_sf, out)
    before = dga.kernel_launch_count()
    ex()
    assert ex.path() == "block_scaled"
    assert dga.kernel_launch_count() > before
    rows = np.random.RandomState(3).choice(m, 16, replace=False)
    rows_t = torch.from_numpy(rows).to(DEV)
    # the device-side quantisation must be the oracle's: same codes and scales on the sampled rows
    x_rows = torch.randn((m, k), device=DEV, generator=torch.Generator(device=DEV).manual_seed(44))[rows_t].cpu().numpy()
    codes, sf = orc.per_token_cast_to_fp8(x_rows, True)
    assert np.array_equal(sf, a_sf[rows_t].cpu().numpy())
    assert np.array_equal(codes, a8[rows_t].view(torch.uint8).cpu().numpy())
    for c0, c1 in ((0, 512), (3072, 3968), (6656, 7168)):   # 128-aligned windows: first tiles, four 224-wide tile seams, last tile
        ref = orc.fp8_block_scaled_golden(a8[rows_t].view(torch.uint8).cpu().numpy(), a_sf[rows_t].cpu().numpy(),
                                          w8[c0:c1].view(torch.uint8).cpu().numpy(), w_sf[c0 // 128:c1 // 128].cpu().numpy())
        got = out[rows_t][:, c0:c1].float().cpu().numpy()
        assert orc.calc_diff(got, ref) < orc.CALC_DIFF_FP8
        assert orc.max_err_over_max(got, ref) < 1e-2
    # and with one scale nudged off a power of two the same executor must take the promotion kernel
    a_sf[0, 0] *= 1.5
    ex()
    assert ex.path() == "promotion"
    ex.close()
