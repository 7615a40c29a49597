"""The arithmetic identities the CUDA kernels rely on (simd_dct_b200/csrc/sdct_math.cuh), proven on
the CPU with exact rational arithmetic / IEEE binary32 emulation.  The GPU side of the same
identities is tests/test_gpu_parity.py::test_dct8_bitwise, ::test_div255_bitwise and the parity tests."""
from fractions import Fraction
import math

import numpy as np
import pytest

f32 = np.float32


def rn32(fr: Fraction, mode="rn") -> np.float32:
    """round an exact rational to binary32 (normal or denormal), nearest-even or toward -inf"""
    if fr == 0:
        return f32(0)
    s = -1 if fr < 0 else 1
    a = abs(fr)
    e = max(math.floor(math.log2(a)) if a >= 1 else -math.ceil(math.log2(1 / a)), -126)
    while e > -126 and Fraction(2) ** e > a:
        e -= 1
    while Fraction(2) ** (e + 1) <= a:
        e += 1
    ulp = Fraction(2) ** (e - 23)
    q = a / ulp
    n = q.numerator // q.denominator
    rem = q - n
    if mode == "rn":
        if rem > Fraction(1, 2) or (rem == Fraction(1, 2) and (n & 1)):
            n += 1
    else:  # toward -inf
        if s < 0 and rem != 0:
            n += 1
    return f32(s * float(Fraction(n) * ulp))


def F(x) -> Fraction:
    return Fraction(float(x))


def test_div255_two_operation_form_is_ieee_division():
    """q = fma(x, r, fl(x * rl)), r = fl(1/255), rl = fl(1/255 - r)  ==  x / 255.f for x = 0..255"""
    r = f32(0.003921568859368562698364257812500)
    rl = f32(-2.319175823606300543e-10)
    assert r == f32(1) / f32(255)
    assert rl == rn32(Fraction(1, 255) - F(r))
    for x in range(256):
        t = rn32(F(f32(x)) * F(rl))
        q = rn32(F(f32(x)) * F(r) + F(t))
        assert q == f32(x) / f32(255), x
    assert sum(1 for x in range(256) if f32(x) * r != f32(x) / f32(255)) == 126  # why one multiply is not enough


def bits(x: np.float32) -> int:
    return int(np.array(x, dtype=np.float32).view(np.int32))


def sat_u8(s32: int) -> int:
    return 0 if s32 < 0 else (255 if s32 > 255 else s32)


def test_denormal_scaled_fma_is_round_to_nearest_even_plus_bias_and_clamp():
    """bits(fma(v, -2^-149, 128 * 2^-149)) saturated to u8 == 255 - clamp(rne(v) + 127, 0, 255): the
    reference's AVX2 quantiser (src/simd_dct.cpp:2224) without an integer add"""
    dmin = Fraction(1, 2 ** 149)
    rng = np.random.default_rng(0)
    vs = list((rng.standard_normal(4000) * 200).astype(f32))
    vs += [f32(k + h) for k in range(-131, 131) for h in (0.0, 0.5, -0.5, 0.25, 0.75)]
    vs += [f32(x) for x in (1e6, -1e6, 8388607.5, -8388608.0, 3e9, -3e9, 2 ** 31 - 128, -(2.0 ** 31))]
    for v in vs:
        w = rn32(-F(v) * dmin + 128 * dmin)
        want = 255 - sat_u8(int(np.rint(np.float64(v))) + 127)
        assert sat_u8(bits(w)) == want, float(v)


def test_roundf_as_two_round_down_operations():
    """roundf(y) for y in [0, 255] == bits(mul_rd(add_rd(y, 0.5), 2^-149)): floor(y + 0.5) cannot be
    fooled by the sum's rounding because round-down never crosses an integer"""
    dmin = Fraction(1, 2 ** 149)
    ys = [f32(k) + f32(d) for k in range(0, 256) for d in (0, 0.5, 0.49999997, 0.50000006, 0.25)]
    ys += [np.nextafter(f32(k + 0.5), f32(0)) for k in range(0, 255)] + [np.nextafter(f32(k + 0.5), f32(300)) for k in range(0, 255)]
    ys = [y for y in ys if 0 <= y <= 255]
    for y in ys:
        u = rn32(F(y) + Fraction(1, 2), "rd")
        n = bits(rn32(F(u) * dmin, "rd"))
        # C roundf: half away from zero (float(y) + 0.5 is exact in binary64)
        assert n == int(math.floor(float(y) + 0.5)), float(y)


def test_pixel_rounding_by_denormal_multiply():
    """bits(x * 2^-149) saturated == clamp(rne(x), 0, 255) (inverse transform output)"""
    dmin = Fraction(1, 2 ** 149)
    for x in [f32(v) for v in (-3.2, -0.4, -0.0, 0.0, 0.4, 0.5, 1.5, 2.5, 254.5, 254.50002, 255.0, 255.6, 1e5, -1e5, 127.49999)]:
        w = rn32(F(x) * dmin)
        want = int(min(255, max(0, np.rint(np.float64(x)))))
        assert sat_u8(bits(w)) == want, float(x)


def test_tensor_form_matrix_is_exact_and_bounded(oracle):
    """mode TENSOR (sdct_tc.cu): the library and the oracle build the same integer matrix; its digits are
    exact in fp16 and every partial sum of 255 * |digit| stays below 2^24, so fp32 accumulation on the
    tensor cores is exact in any order; K / 2^22 is the true C (x) C to 2^-23."""
    import ctypes as C
    from simd_dct_b200 import _lib
    k_lib = np.zeros(64 * 64, dtype=np.int32)
    _lib.load().sdct_tensor_matrix(k_lib.ctypes.data_as(C.POINTER(C.c_int32)))
    k = oracle.tensor_matrix()
    assert np.array_equal(k_lib.reshape(64, 64), k)
    hi = (k + 512) >> 10
    lo = k - 1024 * hi
    assert np.abs(hi).max() <= 1024 and lo.min() >= -512 and lo.max() <= 511
    assert 255 * np.abs(hi).sum(axis=1).max() < 2 ** 24 and 255 * np.abs(lo).sum(axis=1).max() < 2 ** 24
    n = np.arange(8)
    c = np.cos((2 * n[None, :] + 1) * n[:, None] * np.pi / 16) * 0.5
    c[0, :] = np.sqrt(1 / 8)
    true = np.einsum("uy,vx->uvyx", c, c).reshape(64, 64)
    assert np.abs(k / 2.0 ** 22 - true).max() <= 2.0 ** -23 + 1e-12


@pytest.mark.parametrize("layout", [0, 1, 2])
def test_tensor_mode_stays_within_one_lsb_of_exact(oracle, layout):
    import oraclelib as ol
    rng = np.random.default_rng(5)
    y, x = np.mgrid[0:256, 0:512]
    smooth = np.clip(128 + 60 * np.sin(.05 * x) + 50 * np.cos(.031 * y) + rng.integers(0, 9, (256, 512)), 0, 255).astype(np.uint8)
    rand = rng.integers(0, 256, (256, 512), dtype=np.uint8)
    for img in (smooth, rand):
        for scale in (1, 8, 255):
            lut = ol.lut(scale)
            t = oracle.forward(layout, img, lut, mode=ol.MODE_TENSOR)
            e = oracle.forward(layout, img, lut)
            d = np.abs(t.astype(int) - e.astype(int))
            assert d.max() <= 1
            assert np.mean(d != 0) <= 2e-3, np.mean(d != 0)


@pytest.mark.parametrize("layout", [0, 1, 2])
def test_tensor_inverse_definition_stays_within_one_level_of_the_fp32_inverse(oracle, layout):
    """oracle.inverse(tensor=True) (dequantiser folded into an exact-integer matrix, two fp16 digits) against the
    fp32 inverse definition: same pixels except for rare one-level rounding differences, for small and large LUTs;
    both are this repo's definitions (the reference has no inverse)."""
    import oraclelib as ol
    from conftest import make_image
    img = make_image("smooth", 64, 128, seed=5)
    for scale in (4, 32):
        t = ol.lut(scale).copy()
        t[0] = 20.0
        lut = (t * np.float32(255)).astype(np.float32) if layout == 1 else t
        coef = oracle.forward(layout, img, lut)
        a = oracle.inverse(layout, coef, 128, 64, lut).astype(int)
        b = oracle.inverse(layout, coef, 128, 64, lut, tensor=True).astype(int)
        d = np.abs(a - b)
        assert d.max() <= 1 and np.mean(d != 0) < 5e-3, (int(d.max()), float(np.mean(d != 0)))
