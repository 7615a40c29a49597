"""Pins the CPU oracle (oracle/sdct_oracle.c) to the reference:
  * against the committed golden vectors, which tools/make_golden.py produced by running the
    unmodified reference (tests/golden/reference_vectors.npz) - runs everywhere;
  * against the reference itself (oracle/_ref, built by oracle/Makefile from /root/reference/src)
    where that library exists, including the facts SURVEY.md section 4 lists.
The reference has no tests, fixtures or known-answer vectors of its own."""
import os

import numpy as np
import pytest

import oraclelib as ol
from conftest import make_image

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")
W, H = 128, 64


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.mark.parametrize("kind", ["rand", "smooth"])
@pytest.mark.parametrize("scale", [1, 8, 255])
def test_oracle_matches_golden_reference_output(oracle, gold, kind, scale):
    img = gold[f"in_{kind}"]
    assert np.array_equal(img, make_image(kind, H, W, seed=42))  # the generator is reproducible
    lut = ol.lut(scale)
    a = oracle.forward(ol.LAYOUT_A, img, lut)
    b = oracle.forward(ol.LAYOUT_B, img, lut, quirks=ol.QUIRK_AVX_X3)
    c = oracle.forward(ol.LAYOUT_C, img, lut, eyes=2)  # top half = left eye, bottom half = right eye
    assert np.array_equal(a, gold[f"A_{kind}_x{scale}"])
    assert np.array_equal(b, gold[f"B_{kind}_x{scale}"])
    assert np.array_equal(c, gold[f"C_{kind}_x{scale}"])
    c_sse = oracle.forward(ol.LAYOUT_C, img, lut, quirks=ol.QUIRK_SSE, eyes=2)   # the reference's SSE4.1 stereo encoder
    assert np.array_equal(c_sse, gold[f"Csse_{kind}_x{scale}"])


@pytest.mark.parametrize("name", ["A", "B", "C"])
def test_oracle_reference_call_semantics_match_golden_windows(oracle, gold, name):
    """half-height processing, inclusive endY, y-vs-2y window comparison"""
    lut = ol.lut(8)
    for kind in ("rand", "smooth"):
        for (s, e) in ((16, 40), (0, 16), (33, 64)):
            rc, got = oracle.ref_call(name, gold[f"in_{kind}"], lut, W, H, s, e, fill=7)
            assert rc == 0
            assert np.array_equal(got, gold[f"win_{name}_{kind}_{s}_{e}"]), (kind, s, e)


def test_layout_b_without_the_x3_quirk_is_the_true_dct(oracle):
    """quirks = none must be an orthonormal DCT (float64 model), quirks = reference must not"""
    img = make_image("smooth", 64, 128, seed=1)
    lut = ol.lut(255)
    n = np.arange(8)
    cm = np.sqrt(2 / 8) * np.cos((2 * n[None, :] + 1) * n[:, None] * np.pi / 16)
    cm[0] /= np.sqrt(2)
    blocks = img.reshape(8, 8, 16, 8).transpose(0, 2, 1, 3).astype(np.float64)       # band, bx, y, x
    coef = np.einsum("uy,nbyx,vx->nbuv", cm, blocks, cm)
    q = 255.0 / (lut.astype(np.float64) * np.float64(np.float32(0.95)))
    model = np.clip(np.rint(coef.reshape(8, 16, 64) * q) + 127, 0, 255).astype(np.uint8)
    model = model.reshape(8, 2, 8, 64).transpose(0, 1, 3, 2).reshape(-1)             # groups of 8 blocks, [i][blk]
    true = oracle.forward(ol.LAYOUT_B, img, lut, quirks=0)
    quirk = oracle.forward(ol.LAYOUT_B, img, lut, quirks=ol.QUIRK_AVX_X3)
    d = np.abs(true.astype(int) - model.astype(int))
    assert d.max() <= 1 and np.mean(d != 0) < 1e-3
    assert np.mean(quirk != model) > 0.01


needs_ref = pytest.mark.skipif(not ol.have_reference("strict"), reason="oracle/_ref not built (no /root/reference here)")


@needs_ref
@pytest.mark.parametrize("kind", ["rand", "smooth"])
@pytest.mark.parametrize("scale", [1, 8])
def test_oracle_is_byte_identical_to_the_compiled_reference_512(oracle, kind, scale):
    """BASELINE config 1: 512x512, default LUT x1 / x8, all three entry points"""
    ref = ol.Reference("strict")
    img = make_image(kind, 512, 512)
    lut = ol.lut(scale)
    for name, isa in (("A", "scalar"), ("C", "scalar"), ("B", "avx2")):
        if isa == "avx2" and "avx2" not in ol.cpu_flags():
            continue
        ref.set_isa(isa)
        rc, want = ref.call(name, img, lut, 512, 512)
        rc2, got = oracle.ref_call(name, img, lut, 512, 512)
        assert rc == rc2 == 0
        assert np.array_equal(got, want), name


@needs_ref
def test_reference_result_codes(oracle):
    """src/simd_dct.cpp:75-76, :97-98, :117-118, :127"""
    ref = ol.Reference("strict")
    img = make_image("rand", 64, 64)
    lut = ol.lut(1)
    ref.set_isa("scalar")
    assert ref.call("A", img, lut, 60, 64)[0] == 2 and oracle.ref_call("A", img, lut, 60, 64)[0] == 2
    assert ref.call("C", img, lut, 64, 60)[0] == 2 and oracle.ref_call("C", img, lut, 64, 60)[0] == 2
    assert ref.call("B", img, lut, 32, 64)[0] == 2 and oracle.ref_call("B", img, lut, 32, 64)[0] == 2
    assert ref.call("B", img, lut, 64, 64)[0] == 2  # no SIMD flag set -> sdr_NotSupported (:127)


@needs_ref
def test_reference_simd_variants_have_the_surveyed_defects(oracle):
    """the SSE stereo path differs from the scalar path (X1 sign), the AVX2 path equals
    the oracle only with the X3 quirk.  Keeps the choice of parity targets honest."""
    flags = ol.cpu_flags()
    ref = ol.Reference("strict")
    img = make_image("rand", 128, 128)
    lut = ol.lut(8)
    ref.set_isa("scalar")
    _, scalar = ref.call("C", img, lut, 128, 128)
    if "sse4_1" in flags:
        ref.set_isa("sse41")
        _, sse = ref.call("C", img, lut, 128, 128)
        assert np.mean(sse != scalar) > 0.01
    if "avx2" in flags:
        ref.set_isa("avx2")
        _, avx = ref.call("B", img, lut, 128, 128)
        assert np.array_equal(avx, oracle.ref_call("B", img, lut, 128, 128, quirks=ol.QUIRK_AVX_X3)[1])
        assert not np.array_equal(avx, oracle.ref_call("B", img, lut, 128, 128, quirks=0)[1])


@needs_ref
@pytest.mark.parametrize("kind,scale", [("rand", 1), ("rand", 8), ("smooth", 1), ("smooth", 255)])
def test_sse_quirk_is_byte_identical_to_the_reference_sse_stereo_encoders(oracle, kind, scale):
    """quirks = SSE (layout C): the restatement of src/simd_dct.cpp:896-1103 / :672-892 against the reference's own SSE4.1,
    SSSE3 and SSE2 stereo encoders (what its dispatcher runs on every x86-64 CPU, :78-83), whole image and a band window,
    strict and fast-math builds."""
    flags = ol.cpu_flags()
    if "sse2" not in flags:
        pytest.skip("no SSE2 on this host")
    x, y = 256, 128
    img = make_image(kind, y, x, seed=5)
    lut = ol.lut(scale)
    isas = [("sse41", "sse4_1"), ("ssse3", "ssse3"), ("sse2", "sse2")]
    for build in ("strict", "fast"):
        if not ol.have_reference(build):
            continue
        ref = ol.Reference(build)
        for isa, flag in isas:
            if flag not in flags:
                continue
            ref.set_isa(isa)
            for (s0, s1) in ((0, y), (32, 80)):
                rc, want = ref.call("C", img, lut, x, y, s0, s1, fill=0x11)
                rc2, got = oracle.ref_call("C", img, lut, x, y, s0, s1, quirks=ol.QUIRK_SSE, fill=0x11)
                assert rc == rc2 == 0
                assert np.array_equal(got, want), (build, isa, s0, s1, int((got != want).sum()))
    ref = ol.Reference("strict")
    ref.set_isa("scalar")
    if kind == "rand":   # ... and they are NOT the scalar path's bytes
        assert np.mean(ref.call("C", img, lut, x, y)[1] != oracle.ref_call("C", img, lut, x, y, quirks=ol.QUIRK_SSE)[1]) > 0.005
    # native semantics (both eyes of every band) = the reference call over the whole image
    native = oracle.forward(ol.LAYOUT_C, img, lut, quirks=ol.QUIRK_SSE, eyes=2)
    assert np.array_equal(native, oracle.ref_call("C", img, lut, x, y, quirks=ol.QUIRK_SSE)[1])


def test_dct8_sse_is_the_paired_form_with_one_sign(oracle):
    rng = np.random.default_rng(3)
    for v in rng.standard_normal((200, 8)).astype(np.float32) * 50:
        a, b = oracle.dct8("dct8_sse", v), oracle.dct8("paired", v, 0)
        assert np.array_equal(np.delete(a, 1).view(np.uint32), np.delete(b, 1).view(np.uint32))
        assert abs(float(a[1]) - float(b[1])) > 0 or v[4] == v[3]
