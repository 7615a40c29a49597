"""sdct_roundtrip_noalloc_device (include/sdct.h): the two-kernel round trip whose scratch is a file-scope device array.
Through the C ABI, against sdct_roundtrip_device on the same inputs - same bytes (u8 outputs: the bar is bit-exact, as in
test_gpu_parity.py) - on two calls in a row, so that what the first call leaves in the shared array is seen not to leak into
the second.  Collected last on purpose."""
import numpy as np
import pytest

import oraclelib as ol
from conftest import make_image

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import simd_dct_b200 as sd  # noqa: E402
from simd_dct_b200 import _lib  # noqa: E402

MAX_BYTES = 32 << 20   # SDCT_NOALLOC_MAX_COEF_BYTES


def lut_for(layout, quality):
    t = ol.lut(quality).copy()
    t[0] = 20.0            # DC must not saturate the u8 levels (the reference applies no level shift)
    return (t * np.float32(255)).astype(np.float32) if layout == sd.LAYOUT_B else t


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


CASES = [  # (name, layout, flags, components, eyes, h, w, frames)
    ("tensor_rgb_B", sd.LAYOUT_B, sd.TENSOR, 3, 1, 256, 512, 2),            # sdct_roundtrip_device: pool scratch between two kernels
    ("tensor_rgb_ycc_C", sd.LAYOUT_C, sd.TENSOR | sd.COLOR_YCBCR, 3, 1, 128, 1024, 1),
    ("tensor_gray_A", sd.LAYOUT_A, sd.TENSOR, 1, 1, 64, 1088, 3),           # ... the fused tensor-core kernel
    ("tensor_stereo_C", sd.LAYOUT_C, sd.TENSOR, 1, 2, 64, 1088, 1),
    ("exact_A", sd.LAYOUT_A, sd.EXACT, 1, 1, 256, 512, 1),                  # ... the fused fp32 kernels
    ("exact_quirk_B", sd.LAYOUT_B, sd.EXACT | sd.QUIRK_AVX_X3, 1, 1, 256, 512, 1),
    ("exact_rgb_C", sd.LAYOUT_C, sd.EXACT, 3, 1, 128, 512, 2),
    ("fast_B", sd.LAYOUT_B, sd.FAST, 1, 1, 256, 512, 1),
]


@pytest.mark.timeout(300, method="thread")
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_noalloc_roundtrip_matches_roundtrip_device(ctx, case):
    _, layout, flags, comps, eyes, h, w, frames = case
    job = sd.dense_job(layout, w, h, frames=frames, components=comps, eyes=eyes, flags=flags)
    shape = (frames * eyes * h, w) if comps == 1 else (frames * eyes * h, w, comps)
    # two calls in a row: different pixels AND a different LUT (the SDCT_TENSOR inverse operand in the second static array changes)
    for call, (kind, quality) in enumerate((("smooth", 4), ("rand", 9))):
        img = make_image(kind, shape[0], w, seed=90 + call, comps=None if comps == 1 else comps)
        lut = lut_for(layout, quality)
        x = dev(img)
        want = torch.full_like(x, 0xa5)
        ctx.roundtrip(x, lut, job, out=want)
        got = torch.full_like(x, 0x5a)
        free0 = torch.cuda.mem_get_info()[0] if call == 1 else None
        ctx.roundtrip_noalloc(x, lut, job, out=got)
        torch.cuda.synchronize()
        if call == 1:   # every module is loaded by now: the second call leaves the device's free memory where it was
            assert torch.cuda.mem_get_info()[0] >= free0
        assert torch.equal(got, want), (case[0], call, int((got != want).sum()))
        if kind == "smooth":
            err = (got.to(torch.int16) - x.to(torch.int16)).abs().float().mean()
            assert float(err) < 8.0, float(err)   # it IS a round trip, not two equal buffers of fill bytes


@pytest.mark.timeout(300, method="thread")
def test_noalloc_band_window_and_repeat(ctx):
    """a band window leaves the other rows alone, and the same call twice gives the same bytes (the array keeps the first
    call's levels; the second overwrites every byte it reads)"""
    layout, h, w = sd.LAYOUT_B, 96, 2112
    lut = lut_for(layout, 4)
    job = sd.dense_job(layout, w, h, components=3, flags=sd.TENSOR, band_begin=4, band_end=10)
    x = dev(make_image("smooth", h, w, seed=93, comps=3))
    want = torch.full_like(x, 0x33)
    ctx.roundtrip(x, lut, job, out=want)
    for _ in range(2):
        got = torch.full_like(x, 0x33)
        ctx.roundtrip_noalloc(x, lut, job, out=got)
        assert torch.equal(got, want)
    assert bool((want[:32] == 0x33).all()) and bool((want[80:] == 0x33).all()) and not bool((want[32:80] == 0x33).all())


@pytest.mark.timeout(300, method="thread")
def test_noalloc_limit(ctx):
    """the largest job that fits (16 frames of 1920 x 1088 = 33 423 360 bytes) runs; one frame more is rejected with both sizes
    in the message and without touching the output"""
    layout, h, w = sd.LAYOUT_C, 1088, 1920
    lut = lut_for(layout, 4)
    frames = MAX_BYTES // (h * w)
    assert frames == 16
    g = torch.Generator(device="cuda").manual_seed(94)
    x = torch.randint(0, 256, (frames + 1, h, w), dtype=torch.uint8, device="cuda", generator=g)
    for flags in (sd.TENSOR, sd.EXACT):
        job = sd.dense_job(layout, w, h, frames=frames, flags=flags)
        want = ctx.roundtrip(x[:frames], lut, job)
        got = torch.zeros_like(want)
        ctx.roundtrip_noalloc(x[:frames], lut, job, out=got)
        assert torch.equal(got, want)
    big = sd.dense_job(layout, w, h, frames=frames + 1, flags=sd.EXACT)
    out = torch.full_like(x, 0x77)
    lib = _lib.load()
    import ctypes as C
    rc = lib.sdct_roundtrip_noalloc_device(ctx._h, C.byref(big), C.c_void_p(x.data_ptr()), lut.ctypes.data_as(C.POINTER(C.c_float)),
                                           C.c_void_p(out.data_ptr()), None)
    assert rc == 2   # SDCT_NOT_SUPPORTED
    msg = lib.sdct_last_error().decode()
    assert str(MAX_BYTES) in msg and "SDCT_NOALLOC_MAX_COEF_BYTES" in msg, msg
    torch.cuda.synchronize()
    assert bool((out == 0x77).all())
    with pytest.raises(sd.SdctError):
        ctx.roundtrip_noalloc(x, lut, big, out=out)
