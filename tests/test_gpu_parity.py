"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on identical inputs.
Bar: byte-identical for every forward mode and for the inverse (integer outputs)."""
import os
import subprocess

import numpy as np
import pytest

import oraclelib as ol
from conftest import make_image

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import simd_dct_b200 as sd  # noqa: E402

LAYOUTS = [sd.LAYOUT_A, sd.LAYOUT_B, sd.LAYOUT_C]


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def gpu_forward(ctx, img, lut, layout, flags=sd.EXACT, eyes=1, comps=1, **kw):
    h = img.shape[0] // eyes
    w = img.shape[1]
    job = sd.dense_job(layout, w, h, frames=1, components=comps, eyes=eyes, flags=flags, **kw)
    out = ctx.forward(dev(img), lut, job)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def rt_lut(layout):
    """a LUT under which a round trip is meaningful: the reference applies no level shift, so
    the DC entry must be large enough for DC (<= 8 normalised) not to saturate the u8 range."""
    t = ol.lut(4).copy()
    t[0] = 20.0
    return (t * np.float32(255)).astype(np.float32) if layout == sd.LAYOUT_B else t


def flags_for(layout, quirk=True):
    return sd.EXACT | (sd.QUIRK_AVX_X3 if (layout == sd.LAYOUT_B and quirk) else 0)


def _dct8_inputs():
    rng = np.random.default_rng(7)
    ints = rng.integers(0, 256, (60000, 8)).astype(np.float32)            # first-pass inputs of layout B
    unit = (rng.integers(0, 256, (60000, 8)).astype(np.float32) / np.float32(255))  # ... of layouts A / C
    wide = (rng.standard_normal((60000, 8)) * 300).astype(np.float32)     # second-pass-like inputs
    edge = np.array([[0] * 8, [255] * 8, [255, 0] * 4, [0, 0, 0, 0, 255, 255, 255, 255], [1e-30] * 8], dtype=np.float32)
    return np.concatenate([ints, unit, wide, edge])


@pytest.mark.parametrize("kind,name,quirk", [(0, "dct8_scalar", 0), (8, "dct8_scalar", 0), (1, "paired", 0), (9, "paired", 0),
                                             (2, "paired", 1), (10, "paired", 1), (3, "dct8_fast", 0), (4, "idct8", 0),
                                             (14, "dct8_sse", 0)])
def test_dct8_bitwise(ctx, oracle, kind, name, quirk):
    """the device 8-point kernels (packed fp32, sdct_math.cuh) reproduce the CPU restatement bit for
    bit - in particular no multiply-add contraction sneaks into the EXACT graphs."""
    x = _dct8_inputs()
    got = ctx.selftest_dct8(kind, x)
    want = np.stack([oracle.dct8(name, v, quirk) for v in x[:4000]])
    assert np.array_equal(got[:4000].view(np.uint32), want.view(np.uint32))
    want_all = oracle.dct8_many(name, x, quirk)
    bad = np.flatnonzero((got.view(np.uint32) != want_all.view(np.uint32)).any(axis=1))
    assert bad.size == 0, (bad[:5], got[bad[:1]], want_all[bad[:1]])


def test_mul_inv255_bitwise(ctx):
    """x * (1.f / 255) (the reference's SSE encoders, src/simd_dct.cpp:955-962): all 256 inputs"""
    x = np.arange(256, dtype=np.float32).reshape(32, 8)
    got = ctx.selftest_dct8(6, x)
    assert np.array_equal(got.view(np.uint32), (x * (np.float32(1) / np.float32(255))).astype(np.float32).view(np.uint32))


@pytest.mark.parametrize("kind", ["rand", "smooth"])
@pytest.mark.parametrize("scale", [1, 8, 255])
def test_layout_c_sse_quirk(ctx, oracle, kind, scale):
    """SDCT_QUIRK_SSE: the bytes of the reference's SSE stereo encoders (restatement pinned against the compiled reference and
    the golden vectors in test_oracle_pins.py): forward for one and two eyes, a band window, the reference-style call, and
    the fused round trip = the fp32 inverse of those levels."""
    flags = sd.EXACT | sd.QUIRK_SSE
    for (h, w, eyes, window) in ((512, 512, 1, None), (128, 2112, 2, None), (96, 1040, 2, (3, 9))):
        img = make_image(kind, h * eyes, w, seed=17)
        lut = ol.lut(scale)
        b0, b1 = window if window else (0, 0)
        job = sd.dense_job(sd.LAYOUT_C, w, h, eyes=eyes, flags=flags, band_begin=b0, band_end=b1)
        out = torch.full((eyes * h * w,), 0x5a, dtype=torch.uint8, device="cuda")
        ctx.forward(dev(img), lut, job, out=out)
        want_all = oracle.forward(ol.LAYOUT_C, img, lut, quirks=ol.QUIRK_SSE, eyes=eyes).reshape(64, h // 8, eyes * (w // 8))
        want = np.full_like(want_all, 0x5a)
        lo, hi = (b0, b1) if window else (0, h // 8)
        want[:, lo:hi] = want_all[:, lo:hi]
        assert np.array_equal(out.cpu().numpy().reshape(want.shape), want), (h, w, eyes, window)
        assert not np.array_equal(want_all, oracle.forward(ol.LAYOUT_C, img, lut, eyes=eyes).reshape(want_all.shape)) or scale == 255
    # round trip: the inverse does not depend on the quirk (levels mean the same coefficients)
    h, w = 256, 512
    lut = rt_lut(sd.LAYOUT_C)
    img = make_image(kind, h, w, seed=18)
    job = sd.dense_job(sd.LAYOUT_C, w, h, flags=flags)
    coef = oracle.forward(ol.LAYOUT_C, img, lut, quirks=ol.QUIRK_SSE)
    rt = ctx.roundtrip(dev(img), lut, job).cpu().numpy().reshape(h, w)
    assert np.array_equal(rt, oracle.inverse(ol.LAYOUT_C, coef, w, h, lut))
    # misuse
    with pytest.raises(sd.SdctError):
        ctx.forward(dev(img), lut, sd.dense_job(sd.LAYOUT_B, w, h, flags=flags))
    with pytest.raises(sd.SdctError):
        ctx.forward(dev(img[:, :504]), lut, sd.dense_job(sd.LAYOUT_C, 504, h, flags=flags))
    with pytest.raises(sd.SdctError):
        ctx.forward(dev(img), lut, sd.dense_job(sd.LAYOUT_C, w, h, flags=sd.TENSOR | sd.QUIRK_SSE))


def test_reference_style_call_with_the_sse_quirk(oracle):
    """sdct_ref_set_flags(SDCT_EXACT | SDCT_QUIRK_AVX_X3 | SDCT_QUIRK_SSE): every simd_dct.h entry point then returns what the
    reference's SIMD dispatch returns on an x86-64 CPU with AVX2 (layout A: its scalar path - the SSE layout-A encoders are
    the broken ones, never reproduced)"""
    x, y = 256, 128
    img = make_image("rand", y, x, seed=19)
    lut = ol.lut(4)
    old = sd.get_reference_flags() if hasattr(sd, "get_reference_flags") else None
    sd.set_reference_flags(sd.EXACT | sd.QUIRK_AVX_X3 | sd.QUIRK_SSE)
    try:
        for name, fn in (("A", sd.simdDCT_EncodeQuantizeBuffer), ("B", sd.simdDCT_EncodeQuantize32ReorderBuffer), ("C", sd.simdDCT_EncodeQuantizeReorderStereoBuffer)):
            for (s0, s1) in ((0, y), (32, 80)):
                got = np.full(x * y, 0x11, dtype=np.uint8)
                assert fn(img, got, lut, x, y, s0, s1) == sd.sdr_Success
                rc, want = oracle.ref_call(name, img, lut, x, y, s0, s1, quirks=ol.QUIRK_AVX_X3 | ol.QUIRK_SSE, fill=0x11)
                assert rc == 0 and np.array_equal(got, want), (name, s0, s1)
    finally:
        sd.set_reference_flags(old if old is not None else (sd.EXACT | sd.QUIRK_AVX_X3))


def test_div255_bitwise(ctx):
    """x / 255.f (src/simd_dct.cpp:217, :340) as two FMA-pipe operations: all 256 inputs"""
    x = np.arange(256, dtype=np.float32).reshape(32, 8)
    got = ctx.selftest_dct8(5, x)
    assert np.array_equal(got.view(np.uint32), (x / np.float32(255)).astype(np.float32).view(np.uint32))


@pytest.mark.parametrize("kind", ["rand", "smooth"])
@pytest.mark.parametrize("scale", [1, 8, 255])
@pytest.mark.parametrize("layout", LAYOUTS)
def test_config1_forward_exact(ctx, oracle, kind, scale, layout):
    """BASELINE config 1: 512x512 single plane, default LUT x1 / x8 / x255."""
    img = make_image(kind, 512, 512)
    lut = ol.lut(scale)
    for quirk in ([True, False] if layout == sd.LAYOUT_B else [True]):
        got = gpu_forward(ctx, img, lut, layout, flags_for(layout, quirk))
        want = oracle.forward(layout, img, lut, quirks=ol.QUIRK_AVX_X3 if quirk else 0)
        assert np.array_equal(got, want), f"{np.count_nonzero(got != want)} bytes differ"


@pytest.mark.parametrize("layout", LAYOUTS)
def test_forward_fast_matches_its_definition_and_stays_close_to_exact(ctx, oracle, layout):
    img = make_image("smooth", 512, 512, seed=3)
    for scale in (8, 64, 255):
        lut = ol.lut(scale)
        got = gpu_forward(ctx, img, lut, layout, sd.FAST)
        want = oracle.forward(layout, img, lut, mode=ol.MODE_FAST)
        assert np.array_equal(got, want)
        exact = oracle.forward(layout, img, lut)  # true DCT (no X3 quirk)
        d = np.abs(got.astype(int) - exact.astype(int))
        assert d.max() <= 1
        assert np.mean(d != 0) <= 1e-3, np.mean(d != 0)


@pytest.mark.parametrize("layout", LAYOUTS)
def test_stereo_and_mono_layouts_small_and_ragged(ctx, oracle, layout):
    """edge shapes: one block, one group, non-multiple-of-CTA tails, eyes = 2 for layout C."""
    shapes = [(8, 64), (16, 64), (24, 192), (8, 1024), (136, 320)]
    if layout != sd.LAYOUT_B:
        shapes += [(8, 8), (24, 40), (72, 1000)]
    lut = ol.lut(16)
    for (h, w) in shapes:
        for eyes in ([1, 2] if layout == sd.LAYOUT_C else [1]):
            img = make_image("rand", h * eyes, w, seed=h + w)
            got = gpu_forward(ctx, img, lut, layout, flags_for(layout), eyes=eyes)
            want = oracle.forward(layout, img, lut, quirks=ol.QUIRK_AVX_X3, eyes=eyes)
            assert np.array_equal(got, want), (h, w, eyes)


@pytest.mark.parametrize("layout", LAYOUTS)
def test_band_window_leaves_other_bands_untouched(ctx, oracle, layout):
    h, w = 128, 256
    img = make_image("smooth", h, w, seed=5)
    lut = ol.lut(32)
    full = oracle.forward(layout, img, lut, quirks=ol.QUIRK_AVX_X3)
    job = sd.dense_job(layout, w, h, flags=flags_for(layout), band_begin=3, band_end=11)
    out = torch.full((h * w,), 0xA5, dtype=torch.uint8, device="cuda")
    ctx.forward(dev(img), lut, job, out=out)
    got = out.cpu().numpy()
    # expected: window bands from the oracle, everything else still 0xA5
    want = np.full(h * w, 0xA5, dtype=np.uint8)
    bpr = w // 8
    n0, n1 = 3 * bpr, 11 * bpr
    if layout == sd.LAYOUT_C:
        ps = h * w // 64
        for i in range(64):
            want[i * ps + n0:i * ps + n1] = full[i * ps + n0:i * ps + n1]
    else:
        want[n0 * 64:n1 * 64] = full[n0 * 64:n1 * 64]
    assert np.array_equal(got, want)


@pytest.mark.parametrize("layout", LAYOUTS)
def test_interleaved_rgb_forward(ctx, oracle, layout):
    """component reorder fused into the load: interleaved RGB -> 3 planar coefficient sets."""
    h, w = 64, 192
    img = make_image("smooth", h, w, seed=9, comps=3)
    lut = ol.lut(24)
    for flags, mode in ((flags_for(layout), ol.MODE_EXACT), (sd.FAST, ol.MODE_FAST)):
        got = gpu_forward(ctx, img, lut, layout, flags, comps=3)
        quirks = ol.QUIRK_AVX_X3 if mode == ol.MODE_EXACT else 0
        want = np.concatenate([oracle.forward(layout, img, lut, mode=mode, quirks=quirks, comp=k) for k in range(3)])
        assert np.array_equal(got, want)
        planes = oracle.deinterleave(img)
        want2 = np.concatenate([oracle.forward(layout, planes[k], lut, mode=mode, quirks=quirks) for k in range(3)])
        assert np.array_equal(want, want2)


@pytest.mark.parametrize("layout", LAYOUTS)
def test_unaligned_pointers_and_padded_pitch(ctx, oracle, layout):
    h, w, pitch = 40, 128, 128 + 13
    img = make_image("rand", h, w, seed=11)
    lut = ol.lut(8)
    buf = torch.zeros(h * pitch + 64, dtype=torch.uint8, device="cuda")
    view = buf[3:3 + h * pitch].view(h, pitch)
    view[:, :w] = dev(img)
    job = sd.dense_job(layout, w, h, flags=flags_for(layout))
    job.pitch = pitch
    outbuf = torch.zeros(h * w + 64, dtype=torch.uint8, device="cuda")
    ctx.forward(buf[3:], lut, job, out=outbuf[1:])
    got = outbuf[1:1 + h * w].cpu().numpy()
    assert np.array_equal(got, oracle.forward(layout, img, lut, quirks=ol.QUIRK_AVX_X3))


@pytest.mark.parametrize("layout", LAYOUTS)
def test_inverse_and_fused_roundtrip(ctx, oracle, layout):
    h, w = 256, 512
    lut = rt_lut(layout)
    for kind in ("smooth", "rand"):
        for eyes in ([1, 2] if layout == sd.LAYOUT_C else [1]):
            img = make_image(kind, h * eyes, w, seed=2)
            coef = oracle.forward(layout, img, lut, eyes=eyes)
            job = sd.dense_job(layout, w, h, eyes=eyes, flags=sd.EXACT)
            got = ctx.inverse(dev(coef), lut, job).cpu().numpy().reshape(h * eyes, w)
            want = oracle.inverse(layout, coef, w, h, lut, eyes=eyes)
            assert np.array_equal(got, want)
            rt = ctx.roundtrip(dev(img), lut, job).cpu().numpy().reshape(h * eyes, w)
            assert np.array_equal(rt, want)
            if kind == "smooth":
                mse = np.mean((rt.astype(float) - img.astype(float)) ** 2)
                assert 10 * np.log10(255 ** 2 / mse) > 30


@pytest.mark.parametrize("layout", LAYOUTS)
@pytest.mark.parametrize("w,window", [(512, None), (2112, (3, 9)), (72, None), (1032, (1, 5))])
def test_interleaved_rgb_inverse_and_roundtrip(ctx, oracle, layout, w, window):
    """the component reorder on the OUTPUT side: three planar coefficient sets -> interleaved pixels.  An even number
    of blocks per row takes the staged interleave (three components per CTA, 16-byte stores; w = 2112 makes tiles
    wrap block rows), an odd one (w = 72, 1032) the byte-store path; band windows leave the other rows untouched."""
    if layout == sd.LAYOUT_B and w % 64:
        pytest.skip("layout B needs width % 64 == 0")
    h = 96
    img = make_image("smooth", h, w, seed=21, comps=3)
    lut = rt_lut(layout)
    b0, b1 = window if window else (0, h // 8)
    job = sd.dense_job(layout, w, h, components=3, flags=sd.EXACT, band_begin=b0, band_end=b1 if window else 0)
    planes = oracle.deinterleave(img)
    coef = np.concatenate([oracle.forward(layout, planes[k], lut) for k in range(3)])
    want = np.full((h, w, 3), 0x5a, dtype=np.uint8)
    for k in range(3):
        want[b0 * 8:b1 * 8, :, k] = oracle.inverse(layout, coef[k * h * w:(k + 1) * h * w], w, h, lut)[b0 * 8:b1 * 8]
    out = torch.full((h, w, 3), 0x5a, dtype=torch.uint8, device="cuda")
    ctx.inverse(dev(coef), lut, job, out=out)
    assert np.array_equal(out.cpu().numpy(), want)
    out.fill_(0x5a)
    ctx.roundtrip(dev(img), lut, job, out=out)
    assert np.array_equal(out.cpu().numpy(), want)


def test_roundtrip_error_no_worse_than_float64_inverse(ctx, oracle):
    """SURVEY 8c acceptance test for the (unpinned) inverse: GPU fwd->inv error must not exceed
    the error of oracle-forward -> float64 textbook inverse on the same input."""
    h, w = 128, 128
    img = make_image("smooth", h, w, seed=4)
    n = np.arange(8)
    Cm = np.sqrt(2 / 8) * np.cos((2 * n[None, :] + 1) * n[:, None] * np.pi / 16)
    Cm[0] /= np.sqrt(2)
    for layout in LAYOUTS:
        lut = rt_lut(layout)
        coef = oracle.forward(layout, img, lut)
        job = sd.dense_job(layout, w, h)
        got = ctx.inverse(dev(coef), lut, job).cpu().numpy().reshape(h, w).astype(float)
        d = (lut.astype(np.float64) * np.float64(np.float32(0.95)))
        ref = np.zeros((h, w))
        bpr = w // 8
        for by in range(h // 8):
            for bx in range(bpr):
                nblk = by * bpr + bx
                if layout == sd.LAYOUT_A:
                    q = coef[nblk * 64:(nblk + 1) * 64].astype(float)
                    c = ((q - 127) / 255 * d).reshape(8, 8).T
                elif layout == sd.LAYOUT_B:
                    g = coef[(nblk // 8) * 512:(nblk // 8 + 1) * 512].reshape(64, 8)[:, nblk % 8].astype(float)
                    c = ((g - 127) * d / 255).reshape(8, 8) / 255
                else:
                    q = coef.reshape(64, -1)[:, nblk].astype(float)
                    c = ((q - 127) / 255 * d).reshape(8, 8)
                ref[by * 8:by * 8 + 8, bx * 8:bx * 8 + 8] = np.clip(np.rint(Cm.T @ c @ Cm * 255), 0, 255)
        err_gpu = np.abs(got - img).max(), np.mean((got - img) ** 2)
        err_ref = np.abs(ref - img).max(), np.mean((ref - img) ** 2)
        assert np.abs(got - ref).max() <= 1
        assert err_gpu[0] <= err_ref[0] + 1 and err_gpu[1] <= err_ref[1] * 1.01 + 1e-3


def test_batch_of_frames_equals_per_frame_and_shards(ctx, oracle):
    """frames are independent: a batch equals the concatenation of its frames / of its shards."""
    f, h, w = 6, 64, 128
    imgs = np.stack([make_image("smooth", h, w, seed=s) for s in range(f)])
    lut = ol.lut(16)
    for layout in LAYOUTS:
        job = sd.dense_job(layout, w, h, frames=f, flags=flags_for(layout))
        got = ctx.forward(dev(imgs), lut, job).cpu().numpy().reshape(f, -1)
        for k in range(f):
            assert np.array_equal(got[k], oracle.forward(layout, imgs[k], lut, quirks=ol.QUIRK_AVX_X3))
        parts = []
        for r in range(4):
            b, e = sd.shard_range(f, r, 4)
            if e > b:
                j = sd.dense_job(layout, w, h, frames=e - b, flags=flags_for(layout))
                parts.append(ctx.forward(dev(imgs[b:e]), lut, j).cpu().numpy())
        assert np.array_equal(np.concatenate(parts), got.reshape(-1))


@pytest.mark.parametrize("name,layout", [("A", sd.LAYOUT_A), ("B", sd.LAYOUT_B), ("C", sd.LAYOUT_C)])
def test_reference_entry_points_host_buffers(oracle, name, layout):
    """the three simd_dct.h symbols on host buffers: same bytes, windows and result codes as
    the reference-semantics oracle (and as the compiled reference when oracle/_ref travelled)."""
    fn = {"A": sd.simdDCT_EncodeQuantizeBuffer, "B": sd.simdDCT_EncodeQuantize32ReorderBuffer,
          "C": sd.simdDCT_EncodeQuantizeReorderStereoBuffer}[name]
    sx, sy = 256, 128
    img = make_image("rand", sy, sx, seed=21)
    lut = ol.lut(8)
    ref = ol.Reference("strict") if ol.have_reference("strict") else None
    for (s, e) in ((0, sy), (0, 64), (64, 128), (16, 16), (40, 100), (300, 400)):
        out = np.full(sx * sy, 7, dtype=np.uint8)
        rc = fn(img.reshape(-1), out, lut, sx, sy, s, e)
        rc2, want = oracle.ref_call(name, img, lut, sx, sy, s, e, fill=7)
        assert rc == rc2 == 0
        assert np.array_equal(out, want), (s, e)
        if ref is not None and (name != "B" or "avx2" in ol.cpu_flags()):
            ref.set_isa("avx2" if name == "B" else "scalar")
            rc3, want3 = ref.call(name, img, lut, sx, sy, s, e, fill=7)
            assert rc3 == 0 and np.array_equal(out, want3), (s, e)
    out = np.zeros(sx * sy, dtype=np.uint8)
    assert fn(None, out, lut, sx, sy, 0, sy) == sd.sdr_InvalidParameter
    assert fn(img.reshape(-1), None, lut, sx, sy, 0, sy) == sd.sdr_InvalidParameter
    assert fn(img.reshape(-1), out, lut, sx - 4, sy, 0, sy) == sd.sdr_NotSupported
    assert fn(img.reshape(-1), out, lut, sx, sy - 4, 0, sy) == sd.sdr_NotSupported


@pytest.mark.parametrize("layout", LAYOUTS)
def test_host_pipeline_matches_device_path(oracle, layout, monkeypatch):
    monkeypatch.setenv("SDCT_HOST_CHUNK_MB", "1")   # the $SDCT_* switches are read once, in sdct_create
    ctx = sd.Context(0)
    lut = ol.lut(12)
    for (f, h, w, eyes) in ((9, 256, 512, 1), (1, 1080 - 1080 % 8, 1920, 1), (2, 128, 256, 2)):
        if eyes == 2 and layout != sd.LAYOUT_C:
            continue
        imgs = np.stack([make_image("smooth", h * eyes, w, seed=s) for s in range(f)])
        job = sd.dense_job(layout, w, h, frames=f, eyes=eyes, flags=flags_for(layout))
        dev_out = ctx.forward(dev(imgs), lut, job).cpu().numpy()
        host_out = ctx.forward_host(imgs, lut, job)
        assert np.array_equal(dev_out, host_out)
        back_dev = ctx.inverse(dev(dev_out), lut, job).cpu().numpy()
        back_host = ctx.inverse_host(host_out, lut, job)
        assert np.array_equal(back_dev, back_host)


def test_full_size_1080p_batch_properties(ctx, oracle):
    """BASELINE config 2 size (1080p luma; 8 of the 64 frames here): oracle parity on two
    frames, and batch == shards on all (size-independent property)."""
    f, h, w = 8, 1080, 1920
    imgs = np.stack([make_image("smooth" if s % 2 else "rand", h, w, seed=s) for s in range(f)])
    lut = ol.lut(8)
    d = dev(imgs)
    for layout in (sd.LAYOUT_B, sd.LAYOUT_A):
        job = sd.dense_job(layout, w, h, frames=f, flags=flags_for(layout))
        got = ctx.forward(d, lut, job)
        for k in (0, 5):
            want = oracle.forward(layout, imgs[k], lut, quirks=ol.QUIRK_AVX_X3)
            assert np.array_equal(got.view(f, -1)[k].cpu().numpy(), want)
        half = sd.dense_job(layout, w, h, frames=f // 2, flags=flags_for(layout))
        lo = ctx.forward(d[:f // 2], lut, half)
        hi = ctx.forward(d[f // 2:], lut, half)
        assert torch.equal(torch.cat([lo, hi]), got)


CLI_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")


@pytest.mark.skipif(not os.path.exists(os.path.join(CLI_DIR, "simd_dct_cli_b200")), reason="oracle/_ref CLIs not built")
@pytest.mark.parametrize("mode,simd", [("enc-quant", "none"), ("enc-quant32", "avx2"), ("enc-quant-stereo", "none")])
def test_reference_cli_linked_against_libsdct(tmp_path, mode, simd):
    """the reference's own CLI (src/main.cpp, unmodified) linked against
    libsdct.so instead of src/simd_dct.cpp writes the same --to dump as the all-CPU binary."""
    if simd == "avx2" and "avx2" not in ol.cpu_flags():
        pytest.skip("host CPU has no AVX2: the reference cannot produce layout B here")
    img = make_image("smooth", 512, 768, seed=5)
    raw = tmp_path / "img.raw"
    raw.write_bytes(img.tobytes())
    outs = []
    for exe in ("simd_dct_cli_cpu", "simd_dct_cli_b200"):
        dump = tmp_path / (exe + ".bin")
        dump.write_bytes(b"")
        r = subprocess.run([os.path.join(CLI_DIR, exe), str(raw), "768", "512", "--runs", "2", "--quality", "8", "--mode", mode,
                            "--max-simd", simd, "--to", str(dump)], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append(np.frombuffer(dump.read_bytes(), dtype=np.uint8))
    assert outs[0].size == 768 * 512
    # the CLI mallocs its output buffer and never clears it: compare only the bytes the call writes
    # (layouts A / B visit the top half of the image, src/simd_dct.cpp:375 / :2243)
    n = 768 * 512 if mode == "enc-quant-stereo" else 768 * 256
    assert np.array_equal(outs[0][:n], outs[1][:n])


@pytest.mark.skipif(not os.path.exists(os.path.join(CLI_DIR, "simd_dct_cli_cpu")), reason="oracle/_ref CLIs not built")
@pytest.mark.parametrize("mode,simd", [("enc-quant", "none"), ("enc-quant32", "avx2"), ("enc-quant-stereo", "none"), ("enc-quant-stereo", "sse4.1")])
def test_python_cli_dump_equals_the_reference_cli(tmp_path, mode, simd):
    """`python -m simd_dct_b200.cli` follows the reference CLI's workflow (src/main.cpp:84-102, :508-606);
    its --to dump equals the all-CPU reference binary's, and --inverse dumps pixels close to the input."""
    import sys
    if simd == "avx2" and "avx2" not in ol.cpu_flags():
        pytest.skip("host CPU has no AVX2: the reference cannot produce layout B here")
    if simd == "sse4.1" and "sse4_1" not in ol.cpu_flags():
        pytest.skip("host CPU has no SSE4.1")
    sse = ["--sse-quirk"] if simd == "sse4.1" else []   # the reference's SSE stereo encoder against SDCT_QUIRK_SSE
    img = make_image("smooth", 512, 768, seed=6)
    raw = tmp_path / "img.raw"
    raw.write_bytes(img.tobytes())
    ref_dump, my_dump, inv_dump = tmp_path / "ref.bin", tmp_path / "mine.bin", tmp_path / "inv.bin"
    common = [str(raw), "768", "512", "--runs", "2", "--quality", "8", "--mode", mode]
    r = subprocess.run([os.path.join(CLI_DIR, "simd_dct_cli_cpu")] + common + ["--max-simd", simd, "--to", str(ref_dump)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    root = os.path.dirname(os.path.dirname(CLI_DIR))
    r = subprocess.run([sys.executable, "-m", "simd_dct_b200.cli"] + common + sse + ["--to", str(my_dump), "--resident"], capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "MiB/s" in r.stdout and "blocks/s" in r.stdout and "kernel only" in r.stdout
    a, b = np.frombuffer(ref_dump.read_bytes(), dtype=np.uint8), np.frombuffer(my_dump.read_bytes(), dtype=np.uint8)
    n = 768 * 512 if mode == "enc-quant-stereo" else 768 * 256
    assert a.size == b.size == 768 * 512 and np.array_equal(a[:n], b[:n])
    # decode: a quality under which the round trip is meaningful (the reference applies no level shift, so the DC entry
    # .17 * quality must reach 8 / 0.5 / 0.95 for DC not to saturate; the unnormalised layout B needs 255 times that)
    q = "25500" if mode == "enc-quant32" else "100"
    r = subprocess.run([sys.executable, "-m", "simd_dct_b200.cli", str(raw), "768", "512", "--runs", "1", "--quality", q, "--mode", mode,
                        "--no-quirk", "--inverse", "--to", str(inv_dump)], capture_output=True, text=True, timeout=300, cwd=root)
    assert r.returncode == 0, r.stdout + r.stderr
    back = np.frombuffer(inv_dump.read_bytes(), dtype=np.uint8).reshape(512, 768)
    rows = 512 if mode == "enc-quant-stereo" else 256
    err = np.abs(back[:rows].astype(int) - img[:rows].astype(int))
    assert err.mean() < 6, float(err.mean())


def test_config3_4k_interleaved_rgb(ctx, oracle):
    """BASELINE config 3 shape (3840x2160x3 interleaved RGB -> 3 planar coefficient sets), 3 frames:
    oracle parity on sampled bands of frame 1 (oracle = de-interleave + reference single-plane
    forward), batch == per-frame on all, inverse returns interleaved pixels within the
    quantiser's error."""
    f, h, w = 3, 2160, 3840
    g = torch.Generator(device="cuda").manual_seed(11)
    base = torch.randint(0, 256, (f, h // 8, w // 8, 3), dtype=torch.uint8, device="cuda", generator=g)
    imgs = base.repeat_interleave(8, 1).repeat_interleave(8, 2)
    imgs = (imgs.to(torch.int16) // 2 + torch.randint(0, 128, (f, h, w, 3), dtype=torch.int16, device="cuda", generator=g)).to(torch.uint8)
    lut = ol.lut(8)
    for layout in (sd.LAYOUT_B, sd.LAYOUT_C):
        job = sd.dense_job(layout, w, h, frames=f, components=3, flags=flags_for(layout))
        got = ctx.forward(imgs, lut, job).view(f, 3, -1)
        one = sd.dense_job(layout, w, h, frames=1, components=3, flags=flags_for(layout))
        for k in range(f):
            assert torch.equal(ctx.forward(imgs[k], lut, one).view(3, -1), got[k])
        # oracle on bands [100, 104) of frame 1, each component
        b0, b1 = 100, 104
        rows = imgs[1, b0 * 8:b1 * 8].cpu().numpy()
        for c in range(3):
            want = oracle.forward(layout, np.ascontiguousarray(rows[:, :, c]), lut, quirks=ol.QUIRK_AVX_X3)
            n0, n1 = b0 * (w // 8), b1 * (w // 8)
            if layout == sd.LAYOUT_B:
                mine = got[1, c, n0 * 64:n1 * 64].cpu().numpy()
            else:
                mine = got[1, c].view(64, -1)[:, n0:n1].contiguous().view(-1).cpu().numpy()
            assert np.array_equal(mine, want), (layout, c)


def test_config4_inverse_only_4k(ctx, oracle):
    """BASELINE config 4: inverse-only decode of 4K coefficient planes; frames sharded as the
    multi-GPU run shards them (contiguous frame ranges, no collective) give the batch result, and
    sampled bands equal the CPU restatement of the inverse."""
    f, h, w = 4, 2160, 3840
    g = torch.Generator(device="cuda").manual_seed(12)
    imgs = torch.randint(0, 256, (f, h, w), dtype=torch.uint8, device="cuda", generator=g)
    for layout in (sd.LAYOUT_B, sd.LAYOUT_C):
        lut = rt_lut(layout)
        job = sd.dense_job(layout, w, h, frames=f, flags=flags_for(layout))
        coef = ctx.forward(imgs, lut, job)
        back = ctx.inverse(coef, lut, job).view(f, h, w)
        parts = []
        for r in range(2):
            b, e = sd.shard_range(f, r, 2)
            sub = sd.dense_job(layout, w, h, frames=e - b, flags=flags_for(layout))
            parts.append(ctx.inverse(coef.view(f, -1)[b:e].contiguous(), lut, sub).view(e - b, h, w))
        assert torch.equal(torch.cat(parts), back)
        # CPU restatement on 2 bands of frame 2 (band-local coefficient window re-packed for the oracle)
        b0, b1 = 17, 19
        n0, n1 = b0 * (w // 8), b1 * (w // 8)
        cf = coef.view(f, -1)[2]
        if layout == sd.LAYOUT_B:
            window = cf[n0 * 64:n1 * 64].cpu().numpy()
        else:
            window = cf.view(64, -1)[:, n0:n1].contiguous().view(-1).cpu().numpy()
        want = oracle.inverse(layout, window, w, (b1 - b0) * 8, lut)
        assert np.array_equal(back[2, b0 * 8:b1 * 8].cpu().numpy(), want)


def test_config5_gigapixel_band_sharded_roundtrip(ctx, oracle):
    """BASELINE config 5: 32768x32768x3 interleaved image generated on the device, forward + inverse
    over band-range shards (8 shards, as 8 GPUs would take them; base pointers stay global so each
    shard writes where the single-device run writes).  Checks: shards == whole, sampled bands ==
    oracle, round-trip error bounded, fused round trip == forward then inverse."""
    w = h = 32768
    world = 8
    g = torch.Generator(device="cuda").manual_seed(13)
    img = torch.randint(0, 256, (h // 16, w // 16, 3), dtype=torch.uint8, device="cuda", generator=g)
    img = img.repeat_interleave(16, 0).repeat_interleave(16, 1)          # 3 GiB, blocky content
    img += torch.randint(0, 4, (1, w, 3), dtype=torch.uint8, device="cuda", generator=g)  # wraps mod 256: fine
    layout = sd.LAYOUT_B
    lut = rt_lut(layout)
    job = sd.dense_job(layout, w, h, components=3, flags=sd.EXACT)
    whole = ctx.forward(img, lut, job)
    sharded = torch.zeros_like(whole)
    from simd_dct_b200 import sharding
    for r in range(world):
        ctx.forward(img, lut, sharding.shard_bands(job, r, world), out=sharded)
    assert torch.equal(whole, sharded)
    del sharded
    for band in (0, 2049, 4095):
        rows = img[band * 8:band * 8 + 8].cpu().numpy()
        n0, n1 = band * (w // 8), (band + 1) * (w // 8)
        for c in (0, 2):
            want = oracle.forward(layout, np.ascontiguousarray(rows[:, :, c]), lut, quirks=0)
            assert np.array_equal(whole.view(3, -1)[c, n0 * 64:n1 * 64].cpu().numpy(), want)
    back = torch.empty_like(img)
    for r in range(world):
        ctx.inverse(whole, lut, sharding.shard_bands(job, r, world), out=back)
    err = (back[::7].to(torch.int16) - img[::7].to(torch.int16)).abs()
    # coefficients saturate at +-127 quantiser steps, so single pixels can be far off; the mean is the bound
    assert float(err.float().mean()) < 3.0, (int(err.max()), float(err.float().mean()))
    fused = ctx.roundtrip(img, lut, job)
    assert torch.equal(fused, back)


@pytest.mark.timeout(300, method="thread")
@pytest.mark.parametrize("layout", LAYOUTS)
def test_tensor_core_fused_roundtrip(ctx, oracle, layout, monkeypatch):
    """the ONE-kernel tensor-core round trip (single planes, TMA pixel path: rows of at least 128 blocks): the levels stay
    in tensor memory between the two products.  Bytes = inverse_tensor(forward_tensor(x)) of the restatement = the two
    kernels back to back; band windows leave the other rows alone; the bench batch (16 200 tiles) under a hard timeout."""
    lut = rt_lut(layout)
    cases = [(64, 1088, 1, None), (96, 2112, 1, (4, 10))] + ([(64, 1088, 2, None)] if layout == sd.LAYOUT_C else [])
    for (h, w, eyes, window) in cases:
        img = make_image("smooth", h * eyes, w, seed=81)
        b0, b1 = window if window else (0, h // 8)
        job = sd.dense_job(layout, w, h, eyes=eyes, flags=sd.TENSOR, band_begin=b0, band_end=b1 if window else 0)
        coef = oracle.forward(layout, img, lut, mode=ol.MODE_TENSOR, eyes=eyes)
        want = np.full((eyes, h, w), 0x5a, dtype=np.uint8)
        want[:, b0 * 8:b1 * 8] = oracle.inverse(layout, coef, w, h, lut, eyes=eyes, tensor=True).reshape(eyes, h, w)[:, b0 * 8:b1 * 8]
        out = torch.full((eyes * h, w), 0x5a, dtype=torch.uint8, device="cuda")
        n0 = ctx.kernel_launches
        ctx.roundtrip(dev(img), lut, job, out=out)
        assert ctx.kernel_launches - n0 == 1, "the fused kernel was not taken"
        assert np.array_equal(out.cpu().numpy().reshape(eyes, h, w), want), (h, w, eyes, window)
    f, h, w = 64, 1080, 1920
    g = torch.Generator(device="cuda").manual_seed(78)
    x = torch.randint(0, 256, (f, h, w), dtype=torch.uint8, device="cuda", generator=g)
    job = sd.dense_job(layout, w, h, frames=f, flags=sd.TENSOR)
    fused = ctx.roundtrip(x, lut, job)
    torch.cuda.synchronize()
    monkeypatch.setenv("SDCT_TC_NO_RT_FUSE", "1")   # read once, in sdct_create
    two = sd.Context(0)
    monkeypatch.delenv("SDCT_TC_NO_RT_FUSE")
    n0 = two.kernel_launches
    unfused = two.roundtrip(x, lut, job)
    torch.cuda.synchronize()
    assert two.kernel_launches - n0 == 2
    two.close()
    assert torch.equal(fused, unfused)


@pytest.mark.timeout(300, method="thread")
@pytest.mark.parametrize("layout", LAYOUTS)
def test_tensor_core_kernels_full_bench_batch(ctx, layout):
    """the bench workload (64 x 1080p: 16 200 tiles, about 110 per CTA) through the tensor-core kernels: the whole batch
    equals its eight 8-frame shards, forward and inverse.  (A build whose mbarrier phases were ambiguous passed every small
    test and hung here: many tiles per CTA are part of the contract.)"""
    f, h, w = 64, 1080, 1920
    g = torch.Generator(device="cuda").manual_seed(77)
    x = torch.randint(0, 256, (f, h, w), dtype=torch.uint8, device="cuda", generator=g)
    lut = rt_lut(layout)
    job = sd.dense_job(layout, w, h, frames=f, flags=sd.TENSOR)
    sub = sd.dense_job(layout, w, h, frames=8, flags=sd.TENSOR)
    coef = ctx.forward(x, lut, job)
    torch.cuda.synchronize()
    parts = torch.cat([ctx.forward(x[i:i + 8], lut, sub) for i in range(0, f, 8)])
    assert torch.equal(coef, parts)
    back = ctx.inverse(coef, lut, job)
    torch.cuda.synchronize()
    parts = torch.cat([ctx.inverse(coef.view(f, -1)[i:i + 8].contiguous(), lut, sub) for i in range(0, f, 8)])
    assert torch.equal(back.view(-1), parts.view(-1))
    rgb = torch.randint(0, 256, (4, 2160, 3840, 3), dtype=torch.uint8, device="cuda", generator=g)
    j3 = sd.dense_job(layout, 3840, 2160, frames=4, components=3, flags=sd.TENSOR)
    one = sd.dense_job(layout, 3840, 2160, frames=1, components=3, flags=sd.TENSOR)
    c3 = ctx.forward(rgb, lut, j3)
    assert torch.equal(c3, torch.cat([ctx.forward(rgb[i], lut, one) for i in range(4)]))
    b3 = ctx.inverse(c3, lut, j3)
    assert torch.equal(b3.view(-1), torch.cat([ctx.inverse(c3.view(4, -1)[i].contiguous(), lut, one).view(-1) for i in range(4)]))


@pytest.mark.parametrize("layout", LAYOUTS)
def test_tensor_core_forward_is_bit_exact_to_its_cpu_restatement(ctx, oracle, layout):
    """SDCT_TENSOR (tcgen05.mma + TMEM, sdct_tc.cu): integer-exact matrix form of the true DCT.
    Byte-identical to oracle mode TENSOR (which proves the fp32 tensor-core accumulation is exact
    for these bounded integer sums), on aligned, ragged (tail tile), stereo and RGB inputs."""
    for (h, w, kind, scale) in ((512, 512, "rand", 1), (512, 512, "smooth", 8), (1080 - 1080 % 8, 1920, "rand", 255),
                                (8, 64, "rand", 8), (24, 192, "smooth", 64), (72, 1088 - 1088 % 64, "rand", 16),
                                (8 * 1080, 1920, "rand", 64)):   # 2025 tiles: every CTA walks ~14 tiles through all stages
        img = make_image(kind, h, w, seed=h + w)
        lut = ol.lut(scale)
        got = gpu_forward(ctx, img, lut, layout, sd.TENSOR)
        want = oracle.forward(layout, img, lut, mode=ol.MODE_TENSOR)
        assert np.array_equal(got, want), (h, w, kind, scale, np.count_nonzero(got != want))
    if layout == sd.LAYOUT_C:
        img = make_image("rand", 2 * 64, 320, seed=9)
        got = gpu_forward(ctx, img, ol.lut(8), layout, sd.TENSOR, eyes=2)
        assert np.array_equal(got, oracle.forward(layout, img, ol.lut(8), mode=ol.MODE_TENSOR, eyes=2))
    rgb = make_image("smooth", 64, 256, seed=4, comps=3)
    job = sd.dense_job(layout, 256, 64, components=3, flags=sd.TENSOR)
    got = ctx.forward(dev(rgb), ol.lut(8), job).cpu().numpy().reshape(3, -1)
    for c in range(3):
        assert np.array_equal(got[c], oracle.forward(layout, rgb, ol.lut(8), mode=ol.MODE_TENSOR, comp=c))
    # unaligned base pointer and padded pitch (byte-granular load / store paths)
    h, w, pitch = 40, 128, 128 + 24
    buf = np.zeros(h * pitch + 3, dtype=np.uint8)
    img = make_image("rand", h, w, seed=77)
    view = buf[3:3 + h * pitch].reshape(h, pitch)
    view[:, :w] = img
    d = dev(buf)
    job = sd.dense_job(layout, w, h, flags=sd.TENSOR)
    job.pitch = pitch
    out = torch.zeros(h * w + 1, dtype=torch.uint8, device="cuda")
    ctx.forward(d[3:], ol.lut(8), job, out=out[1:])
    assert np.array_equal(out[1:].cpu().numpy(), oracle.forward(layout, img, ol.lut(8), mode=ol.MODE_TENSOR))


@pytest.mark.parametrize("layout", LAYOUTS)
def test_tensor_core_wide_images_tma_pixel_path(ctx, oracle, layout, monkeypatch):
    """Images at least 128 blocks wide with 16-byte-aligned rows take the TMA pixel path of sdct_tc.cu (bulk
    tensor copies of 8 rows x 1024 bytes, a second box where a tile runs on into the next block row); layout B
    the direct byte-pair stores, layout C one tensor-map store per tile (clipped at the window's end): batch of frames with a ragged tail tile, band
    window whose first block is not a multiple of 8, padded pitch, two eyes."""
    f, h, w = 3, 40, 1088            # 136 blocks per row (layout B needs width % 64 == 0): every tile straddles block rows; 680 = 5 * 128 + 40 blocks per frame
    imgs = np.stack([make_image("rand", h, w, seed=40 + s) for s in range(f)])
    lut = ol.lut(16)
    job = sd.dense_job(layout, w, h, frames=f, flags=sd.TENSOR)
    got = ctx.forward(dev(imgs), lut, job).cpu().numpy().reshape(f, -1)
    want = [oracle.forward(layout, imgs[k], lut, mode=ol.MODE_TENSOR) for k in range(f)]
    for k in range(f):
        assert np.array_equal(got[k], want[k]), (k, np.count_nonzero(got[k] != want[k]))
    # the same through the thread-copy paths (what a geometry that fits no tensor map gets)
    monkeypatch.setenv("SDCT_TC_NO_TMA", "1")   # read once, in sdct_create
    ctx_no_tma = sd.Context(0)
    monkeypatch.delenv("SDCT_TC_NO_TMA")
    got = ctx_no_tma.forward(dev(imgs), lut, job).cpu().numpy().reshape(f, -1)
    torch.cuda.synchronize()
    ctx_no_tma.close()
    for k in range(f):
        assert np.array_equal(got[k], want[k]), (k, np.count_nonzero(got[k] != want[k]))
    # band window: bands 1..3 of frame 0 only, the rest of the output untouched
    job = sd.dense_job(layout, w, h, flags=sd.TENSOR, band_begin=1, band_end=4)
    out = torch.full((h * w,), 0x5A, dtype=torch.uint8, device="cuda")
    ctx.forward(dev(imgs[0]), lut, job, out=out)
    expect = np.full(h * w, 0x5A, dtype=np.uint8)
    bpr = w // 8
    n0, n1 = 1 * bpr, 4 * bpr
    if layout == sd.LAYOUT_C:
        ps = h * w // 64
        for i in range(64):
            expect[i * ps + n0:i * ps + n1] = want[0][i * ps + n0:i * ps + n1]
    else:
        expect[n0 * 64:n1 * 64] = want[0][n0 * 64:n1 * 64]
    assert np.array_equal(out.cpu().numpy(), expect)
    # padded pitch (still 16-byte aligned)
    pitch = w + 48
    buf = np.zeros((h, pitch), dtype=np.uint8)
    buf[:, :w] = imgs[1]
    job = sd.dense_job(layout, w, h, flags=sd.TENSOR)
    job.pitch = pitch
    assert np.array_equal(ctx.forward(dev(buf), lut, job).cpu().numpy(), want[1])
    if layout == sd.LAYOUT_C:   # two eyes (block rows alternate between the eyes; only layout C has them)
        img2 = make_image("smooth", 2 * 24, w, seed=91)
        got2 = gpu_forward(ctx, img2, ol.lut(8), layout, sd.TENSOR, eyes=2)
        assert np.array_equal(got2, oracle.forward(layout, img2, ol.lut(8), mode=ol.MODE_TENSOR, eyes=2))


@pytest.mark.parametrize("layout", LAYOUTS)
def test_tensor_core_levels_stay_within_one_of_the_reference_arithmetic(ctx, oracle, layout):
    """SDCT_TENSOR is not the reference's fp32 rounding sequence: against the EXACT levels (true DCT, no X3
    quirk) of a smooth image it may differ by one level in rare rounding cases, never by more."""
    img = make_image("smooth", 512, 512, seed=3)
    for scale in (8, 64, 255):
        lut = ol.lut(scale)
        got = gpu_forward(ctx, img, lut, layout, sd.TENSOR)
        exact = oracle.forward(layout, img, lut)
        d = np.abs(got.astype(int) - exact.astype(int))
        assert d.max() <= 1
        assert np.mean(d != 0) <= 1e-3, np.mean(d != 0)


@pytest.mark.parametrize("layout", LAYOUTS)
def test_tensor_core_fused_rgb_components(ctx, oracle, layout):
    """Interleaved RGB with 16-byte-aligned rows: one copy of a pixel tile feeds the three components' tiles
    (sdct_tc.cu, VEC_RGB).  Tiles that run on into the next block row, a ragged tail, several frames, a CTA
    walking several pixel tiles, and the unaligned fallback must all give the per-plane oracle bytes."""
    lut = ol.lut(16)
    for (f, h, w) in ((2, 40, 1088), (1, 1080 - 1080 % 8, 1920)):
        rgb = np.stack([make_image("rand", h, w, seed=60 + k, comps=3) for k in range(f)])
        job = sd.dense_job(layout, w, h, frames=f, components=3, flags=sd.TENSOR)
        got = ctx.forward(dev(rgb), lut, job).cpu().numpy().reshape(f, 3, -1)
        for k in range(f):
            for c in range(3):
                want = oracle.forward(layout, rgb[k], lut, mode=ol.MODE_TENSOR, comp=c)
                assert np.array_equal(got[k, c], want), (f, h, w, k, c, np.count_nonzero(got[k, c] != want))
    # rows that are only 8-byte aligned (pitch = 3 * 1080 + 8): register-prefetch path, same bytes
    h, w = 24, 1088
    rgb = make_image("smooth", h, w, seed=70, comps=3)
    pitch = 3 * w + 8
    buf = np.zeros((h, pitch), dtype=np.uint8)
    buf[:, :3 * w] = rgb.reshape(h, 3 * w)
    job = sd.dense_job(layout, w, h, components=3, flags=sd.TENSOR)
    job.pitch = pitch
    got = ctx.forward(dev(buf), lut, job).cpu().numpy().reshape(3, -1)
    for c in range(3):
        assert np.array_equal(got[c], oracle.forward(layout, rgb, lut, mode=ol.MODE_TENSOR, comp=c))


@pytest.mark.parametrize("layout", LAYOUTS)
def test_tensor_core_inverse_is_bit_exact_to_its_cpu_restatement(ctx, oracle, layout, monkeypatch):
    """SDCT_TENSOR inverse (sdct_tc.cu, inverse_tc_kernel): dequantiser folded into an exact-integer matrix,
    A operand in shared memory (K-major for layout A, MN-major for B and C).  Byte-identical to
    oracle.inverse(tensor=True); within one level of the fp32 inverse; unsupported geometries are refused."""
    for (h, w, kind, seed) in ((512, 512, "smooth", 1), (48, 1088, "rand", 2), (8 * 1080, 1920, "smooth", 3)):
        img = make_image(kind, h, w, seed=seed)
        lut = rt_lut(layout)
        coef = oracle.forward(layout, img, lut)
        job = sd.dense_job(layout, w, h, flags=sd.TENSOR)
        got = ctx.inverse(dev(coef), lut, job).cpu().numpy().reshape(h, w)
        want = oracle.inverse(layout, coef, w, h, lut, tensor=True)
        assert np.array_equal(got, want), (h, w, np.count_nonzero(got != want))
        if h <= 512:
            fp32 = oracle.inverse(layout, coef, w, h, lut)
            d = np.abs(got.astype(int) - fp32.astype(int))
            assert d.max() <= 1 and np.mean(d != 0) < 5e-3, (int(d.max()), float(np.mean(d != 0)))
    # a second LUT on the same context (the LUT-dependent operand is rebuilt), several frames
    f, h, w = 3, 64, 256
    imgs = np.stack([make_image("smooth", h, w, seed=20 + k) for k in range(f)])
    lut = (rt_lut(layout) * np.float32(1.5)).astype(np.float32)
    coefs = np.stack([oracle.forward(layout, imgs[k], lut) for k in range(f)])
    job = sd.dense_job(layout, w, h, frames=f, flags=sd.TENSOR)
    got = ctx.inverse(dev(coefs), lut, job).cpu().numpy().reshape(f, h, w)
    for k in range(f):
        assert np.array_equal(got[k], oracle.inverse(layout, coefs[k], w, h, lut, tensor=True))
    # narrow images (fewer than 8 blocks per row: the generic address walk) and, for layout C, two eyes with a tile
    # that runs across block rows of both
    for (h2, w2, eyes) in ((128, 32, 1), (64, 1088, 2)):
        if (layout == sd.LAYOUT_B and w2 % 64) or (eyes == 2 and layout != sd.LAYOUT_C):
            continue
        img = make_image("rand", h2 * eyes, w2, seed=7 + eyes)
        coef = oracle.forward(layout, img, rt_lut(layout), eyes=eyes)
        job2 = sd.dense_job(layout, w2, h2, eyes=eyes, flags=sd.TENSOR)
        got = ctx.inverse(dev(coef), rt_lut(layout), job2).cpu().numpy().reshape(h2 * eyes, w2)
        assert np.array_equal(got, oracle.inverse(layout, coef, w2, h2, rt_lut(layout), eyes=eyes, tensor=True)), (h2, w2, eyes)
    # the same with the tensor-map path of layout C switched off (the thread-copy fallback for geometries no tensor
    # map describes; layouts A and B always use the 1-D bulk copy, which needs no map)
    monkeypatch.setenv("SDCT_TC_NO_TMA", "1")   # read once, in sdct_create
    ctx_no_tma = sd.Context(0)
    monkeypatch.delenv("SDCT_TC_NO_TMA")
    got = ctx_no_tma.inverse(dev(coefs), lut, job).cpu().numpy().reshape(f, h, w)
    torch.cuda.synchronize()
    ctx_no_tma.close()
    for k in range(f):
        assert np.array_equal(got[k], oracle.inverse(layout, coefs[k], w, h, lut, tensor=True))
    # a geometry the tensor-core inverse has no path for (an odd number of blocks per row of interleaved pixels) is refused,
    # not silently replaced by the fp32 arithmetic
    if layout != sd.LAYOUT_B:
        job = sd.dense_job(layout, 72, 64, components=3, flags=sd.TENSOR)
        with pytest.raises(sd.SdctError):
            ctx.inverse(torch.zeros(3 * 72 * 64, dtype=torch.uint8, device="cuda"), lut, job)


@pytest.mark.parametrize("layout", LAYOUTS)
@pytest.mark.parametrize("w,h,window", [(512, 96, None), (2112, 96, (4, 10)), (3840, 64, None)])
def test_tensor_core_rgb_inverse_and_roundtrip(ctx, oracle, layout, w, h, window):
    """SDCT_TENSOR on the output side of interleaved RGB (three components per pixel tile staged planar, written
    interleaved) and the SDCT_TENSOR round trip (forward + inverse kernels back to back): byte-identical to the CPU
    restatement inverse_tensor(forward_tensor(x)) per component; bytes outside the band window untouched."""
    img = make_image("smooth", h, w, seed=31, comps=3)
    lut = rt_lut(layout)
    b0, b1 = window if window else (0, h // 8)
    job = sd.dense_job(layout, w, h, components=3, flags=sd.TENSOR, band_begin=b0, band_end=b1 if window else 0)
    planes = oracle.deinterleave(img)
    coef = np.concatenate([oracle.forward(layout, planes[k], lut, mode=ol.MODE_TENSOR) for k in range(3)])
    want = np.full((h, w, 3), 0x5a, dtype=np.uint8)
    for k in range(3):
        want[b0 * 8:b1 * 8, :, k] = oracle.inverse(layout, coef[k * h * w:(k + 1) * h * w], w, h, lut, tensor=True)[b0 * 8:b1 * 8]
    out = torch.full((h, w, 3), 0x5a, dtype=torch.uint8, device="cuda")
    ctx.inverse(dev(coef), lut, job, out=out)
    assert np.array_equal(out.cpu().numpy(), want)
    out.fill_(0x5a)
    n0 = ctx.kernel_launches
    ctx.roundtrip(dev(img), lut, job, out=out)
    assert ctx.kernel_launches - n0 == 2
    assert np.array_equal(out.cpu().numpy(), want)
    # single plane too
    luma = np.ascontiguousarray(img[:, :, 1])
    j1 = sd.dense_job(layout, w, h, flags=sd.TENSOR)
    c1 = oracle.forward(layout, luma, lut, mode=ol.MODE_TENSOR)
    rt = ctx.roundtrip(dev(luma), lut, j1).cpu().numpy().reshape(h, w)
    assert np.array_equal(rt, oracle.inverse(layout, c1, w, h, lut, tensor=True))
