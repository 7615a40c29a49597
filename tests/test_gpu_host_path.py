"""GPU tests of the host-buffer path (sdct_host.cpp): what an existing caller of src/simd_dct.h:29-31 sees.
Bytes outside the job's window stay untouched (row padding included), caller threads with disjoint windows
overlap and agree with one call, one call can be sharded over several devices, and a second context on another
device works in the same process."""
import threading

import numpy as np
import pytest

import oraclelib as ol
from conftest import make_image

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import simd_dct_b200 as sd  # noqa: E402

LAYOUTS = [sd.LAYOUT_A, sd.LAYOUT_B, sd.LAYOUT_C]


def flags_for(layout):
    return sd.EXACT | (sd.QUIRK_AVX_X3 if layout == sd.LAYOUT_B else 0)


def rt_lut(layout):
    t = ol.lut(4).copy()
    t[0] = 20.0
    return (t * np.float32(255)).astype(np.float32) if layout == sd.LAYOUT_B else t


@pytest.mark.parametrize("engine_streams", ["0", "1"])
def test_pinned_batch_through_both_stream_plans(oracle, monkeypatch, engine_streams):
    """the chunked pipeline on pinned caller memory with one stream per engine (default) and with one stream per slot
    ($SDCT_HOST_ENGINE_STREAMS=0, read in sdct_create): many chunks (1 MB), every slot reused several times, forward and
    inverse, twice in a row on the same context (buffers and events of the first call are reused by the second)"""
    monkeypatch.setenv("SDCT_HOST_ENGINE_STREAMS", engine_streams)
    monkeypatch.setenv("SDCT_HOST_CHUNK_MB", "1")
    c = sd.Context(0)
    try:
        f, h, w = 24, 256, 512
        layout = sd.LAYOUT_B
        lut = rt_lut(layout)
        job = sd.dense_job(layout, w, h, frames=f, flags=flags_for(layout))
        for rep in range(2):
            img = make_image("smooth", f * h, w, seed=60 + rep)
            h_in = torch.from_numpy(img).pin_memory()
            h_coef = torch.zeros(f * h * w, dtype=torch.uint8).pin_memory()
            c.forward_host(h_in, lut, job, out=h_coef)
            want = np.concatenate([oracle.forward(layout, img[k * h:(k + 1) * h], lut, quirks=ol.QUIRK_AVX_X3) for k in range(f)])
            assert np.array_equal(h_coef.numpy(), want), (engine_streams, rep)
            h_pix = torch.zeros(f * h, w, dtype=torch.uint8).pin_memory()
            c.inverse_host(h_coef, lut, job, out=h_pix)
            back = np.concatenate([oracle.inverse(layout, want[k * h * w:(k + 1) * h * w], w, h, lut) for k in range(f)])
            assert np.array_equal(h_pix.numpy(), back), (engine_streams, rep)
    finally:
        c.close()


@pytest.mark.parametrize("pinned", [False, True])
@pytest.mark.parametrize("layout", LAYOUTS)
def test_inverse_host_padded_pitch_leaves_padding_untouched(ctx, oracle, layout, pinned):
    """the pixel side of the inverse is written row by row - the bytes between the rows of a
    padded pitch, and the pixels outside a band window, keep the caller's values."""
    h, w, pitch = 72, 128, 128 + 40
    img = make_image("smooth", h, w, seed=31)
    lut = rt_lut(layout)
    coef = oracle.forward(layout, img, lut)
    want = oracle.inverse(layout, coef, w, h, lut)
    job = sd.dense_job(layout, w, h, flags=sd.EXACT, band_begin=2, band_end=7)
    job.pitch = pitch
    if pinned:
        out_t = torch.full((h * pitch,), 0xA5, dtype=torch.uint8).pin_memory()
        coef_t = torch.from_numpy(coef).pin_memory()
        ctx.inverse_host(coef_t, lut, job, out=out_t)
        out = out_t.numpy().reshape(h, pitch)
    else:
        out = np.full((h, pitch), 0xA5, dtype=np.uint8)
        ctx.inverse_host(coef, lut, job, out=out.reshape(-1))
    expect = np.full((h, pitch), 0xA5, dtype=np.uint8)
    expect[16:56, :w] = want[16:56]
    assert np.array_equal(out, expect)


@pytest.mark.parametrize("layout", LAYOUTS)
def test_forward_host_unaligned_padded_input_and_window(ctx, oracle, layout):
    """pageable input with an odd pitch and an odd base offset; only the window's coefficient bytes are written"""
    h, w, pitch = 64, 192, 192 + 13
    img = make_image("rand", h, w, seed=32)
    buf = np.zeros(h * pitch + 5, dtype=np.uint8)
    buf[5:5 + h * pitch].reshape(h, pitch)[:, :w] = img
    lut = ol.lut(8)
    job = sd.dense_job(layout, w, h, flags=flags_for(layout), band_begin=1, band_end=6)
    job.pitch = pitch
    out = np.full(h * w + 3, 0x3C, dtype=np.uint8)
    ctx.forward_host(buf[5:], lut, job, out=out[3:])
    full = oracle.forward(layout, img, lut, quirks=ol.QUIRK_AVX_X3)
    expect = np.full(h * w, 0x3C, dtype=np.uint8)
    n0, n1 = 1 * (w // 8), 6 * (w // 8)
    if layout == sd.LAYOUT_C:
        ps = h * w // 64
        for i in range(64):
            expect[i * ps + n0:i * ps + n1] = full[i * ps + n0:i * ps + n1]
    else:
        expect[n0 * 64:n1 * 64] = full[n0 * 64:n1 * 64]
    assert np.array_equal(out[3:], expect) and (out[:3] == 0x3C).all()


def test_small_window_over_many_large_frames_is_compacted(ctx, oracle):
    """a one-band window over a batch of frames is staged as the window only (the device buffers
    of a chunk are the dense sub-job, not frames x frame stride)"""
    f, h, w = 24, 1080 - 1080 % 8, 1920
    rng = np.random.default_rng(5)
    imgs = rng.integers(0, 256, (f, h, w), dtype=np.uint8)
    lut = ol.lut(8)
    job = sd.dense_job(sd.LAYOUT_B, w, h, frames=f, flags=flags_for(sd.LAYOUT_B), band_begin=50, band_end=51)
    out = np.zeros(f * h * w, dtype=np.uint8)
    before = torch.cuda.memory_reserved()
    free0, _ = torch.cuda.mem_get_info()
    ctx.forward_host(imgs, lut, job, out=out)
    free1, _ = torch.cuda.mem_get_info()
    assert free0 - free1 < 64 << 20, "a 15 KiB-per-frame window must not allocate frame-sized device buffers"
    del before
    for k in (0, 23):
        want = oracle.forward(sd.LAYOUT_B, imgs[k, 400:408], lut, quirks=ol.QUIRK_AVX_X3)
        n0 = 50 * (w // 8)
        assert np.array_equal(out.reshape(f, -1)[k, n0 * 64:(n0 + w // 8) * 64], want)
        assert not out.reshape(f, -1)[k, :n0 * 64].any()


@pytest.mark.parametrize("name,layout", [("A", sd.LAYOUT_A), ("B", sd.LAYOUT_B), ("C", sd.LAYOUT_C)])
def test_four_caller_threads_with_disjoint_windows(oracle, name, layout):
    """the reference's concurrency model (src/simd_dct.cpp:1073-1087): N threads call the same entry point on one
    image with disjoint startY/endY windows.  The shim runs them on separate pipelines; together they must write
    exactly what one full call writes."""
    fn = {"A": sd.simdDCT_EncodeQuantizeBuffer, "B": sd.simdDCT_EncodeQuantize32ReorderBuffer,
          "C": sd.simdDCT_EncodeQuantizeReorderStereoBuffer}[name]
    sx, sy = 1024, 2048                      # half height 1024 rows = 128 bands
    img = make_image("smooth", sy, sx, seed=33).reshape(-1)
    lut = ol.lut(8)
    one = np.full(sx * sy, 9, dtype=np.uint8)
    assert fn(img, one, lut, sx, sy, 0, sy) == 0
    rc_want, want = oracle.ref_call(name, img, lut, sx, sy, 0, sy, fill=9)
    assert rc_want == 0 and np.array_equal(one, want)
    many = np.full(sx * sy, 9, dtype=np.uint8)
    scale = 1 if name == "A" else 2           # window units: y for layout A's scalar loop, 2y otherwise
    bands = sy // 16
    cuts = [0, 32, 64, 96, bands]
    rcs = [None] * 4

    def work(t):
        start = cuts[t] * 8 * scale
        end = (cuts[t + 1] - 1) * 8 * scale   # inclusive end: the last band of this thread's range
        rcs[t] = fn(img, many, lut, sx, sy, start, end)
    for rep in range(3):
        many[:] = 9
        threads = [threading.Thread(target=work, args=(t,)) for t in range(4)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert rcs == [0, 0, 0, 0]
        assert np.array_equal(many, one), rep


@pytest.mark.parametrize("layout", LAYOUTS)
def test_one_call_sharded_over_several_contexts(ctx, oracle, layout):
    """sdct_forward_host_multi / sdct_inverse_host_multi: frame ranges (batch) and band ranges (single frame).
    Uses every visible device; on a one-GPU box two contexts on device 0 exercise the same code path."""
    ndev = torch.cuda.device_count()
    ctxs = [sd.Context(d) for d in range(ndev)] if ndev > 1 else [sd.Context(0), sd.Context(0), sd.Context(0)]
    lut = rt_lut(layout)
    for (f, h, w) in ((7, 64, 256), (1, 1080 - 1080 % 8, 1920)):
        imgs = np.stack([make_image("smooth", h, w, seed=40 + k) for k in range(f)])
        job = sd.dense_job(layout, w, h, frames=f, flags=flags_for(layout))
        want = ctx.forward_host(imgs, lut, job)
        got = sd.forward_host_multi(ctxs, imgs, lut, job)
        assert np.array_equal(got, want), (f, h, w)
        back_want = ctx.inverse_host(want, lut, job)
        back = sd.inverse_host_multi(ctxs, want, lut, job)
        assert np.array_equal(back, back_want)
    for c in ctxs:
        c.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_devices_in_one_process():
    """kernel attributes (opt-in shared memory) are per device - a context on device 1 created after
    device 0 has run must launch the fp32 and the tensor-core kernels."""
    lut = ol.lut(8)
    img = make_image("rand", 256, 512, seed=50)
    outs = []
    for d in (0, 1):
        c = sd.Context(d)
        with torch.cuda.device(d):
            x = torch.from_numpy(img).cuda(d)
            res = []
            for layout in LAYOUTS:
                for flags in (flags_for(layout), sd.TENSOR):
                    job = sd.dense_job(layout, 512, 256, flags=flags)
                    coef = c.forward(x, lut, job)
                    res.append(coef.cpu().numpy())
                    res.append(c.inverse(coef, lut, job).cpu().numpy())
            torch.cuda.synchronize(d)
        outs.append(res)
        c.close()
    for a, b in zip(*outs):
        assert np.array_equal(a, b)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_library_leaves_the_callers_current_device_alone():
    """contexts on other devices (creation, launches, host-buffer calls, destruction) must not change the calling
    thread's current device - bench.py at N > 1 once compared tensors on the wrong device because of that"""
    torch.cuda.set_device(0)
    lut = ol.lut(8)
    img = make_image("rand", 64, 128, seed=51)
    c = sd.Context(1)
    assert torch.cuda.current_device() == 0
    job = sd.dense_job(sd.LAYOUT_B, 128, 64, flags=flags_for(sd.LAYOUT_B))
    x = torch.from_numpy(img).cuda(1)
    c.forward(x, lut, job)
    c.forward_host(img, lut, job)
    c.synchronize()
    assert torch.cuda.current_device() == 0
    c.close()
    assert torch.cuda.current_device() == 0
