"""per-plane statistics of layout-C coefficient sets (sdct_plane_stats_device).  The CPU restatement
(oracle/sdct_oracle.c: sdct_oracle_plane_stats) is pinned against a numpy one-liner here; the GPU kernel against the
restatement on sets produced by the forward kernel."""
import numpy as np
import pytest

import oraclelib as ol
from conftest import make_image


def numpy_stats(plane):
    a = np.abs(plane.astype(np.int64) - 127)
    z = a == 0
    starts = z & np.concatenate([[True], ~z[:-1]])
    return np.array([z.sum(), starts.sum(), a.sum(), np.bitwise_or.reduce(a) if a.size else 0], dtype=np.uint64)


def test_oracle_plane_stats_equals_numpy(oracle):
    rng = np.random.default_rng(3)
    for n in (0, 1, 15, 16, 17, 4096, 40000):
        for p_zero in (0.0, 0.5, 0.97, 1.0):
            plane = rng.integers(0, 256, n, dtype=np.uint8)
            plane[rng.random(n) < p_zero] = 127
            assert np.array_equal(oracle.plane_stats(plane), numpy_stats(plane)), (n, p_zero)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["smooth", "rand"])
def test_plane_stats_on_the_gpu(ctx, oracle, kind):
    torch = pytest.importorskip("torch")
    import simd_dct_b200 as sd
    f, h, w = 2, 264, 1920     # 33 bands x 240 blocks: planes of 7920 bytes (not a multiple of the kernel's pieces)
    imgs = np.stack([make_image(kind, h, w, seed=40 + k, comps=3) for k in range(f)])
    lut = ol.lut(24)
    job = sd.dense_job(sd.LAYOUT_C, w, h, frames=f, components=3, flags=sd.EXACT)
    coef = ctx.forward(torch.from_numpy(imgs).cuda(), lut, job)
    got = ctx.plane_stats(coef, job).cpu().numpy().astype(np.uint64)
    planes = coef.cpu().numpy().reshape(f, 3, 64, -1)
    for k in range(f):
        for c in range(3):
            for i in (0, 1, 8, 27, 63):
                assert np.array_equal(got[k, c, i], oracle.plane_stats(planes[k, c, i])), (k, c, i)
    want = np.array([[[numpy_stats(planes[k, c, i]) for i in range(64)] for c in range(3)] for k in range(f)])
    assert np.array_equal(got, want)
    # a band window (16-byte aligned start: 240 blocks per band) and an unaligned one-plane job
    win = sd.dense_job(sd.LAYOUT_C, w, h, frames=f, components=3, flags=sd.EXACT, band_begin=5, band_end=12)
    gw = ctx.plane_stats(coef, win).cpu().numpy().astype(np.uint64)
    assert np.array_equal(gw[1, 2, 9], oracle.plane_stats(planes[1, 2, 9, 5 * 240:12 * 240]))
    h2, w2 = 40, 72            # 9 blocks per row: every plane starts at an odd offset
    img = make_image(kind, h2, w2, seed=44)
    j2 = sd.dense_job(sd.LAYOUT_C, w2, h2, flags=sd.EXACT)
    c2 = ctx.forward(torch.from_numpy(img).cuda(), lut, j2)
    g2 = ctx.plane_stats(c2, j2).cpu().numpy().astype(np.uint64)
    p2 = c2.cpu().numpy().reshape(64, -1)
    assert np.array_equal(g2[0, 0], np.array([numpy_stats(p2[i]) for i in range(64)]))
    with pytest.raises(sd.SdctError):
        ctx.plane_stats(c2, sd.dense_job(sd.LAYOUT_A, w2, h2))
