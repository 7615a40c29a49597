"""the colour front-end (SDCT_COLOR_YCBCR).  The CPU restatement (oracle/sdct_oracle.c) is pinned against
the textbook float formulas here; the GPU paths - conversion fused into the forward load (fp32 and tensor-core kernels)
and into the inverse store - against the restatement composed with the pinned single-plane oracle."""
import numpy as np
import pytest

import oraclelib as ol
from conftest import make_image


def test_oracle_colour_conversion_matches_jfif(oracle):
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (64, 96, 3), dtype=np.uint8)
    img[0, :8] = [[0, 0, 0], [255, 255, 255], [255, 0, 0], [0, 255, 0], [0, 0, 255], [255, 255, 0], [0, 255, 255], [255, 0, 255]]
    ycc = oracle.rgb_to_ycc(img)
    r, g, b = [img[..., k].astype(np.float64) for k in range(3)]
    want = np.stack([0.299 * r + 0.587 * g + 0.114 * b, 128 - 0.168736 * r - 0.331264 * g + 0.5 * b, 128 + 0.5 * r - 0.418688 * g - 0.081312 * b])
    assert np.abs(ycc.astype(np.float64) - want).max() <= 1.51          # 8-bit coefficients: within 1.5 levels of the exact matrix
    assert ycc[1:, 0, 0].tolist() == [128, 128] and ycc[0, 0, 0] == 0 and ycc[0, 0, 1] == 255   # grey stays grey
    back = oracle.ycc_to_rgb(ycc)
    assert np.abs(back.astype(int) - img.astype(int)).max() <= 3          # the 8-bit round trip
    # exhaustive check of ranges: every sum of the forward form is a byte
    grid = np.stack(np.meshgrid(*[np.arange(0, 256, 5)] * 3, indexing="ij"), -1).reshape(-1, 1, 3).astype(np.uint8)
    assert oracle.rgb_to_ycc(grid).shape == (3, grid.shape[0], 1)


@pytest.mark.gpu
@pytest.mark.parametrize("layout_name", ["A", "B", "C"])
def test_colour_front_end_on_the_gpu(ctx, oracle, layout_name):
    torch = pytest.importorskip("torch")
    import simd_dct_b200 as sd
    layout = {"A": sd.LAYOUT_A, "B": sd.LAYOUT_B, "C": sd.LAYOUT_C}[layout_name]
    for (h, w) in ((96, 512), (64, 2112), (64, 3840)):
        img = make_image("smooth", h, w, seed=61, comps=3)
        img[::3, ::5, 0] = 255 - img[::3, ::5, 0]     # some saturated colours
        ycc = oracle.rgb_to_ycc(img)
        lut = ol.lut(8)
        x = torch.from_numpy(img).cuda()
        quirk = sd.QUIRK_AVX_X3 if layout == sd.LAYOUT_B else 0
        for flags, mode, q in ((sd.EXACT | quirk, ol.MODE_EXACT, ol.QUIRK_AVX_X3), (sd.FAST, ol.MODE_FAST, 0), (sd.TENSOR, ol.MODE_TENSOR, 0)):
            job = sd.dense_job(layout, w, h, components=3, flags=flags | sd.COLOR_YCBCR)
            got = ctx.forward(x, lut, job).cpu().numpy()
            want = np.concatenate([oracle.forward(layout, ycc[k], lut, mode=mode, quirks=q) for k in range(3)])
            assert np.array_equal(got, want), (h, w, flags)
        # inverse side: decode + YCbCr -> RGB fused into the store (fp32 and tensor-core), and the round trips
        from test_gpu_parity import rt_lut
        lut = rt_lut(layout)
        coef = np.concatenate([oracle.forward(layout, ycc[k], lut) for k in range(3)])
        dec = np.stack([oracle.inverse(layout, coef[k * h * w:(k + 1) * h * w], w, h, lut) for k in range(3)])
        job = sd.dense_job(layout, w, h, components=3, flags=sd.EXACT | sd.COLOR_YCBCR)
        got = ctx.inverse(torch.from_numpy(coef).cuda(), lut, job).cpu().numpy().reshape(h, w, 3)
        assert np.array_equal(got, oracle.ycc_to_rgb(dec))
        assert np.array_equal(ctx.roundtrip(x, lut, job).cpu().numpy().reshape(h, w, 3), got)
        assert np.abs(got.astype(int) - img.astype(int)).mean() < 4
        dect = np.stack([oracle.inverse(layout, coef[k * h * w:(k + 1) * h * w], w, h, lut, tensor=True) for k in range(3)])
        jt = sd.dense_job(layout, w, h, components=3, flags=sd.TENSOR | sd.COLOR_YCBCR)
        gt = ctx.inverse(torch.from_numpy(coef).cuda(), lut, jt).cpu().numpy().reshape(h, w, 3)
        assert np.array_equal(gt, oracle.ycc_to_rgb(dect))
    # an odd number of blocks per row: the forward converts (byte-load path), the output side is refused
    h, w = 40, 72
    if layout != sd.LAYOUT_B:
        img = make_image("rand", h, w, seed=62, comps=3)
        ycc = oracle.rgb_to_ycc(img)
        job = sd.dense_job(layout, w, h, components=3, flags=sd.EXACT | sd.COLOR_YCBCR)
        buf = torch.zeros(h * w * 3 + 16, dtype=torch.uint8, device="cuda")
        buf[1:1 + h * w * 3] = torch.from_numpy(img).cuda().view(-1)      # unaligned base: byte loads
        got = ctx.forward(buf[1:], ol.lut(8), job).cpu().numpy()
        assert np.array_equal(got, np.concatenate([oracle.forward(layout, ycc[k], ol.lut(8)) for k in range(3)]))
        with pytest.raises(sd.SdctError):
            ctx.inverse(torch.zeros(h * w * 3, dtype=torch.uint8, device="cuda"), ol.lut(8), job)
    with pytest.raises(sd.SdctError):
        ctx.forward(torch.zeros(64 * 64, dtype=torch.uint8, device="cuda"), ol.lut(8), sd.dense_job(layout, 64, 64, flags=sd.COLOR_YCBCR))
