#!/usr/bin/env python
"""Profiling only: the device entry points called straight on PINNED HOST buffers (unified addressing: the kernel itself reads
and writes host memory over PCIe, no copy engine, no chunks), against the chunked host pipeline (bench.py's e2e leg)."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import simd_dct_b200 as sd
F, H, W = 64, 1080, 1920
rng = np.random.default_rng(0)
h_in = torch.from_numpy(rng.integers(0, 256, (F, H, W), dtype=np.uint8)).pin_memory()
h_out = torch.empty(F * H * W, dtype=torch.uint8).pin_memory()
d_in = h_in.cuda()
d_out = torch.empty(F * H * W, dtype=torch.uint8, device="cuda")
lut = sd.quality_lut(1.0)
ctx = sd.Context(0)


def wall(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
        torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n


def shim():
    assert sd.simdDCT_EncodeQuantize32ReorderBuffer(h_in, h_out, lut, W, 2 * F * H, 0, 2 * F * H) == sd.sdr_Success


t = wall(shim)
print(f"chunked pipeline (shim): {t*1e3:.3f} ms  {F*H*W/t/1e9:.1f} GB/s each way", flush=True)
ref = h_out.clone()
for name, flags in (("exact+x3", sd.EXACT | sd.QUIRK_AVX_X3), ("tensor", sd.TENSOR), ("fast", sd.FAST)):
    job = sd.dense_job(sd.LAYOUT_B, W, H, frames=F, flags=flags)
    for src, dst, label in ((h_in, h_out, "host -> host"), (h_in, d_out, "host -> device"), (d_in, h_out, "device -> host"), (d_in, d_out, "device -> device")):
        try:
            dst.zero_()
            t = wall(lambda: ctx.forward(src, lut, job, out=dst))
            ok = ""
            if name == "exact+x3":
                ok = " bytes == pipeline's: %s" % bool(torch.equal(dst.cpu(), ref))
            print(f"{name:9s} {label:16s}: {t*1e3:.3f} ms  {F*H*W/t/1e9:.1f} GB/s each way{ok}", flush=True)
        except Exception as e:  # noqa: BLE001
            print(f"{name:9s} {label:16s}: failed: {e}", flush=True)
