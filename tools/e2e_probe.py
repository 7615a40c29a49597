#!/usr/bin/env python
"""Profiling only: wall-clock time of the reference-facing host call (bench.py's e2e leg) for the current
$SDCT_HOST_CHUNK_MB."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import simd_dct_b200 as sd
F, H, W = 64, 1080, 1920
rng = np.random.default_rng(0)
h_in = torch.from_numpy(rng.integers(0, 256, (F, H, W), dtype=np.uint8)).pin_memory()
h_out = torch.empty(F * H * W, dtype=torch.uint8).pin_memory()
lut = sd.quality_lut(1.0)
def step():
    rc = sd.simdDCT_EncodeQuantize32ReorderBuffer(h_in, h_out, lut, W, 2 * F * H, 0, 2 * F * H)
    assert rc == sd.sdr_Success
for _ in range(3): step()
t0 = time.perf_counter()
for _ in range(10): step()
dt = (time.perf_counter() - t0) / 10
print(f"chunk_mb={os.environ.get('SDCT_HOST_CHUNK_MB','16')}: {dt*1e3:.3f} ms  {F*H*W/dt/1e9:.1f} GB/s each way", flush=True)
