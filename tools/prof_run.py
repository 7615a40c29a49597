#!/usr/bin/env python
"""Runs selected kernels a few times on the bench workload (for ncu / compute-sanitizer).
usage: prof_run.py [--frames 64] [--w 1920] [--h 1080] [--reps 4] name [name ...]
names: fwd_{A,B,C}_{exact,fast,tensor}, inv_{A,B,C}[_tensor], rt_{A,B,C}_{exact,fast,tensor}, rgb_{A,B,C}_{exact,fast,tensor} (interleaved RGB
forward), irgb_* (inverse to interleaved RGB), rtrgb_* (round trip)"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import simd_dct_b200 as sd  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=64)
ap.add_argument("--w", type=int, default=1920)
ap.add_argument("--h", type=int, default=1080)
ap.add_argument("--reps", type=int, default=4)
ap.add_argument("names", nargs="+")
a = ap.parse_args()
L = {"A": sd.LAYOUT_A, "B": sd.LAYOUT_B, "C": sd.LAYOUT_C}
ctx = sd.Context(0)
lut = sd.quality_lut(1.0)
rng = np.random.default_rng(0)
for name in a.names:
    parts = name.split("_")
    op, lay = parts[0], L[parts[1]]
    mode = parts[2] if len(parts) > 2 else "exact"
    comps = 3 if op in ("rgb", "irgb", "rtrgb") else 1
    flags = sd.TENSOR if mode == "tensor" else sd.FAST if mode == "fast" else (sd.EXACT | (sd.QUIRK_AVX_X3 if lay == sd.LAYOUT_B else 0))
    if len(parts) > 3 and parts[3] == "ycc":   # colour front-end (interleaved pixels)
        flags |= sd.COLOR_YCBCR
    job = sd.dense_job(lay, a.w, a.h, frames=a.frames, components=comps, flags=flags)
    shape = (a.frames, a.h, a.w) if comps == 1 else (a.frames, a.h, a.w, 3)
    pix = torch.from_numpy(rng.integers(0, 256, shape, dtype=np.uint8)).cuda()
    coef = torch.empty(a.frames * a.h * a.w * comps, dtype=torch.uint8, device="cuda")
    back = torch.empty_like(pix)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.forward(pix, lut, job, out=coef)
    for rep in range(a.reps):
        if rep == 1:
            e0.record()
        if op in ("fwd", "rgb"):
            ctx.forward(pix, lut, job, out=coef)
        elif op in ("inv", "irgb"):
            ctx.inverse(coef, lut, job, out=back)
        else:
            ctx.roundtrip(pix, lut, job, out=back)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / max(1, a.reps - 1)
    blocks = a.frames * (a.h // 8) * (a.w // 8) * comps
    print(f"{name}: {ms*1e3:.1f} us/launch, {blocks*128/ms/1e6:.0f} GB/s algorithmic", flush=True)
