#!/usr/bin/env python
"""Profiling only (build with -DSDCT_TRACE_CTA=1): when every warp of the headline kernel (forward, layout B, exact + X3) enters and
leaves, by globaltimer - how many warps are resident over the kernel's duration, per SM spread of the finishing times."""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import simd_dct_b200 as sd
from simd_dct_b200 import _lib
ctx = sd.Context(0)
job = sd.dense_job(sd.LAYOUT_B, 1920, 1080, frames=64, flags=sd.EXACT | sd.QUIRK_AVX_X3)
pix = torch.randint(0, 256, (64, 1080, 1920), dtype=torch.uint8, device="cuda")
out = torch.empty(64 * 1080 * 1920, dtype=torch.uint8, device="cuda")
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for it in range(30):
    if it == 29:
        ev0.record()
    ctx.forward(pix, sd.quality_lut(1), job, out=out)
ev1.record()
torch.cuda.synchronize()
print(f"last launch (back to back): {ev0.elapsed_time(ev1) * 1e3:.1f} us by CUDA events")
t = np.zeros(1024 * 4 * 3, dtype=np.uint64)
assert _lib.load().sdct_k_warp_times(t.ctypes.data_as(C.c_void_p)) == 0
t = t.reshape(1024, 4, 3).astype(np.int64)
n = int((t[:, 0, 0] > 0).sum())
t = t[:n]
t0 = t[:, :, 0].min()
start, end, sm = (t[:, :, 0] - t0) / 1e3, (t[:, :, 1] - t0) / 1e3, t[:, 0, 2]
q = lambda a: "min %.1f  p10 %.1f  median %.1f  p90 %.1f  max %.1f" % (a.min(), np.percentile(a, 10), np.median(a), np.percentile(a, 90), a.max())
print(f"{n} CTAs on {len(set(sm.tolist()))} SMs; us after the first warp's entry:")
print("  warp entry:", q(start))
print("  warp exit: ", q(end))
print("  warp lifetime:", q(end - start))
span = end.max()
grid = np.linspace(0, span, 41)
res = [(int(((start <= g) & (end > g)).sum())) for g in grid]
print("  resident warps over time (41 samples):", res)
print("  mean resident warps / SM: %.2f of 12" % (float(np.sum(end - start)) / span / 148))
per_sm = {}
for c in range(n):
    per_sm.setdefault(int(sm[c]), []).append(float(end[c].max()))
last = np.array([max(v) for v in per_sm.values()])
cnt = np.array([len(v) for v in per_sm.values()])
print("  CTAs per SM: min %d max %d; last exit per SM:" % (cnt.min(), cnt.max()), q(last))
