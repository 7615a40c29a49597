#!/usr/bin/env python
"""Profiling only: per-tile clock64 stamps of CTA 0's producer / MMA / consumer roles (SDCT_TC_ABLATE bit 5)."""
import ctypes as C, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
abl = int(sys.argv[1]) if len(sys.argv) > 1 else 0
layout_name = sys.argv[2] if len(sys.argv) > 2 else "A"
inverse = len(sys.argv) > 3 and sys.argv[3] == "inverse"
roundtrip = len(sys.argv) > 3 and sys.argv[3] == "roundtrip"
os.environ["SDCT_TC_ABLATE"] = str(abl | 32)
import simd_dct_b200 as sd
from simd_dct_b200 import _lib
ctx = sd.Context(0)
L = {"A": sd.LAYOUT_A, "B": sd.LAYOUT_B, "C": sd.LAYOUT_C}[layout_name]
job = sd.dense_job(L, 1920, 1080, frames=64, flags=sd.TENSOR)
pix = torch.randint(0, 256, (64, 1080, 1920), dtype=torch.uint8, device="cuda")
out = torch.empty(64 * 1080 * 1920, dtype=torch.uint8, device="cuda")
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for it in range(3):
    if it == 2:
        torch.cuda.synchronize()
        ev0.record()
    if roundtrip:
        ctx.roundtrip(pix, sd.quality_lut(1), job, out=out.view(64, 1080, 1920))
    elif inverse:
        ctx.inverse(out, sd.quality_lut(1), job, out=pix.view(-1))
    else:
        ctx.forward(pix, sd.quality_lut(1), job, out=out)
ev1.record()
torch.cuda.synchronize()
print(f"last launch: {ev0.elapsed_time(ev1) * 1e3:.1f} us by CUDA events")
ct = np.zeros(3 * 512, dtype=np.uint64)
lib = _lib.load()
if hasattr(lib, "sdct_k_tc_cta_times") and lib.sdct_k_tc_cta_times(ct.ctypes.data_as(C.c_void_p)) == 0:
    ct = ct.reshape(3, 512).astype(np.int64)
    n = int((ct[0] > 0).sum())
    e0 = ct[0, :n].min()
    q = lambda a: "min %.1f  median %.1f  max %.1f" % (a.min() / 1e3, np.median(a) / 1e3, a.max() / 1e3)
    print(f"{n} CTAs, us after the first CTA's entry: entry {q(ct[0, :n] - e0)} | prologue done {q(ct[1, :n] - e0)} | exit {q(ct[2, :n] - e0)}")
    print("   prologue duration: " + q(ct[1, :n] - ct[0, :n]) + " | CTA 0: entry %.1f prologue done %.1f exit %.1f" % tuple((ct[i, 0] - e0) / 1e3 for i in range(3)))
tr = np.zeros(3 * 256 * 8, dtype=np.uint64)
assert _lib.load().sdct_k_tc_trace(tr.ctypes.data_as(C.c_void_p)) == 0
tr = tr.reshape(3, 256, 8).astype(np.int64)
t0 = tr[0, 0, 0]
names = ["producer", "mma", "consumer0"]
print(f"ablate={abl} layout={layout_name} {'roundtrip' if roundtrip else 'inverse' if inverse else 'forward'}; cycles relative to producer tile 0")
for role in range(3):
    print(names[role])
    last = max(k for k in range(256) if tr[role, k, 0])
    ks = list(range(0, 4)) + list(range(12, 30)) + list(range(last - 3, last + 1)) if role < 2 else list(range(0, 4, 2)) + list(range(12, 30, 2)) + list(range(last - 2, last + 1, 2))
    prev = None
    for k in ks:
        row = tr[role, k]
        if row[0] == 0:
            continue
        rel = [int(x - t0) if x else -1 for x in row[:8]]
        d = "" if prev is None else f"  (+{rel[0] - prev} since previous)"
        prev = rel[0]
        print(f"  k={k:3d} " + " ".join(f"{x:8d}" for x in rel) + d)
span = int(tr[tr > 0].max() - t0)
print(f"CTA 0: first producer stamp to last stamp of any role: {span} clocks = {span / 1.965e3:.1f} us at 1965 MHz")
