#!/usr/bin/env python
"""Generates tests/golden/reference_vectors.npz by RUNNING THE UNMODIFIED REFERENCE
(oracle/_ref/libsimd_dct_ref_strict.so = /root/reference/src/simd_dct.cpp + simd_platform.c compiled
by oracle/Makefile with -O2 -fno-fast-math -ffp-contract=off).  Only works where /root/reference is
mounted (the build container); the vectors travel with the repo, the reference does not.

Contents: two seeded 64x128 u8 inputs, and for each input / LUT scale / entry point the bytes the
reference wrote (through its public API, with its own call semantics: sizeY = stereo height, only
rows below sizeY/2 visited for A and B).  Layout A and C come from the scalar path (all CPU feature
bools false), layout B from the AVX2 path (avx2Supported = true only) - SURVEY.md section 8c; Csse_* are layout C from
the SSE4.1 stereo encoder (the quirks = SSE restatement's pin)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oraclelib as ol  # noqa: E402
from conftest import make_image  # noqa: E402

ref = ol.Reference("strict")
assert "avx2" in ol.cpu_flags(), "the AVX2 path of the reference must be executable to make the layout-B vectors"
assert "sse4_1" in ol.cpu_flags()
out = {}
W, H = 128, 64
for kind in ("rand", "smooth"):
    img = make_image(kind, H, W, seed=42)
    out[f"in_{kind}"] = img
    for scale in (1, 8, 255):
        lut = ol.lut(scale)
        # A and B: sizeY = 2H with an H-row buffer -> the whole plane is encoded
        ref.set_isa("scalar")
        rc, a = ref.call("A", img, lut, W, 2 * H, 0, 2 * H, out=np.zeros(W * 2 * H + 4096, np.uint8))
        assert rc == 0
        out[f"A_{kind}_x{scale}"] = a[: W * H].copy()
        ref.set_isa("avx2")
        rc, b = ref.call("B", img, lut, W, 2 * H, 0, 2 * H, out=np.zeros(W * 2 * H + 4096, np.uint8))
        assert rc == 0
        out[f"B_{kind}_x{scale}"] = b[: W * H].copy()
        # C: the image is the stereo pair (left eye = top half), sizeY = H
        ref.set_isa("scalar")
        rc, c = ref.call("C", img, lut, W, H, 0, H)
        assert rc == 0
        out[f"C_{kind}_x{scale}"] = c.copy()
        # ... and the bytes of its SSE4.1 stereo encoder (the variant its dispatcher picks on x86-64; flag SDCT_QUIRK_SSE)
        ref.set_isa("sse41")
        rc, c = ref.call("C", img, lut, W, H, 0, H)
        assert rc == 0
        out[f"Csse_{kind}_x{scale}"] = c.copy()
    # window semantics (inclusive endY, 2y vs y comparison), scale 8
    lut = ol.lut(8)
    for name, isa in (("A", "scalar"), ("B", "avx2"), ("C", "scalar")):
        ref.set_isa(isa)
        for (s, e) in ((16, 40), (0, 16), (33, 64)):
            rc, o = ref.call(name, img, lut, W, H, s, e, fill=7)
            assert rc == 0
            out[f"win_{name}_{kind}_{s}_{e}"] = o.copy()
path = os.path.join(ROOT, "tests", "golden", "reference_vectors.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), "bytes,", len(out), "arrays")
