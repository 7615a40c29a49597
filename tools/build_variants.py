#!/usr/bin/env python
"""Experiment builds of libsdct.so: tools/build_variants.py name=DEF1,DEF2 [name2=...] writes
simd_dct_b200/variants/libsdct_<name>.so (git-ignored, travels to the GPU box); select one with
SDCT_LIB=<path>.  Used to time kernel alternatives side by side in one gpurun call."""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simd_dct_b200 import build as b  # noqa: E402

outdir = os.path.join(ROOT, "simd_dct_b200", "variants")
os.makedirs(outdir, exist_ok=True)


def one(arg):
    name, _, defs = arg.partition("=")
    out = os.path.join(outdir, f"libsdct_{name}.so")
    b.build(force=True, defs=[d for d in defs.split(",") if d], out=out)
    return out


with ThreadPoolExecutor(2) as ex:
    for path in ex.map(one, sys.argv[1:]):
        print(path)
