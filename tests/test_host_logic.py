"""Host-side logic that needs no GPU: the C-ABI library loads and exports every symbol the public
headers declare, job geometry / strides / shard ranges, and the error paths (including the loud
failure when no device is present - there is no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import simd_dct_b200 as sd
from simd_dct_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//.*", "", txt)
    return sorted(set(re.findall(r"\b((?:sdct|simdDCT)_\w+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = declared_functions("sdct.h")
    assert len(names) >= 20
    assert set(names) == set(_lib.SYMBOLS), set(names) ^ set(_lib.SYMBOLS)
    for n in names:
        assert getattr(lib, n) is not None
    # the reference's three C++-linkage entry points keep their Itanium-mangled names (src/simd_dct.h:29-31)
    assert declared_functions("simd_dct.h") == sorted(_lib.SHIM_SYMBOLS)
    for mangled in _lib.SHIM_SYMBOLS.values():
        assert getattr(lib, mangled) is not None
    assert lib.sdct_abi_version() == 1


def test_library_does_not_link_the_oracle():
    out = os.popen("ldd %s" % _lib.LIB_PATH).read()
    assert "oracle" not in out and "simd_dct_ref" not in out


def test_dense_job_strides():
    j = sd.dense_job(sd.LAYOUT_C, 1920, 1080 - 1080 % 8, frames=3, components=3, eyes=2)
    h = 1080 - 1080 % 8
    assert (j.pitch, j.eye_stride, j.pixel_frame_stride) == (1920 * 3, 1920 * 3 * h, 1920 * 3 * h * 2)
    assert j.coef_plane_stride == 1920 * h * 2 // 64
    assert j.coef_comp_stride == 1920 * h * 2 and j.coef_frame_stride == 3 * j.coef_comp_stride
    j = sd.dense_job(sd.LAYOUT_B, 128, 64, eyes=2)
    assert j.eyes == 1  # eyes only exist for layout C


def test_shard_range_partitions_without_gaps():
    for total in (0, 1, 7, 64, 135, 4096):
        for world in (1, 2, 3, 4, 8):
            cuts = [sd.shard_range(total, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in cuts]
            assert max(sizes) - min(sizes) <= 1


def test_validation_and_error_strings_without_a_device():
    lib = _lib.load()
    job = sd.dense_job(sd.LAYOUT_B, 96, 64)           # width % 64 != 0
    buf = np.zeros(64, np.uint8)
    lut = sd.quality_lut(1)
    fp = lut.ctypes.data_as(C.POINTER(C.c_float))
    rc = lib.sdct_forward_device(None, C.byref(job), buf.ctypes.data, fp, buf.ctypes.data, None)
    assert rc == _lib.NOT_SUPPORTED and b"64" in lib.sdct_last_error()
    job = sd.dense_job(sd.LAYOUT_A, 64, 64, components=2)
    assert lib.sdct_forward_device(None, C.byref(job), buf.ctypes.data, fp, buf.ctypes.data, None) == _lib.NOT_SUPPORTED
    job = sd.dense_job(sd.LAYOUT_A, 64, 64)
    assert lib.sdct_forward_device(None, C.byref(job), buf.ctypes.data, fp, buf.ctypes.data, None) == _lib.INVALID_PARAMETER
    assert lib.sdct_forward_device(None, None, None, None, None, None) == _lib.INVALID_PARAMETER


def test_reference_entry_points_validate_like_the_reference():
    """src/simd_dct.cpp:75-76, :97-98, :117-118 - checked before any device is touched"""
    img = np.zeros(64 * 64, np.uint8)
    out = np.zeros_like(img)
    lut = sd.quality_lut(1)
    for fn, bad_x in ((sd.simdDCT_EncodeQuantizeBuffer, 60), (sd.simdDCT_EncodeQuantizeReorderStereoBuffer, 60),
                      (sd.simdDCT_EncodeQuantize32ReorderBuffer, 32)):
        assert fn(None, out, lut, 64, 64, 0, 64) == sd.sdr_InvalidParameter
        assert fn(img, None, lut, 64, 64, 0, 64) == sd.sdr_InvalidParameter
        assert fn(img, out, lut, bad_x, 64, 0, 64) == sd.sdr_NotSupported
        assert fn(img, out, lut, 64, 60, 0, 64) == sd.sdr_NotSupported
        assert fn(img, out, lut, 64, 64, 100, 200) == sd.sdr_Success   # empty window: nothing to do, no device needed


def test_no_cpu_fallback_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(sd.SdctError) as e:
        sd.Context(0)
    assert e.value.code == _lib.NO_DEVICE and "no CPU fallback" in str(e.value)
    img = np.zeros(64 * 64, np.uint8)
    out = np.full_like(img, 9)
    assert sd.simdDCT_EncodeQuantizeBuffer(img, out, sd.quality_lut(1), 64, 64, 0, 64) == sd.sdr_NotSupported
    assert (out == 9).all() and "no CPU fallback" in sd.last_error()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "simd_dct_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in txt and "oraclelib" not in txt and "simd_dct_ref" not in txt and "dlopen" not in txt, f
                assert not re.search(r'#\s*include[^\n]*oracle', txt), f   # comments may cite the oracle, code may not use it
