"""ctypes access to the CPU oracle (oracle/liboracle.so) and, when it was built, to the
unmodified reference (oracle/_ref/libsimd_dct_ref_{strict,fast}.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing under simd_dct_b200/ imports this module.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

LAYOUT_A, LAYOUT_B, LAYOUT_C = 0, 1, 2
MODE_EXACT, MODE_FAST, MODE_TENSOR = 0, 1, 2
QUIRK_NONE, QUIRK_AVX_X3, QUIRK_SSE = 0, 1, 2
TAB_SCALAR, TAB_AVX, TAB_FAST, TAB_INVERSE = 0, 1, 2, 3

# the reference CLI's default quantisation LUT (src/main.cpp:179-189): the JPEG luminance
# table / 100 with entry [0] = .17; `--quality N` multiplies every entry by N (:214-217)
DEFAULT_LUT = np.array(
    [.17, .11, .10, .16, .24, .40, .51, .61,
     .12, .12, .14, .19, .26, .58, .60, .55,
     .14, .13, .16, .24, .40, .57, .69, .56,
     .14, .17, .22, .29, .51, .87, .80, .62,
     .18, .22, .37, .56, .68, 1.09, 1.03, .77,
     .24, .35, .55, .64, .81, 1.04, 1.13, .92,
     .49, .64, .78, .87, 1.03, 1.21, 1.20, 1.01,
     .72, .92, .95, .98, 1.12, 1.00, 1.03, .99], dtype=np.float32)

_u8p = C.POINTER(C.c_uint8)
_f32p = C.POINTER(C.c_float)
_sz = C.c_size_t


def _p8(a):
    return a.ctypes.data_as(_u8p)


def _pf(a):
    return a.ctypes.data_as(_f32p)


def lut(scale: float = 1.0) -> np.ndarray:
    return (DEFAULT_LUT * np.float32(scale)).astype(np.float32)


class Oracle:
    def __init__(self, path: str = ORACLE_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing - run `make -C oracle` (or __graft_entry__.build())")
        L = self.lib = C.CDLL(path)
        L.sdct_oracle_forward.argtypes = [C.c_int, C.c_int, C.c_int, _u8p, _sz, _sz, _sz, _sz, C.c_int, _sz,
                                          _f32p, _u8p, _sz]
        L.sdct_oracle_forward.restype = None
        L.sdct_oracle_inverse.argtypes = [C.c_int, _u8p, _sz, _sz, _sz, C.c_int, _sz, _f32p, _u8p, _sz, _sz]
        L.sdct_oracle_inverse.restype = None
        L.sdct_oracle_inverse_tensor.argtypes = L.sdct_oracle_inverse.argtypes
        L.sdct_oracle_inverse_tensor.restype = None
        L.sdct_oracle_tables.argtypes = [_f32p, C.c_int, _f32p, C.c_int]
        L.sdct_oracle_tables.restype = None
        for n in ("dct8_scalar", "dct8_fast", "idct8", "dct8_sse"):
            getattr(L, "sdct_oracle_" + n).argtypes = [_f32p]
            getattr(L, "sdct_oracle_" + n).restype = None
        L.sdct_oracle_dct8_paired.argtypes = [_f32p, C.c_int]
        L.sdct_oracle_dct8_paired.restype = None
        L.sdct_oracle_rgb_to_ycc.argtypes = [_u8p, _sz, _sz, _sz, _u8p]
        L.sdct_oracle_ycc_to_rgb.argtypes = [_u8p, _sz, _sz, _u8p, _sz]
        L.sdct_oracle_plane_stats.argtypes = [_u8p, _sz, C.POINTER(C.c_uint64)]
        L.sdct_oracle_plane_stats.restype = None
        L.sdct_oracle_deinterleave.argtypes = [_u8p, _sz, _sz, _sz, C.c_int, _u8p]
        L.sdct_oracle_interleave.argtypes = [_u8p, _sz, _sz, C.c_int, _u8p, _sz]
        sig = [_u8p, _u8p, _f32p, _sz, _sz, _sz, _sz]
        L.sdct_oracle_ref_encode_quantize_buffer.argtypes = sig
        L.sdct_oracle_ref_encode_quantize32_reorder_buffer.argtypes = sig + [C.c_int]
        L.sdct_oracle_ref_encode_quantize_reorder_stereo_buffer.argtypes = sig
        L.sdct_oracle_ref_encode_quantize_reorder_stereo_buffer_q.argtypes = sig + [C.c_int]

    # ---- native semantics -------------------------------------------------------------
    def forward(self, layout, img, lut_, mode=MODE_EXACT, quirks=QUIRK_NONE, eyes=1, comp=None):
        """img: (H, W) u8 plane, or (eyes*H, W) for layout C with eyes=2, or (H, W, ncomp)
        interleaved with comp = component index.  Returns the flat coefficient buffer."""
        img = np.ascontiguousarray(img)
        if img.ndim == 3:
            h, w, nc = img.shape
            pitch, xstep, off = w * nc, nc, int(comp)
        else:
            h, w = img.shape
            pitch, xstep, off = w, 1, 0
        h //= eyes
        out = np.zeros(eyes * h * w, dtype=np.uint8)
        base = C.cast(C.c_void_p(img.ctypes.data + off), _u8p)
        self.lib.sdct_oracle_forward(layout, mode, quirks, base, pitch, xstep, w, h, eyes, h * pitch,
                                     _pf(lut_), _p8(out), eyes * h * w // 64)
        return out

    def inverse(self, layout, coef, w, h, lut_, eyes=1, tensor=False):
        coef = np.ascontiguousarray(coef)
        out = np.zeros((eyes * h, w), dtype=np.uint8)
        fn = self.lib.sdct_oracle_inverse_tensor if tensor else self.lib.sdct_oracle_inverse
        fn(layout, _p8(coef), eyes * h * w // 64, w, h, eyes, h * w, _pf(lut_), _p8(out), w, 1)
        return out

    def tensor_matrix(self):
        k = np.zeros(64 * 64, dtype=np.int32)
        self.lib.sdct_oracle_tensor_matrix(k.ctypes.data_as(C.POINTER(C.c_int32)))
        return k.reshape(64, 64)

    def tables(self, lut_, layout, which):
        out = np.zeros(64, dtype=np.float32)
        self.lib.sdct_oracle_tables(_pf(lut_), layout, _pf(out), which)
        return out

    def dct8(self, kind, v, quirk=0):
        v = np.array(v, dtype=np.float32)
        if kind == "paired":
            self.lib.sdct_oracle_dct8_paired(_pf(v), quirk)
        else:
            getattr(self.lib, "sdct_oracle_" + kind)(_pf(v))
        return v

    def dct8_many(self, kind, vs, quirk=0):
        vs = np.ascontiguousarray(vs, dtype=np.float32).copy()
        fn = self.lib.sdct_oracle_dct8_paired if kind == "paired" else getattr(self.lib, "sdct_oracle_" + kind)
        base = vs.ctypes.data
        for k in range(vs.shape[0]):
            ptr = C.cast(C.c_void_p(base + 32 * k), _f32p)
            if kind == "paired":
                fn(ptr, quirk)
            else:
                fn(ptr)
        return vs

    def rgb_to_ycc(self, img):
        """(H, W, 3) interleaved RGB -> (3, H, W) planar Y, Cb, Cr (the colour front-end's definition)"""
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w, _ = img.shape
        out = np.empty((3, h, w), dtype=np.uint8)
        self.lib.sdct_oracle_rgb_to_ycc(_p8(img), 3 * w, w, h, _p8(out))
        return out

    def ycc_to_rgb(self, planes):
        planes = np.ascontiguousarray(planes, dtype=np.uint8)
        _, h, w = planes.shape
        out = np.empty((h, w, 3), dtype=np.uint8)
        self.lib.sdct_oracle_ycc_to_rgb(_p8(planes), w, h, _p8(out), 3 * w)
        return out

    def plane_stats(self, plane):
        """{zeros, zero_runs, sum_abs, or_abs} of one coefficient plane (bytes in block order)"""
        plane = np.ascontiguousarray(plane, dtype=np.uint8)
        out = np.zeros(4, dtype=np.uint64)
        self.lib.sdct_oracle_plane_stats(_p8(plane), plane.size, out.ctypes.data_as(C.POINTER(C.c_uint64)))
        return out

    def deinterleave(self, img):
        h, w, nc = img.shape
        img = np.ascontiguousarray(img)
        out = np.zeros((nc, h, w), dtype=np.uint8)
        self.lib.sdct_oracle_deinterleave(_p8(img), w * nc, w, h, nc, _p8(out))
        return out

    # ---- the reference's call semantics -------------------------------------------------
    def ref_call(self, name, buf, lut_, size_x, size_y, start_y=0, end_y=None, quirks=QUIRK_AVX_X3, fill=0):
        end_y = size_y if end_y is None else end_y
        buf = np.ascontiguousarray(buf).reshape(-1)
        out = np.full(size_x * size_y, fill, dtype=np.uint8)
        if name == "A":
            rc = self.lib.sdct_oracle_ref_encode_quantize_buffer(_p8(buf), _p8(out), _pf(lut_), size_x, size_y,
                                                                 start_y, end_y)
        elif name == "B":
            rc = self.lib.sdct_oracle_ref_encode_quantize32_reorder_buffer(_p8(buf), _p8(out), _pf(lut_), size_x,
                                                                           size_y, start_y, end_y, quirks)
        else:
            rc = self.lib.sdct_oracle_ref_encode_quantize_reorder_stereo_buffer_q(_p8(buf), _p8(out), _pf(lut_), size_x, size_y,
                                                                                  start_y, end_y, quirks & QUIRK_SSE)
        return rc, out


_REF_SYMS = {
    "A": b"_Z28simdDCT_EncodeQuantizeBufferPKhPhPKfmmmm",
    "B": b"_Z37simdDCT_EncodeQuantize32ReorderBufferPKhPhPKfmmmm",
    "C": b"_Z41simdDCT_EncodeQuantizeReorderStereoBufferPKhPhPKfmmmm",
}
_FLAGS = ["sseSupported", "sse2Supported", "sse3Supported", "ssse3Supported", "sse41Supported", "sse42Supported",
          "avxSupported", "avx2Supported", "fma3Supported", "avx512FSupported", "avx512VLSupported"]


class Reference:
    """The unmodified reference (src/simd_dct.cpp + src/simd_platform.c) compiled by
    oracle/Makefile.  `isa` picks the variant through the reference's own feature bools
    (src/simd_platform.c:30-55): 'scalar' (all false), 'sse41', 'avx2', 'avx512vl', 'native'."""

    def __init__(self, build: str = "strict"):
        path = os.path.join(REF_DIR, f"libsimd_dct_ref_{build}.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.fn = {}
        for k, sym in _REF_SYMS.items():
            f = getattr(self.lib, sym.decode())
            f.argtypes = [_u8p, _u8p, _f32p, _sz, _sz, _sz, _sz]
            f.restype = C.c_int
            self.fn[k] = f
        self.set_isa("scalar")

    def _flag(self, name):
        return C.c_bool.in_dll(self.lib, name)

    def host_supports(self, name):
        return name in open("/proc/cpuinfo").read().split("flags", 1)[-1].split("\n", 1)[0].split()

    def set_isa(self, isa: str):
        for n in _FLAGS:
            self._flag(n).value = False
        if isa == "scalar":
            pass
        elif isa == "sse41":
            for n in ("sseSupported", "sse2Supported", "sse3Supported", "ssse3Supported", "sse41Supported"):
                self._flag(n).value = True
        elif isa == "ssse3":
            for n in ("sseSupported", "sse2Supported", "sse3Supported", "ssse3Supported"):
                self._flag(n).value = True
        elif isa == "sse2":
            for n in ("sseSupported", "sse2Supported"):
                self._flag(n).value = True
        elif isa == "avx2":
            self._flag("avx2Supported").value = True
        elif isa == "avx512vl":
            self._flag("avx512VLSupported").value = True
        elif isa == "native":
            self.lib._DetectCPUFeatures()
        else:
            raise ValueError(isa)
        self.isa = isa

    def call(self, name, buf, lut_, size_x, size_y, start_y=0, end_y=None, fill=0, out=None):
        end_y = size_y if end_y is None else end_y
        buf = np.ascontiguousarray(buf).reshape(-1)
        if out is None:
            # + slack: the reference's layout-A SIMD variants write past the half image
            out = np.full(size_x * size_y + 4096, fill, dtype=np.uint8)
        rc = self.fn[name](_p8(buf), _p8(out), _pf(lut_), size_x, size_y, start_y, end_y)
        return rc, out[: size_x * size_y]


def have_reference(build="strict"):
    return os.path.exists(os.path.join(REF_DIR, f"libsimd_dct_ref_{build}.so"))


def cpu_flags():
    try:
        txt = open("/proc/cpuinfo").read()
        return set(txt.split("flags", 1)[1].split(":", 1)[1].split("\n", 1)[0].split())
    except Exception:
        return set()
