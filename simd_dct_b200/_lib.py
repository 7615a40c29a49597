"""ctypes binding of libsdct.so (include/sdct.h).  Loading fails loudly when the library has not
been built; there is no Python or CPU fallback for any compute entry point."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# $SDCT_LIB: an experiment build of the same library (tools/build_variants.py); never a different implementation
LIB_PATH = os.environ.get("SDCT_LIB") or os.path.join(HERE, "libsdct.so")

OK, INVALID_PARAMETER, NOT_SUPPORTED, CUDA_ERROR, NO_DEVICE = range(5)
LAYOUT_A, LAYOUT_B, LAYOUT_C = 0, 1, 2
EXACT, FAST, QUIRK_AVX_X3, TENSOR, COLOR_YCBCR, QUIRK_SSE = 0, 1, 2, 4, 8, 16


class Job(C.Structure):
    """struct sdct_job (include/sdct.h)"""
    _fields_ = [("layout", C.c_int32), ("flags", C.c_int32), ("components", C.c_int32), ("eyes", C.c_int32),
                ("width", C.c_uint64), ("height", C.c_uint64), ("pitch", C.c_uint64), ("eye_stride", C.c_uint64),
                ("frames", C.c_uint64), ("pixel_frame_stride", C.c_uint64), ("coef_frame_stride", C.c_uint64),
                ("coef_comp_stride", C.c_uint64), ("coef_plane_stride", C.c_uint64),
                ("band_begin", C.c_uint64), ("band_end", C.c_uint64)]


# every symbol include/sdct.h declares: name -> (restype, argtypes)
_vp, _sz, _fp = C.c_void_p, C.c_size_t, C.POINTER(C.c_float)
_JP = C.POINTER(Job)
SYMBOLS = {
    "sdct_job_dense": (None, [_JP, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64]),
    "sdct_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "sdct_destroy": (None, [_vp]),
    "sdct_last_error": (C.c_char_p, []),
    "sdct_abi_version": (C.c_int, []),
    "sdct_device_count": (C.c_int, []),
    "sdct_sm_count": (C.c_int, [_vp]),
    "sdct_kernel_launches": (C.c_uint64, [_vp]),
    "sdct_forward_device": (C.c_int, [_vp, _JP, _vp, _fp, _vp, _vp]),
    "sdct_inverse_device": (C.c_int, [_vp, _JP, _vp, _fp, _vp, _vp]),
    "sdct_roundtrip_device": (C.c_int, [_vp, _JP, _vp, _fp, _vp, _vp]),
    "sdct_roundtrip_noalloc_device": (C.c_int, [_vp, _JP, _vp, _fp, _vp, _vp]),
    "sdct_forward_host": (C.c_int, [_vp, _JP, _vp, _fp, _vp]),
    "sdct_inverse_host": (C.c_int, [_vp, _JP, _vp, _fp, _vp]),
    "sdct_forward_host_multi": (C.c_int, [C.POINTER(_vp), C.c_int, _JP, _vp, _fp, _vp]),
    "sdct_inverse_host_multi": (C.c_int, [C.POINTER(_vp), C.c_int, _JP, _vp, _fp, _vp]),
    "sdct_plane_stats_device": (C.c_int, [_vp, _JP, _vp, _vp, _vp]),
    "sdct_selftest_dct8": (C.c_int, [_vp, C.c_int, _fp, _sz]),
    "sdct_tensor_matrix": (None, [C.POINTER(C.c_int32)]),
    "sdct_synchronize": (C.c_int, [_vp]),
    "sdct_host_alloc": (_vp, [_sz]),
    "sdct_host_free": (None, [_vp]),
    "sdct_ref_encode_quantize_buffer": (C.c_int, [_vp, _vp, _fp, _sz, _sz, _sz, _sz]),
    "sdct_ref_encode_quantize_reorder_stereo_buffer": (C.c_int, [_vp, _vp, _fp, _sz, _sz, _sz, _sz]),
    "sdct_ref_encode_quantize32_reorder_buffer": (C.c_int, [_vp, _vp, _fp, _sz, _sz, _sz, _sz]),
    "sdct_ref_set_flags": (None, [C.c_int]),
    "sdct_ref_get_flags": (C.c_int, []),
    "sdct_ref_device_count": (C.c_int, []),
    "sdct_shard_range": (None, [C.c_uint64, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
}
# the reference's C++-linkage entry points (include/simd_dct.h), Itanium-mangled
SHIM_SYMBOLS = {
    "simdDCT_EncodeQuantizeBuffer": "_Z28simdDCT_EncodeQuantizeBufferPKhPhPKfmmmm",
    "simdDCT_EncodeQuantizeReorderStereoBuffer": "_Z41simdDCT_EncodeQuantizeReorderStereoBufferPKhPhPKfmmmm",
    "simdDCT_EncodeQuantize32ReorderBuffer": "_Z37simdDCT_EncodeQuantize32ReorderBufferPKhPhPKfmmmm",
}

_lib = None


class SdctError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"sdct status {code}: {msg}")
        self.code = code


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -m simd_dct_b200.build` "
                          "(or __graft_entry__.build()); simd_dct_b200 has no fallback path")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    for name, mangled in SHIM_SYMBOLS.items():
        fn = getattr(lib, mangled)
        fn.restype, fn.argtypes = C.c_int, [_vp, _vp, _fp, _sz, _sz, _sz, _sz]
    if lib.sdct_abi_version() != 1:
        raise ImportError("libsdct.so ABI version mismatch")
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != OK:
        raise SdctError(rc, load().sdct_last_error().decode(errors="replace"))
