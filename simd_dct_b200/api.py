"""Host-side mirror of the reference interface for the 8x8 block-DCT hot path.

Two levels, both thin over the C ABI (include/sdct.h):

* the reference's own three calls, same names / argument order / result codes as
  src/simd_dct.h:29-31, taking numpy (host) buffers and going through the exported C++ shim
  symbols of libsdct.so - what an existing caller of the reference sees;
* `Context`: batched, device-resident forward / inverse / fused round trip on torch CUDA tensors
  (torch is only the owner of device memory and streams here), plus the pipelined host path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib
from ._lib import COLOR_YCBCR, EXACT, FAST, LAYOUT_A, LAYOUT_B, LAYOUT_C, QUIRK_AVX_X3, QUIRK_SSE, TENSOR, Job, SdctError, check  # noqa: F401

sdr_Success, sdr_InvalidParameter, sdr_NotSupported = 0, 1, 2

# the reference CLI's default table (src/main.cpp:179-189); `--quality N` scales it (:214-217)
DEFAULT_LUT = np.array(
    [.17, .11, .10, .16, .24, .40, .51, .61, .12, .12, .14, .19, .26, .58, .60, .55,
     .14, .13, .16, .24, .40, .57, .69, .56, .14, .17, .22, .29, .51, .87, .80, .62,
     .18, .22, .37, .56, .68, 1.09, 1.03, .77, .24, .35, .55, .64, .81, 1.04, 1.13, .92,
     .49, .64, .78, .87, 1.03, 1.21, 1.20, 1.01, .72, .92, .95, .98, 1.12, 1.00, 1.03, .99], dtype=np.float32)


def quality_lut(quality: float = 1.0) -> np.ndarray:
    return (DEFAULT_LUT * np.float32(quality)).astype(np.float32)


def _fp(lut) -> "C.POINTER(C.c_float)":
    if lut is None:
        return None
    lut = np.ascontiguousarray(lut, dtype=np.float32)
    if lut.size != 64:
        raise ValueError("quantisation LUT must have 64 entries")
    _fp.keep = lut  # keep alive for the duration of the (synchronous) call
    return lut.ctypes.data_as(C.POINTER(C.c_float))


def _host_ptr(a) -> Optional[int]:
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        if a.dtype != np.uint8 or not a.flags.c_contiguous:
            raise ValueError("buffers must be C-contiguous uint8")
        return a.ctypes.data
    return int(a.data_ptr())  # CPU torch tensor (possibly pinned)


# ---- the reference's call surface ---------------------------------------------------------------
def _ref(name, pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY) -> int:
    fn = getattr(_lib.load(), _lib.SHIM_SYMBOLS[name])
    return fn(_host_ptr(pFrom), _host_ptr(pTo), _fp(pQuantizeLUT), sizeX, sizeY, startY, endY)


def simdDCT_EncodeQuantizeBuffer(pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY) -> int:
    """src/simd_dct.h:29 - layout A."""
    return _ref("simdDCT_EncodeQuantizeBuffer", pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY)


def simdDCT_EncodeQuantizeReorderStereoBuffer(pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY) -> int:
    """src/simd_dct.h:30 - layout C."""
    return _ref("simdDCT_EncodeQuantizeReorderStereoBuffer", pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY)


def simdDCT_EncodeQuantize32ReorderBuffer(pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY) -> int:
    """src/simd_dct.h:31 - layout B."""
    return _ref("simdDCT_EncodeQuantize32ReorderBuffer", pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY)


def set_reference_flags(flags: int) -> None:
    _lib.load().sdct_ref_set_flags(flags)


def get_reference_flags() -> int:
    return int(_lib.load().sdct_ref_get_flags())


def last_error() -> str:
    return _lib.load().sdct_last_error().decode(errors="replace")


def shard_range(total: int, rank: int, world: int):
    b, e = C.c_uint64(), C.c_uint64()
    _lib.load().sdct_shard_range(total, rank, world, C.byref(b), C.byref(e))
    return b.value, e.value


def dense_job(layout, width, height, frames=1, components=1, eyes=1, flags=EXACT, band_begin=0, band_end=0) -> Job:
    j = Job()
    _lib.load().sdct_job_dense(C.byref(j), layout, flags, components, eyes, width, height, frames)
    j.band_begin, j.band_end = band_begin, band_end
    return j


# ---- batched device / host API -----------------------------------------------------------------
class Context:
    """One GPU: streams, staging buffers, launch counter (sdct_context)."""

    def __init__(self, device: int = 0):
        self._lib = _lib.load()
        h = C.c_void_p()
        check(self._lib.sdct_create(device, C.byref(h)))
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.sdct_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def sm_count(self) -> int:
        return self._lib.sdct_sm_count(self._h)

    @property
    def kernel_launches(self) -> int:
        return self._lib.sdct_kernel_launches(self._h)

    def selftest_dct8(self, kind: int, vectors):
        """runs one device 8-point kernel on (n, 8) float32 vectors; returns the transformed copy"""
        v = np.ascontiguousarray(vectors, dtype=np.float32).copy()
        if v.ndim != 2 or v.shape[1] != 8:
            raise ValueError("vectors must have shape (n, 8)")
        check(self._lib.sdct_selftest_dct8(self._h, kind, v.ctypes.data_as(C.POINTER(C.c_float)), v.shape[0]))
        return v

    def synchronize(self):
        check(self._lib.sdct_synchronize(self._h))

    @staticmethod
    def _stream():
        """torch's current stream, so that torch events and tensors order with the launch.  The
        legacy default stream has handle 0, which the C ABI reads as "the context's own stream";
        pass cudaStreamLegacy (handle 1) for it instead."""
        import torch
        return C.c_void_p(torch.cuda.current_stream().cuda_stream or 1)

    def _coef_bytes(self, job: Job) -> int:
        return int(job.frames * job.coef_frame_stride)

    # device-resident: tensors are torch uint8 CUDA tensors
    def forward(self, pixels, lut, job: Job, out=None):
        import torch
        if out is None:
            out = torch.empty(self._coef_bytes(job), dtype=torch.uint8, device=pixels.device)
        check(self._lib.sdct_forward_device(self._h, C.byref(job), C.c_void_p(pixels.data_ptr()), _fp(lut),
                                            C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def inverse(self, coefs, lut, job: Job, out=None):
        import torch
        if out is None:
            out = torch.empty(int(job.frames * job.pixel_frame_stride), dtype=torch.uint8, device=coefs.device)
        check(self._lib.sdct_inverse_device(self._h, C.byref(job), C.c_void_p(coefs.data_ptr()), _fp(lut),
                                            C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def roundtrip(self, pixels, lut, job: Job, out=None):
        import torch
        if out is None:
            out = torch.empty_like(pixels)
        check(self._lib.sdct_roundtrip_device(self._h, C.byref(job), C.c_void_p(pixels.data_ptr()), _fp(lut),
                                              C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def roundtrip_noalloc(self, pixels, lut, job: Job, out):
        """sdct_roundtrip_noalloc_device: the two-kernel round trip with the levels in the library's file-scope device array
        (at most NOALLOC_MAX_COEF_BYTES); `out` is the caller's - nothing is allocated on this path, here or in the library"""
        check(self._lib.sdct_roundtrip_noalloc_device(self._h, C.byref(job), C.c_void_p(pixels.data_ptr()), _fp(lut),
                                                      C.c_void_p(out.data_ptr()), self._stream()))
        return out

    def plane_stats(self, coefs, job: Job):
        """layout C: per coefficient plane {zero levels, zero runs, sum |level - 127|, OR of |level - 127|} over the job's
        band window -> int64 tensor [frames, components, 64, 4] on the device (sdct_plane_stats_device)"""
        import torch
        out = torch.empty((int(job.frames), int(job.components), 64, 4), dtype=torch.int64, device=coefs.device)
        check(self._lib.sdct_plane_stats_device(self._h, C.byref(job), C.c_void_p(coefs.data_ptr()), C.c_void_p(out.data_ptr()), self._stream()))
        return out

    # host buffers (numpy arrays or CPU torch tensors, pinned or pageable); synchronous
    def forward_host(self, pixels, lut, job: Job, out=None):
        if out is None:
            out = np.empty(self._coef_bytes(job), dtype=np.uint8)
        check(self._lib.sdct_forward_host(self._h, C.byref(job), C.c_void_p(_host_ptr(pixels)), _fp(lut),
                                          C.c_void_p(_host_ptr(out))))
        return out

    def inverse_host(self, coefs, lut, job: Job, out=None):
        if out is None:
            out = np.empty(int(job.frames * job.pixel_frame_stride), dtype=np.uint8)
        check(self._lib.sdct_inverse_host(self._h, C.byref(job), C.c_void_p(_host_ptr(coefs)), _fp(lut),
                                          C.c_void_p(_host_ptr(out))))
        return out


def _ctx_array(ctxs):
    arr = (C.c_void_p * len(ctxs))(*[c._h for c in ctxs])
    return arr, len(ctxs)


def forward_host_multi(ctxs, pixels, lut, job: Job, out=None):
    """one host-buffer call sharded over several Contexts (= devices): frame ranges or band ranges, no collective"""
    if out is None:
        out = np.empty(int(job.frames * job.coef_frame_stride), dtype=np.uint8)
    arr, n = _ctx_array(ctxs)
    check(_lib.load().sdct_forward_host_multi(arr, n, C.byref(job), C.c_void_p(_host_ptr(pixels)), _fp(lut), C.c_void_p(_host_ptr(out))))
    return out


def inverse_host_multi(ctxs, coefs, lut, job: Job, out=None):
    if out is None:
        out = np.empty(int(job.frames * job.pixel_frame_stride), dtype=np.uint8)
    arr, n = _ctx_array(ctxs)
    check(_lib.load().sdct_inverse_host_multi(arr, n, C.byref(job), C.c_void_p(_host_ptr(coefs)), _fp(lut), C.c_void_p(_host_ptr(out))))
    return out
