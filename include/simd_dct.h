/*
 * simd_dct.h - drop-in C++ call surface of rainerzufalldererste/simd_dct, served by libsdct.so.
 *
 * Same three entry points, argument meaning and result codes as the reference's public header
 * (src/simd_dct.h:22-31); a caller built against that header links against libsdct.so
 * unchanged, because the functions keep C++ linkage and therefore the same mangled names
 * (_Z28simdDCT_EncodeQuantizeBufferPKhPhPKfmmmm and friends).  The work runs on a B200 through
 * the C ABI in sdct.h; there is no CPU path behind these symbols.
 *
 * Differences a caller can observe (DESIGN.md "Boundary"):
 *   - no _DetectCPUFeatures() call is needed (src/simd_platform.h:131); the device context is
 *     created on first use ($SDCT_DEVICE selects the GPU);
 *   - simdDCT_EncodeQuantize32ReorderBuffer never returns sdr_NotSupported for a missing ISA;
 *   - a null pQuantizeLUT yields sdr_InvalidParameter instead of a crash;
 *   - CUDA failures surface as sdr_NotSupported, text via sdct_last_error().
 */
#pragma once

#include <stddef.h>
#include <stdint.h>

enum simdDctResult
{
  sdr_Success = 0,
  sdr_InvalidParameter = 1,
  sdr_NotSupported = 2,
};

#define _SUCCEEDED(errorCode) (sdr_Success == (errorCode))
#define _FAILED(errorCode) (!(_SUCCEEDED(errorCode)))

/* layout A: 64 contiguous bytes per 8x8 block (src/simd_dct.cpp:93-111) */
simdDctResult simdDCT_EncodeQuantizeBuffer(const uint8_t *pFrom, uint8_t *pTo, const float *pQuantizeLUT,
                                           const size_t sizeX, const size_t sizeY,
                                           const size_t startY, const size_t endY);

/* layout C: 64 coefficient planes, left/right eye interleaved per band (src/simd_dct.cpp:71-91) */
simdDctResult simdDCT_EncodeQuantizeReorderStereoBuffer(const uint8_t *pFrom, uint8_t *pTo, const float *pQuantizeLUT,
                                                        const size_t sizeX, const size_t sizeY,
                                                        const size_t startY, const size_t endY);

/* layout B: groups of 8 blocks, coefficient-major inside a group (src/simd_dct.cpp:113-133) */
simdDctResult simdDCT_EncodeQuantize32ReorderBuffer(const uint8_t *pFrom, uint8_t *pTo, const float *pQuantizeLUT,
                                                    const size_t sizeX, const size_t sizeY,
                                                    const size_t startY, const size_t endY);
