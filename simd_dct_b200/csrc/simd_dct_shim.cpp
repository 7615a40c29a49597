// simd_dct_shim.cpp - the reference's three C++-linkage entry points (src/simd_dct.h:29-31),
// forwarding to the C ABI.  Status codes 0..2 pass through; anything else (CUDA error, no
// device) is reported as sdr_NotSupported, the reference's "cannot run here" code
// (src/simd_dct.cpp:127).
#include "../../include/simd_dct.h"
#include "../../include/sdct.h"

static simdDctResult to_result(int rc)
{
  if (rc == SDCT_OK) return sdr_Success;
  if (rc == SDCT_INVALID_PARAMETER) return sdr_InvalidParameter;
  return sdr_NotSupported;
}

simdDctResult simdDCT_EncodeQuantizeBuffer(const uint8_t *pFrom, uint8_t *pTo, const float *pQuantizeLUT,
                                           const size_t sizeX, const size_t sizeY, const size_t startY, const size_t endY)
{
  return to_result(sdct_ref_encode_quantize_buffer(pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY));
}

simdDctResult simdDCT_EncodeQuantizeReorderStereoBuffer(const uint8_t *pFrom, uint8_t *pTo, const float *pQuantizeLUT,
                                                        const size_t sizeX, const size_t sizeY, const size_t startY, const size_t endY)
{
  return to_result(sdct_ref_encode_quantize_reorder_stereo_buffer(pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY));
}

simdDctResult simdDCT_EncodeQuantize32ReorderBuffer(const uint8_t *pFrom, uint8_t *pTo, const float *pQuantizeLUT,
                                                    const size_t sizeX, const size_t sizeY, const size_t startY, const size_t endY)
{
  return to_result(sdct_ref_encode_quantize32_reorder_buffer(pFrom, pTo, pQuantizeLUT, sizeX, sizeY, startY, endY));
}
