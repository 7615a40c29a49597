"""simd_dct_b200 - B200 (sm_100a) implementation of simd_dct's hot path: the 8x8 block DCT with
quantisation and coefficient/component reordering, and its inverse.  See DESIGN.md."""
from .api import (COLOR_YCBCR, DEFAULT_LUT, EXACT, FAST, LAYOUT_A, LAYOUT_B, LAYOUT_C, QUIRK_AVX_X3, QUIRK_SSE, TENSOR, Context, Job, SdctError,  # noqa: F401
                  dense_job, forward_host_multi, inverse_host_multi, get_reference_flags, last_error, quality_lut, set_reference_flags, shard_range,
                  simdDCT_EncodeQuantize32ReorderBuffer, simdDCT_EncodeQuantizeBuffer,
                  simdDCT_EncodeQuantizeReorderStereoBuffer, sdr_InvalidParameter, sdr_NotSupported, sdr_Success)
