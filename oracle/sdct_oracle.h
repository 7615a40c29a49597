/*
 * sdct_oracle.h - CPU restatement of the 8x8 block-DCT hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Nothing under oracle/ is part of the product: only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load liboracle.so or oracle/_ref/.
 * The shipped library (simd_dct_b200/csrc -> libsdct.so) never links, loads or calls it.
 *
 * Parity status
 *   forward, layouts A / B / C : PINNED  - byte-identical to the reference itself
 *       (oracle/_ref/libsimd_dct_ref_strict.so, built from /root/reference/src by
 *       oracle/Makefile) and to the committed fixtures under tests/golden/
 *       (tests/test_oracle_pins.py).
 *   inverse, layouts A / B / C : PARITY UNPINNED - the reference has no inverse of any kind
 *       (src/simd_dct.h:29-31 declares only three Encode functions).  The op order below is
 *       this repo's definition (transposed flow graph of src/simd_dct.cpp:148-171); it is
 *       validated by round-trip error against a float64 model, not by reference bytes.
 *   forward, mode FAST : this repo's FMA-contracted arithmetic (no reference counterpart);
 *       pinned only relative to mode EXACT (<= 1 LSB on a measured fraction of bytes).
 *   forward, mode TENSOR : this repo's integer-matrix form of the true DCT (exact integer sums, one
 *       fma + one multiply + round); no reference counterpart, pinned relative to mode EXACT.
 *   inverse, mode TENSOR : this repo's integer-matrix form of the inverse with the dequantiser folded into the
 *       matrix (sdct_oracle_inverse_tensor); PARITY UNPINNED like every inverse, pinned relative to the fp32
 *       inverse definition (<= 1 level on a measured fraction of pixels).
 *   interleaved 3-component input : extension, defined by composition (deinterleave, then
 *       the pinned single-plane forward per component).
 *
 * Conventions ("native" semantics; the reference's half-height/window quirks live only in
 * the sdct_oracle_ref_* wrappers at the bottom):
 *   - a plane is `width` x `height` u8, row pitch `pitch` bytes, width%8==0, height%8==0
 *     (layout B: width%64==0); every band (8 rows) of the plane is processed;
 *   - coefficient index i = u*8+v with u = vertical, v = horizontal frequency, except
 *     layout A which stores (and quantises) at the transposed position v*8+u
 *     (src/simd_dct.cpp:343-362: row pass, transpose, row pass, no transpose back).
 */
#ifndef SDCT_ORACLE_H
#define SDCT_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { SDCT_O_LAYOUT_A = 0, SDCT_O_LAYOUT_B = 1, SDCT_O_LAYOUT_C = 2 };
enum { SDCT_O_MODE_EXACT = 0, SDCT_O_MODE_FAST = 1, SDCT_O_MODE_TENSOR = 2 };
enum { SDCT_O_QUIRK_NONE = 0, SDCT_O_QUIRK_AVX_X3 = 1, SDCT_O_QUIRK_SSE = 2 /* layout C: the SSE encoders' arithmetic incl. their X1 sign, :896-1103 */ };
enum { SDCT_O_TAB_SCALAR = 0, SDCT_O_TAB_AVX = 1, SDCT_O_TAB_FAST = 2, SDCT_O_TAB_INVERSE = 3, SDCT_O_TAB_TENSOR = 4 };

/* 8-point kernels --------------------------------------------------------------------- */
void sdct_oracle_dct8_scalar(float v[8]);                 /* src/simd_dct.cpp:138-172 */
void sdct_oracle_dct8_paired(float v[8], int quirk_x3);   /* src/simd_dct.cpp:2160-2183 */
void sdct_oracle_dct8_sse(float v[8]);                     /* src/simd_dct.cpp:672-892 (with its X1 sign) */
void sdct_oracle_dct8_fast(float v[8]);                   /* this repo: FMA form, unnormalised */
void sdct_oracle_idct8(float v[8]);                       /* this repo, unnormalised */

/* mode TENSOR (this repo): the integer matrix K[c*64+p] = round(2^22 C[u][y] C[v][x]), c = u*8+v,
 * p = y*8+x, of the tensor-core form (simd_dct_b200/csrc/sdct_tc.cu); its two digits are
 * hi = floor((K + 512) / 1024), lo = K - 1024 hi */
void sdct_oracle_tensor_matrix(int32_t K[64 * 64]);

/* quantiser / dequantiser scale tables (64 floats) derived from the caller's LUT */
void sdct_oracle_tables(const float lut[64], int layout, float out[64], int which);

/* forward -----------------------------------------------------------------------------
 * in: pixel (y, x) of eye e at in[e*eye_stride + y*pitch + x*xstep]  (xstep = 1 for a planar
 *     plane, = ncomp for one component of an interleaved image).
 * out: layout A  block (band,bx) -> out + (band*width/8 + bx)*64, byte [v*8+u]       (:300-395)
 *      layout B  group g of 8 blocks -> out + band*8*width + g*512, byte [i*8+blk]   (:2064-2262)
 *      layout C  plane i at out + i*plane_stride; inside, block (band,eye,bx) at
 *                (band*eyes + eye)*(width/8) + bx                                    (:177-298)
 *      eyes / eye_stride / plane_stride are only used by layout C; the reference's stereo call
 *      is eyes = 2, eye_stride = width*height, plane_stride = 2*width*height/64. */
void sdct_oracle_forward(int layout, int mode, int quirks, const uint8_t *in, size_t pitch, size_t xstep,
                         size_t width, size_t height, int eyes, size_t eye_stride, const float lut[64],
                         uint8_t *out, size_t plane_stride);

/* inverse (PARITY UNPINNED, see header): coefficient layouts as above -> pixels */
void sdct_oracle_inverse(int layout, const uint8_t *coef, size_t plane_stride, size_t width, size_t height,
                         int eyes, size_t eye_stride, const float lut[64], uint8_t *out, size_t pitch, size_t xstep);
/* inverse, mode TENSOR (exact-integer matrix form with the dequantiser folded in; parity unpinned) */
void sdct_oracle_inverse_tensor(int layout, const uint8_t *coef, size_t plane_stride, size_t width, size_t height,
                         int eyes, size_t eye_stride, const float lut[64], uint8_t *out, size_t pitch, size_t xstep);

/* component reorder: interleaved (width*ncomp bytes per row) <-> ncomp planar planes ---- */
/* this repo: interleaved RGB rows -> planar Y, Cb, Cr planes and back (8-bit fixed point, exact) */
void sdct_oracle_rgb_to_ycc(const uint8_t *in, size_t pitch, size_t width, size_t height, uint8_t *out);
void sdct_oracle_ycc_to_rgb(const uint8_t *planes, size_t width, size_t height, uint8_t *out, size_t pitch);
void sdct_oracle_plane_stats(const uint8_t *plane, size_t n, uint64_t out[4]);   /* this repo */
void sdct_oracle_deinterleave(const uint8_t *in, size_t pitch, size_t width, size_t height,
                              int ncomp, uint8_t *out /* ncomp planes of width*height */);
void sdct_oracle_interleave(const uint8_t *planes, size_t width, size_t height, int ncomp,
                            uint8_t *out, size_t pitch);

/* reference call semantics (sizeY = double height, inclusive endY window) --------------- */
/* returns the reference's simdDctResult values: 0 ok, 1 invalid parameter, 2 not supported */
int sdct_oracle_ref_encode_quantize_buffer(const uint8_t *from, uint8_t *to, const float *lut,
                                           size_t size_x, size_t size_y, size_t start_y, size_t end_y);
int sdct_oracle_ref_encode_quantize32_reorder_buffer(const uint8_t *from, uint8_t *to, const float *lut,
                                                     size_t size_x, size_t size_y, size_t start_y,
                                                     size_t end_y, int quirks);
int sdct_oracle_ref_encode_quantize_reorder_stereo_buffer_q(const uint8_t *from, uint8_t *to, const float *lut, size_t size_x,
                                                            size_t size_y, size_t start_y, size_t end_y, int quirks);
int sdct_oracle_ref_encode_quantize_reorder_stereo_buffer(const uint8_t *from, uint8_t *to,
                                                          const float *lut, size_t size_x, size_t size_y,
                                                          size_t start_y, size_t end_y);

#ifdef __cplusplus
}
#endif
#endif
