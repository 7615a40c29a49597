/*
 * sdct_oracle.c - CPU restatement of the 8x8 block-DCT hot path.  TEST INFRASTRUCTURE ONLY.
 * See sdct_oracle.h for the parity status of every entry point.  Build with
 *   gcc -std=c11 -O2 -fno-fast-math -ffp-contract=off     (oracle/Makefile)
 * so that every float operation below is one IEEE-754 binary32 operation, rounded to nearest
 * even, in exactly the order written (fmaf() calls are explicit single-rounding FMAs).
 *
 * All file:line citations are into /root/reference/src/simd_dct.cpp unless stated otherwise.
 */
#include "sdct_oracle.h"

#include <limits.h>
#include <math.h>
#include <string.h>

/* 8-point constants, :140-146 (narrowed to float exactly as the reference's literals are). */
static const float K_A = 1.3870398453221474618216191915664f;  /* sqrt2 cos(1pi/16) */
static const float K_B = 1.3065629648763765278566431734272f;  /* sqrt2 cos(2pi/16) */
static const float K_C = 1.1758756024193587169744671046113f;  /* sqrt2 cos(3pi/16) */
static const float K_D = 0.78569495838710218127789736765722f; /* sqrt2 cos(5pi/16) */
static const float K_E = 0.54119610014619698439972320536639f; /* sqrt2 cos(6pi/16) */
static const float K_F = 0.27589937928294301233595756366937f; /* sqrt2 cos(7pi/16) */
static const float K_N = 0.35355339059327376220042218105242f; /* 1/sqrt8 */
static const float K_VR = 0.95f;                              /* :191, :906, :2066 */

/* ---------------------------------------------------------------------------------------
 * 8-point forward kernels
 * ------------------------------------------------------------------------------------- */

/* :138-172.  Odd outputs are summed left to right: ((t0 - t1) + t2) - t3. */
void sdct_oracle_dct8_scalar(float v[8])
{
    const float s07 = v[0] + v[7], s16 = v[1] + v[6], s25 = v[2] + v[5], s34 = v[3] + v[4];
    const float d07 = v[0] - v[7], d61 = v[6] - v[1], d25 = v[2] - v[5], d43 = v[4] - v[3];
    const float pp = s07 + s34, pm = s07 - s34, qq = s16 + s25, qm = s16 - s25;

    v[0] = K_N * (pp + qq);
    v[2] = K_N * (K_B * pm + K_E * qm);
    v[4] = K_N * (pp - qq);
    v[6] = K_N * (K_E * pm - K_B * qm);
    v[1] = K_N * (((K_A * d07 - K_C * d61) + K_D * d25) - K_F * d43);
    v[3] = K_N * (((K_C * d07 + K_F * d61) - K_A * d25) + K_D * d43);
    v[5] = K_N * (((K_D * d07 + K_A * d61) + K_F * d25) - K_C * d43);
    v[7] = K_N * (((K_F * d07 + K_D * d61) + K_C * d25) + K_A * d43);
}

/* :2160-2183 (AVX2) == :1965-1988 (AVX-512VL) lane arithmetic.  Odd outputs are summed
 * pairwise, (t0 op t1) op (t2 op t3).  quirk_x3 != 0 reproduces the reference's X3 line
 * (:2181), which subtracts (A*d25 + D*d43) where the true DCT needs (A*d25 - D*d43). */
void sdct_oracle_dct8_paired(float v[8], int quirk_x3)
{
    const float s07 = v[0] + v[7], s16 = v[1] + v[6], s25 = v[2] + v[5], s34 = v[3] + v[4];
    const float d07 = v[0] - v[7], d61 = v[6] - v[1], d25 = v[2] - v[5], d43 = v[4] - v[3];
    const float pp = s07 + s34, pm = s07 - s34, qq = s16 + s25, qm = s16 - s25;

    v[0] = K_N * (pp + qq);
    v[4] = K_N * (pp - qq);
    v[2] = K_N * (K_B * pm + K_E * qm);
    v[6] = K_N * (K_E * pm - K_B * qm);
    v[1] = K_N * ((K_A * d07 - K_C * d61) + (K_D * d25 - K_F * d43));
    if (quirk_x3)
        v[3] = K_N * ((K_C * d07 + K_F * d61) - (K_A * d25 + K_D * d43));
    else
        v[3] = K_N * ((K_C * d07 + K_F * d61) - (K_A * d25 - K_D * d43));
    v[5] = K_N * ((K_D * d07 + K_A * d61) + (K_F * d25 - K_C * d43));
    v[7] = K_N * ((K_F * d07 + K_D * d61) + (K_C * d25 + K_A * d43));
}

/* :672-892 (SSE4.1) == :434-654 (SSE2) lane arithmetic, one row of two adjacent blocks per call.  The sums, differences,
 * even outputs and the pairwise summation of the odd outputs are those of the AVX2 form above (the sign flips the SSE code
 * applies by XOR, :717-721, and its negative factors -C_c, -C_a, -C_b are exact negations, so a + (-b) = a - b bit for
 * bit); what differs is the defect: X1 adds C_f * x43m where the DCT subtracts it (:762 "xf14 = (C_f * x34m)", :788). */
void sdct_oracle_dct8_sse(float v[8])
{
    const float s07 = v[0] + v[7], s16 = v[1] + v[6], s25 = v[2] + v[5], s34 = v[3] + v[4];
    const float d07 = v[0] - v[7], d61 = v[6] - v[1], d25 = v[2] - v[5], d43 = v[4] - v[3];
    const float pp = s07 + s34, pm = s07 - s34, qq = s16 + s25, qm = s16 - s25;

    v[0] = K_N * (pp + qq);
    v[4] = K_N * (pp - qq);
    v[2] = K_N * (K_B * pm + K_E * qm);
    v[6] = K_N * (K_E * pm - K_B * qm);
    v[1] = K_N * ((K_A * d07 - K_C * d61) + (K_D * d25 + K_F * d43));   /* the defect */
    v[3] = K_N * ((K_C * d07 + K_F * d61) - (K_A * d25 - K_D * d43));
    v[5] = K_N * ((K_D * d07 + K_A * d61) + (K_F * d25 - K_C * d43));
    v[7] = K_N * ((K_F * d07 + K_D * d61) + (K_C * d25 + K_A * d43));
}

/* ---------------------------------------------------------------------------------------
 * This repo's own arithmetic definitions (no reference counterpart): the FMA-contracted
 * "fast" forward and the inverse.  Both are *defined* here and the CUDA kernels reproduce
 * them operation for operation, so GPU-vs-oracle is bit-exact for them as well.
 * ------------------------------------------------------------------------------------- */

/* Unnormalised forward (the two 1/sqrt8 factors = 1/8 per 2-D transform are folded into the
 * quantiser scale).  34 operations instead of 56. */
void sdct_oracle_dct8_fast(float v[8])
{
    const float s07 = v[0] + v[7], s16 = v[1] + v[6], s25 = v[2] + v[5], s34 = v[3] + v[4];
    const float d07 = v[0] - v[7], d61 = v[6] - v[1], d25 = v[2] - v[5], d43 = v[4] - v[3];
    const float pp = s07 + s34, pm = s07 - s34, qq = s16 + s25, qm = s16 - s25;

    v[0] = pp + qq;
    v[4] = pp - qq;
    v[2] = fmaf(K_B, pm, K_E * qm);
    v[6] = fmaf(K_E, pm, -(K_B * qm));
    v[1] = fmaf(-K_F, d43, fmaf(K_D, d25, fmaf(-K_C, d61, K_A * d07)));
    v[3] = fmaf(K_D, d43, fmaf(-K_A, d25, fmaf(K_F, d61, K_C * d07)));
    v[5] = fmaf(-K_C, d43, fmaf(K_F, d25, fmaf(K_A, d61, K_D * d07)));
    v[7] = fmaf(K_A, d43, fmaf(K_C, d25, fmaf(K_D, d61, K_F * d07)));
}

/* Unnormalised inverse 8-point (transposed flow graph of :148-171, SURVEY A.6, with the
 * scalar factors folded into the dequantiser scale).  x = 2*sqrt2 * C^T X. */
void sdct_oracle_idct8(float v[8])
{
    const float e0 = v[0] + v[4], e1 = v[0] - v[4];
    const float e2 = fmaf(K_B, v[2], K_E * v[6]);
    const float e3 = fmaf(K_E, v[2], -(K_B * v[6]));
    const float a0 = e0 + e2, a3 = e0 - e2, a1 = e1 + e3, a2 = e1 - e3;
    const float o0 = fmaf(K_F, v[7], fmaf(K_D, v[5], fmaf(K_C, v[3], K_A * v[1])));
    const float o1 = fmaf(K_D, v[7], fmaf(K_A, v[5], fmaf(K_F, v[3], -(K_C * v[1]))));
    const float o2 = fmaf(K_C, v[7], fmaf(K_F, v[5], fmaf(-K_A, v[3], K_D * v[1])));
    const float o3 = fmaf(K_A, v[7], fmaf(-K_C, v[5], fmaf(K_D, v[3], -(K_F * v[1]))));
    v[0] = a0 + o0; v[7] = a0 - o0;
    v[1] = a1 - o1; v[6] = a1 + o1;
    v[2] = a2 + o2; v[5] = a2 - o2;
    v[3] = a3 - o3; v[4] = a3 + o3;
}

/* ---------------------------------------------------------------------------------------
 * Quantisers
 * ------------------------------------------------------------------------------------- */

/* the reference's _clamp template, :45-49: NaN falls to `lo` */
static float clamp_ref(float x, float lo, float hi) { return x > lo ? (x < hi ? x : hi) : lo; }

/* scalar paths, :245 / :362 : reciprocal table 1/(lut*vr), bias 127/255, roundf (half away) */
static uint8_t quant_scalar(float c, float rq)
{
    const float bias = 127.0f / 255.0f;
    return (uint8_t)roundf(clamp_ref(c * rq + bias, 0.f, 1.f) * 255.f);
}

/* _mm256_cvtps_epi32 semantics: round to nearest even; NaN / out of range -> INT_MIN */
static int32_t cvtps_epi32(float x)
{
    if (!(x >= -2147483648.0f && x < 2147483648.0f)) return INT_MIN;
    return (int32_t)lrintf(x);     /* default rounding mode = nearest even */
}

/* AVX2 / AVX-512VL path, :2224 : scale 255/(lut*vr), integer bias after the conversion */
static uint8_t quant_avx(float c, float q)
{
    int64_t r = (int64_t)cvtps_epi32(c * q) + 127;
    return (uint8_t)(r > 255 ? 255 : (r < 0 ? 0 : r));
}

/* this repo's fast quantiser: the AVX2 path's form with a single pre-folded scale,
 * clamp(rne(fl(c * qs)) + 127, 0, 255).  |c * qs| < 2^31 is guaranteed by the library's table check. */
static uint8_t quant_fast(float c, float qs)
{
    long r = lrintf(c * qs) + 127;     /* default rounding mode = nearest even */
    return (uint8_t)(r > 255 ? 255 : (r < 0 ? 0 : r));
}

/* this repo's pixel rounding for the inverse: clamp(rne(x), 0, 255), x in pixel units */
static uint8_t pixel_u8(float x)
{
    if (!(x > 0.f)) return 0;
    if (x >= 255.f) return 255;
    return (uint8_t)lrintf(x);
}

void sdct_oracle_tables(const float lut[64], int layout, float out[64], int which)
{
    for (int i = 0; i < 64; i++) {
        const float d = lut[i] * K_VR;
        switch (which) {
        case SDCT_O_TAB_SCALAR: out[i] = 1.f / d; break;                 /* :194-212 */
        case SDCT_O_TAB_AVX:    out[i] = 255.0f / d; break;              /* :2236-2239 */
        case SDCT_O_TAB_FAST:   /* fwd fast on raw pixels: 1/8 of the two unnormalised passes folded in;
                                   A/C: 255 * (1/255) cancels, B: the AVX path's 255/d */
            out[i] = (layout == SDCT_O_LAYOUT_B) ? 0.125f * (255.0f / d) : 0.125f * (1.f / d);
            break;
        case SDCT_O_TAB_TENSOR: /* the matrix carries 2^22; A/C: 255 * (1/255) cancels, B: the AVX path's 255/d */
            out[i] = ldexpf((layout == SDCT_O_LAYOUT_B) ? 255.0f / d : 1.f / d, -22);
            break;
        default:                /* SDCT_O_TAB_INVERSE: level -> coefficient in pixel units, 1/8 folded in */
            out[i] = (layout == SDCT_O_LAYOUT_B) ? (d / 255.f) * 0.125f : d * 0.125f;
            break;
        }
    }
}

/* ---------------------------------------------------------------------------------------
 * Per-block 2-D transforms.  `px` points at the block's top-left pixel, row pitch `pitch`.
 * Results are written to o[64] in the index order of the respective layout.
 * ------------------------------------------------------------------------------------- */

static void transpose8(float b[64])
{
    for (int i = 0; i < 8; i++)
        for (int j = i + 1; j < 8; j++) { float t = b[i * 8 + j]; b[i * 8 + j] = b[j * 8 + i]; b[j * 8 + i] = t; }
}

/* layout A, :343-362 : x/255.f, row pass, swap, row pass; NOT swapped back */
static void block_fwd_a(const uint8_t *px, size_t pitch, const float rq[64], uint8_t o[64])
{
    float b[64];
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++) b[r * 8 + c] = px[r * pitch + c] / 255.f;
    for (int r = 0; r < 8; r++) sdct_oracle_dct8_scalar(b + r * 8);
    transpose8(b);
    for (int r = 0; r < 8; r++) sdct_oracle_dct8_scalar(b + r * 8);
    for (int i = 0; i < 64; i++) o[i] = quant_scalar(b[i], rq[i]);
}

/* layout C, :220-247 : x/255.f, swap, row pass (= vertical), swap, row pass (= horizontal) */
static void block_fwd_c(const uint8_t *px, size_t pitch, const float rq[64], uint8_t o[64])
{
    float b[64];
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++) b[r * 8 + c] = px[r * pitch + c] / 255.f;
    transpose8(b);
    for (int r = 0; r < 8; r++) sdct_oracle_dct8_scalar(b + r * 8);
    transpose8(b);
    for (int r = 0; r < 8; r++) sdct_oracle_dct8_scalar(b + r * 8);
    for (int i = 0; i < 64; i++) o[i] = quant_scalar(b[i], rq[i]);
}

/* layout C as the reference's SSE4.1 / SSSE3 / SSE2 stereo encoders compute it (:896-1103; quirk SDCT_O_QUIRK_SSE):
 * x * (1.f / 255) instead of x / 255.f (:955-962), the transposes and passes of the scalar form with inplace_dct8_sse41
 * (:964-1007), and the quantiser min(255, max(0, cvtps_epi32(c * (255 / (lut * .95)) + 127.f))) (:1018-1024): product and
 * sum are rounded separately, the conversion rounds to nearest even, out-of-range and NaN convert to INT_MIN and clamp to 0 */
static uint8_t quant_sse(float c, float q)
{
    const float p = c * q;
    const float y = p + 127.0f;
    if (!(y >= -2147483648.f && y < 2147483648.f)) return 0;
    const long r = lrintf(y);
    return (uint8_t)(r > 255 ? 255 : (r < 0 ? 0 : r));
}

static void block_fwd_c_sse(const uint8_t *px, size_t pitch, const float q[64], uint8_t o[64])
{
    const float inv255 = 1.f / (float)0xFF;
    float b[64];
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++) b[r * 8 + c] = inv255 * (float)px[r * pitch + c];
    transpose8(b);
    for (int r = 0; r < 8; r++) sdct_oracle_dct8_sse(b + r * 8);
    transpose8(b);
    for (int r = 0; r < 8; r++) sdct_oracle_dct8_sse(b + r * 8);
    for (int i = 0; i < 64; i++) o[i] = quant_sse(b[i], q[i]);
}

/* layout B, :2107-2228 : (float)x, horizontal pass, vertical pass, natural order */
static void block_fwd_b(const uint8_t *px, size_t pitch, const float q[64], int quirks, uint8_t o[64])
{
    float b[64], col[8];
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++) b[r * 8 + c] = (float)px[r * pitch + c];
    for (int r = 0; r < 8; r++) sdct_oracle_dct8_paired(b + r * 8, quirks & SDCT_O_QUIRK_AVX_X3);
    for (int c = 0; c < 8; c++) {
        for (int r = 0; r < 8; r++) col[r] = b[r * 8 + c];
        sdct_oracle_dct8_paired(col, quirks & SDCT_O_QUIRK_AVX_X3);
        for (int r = 0; r < 8; r++) b[r * 8 + c] = col[r];
    }
    for (int i = 0; i < 64; i++) o[i] = quant_avx(b[i], q[i]);
}

/* fast mode (this repo): (float)x, horizontal pass, vertical pass, unnormalised; the output
 * order is natural (u*8+v) or, for layout A, transposed (v*8+u) with the LUT at that index */
static void block_fwd_fast(const uint8_t *px, size_t pitch, const float qs[64], int transposed, uint8_t o[64])
{
    float b[64], col[8];
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++) b[r * 8 + c] = (float)px[r * pitch + c];
    for (int r = 0; r < 8; r++) sdct_oracle_dct8_fast(b + r * 8);
    for (int c = 0; c < 8; c++) {
        for (int r = 0; r < 8; r++) col[r] = b[r * 8 + c];
        sdct_oracle_dct8_fast(col);
        for (int r = 0; r < 8; r++) b[r * 8 + c] = col[r];
    }
    for (int u = 0; u < 8; u++)
        for (int v = 0; v < 8; v++) {
            const int i = transposed ? v * 8 + u : u * 8 + v;
            o[i] = quant_fast(b[u * 8 + v], qs[i]);
        }
}

/* mode TENSOR (this repo; simd_dct_b200/csrc/sdct_tc.cu): K = round(2^22 C (x) C) as an integer
 * matrix, entries of C = the reference's constants (:140-146) as doubles.  Per coefficient:
 * exact integer sums D1 = sum x*hi, D0 = sum x*lo (|.| < 2^24, so the GPU's fp32 tensor-core
 * accumulation is exact too), t = fma(D1, 1024, D0), level = clamp(rne(fl(t * qs) + 127)). */
/* mode TENSOR quantiser: clamp(rne(fl(t * qs) + 127), 0, 255) - the bias is added exactly before the
 * single round-to-nearest-even (the GPU does both in one FMA onto the denormal grid) */
static uint8_t quant_tensor(float t, float qs)
{
    const double r = rint((double)(t * qs) + 127.0);     /* float product, exact double sum, ties to even */
    return (uint8_t)(r > 255.0 ? 255 : (r < 0.0 ? 0 : (int)r));
}

static double dct_entry(int u, int y)
{
    static const double V[9] = {1.4142135623730950488016887242097, 1.3870398453221474618216191915664,
                                1.3065629648763765278566431734272, 1.1758756024193587169744671046113, 1.0,
                                0.78569495838710218127789736765722, 0.54119610014619698439972320536639,
                                0.27589937928294301233595756366937, 0.0};
    const double N = 0.35355339059327376220042218105242;
    if (u == 0) return N;
    int m = ((2 * y + 1) * u) % 32;
    double sign = 1.0;
    if (m > 16) m = 32 - m;
    if (m > 8) { m = 16 - m; sign = -1.0; }
    return N * (sign * V[m]);
}

void sdct_oracle_tensor_matrix(int32_t K[64 * 64])
{
    for (int u = 0; u < 8; u++)
        for (int v = 0; v < 8; v++)
            for (int y = 0; y < 8; y++)
                for (int x = 0; x < 8; x++)
                    K[(u * 8 + v) * 64 + y * 8 + x] = (int32_t)llrint(ldexp(dct_entry(u, y) * dct_entry(v, x), 22));
}

static void block_fwd_tensor(const uint8_t *px, size_t pitch, const float qs[64], int transposed, uint8_t o[64])
{
    static int32_t K[64 * 64];
    static int have = 0;
    if (!have) { sdct_oracle_tensor_matrix(K); have = 1; }
    for (int u = 0; u < 8; u++)
        for (int v = 0; v < 8; v++) {
            int32_t d1 = 0, d0 = 0;
            for (int p = 0; p < 64; p++) {
                const int32_t full = K[(u * 8 + v) * 64 + p];
                const int32_t hi = (full + 512) >> 10, lo = full - hi * 1024;
                const int32_t x = px[(p >> 3) * pitch + (p & 7)];
                d1 += x * hi;
                d0 += x * lo;
            }
            const float t = fmaf((float)d1, 1024.f, (float)d0);
            const int i = transposed ? v * 8 + u : u * 8 + v;
            o[i] = quant_tensor(t, qs[i]);
        }
}

/* inverse (this repo): dequantise with one FMA, (q - 127) * tab = fma(q, tab, -(127 * tab)),
 * vertical pass, horizontal pass, clamp(rne(.), 0, 255) */
static void block_inv(const uint8_t q[64], const float tab[64], int transposed, uint8_t *px, size_t pitch)
{
    float b[64], col[8];
    for (int u = 0; u < 8; u++)
        for (int v = 0; v < 8; v++) {
            const int i = transposed ? v * 8 + u : u * 8 + v;
            b[u * 8 + v] = fmaf((float)q[i], tab[i], -(127.f * tab[i]));
        }
    for (int c = 0; c < 8; c++) {
        for (int r = 0; r < 8; r++) col[r] = b[r * 8 + c];
        sdct_oracle_idct8(col);
        for (int r = 0; r < 8; r++) b[r * 8 + c] = col[r];
    }
    for (int r = 0; r < 8; r++) sdct_oracle_idct8(b + r * 8);
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++) px[r * pitch + c] = pixel_u8(b[r * 8 + c]);
}

/* ---------------------------------------------------------------------------------------
 * Plane-level entry points ("native" semantics: every band of the plane is processed)
 * ------------------------------------------------------------------------------------- */

static void gather_block(const uint8_t *src, size_t pitch, size_t xstep, uint8_t tile[64])
{
    for (int r = 0; r < 8; r++)
        for (int c = 0; c < 8; c++) tile[r * 8 + c] = src[r * pitch + (size_t)c * xstep];
}

/* where block (band, eye, bx) of an image `width` wide puts coefficient i */
static uint8_t *coef_addr(int layout, uint8_t *out, size_t plane_stride, size_t width, int eyes,
                          size_t band, int eye, size_t bx, int i)
{
    const size_t bpr = width / 8;
    switch (layout) {
    case SDCT_O_LAYOUT_A: return out + (band * bpr + bx) * 64 + i;                        /* :361-365 */
    case SDCT_O_LAYOUT_B: return out + band * 8 * width + (bx / 8) * 512 + i * 8 + (bx % 8); /* :2221-2230 */
    default:              return out + (size_t)i * plane_stride + (band * eyes + eye) * bpr + bx; /* :256-264, :289-296 */
    }
}

static void forward_band(int layout, int mode, int quirks, const uint8_t *in, size_t pitch, size_t xstep,
                         size_t width, size_t band, int eyes, size_t eye_stride, const float tab[64],
                         uint8_t *out, size_t plane_stride)
{
    uint8_t tile[64], o[64];
    for (int eye = 0; eye < eyes; eye++)
        for (size_t bx = 0; bx < width / 8; bx++) {
            gather_block(in + (size_t)eye * eye_stride + band * 8 * pitch + bx * 8 * xstep, pitch, xstep, tile);
            if (mode == SDCT_O_MODE_TENSOR) block_fwd_tensor(tile, 8, tab, layout == SDCT_O_LAYOUT_A, o);
            else if (mode == SDCT_O_MODE_FAST) block_fwd_fast(tile, 8, tab, layout == SDCT_O_LAYOUT_A, o);
            else if (layout == SDCT_O_LAYOUT_A) block_fwd_a(tile, 8, tab, o);
            else if (layout == SDCT_O_LAYOUT_B) block_fwd_b(tile, 8, tab, quirks, o);
            else if (quirks & SDCT_O_QUIRK_SSE) block_fwd_c_sse(tile, 8, tab, o);
            else block_fwd_c(tile, 8, tab, o);
            for (int i = 0; i < 64; i++) *coef_addr(layout, out, plane_stride, width, eyes, band, eye, bx, i) = o[i];
        }
}

static void forward_table(int layout, int mode, int quirks, const float lut[64], float tab[64])
{
    const int which = mode == SDCT_O_MODE_TENSOR ? SDCT_O_TAB_TENSOR : mode == SDCT_O_MODE_FAST ? SDCT_O_TAB_FAST
                    : (layout == SDCT_O_LAYOUT_B || (layout == SDCT_O_LAYOUT_C && (quirks & SDCT_O_QUIRK_SSE))) ? SDCT_O_TAB_AVX
                    : SDCT_O_TAB_SCALAR;   /* (the SSE encoders build the AVX path's 255 / (lut * .95), :908-923) */
    sdct_oracle_tables(lut, layout, tab, which);
}

void sdct_oracle_forward(int layout, int mode, int quirks, const uint8_t *in, size_t pitch, size_t xstep,
                         size_t width, size_t height, int eyes, size_t eye_stride, const float lut[64],
                         uint8_t *out, size_t plane_stride)
{
    float tab[64];
    forward_table(layout, mode, quirks, lut, tab);
    if (layout != SDCT_O_LAYOUT_C) eyes = 1;
    for (size_t band = 0; band < height / 8; band++)
        forward_band(layout, mode, quirks, in, pitch, xstep, width, band, eyes, eye_stride, tab, out, plane_stride);
}

void sdct_oracle_inverse(int layout, const uint8_t *coef, size_t plane_stride, size_t width, size_t height,
                         int eyes, size_t eye_stride, const float lut[64], uint8_t *out, size_t pitch, size_t xstep)
{
    float tab[64];
    uint8_t q[64], tile[64];
    sdct_oracle_tables(lut, layout, tab, SDCT_O_TAB_INVERSE);
    if (layout != SDCT_O_LAYOUT_C) eyes = 1;
    for (size_t band = 0; band < height / 8; band++)
        for (int eye = 0; eye < eyes; eye++)
            for (size_t bx = 0; bx < width / 8; bx++) {
                for (int i = 0; i < 64; i++)
                    q[i] = *coef_addr(layout, (uint8_t *)coef, plane_stride, width, eyes, band, eye, bx, i);
                block_inv(q, tab, layout == SDCT_O_LAYOUT_A, tile, 8);
                uint8_t *dst = out + (size_t)eye * eye_stride + band * 8 * pitch + bx * 8 * xstep;
                for (int r = 0; r < 8; r++)
                    for (int c = 0; c < 8; c++) dst[r * pitch + (size_t)c * xstep] = tile[r * 8 + c];
            }
}

/* Inverse, mode TENSOR (this repo; simd_dct_b200/csrc/sdct_tc.cu, inverse_tc_kernel).  PARITY UNPINNED like every
 * inverse here (the reference has none).  The dequantiser scale is folded into the matrix:
 *   E[p][i]   = W[p][i] * tab[i]            (double; W = unnormalised inverse basis, tab = SDCT_O_TAB_INVERSE as float,
 *                                            i = the layout's own coefficient index, p = y*8+x)
 *   S         = the largest integer with 2^S * max|E| < 2^20 - 512
 *   full      = llrint(2^S * E), hi = (full + 512) >> 10, lo = full - 1024 * hi
 *   per pixel : D1 = sum q_i * hi, D0 = sum q_i * lo (exact integers, |.| < 2^24: the tensor core's fp32
 *               accumulation is exact), t = fma(D1, 1024, D0), y = fma(t, 2^-S, off[p]) with
 *               off[p] = (float)(-127 * sum_i E[p][i]), pixel = clamp(rne(y), 0, 255). */
static int inverse_tensor_matrix(int layout, const float tab[64], int32_t full[64 * 64], float off[64])
{
    static double E[64 * 64];
    double emax = 0.0;
    const double N = 0.35355339059327376220042218105242;
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++)
            for (int u = 0; u < 8; u++)
                for (int v = 0; v < 8; v++) {
                    const int i = layout == SDCT_O_LAYOUT_A ? v * 8 + u : u * 8 + v;
                    const double w = (dct_entry(u, y) / N) * (dct_entry(v, x) / N);
                    const double e = w * (double)tab[i];
                    E[(y * 8 + x) * 64 + i] = e;
                    if (fabs(e) > emax) emax = fabs(e);
                }
    int S = 0;
    if (emax > 0.0) {
        S = (int)floor(log2((1048576.0 - 512.0) / emax));
        while (ldexp(emax, S) >= 1048576.0 - 512.0) S--;
        while (ldexp(emax, S + 1) < 1048576.0 - 512.0) S++;
    }
    for (int p = 0; p < 64; p++) {
        /* the level offset -127 goes through the ROUNDED matrix too, so that the all-127 block decodes to exactly 0 */
        int64_t sum = 0;
        for (int i = 0; i < 64; i++) {
            full[p * 64 + i] = (int32_t)llrint(ldexp(E[p * 64 + i], S));
            sum += full[p * 64 + i];
        }
        off[p] = (float)(-127.0 * ldexp((double)sum, -S));
    }
    return S;
}

void sdct_oracle_inverse_tensor(int layout, const uint8_t *coef, size_t plane_stride, size_t width, size_t height,
                                int eyes, size_t eye_stride, const float lut[64], uint8_t *out, size_t pitch, size_t xstep)
{
    float tab[64], off[64];
    static int32_t full[64 * 64];
    sdct_oracle_tables(lut, layout, tab, SDCT_O_TAB_INVERSE);
    const int S = inverse_tensor_matrix(layout, tab, full, off);
    const float unscale = ldexpf(1.f, -S);
    if (layout != SDCT_O_LAYOUT_C) eyes = 1;
    for (size_t band = 0; band < height / 8; band++)
        for (int eye = 0; eye < eyes; eye++)
            for (size_t bx = 0; bx < width / 8; bx++) {
                int32_t q[64];
                for (int i = 0; i < 64; i++)
                    q[i] = *coef_addr(layout, (uint8_t *)coef, plane_stride, width, eyes, band, eye, bx, i);
                uint8_t *dst = out + (size_t)eye * eye_stride + band * 8 * pitch + bx * 8 * xstep;
                for (int p = 0; p < 64; p++) {
                    int32_t d1 = 0, d0 = 0;
                    for (int i = 0; i < 64; i++) {
                        const int32_t f = full[p * 64 + i];
                        const int32_t hi = (f + 512) >> 10, lo = f - hi * 1024;
                        d1 += q[i] * hi;
                        d0 += q[i] * lo;
                    }
                    const float t = fmaf((float)d1, 1024.f, (float)d0);
                    dst[(p >> 3) * pitch + (size_t)(p & 7) * xstep] = pixel_u8(fmaf(t, unscale, off[p]));
                }
            }
}

/* colour front-end (this repo's definition; include/sdct.h SDCT_COLOR_YCBCR): JFIF / BT.601 full-range
 * RGB <-> YCbCr in 8-bit fixed point.  in: interleaved rows (3 bytes per pixel), out: three planar planes Y, Cb, Cr. */
static int floor_shift8(int v) { return v >= 0 ? v >> 8 : -((-v + 255) >> 8); }
static uint8_t clamp_u8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

void sdct_oracle_rgb_to_ycc(const uint8_t *in, size_t pitch, size_t width, size_t height, uint8_t *out)
{
    for (size_t y = 0; y < height; y++)
        for (size_t x = 0; x < width; x++) {
            const int r = in[y * pitch + 3 * x], g = in[y * pitch + 3 * x + 1], b = in[y * pitch + 3 * x + 2];
            out[(0 * height + y) * width + x] = (uint8_t)((77 * r + 150 * g + 29 * b + 128) >> 8);
            out[(1 * height + y) * width + x] = (uint8_t)((-43 * r - 85 * g + 128 * b + 127 + 32768) >> 8);
            out[(2 * height + y) * width + x] = (uint8_t)((128 * r - 107 * g - 21 * b + 127 + 32768) >> 8);
        }
}

void sdct_oracle_ycc_to_rgb(const uint8_t *planes, size_t width, size_t height, uint8_t *out, size_t pitch)
{
    for (size_t y = 0; y < height; y++)
        for (size_t x = 0; x < width; x++) {
            const int yy = planes[(0 * height + y) * width + x], cb = planes[(1 * height + y) * width + x] - 128,
                      cr = planes[(2 * height + y) * width + x] - 128;
            out[y * pitch + 3 * x] = clamp_u8(floor_shift8(256 * yy + 359 * cr + 128));
            out[y * pitch + 3 * x + 1] = clamp_u8(floor_shift8(256 * yy - 88 * cb - 183 * cr + 128));
            out[y * pitch + 3 * x + 2] = clamp_u8(floor_shift8(256 * yy + 454 * cb + 128));
        }
}

/* per-plane statistics of a layout-C coefficient plane (this repo's definition, include/sdct.h sdct_plane_stat; the
 * reference stops at the quantised planes, src/simd_dct.cpp:1061-1069): zero levels (byte 127), zero runs, sum and
 * bitwise OR of |level - 127| over bytes [0, n) */
void sdct_oracle_plane_stats(const uint8_t *plane, size_t n, uint64_t out[4])
{
    uint64_t zeros = 0, runs = 0, sum = 0, mx = 0;
    for (size_t k = 0; k < n; k++) {
        const uint64_t a = plane[k] >= 127 ? (uint64_t)(plane[k] - 127) : (uint64_t)(127 - plane[k]);
        zeros += a == 0;
        runs += a == 0 && (k == 0 || plane[k - 1] != 127);
        sum += a;
        mx |= a;
    }
    out[0] = zeros; out[1] = runs; out[2] = sum; out[3] = mx;
}

/* component reorder: interleaved rows (ncomp bytes per pixel) <-> ncomp planar planes */
void sdct_oracle_deinterleave(const uint8_t *in, size_t pitch, size_t width, size_t height, int ncomp, uint8_t *out)
{
    for (int k = 0; k < ncomp; k++)
        for (size_t y = 0; y < height; y++)
            for (size_t x = 0; x < width; x++)
                out[((size_t)k * height + y) * width + x] = in[y * pitch + x * ncomp + k];
}

void sdct_oracle_interleave(const uint8_t *planes, size_t width, size_t height, int ncomp, uint8_t *out, size_t pitch)
{
    for (int k = 0; k < ncomp; k++)
        for (size_t y = 0; y < height; y++)
            for (size_t x = 0; x < width; x++)
                out[y * pitch + x * ncomp + k] = planes[((size_t)k * height + y) * width + x];
}

/* ---------------------------------------------------------------------------------------
 * The reference's call semantics: sizeY is the stereo (double) height, only rows below
 * sizeY/2 are visited (:268, :375, :2243), and [startY, endY] is an inclusive window.
 * Return values follow simdDctResult (src/simd_dct.h:22-27).
 * ------------------------------------------------------------------------------------- */

int sdct_oracle_ref_encode_quantize_buffer(const uint8_t *from, uint8_t *to, const float *lut,
                                           size_t size_x, size_t size_y, size_t start_y, size_t end_y)
{
    float tab[64];
    if (!from || !to) return 1;                               /* :97 */
    if ((size_x & 7) || (size_y & 7)) return 2;               /* :98 */
    forward_table(SDCT_O_LAYOUT_A, SDCT_O_MODE_EXACT, 0, lut, tab);
    for (size_t y = 0; y < size_y / 2; y += 8) {
        if (y < start_y) continue;                            /* :377 - compares y, not 2y */
        if (y > end_y) break;                                 /* :384 */
        forward_band(SDCT_O_LAYOUT_A, SDCT_O_MODE_EXACT, 0, from, size_x, 1, size_x, y / 8, 1, 0, tab, to, 0);
    }
    return 0;
}

int sdct_oracle_ref_encode_quantize32_reorder_buffer(const uint8_t *from, uint8_t *to, const float *lut,
                                                     size_t size_x, size_t size_y, size_t start_y, size_t end_y,
                                                     int quirks)
{
    float tab[64];
    if (!from || !to) return 1;                               /* :117 */
    if ((size_x & 63) || (size_y & 7)) return 2;              /* :118 */
    forward_table(SDCT_O_LAYOUT_B, SDCT_O_MODE_EXACT, quirks, lut, tab);
    for (size_t y = 0; y < size_y / 2; y += 8) {
        if (y * 2 < start_y) continue;                        /* :2245 */
        if (y * 2 > end_y) break;                             /* :2252 */
        forward_band(SDCT_O_LAYOUT_B, SDCT_O_MODE_EXACT, quirks, from, size_x, 1, size_x, y / 8, 1, 0, tab, to, 0);
    }
    return 0;
}

/* quirks = SDCT_O_QUIRK_SSE: the bytes of the SSE4.1 / SSSE3 / SSE2 variants the dispatcher picks on every x86-64 CPU
 * (:78-83); they step through a row 16 pixels at a time (:945), so size_x must be a multiple of 16 here */
int sdct_oracle_ref_encode_quantize_reorder_stereo_buffer_q(const uint8_t *from, uint8_t *to, const float *lut,
                                                            size_t size_x, size_t size_y, size_t start_y, size_t end_y,
                                                            int quirks)
{
    float tab[64];
    if (!from || !to) return 1;                               /* :75 */
    if ((size_x & 7) || (size_y & 7)) return 2;               /* :76 */
    if ((quirks & SDCT_O_QUIRK_SSE) && (size_x & 15)) return 2;
    forward_table(SDCT_O_LAYOUT_C, SDCT_O_MODE_EXACT, quirks, lut, tab);
    for (size_t y = 0; y < size_y / 2; y += 8) {
        if (y * 2 < start_y) continue;                        /* :270, :1070 */
        if (y * 2 > end_y) break;                             /* :280, :1080 */
        forward_band(SDCT_O_LAYOUT_C, SDCT_O_MODE_EXACT, quirks, from, size_x, 1, size_x, y / 8, 2,
                     size_x * size_y / 2, tab, to, size_x * size_y / 64);
    }
    return 0;
}

int sdct_oracle_ref_encode_quantize_reorder_stereo_buffer(const uint8_t *from, uint8_t *to, const float *lut,
                                                          size_t size_x, size_t size_y, size_t start_y, size_t end_y)
{
    return sdct_oracle_ref_encode_quantize_reorder_stereo_buffer_q(from, to, lut, size_x, size_y, start_y, end_y, 0);
}
