// sdct_io.cuh - tile bookkeeping and pixel-block loads shared by the fp32 kernels (sdct_kernels.cu)
// and the tensor-core kernels (sdct_tc.cu).
#pragma once
#include "sdct_kernels.h"
#include "sdct_math.cuh"

namespace sdct {

// one tile = the CB consecutive blocks [n0, n0 + count) of one coefficient set; CTA-uniform
struct Tile {
    const uint8_t *pix;     // pixel base of the frame
    uint8_t *coef;          // start of the coefficient set of (frame, comp)
    uint32_t comp;
    uint32_t n0;            // first n_local of this tile
    uint32_t count;         // valid blocks in this tile
};

// ---- division-free tile walk -------------------------------------------------------------------------
// n / d for any 32-bit n (Granlund-Montgomery, magic computed by make_fastdiv on the host)
SDCT_DEV uint32_t fastdiv(uint32_t n, const KFastDiv &d)
{
    const uint32_t t = __umulhi(n, d.m);
    return (t + ((n - t) >> d.sh1)) >> d.sh2;
}
static inline KFastDiv make_fastdiv(uint32_t d)
{
    KFastDiv r;
    if (d == 0) d = 1;
    uint32_t l = 0;
    while ((1ull << l) < d) l++;   // ceil(log2 d)
    r.m = (uint32_t)((((1ull << l) - d) << 32) / d + 1);
    r.sh1 = l < 1 ? l : 1;
    r.sh2 = l > 0 ? l - 1 : 0;
    return r;
}

// A persistent CTA's tiles k = first, first + stride, ... are global tiles blockIdx.x + k * gridDim.x.  A walk
// carries (tile / chunks, tile % chunks) from tile to tile, so that the per-tile bookkeeping has no division.
// (the launcher rejects tile counts above 0xffff0000, so a walk may step past the end without wrapping)
struct TileWalk {
    uint32_t tile, fc, chunk, step, dq, dr;
    SDCT_DEV void init(const KParams &p, uint32_t first, uint32_t stride)
    {
        tile = blockIdx.x + first * gridDim.x;
        fc = fastdiv(tile, p.div_chunks);
        chunk = tile - fc * p.chunks;
        step = stride * gridDim.x;
        dq = fastdiv(step, p.div_chunks);
        dr = step - dq * p.chunks;
    }
    SDCT_DEV void next(const KParams &p)
    {
        tile += step; fc += dq; chunk += dr;
        if (chunk >= p.chunks) { chunk -= p.chunks; fc++; }
    }
};

template <int COMPS, int CB, class Walk = TileWalk>
SDCT_DEV Tile tile_at(const KParams &p, const Walk &w)
{
    Tile t;
    uint32_t frame = w.fc;
    t.comp = 0;
    if (COMPS != 1) { frame = w.fc / COMPS; t.comp = w.fc - frame * COMPS; }
    t.n0 = w.chunk * CB;
    t.count = min((uint32_t)CB, p.rows * p.bpr - t.n0);
    t.pix = p.pixels + (uint64_t)frame * p.pixel_frame_stride;
    t.coef = p.coefs + (uint64_t)frame * p.coef_frame_stride + (uint64_t)t.comp * p.coef_comp_stride;
    return t;
}

// ---- colour front-end (flag SDCT_COLOR_YCBCR) -------------------------------------------------------------
// JFIF / BT.601 full-range RGB <-> YCbCr in 8-bit fixed point, exact integer arithmetic (restated in oracle/sdct_oracle.c):
//   Y  = (77 R + 150 G + 29 B + 128) >> 8
//   Cb = (-43 R - 85 G + 128 B + 127 + 32768) >> 8        Cr = (128 R - 107 G - 21 B + 127 + 32768) >> 8
//   R = clamp((256 Y + 359 (Cr - 128) + 128) >> 8)         B = clamp((256 Y + 454 (Cb - 128) + 128) >> 8)
//   G = clamp((256 Y - 88 (Cb - 128) - 183 (Cr - 128) + 128) >> 8)
// Forward: one DP4A per pixel and component on the pixel's bytes gathered into one word [R, G, B, .]; the coefficient
// 128 does not fit a signed byte, so the chroma forms take the complemented byte: 128 B = -128 (255 - B) + 128 * 255.
// Every sum lies in [0, 65535]; the result is byte 1 of the sum.
SDCT_DEV uint32_t dp4a_u8_s8(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// the 16-bit sum whose byte 1 is component k (0 Y, 1 Cb, 2 Cr) of the pixel gathered in g = [R, G, B, .]
SDCT_DEV uint32_t ycc_sum(uint32_t g, uint32_t k)
{
    const uint32_t coef = k == 0 ? 0x001d964du : (k == 1 ? 0x0080abd5u : 0x00eb9580u);   // [77, 150, 29] / [-43, -85, -128] / [-128, -107, -21]
    const uint32_t flip = k == 0 ? 0u : (k == 1 ? 0x00ff0000u : 0x000000ffu);
    const uint32_t bias = k == 0 ? 128u : 65535u;                                          // 128 * 255 + 127 + 32768
    if (k == 0) return __dp4a(g, coef, bias);            // (150 needs the unsigned form; byte 3 meets a zero coefficient)
    return dp4a_u8_s8(g ^ flip, coef, bias);
}
// interleaved pixel x of a 24-byte row held in six words -> [R, G, B, .]
template <int X>
SDCT_DEV uint32_t gather_rgb(const uint32_t (&s)[6])
{
    constexpr int b0 = 3 * X, w0 = b0 >> 2, o = b0 & 3;
    if (o == 0) return s[w0];
    constexpr int w1 = w0 + 1 < 6 ? w0 + 1 : 5;
    return __byte_perm(s[w0], s[w1], (uint32_t)(o | ((o + 1) << 4) | ((o + 2) << 8) | ((o + 3 > 7 ? 7 : o + 3) << 12)));
}
// component k of the eight pixels of a row -> two words of four bytes
SDCT_DEV void ycc_row(const uint32_t (&s)[6], uint32_t k, uint32_t &lo, uint32_t &hi)
{
    const uint32_t v0 = ycc_sum(gather_rgb<0>(s), k), v1 = ycc_sum(gather_rgb<1>(s), k), v2 = ycc_sum(gather_rgb<2>(s), k), v3 = ycc_sum(gather_rgb<3>(s), k);
    const uint32_t v4 = ycc_sum(gather_rgb<4>(s), k), v5 = ycc_sum(gather_rgb<5>(s), k), v6 = ycc_sum(gather_rgb<6>(s), k), v7 = ycc_sum(gather_rgb<7>(s), k);
    lo = __byte_perm(__byte_perm(v0, v1, 0x0051), __byte_perm(v2, v3, 0x0051), 0x5410);
    hi = __byte_perm(__byte_perm(v4, v5, 0x0051), __byte_perm(v6, v7, 0x0051), 0x5410);
}
// the inverse for one pixel: levels (y, cb, cr) as integers -> the three sums whose >> 8, clamped, are R, G, B
SDCT_DEV void ycc_to_rgb_sums(int y, int cb, int cr, int &r, int &g, int &b)
{
    const int y8 = (y << 8) + 128;
    r = (y8 + 359 * cr - 359 * 128) >> 8;
    g = (y8 - 88 * cb - 183 * cr + (88 + 183) * 128) >> 8;
    b = (y8 + 454 * cb - 454 * 128) >> 8;
}

// four pixels' Y, Cb, Cr bytes (one word each) -> the 12 interleaved RGB bytes; the saturating pack does the clamp
SDCT_DEV void ycc_words_to_rgb(uint32_t yw, uint32_t cbw, uint32_t crw, uint32_t *o)
{
    int r[4], g[4], b[4];
#pragma unroll
    for (int i = 0; i < 4; i++)
        ycc_to_rgb_sums((int)((yw >> (8 * i)) & 0xffu), (int)((cbw >> (8 * i)) & 0xffu), (int)((crw >> (8 * i)) & 0xffu), r[i], g[i], b[i]);
    // cvt.pack.sat.u8.s32.b32 d, a, b, c: d = c << 16 | sat(a) << 8 | sat(b)
#define SDCT_PK(b0, b1, b2, b3) pack_sat((uint32_t)(b1), (uint32_t)(b0), pack_sat((uint32_t)(b3), (uint32_t)(b2), 0u))
    o[0] = SDCT_PK(r[0], g[0], b[0], r[1]);
    o[1] = SDCT_PK(g[1], b[1], r[2], g[2]);
    o[2] = SDCT_PK(b[2], r[3], g[3], b[3]);
#undef SDCT_PK
}

// ---- pixel block -> 16 words (row r = words 2r, 2r+1), component extraction fused ----------------
// VEC: 0 = byte loads, 1 = 8-byte loads, 2 = one 16-byte load per row for the adjacent pair (P, Q)
// color (interleaved pixels only): component `comp` of the pixels converted to YCbCr instead of the raw component
template <int COMPS, int VEC>
SDCT_DEV void load_block(const uint8_t *px, uint64_t pitch, uint32_t comp, uint32_t w[16], uint32_t color = 0)
{
    if (COMPS == 1) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
            if (VEC) {
                const uint2 v = __ldg(reinterpret_cast<const uint2 *>(px + r * pitch));
                w[2 * r] = v.x; w[2 * r + 1] = v.y;
            } else {
                uint32_t lo = 0, hi = 0;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    lo |= (uint32_t)__ldg(px + r * pitch + c) << (8 * c);
                    hi |= (uint32_t)__ldg(px + r * pitch + 4 + c) << (8 * c);
                }
                w[2 * r] = lo; w[2 * r + 1] = hi;
            }
        }
    } else {
        // 8 interleaved pixels = 24 bytes; component k sits at bytes k, k+3, ..., k+21.
        auto row_words = [&](int r, uint32_t (&s)[6]) {
            if (VEC) {
                const uint2 *q = reinterpret_cast<const uint2 *>(px + r * pitch);
                const uint2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
                s[0] = a.x; s[1] = a.y; s[2] = b.x; s[3] = b.y; s[4] = c.x; s[5] = c.y;
            } else {
#pragma unroll
                for (int k = 0; k < 6; k++) {
                    s[k] = 0;
#pragma unroll
                    for (int c = 0; c < 4; c++) s[k] |= (uint32_t)__ldg(px + r * pitch + 4 * k + c) << (8 * c);
                }
            }
        };
        if (color) {   // (uniform; its own loop, so that the plain path below keeps its schedule and registers)
#pragma unroll
            for (int r = 0; r < 8; r++) {
                uint32_t s[6];
                row_words(r, s);
                ycc_row(s, comp, w[2 * r], w[2 * r + 1]);
            }
            return;
        }
        // Two PRMTs per output word; the selectors repeat every 12 bytes (= 3 words).
        const uint32_t selT = comp == 0 ? 0x0630u : (comp == 1 ? 0x0741u : 0x0052u);
        const uint32_t selU = comp == 0 ? 0x5210u : (comp == 1 ? 0x6210u : 0x7410u);
#pragma unroll
        for (int r = 0; r < 8; r++) {
            uint32_t s[6];
            row_words(r, s);
            w[2 * r] = __byte_perm(__byte_perm(s[0], s[1], selT), s[2], selU);
            w[2 * r + 1] = __byte_perm(__byte_perm(s[3], s[4], selT), s[5], selU);
        }
    }
}

}  // namespace sdct
