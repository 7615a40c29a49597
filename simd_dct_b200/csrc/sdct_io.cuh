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

// ---- pixel block -> 16 words (row r = words 2r, 2r+1), component extraction fused ----------------
// VEC: 0 = byte loads, 1 = 8-byte loads, 2 = one 16-byte load per row for the adjacent pair (P, Q)
template <int COMPS, int VEC>
SDCT_DEV void load_block(const uint8_t *px, uint64_t pitch, uint32_t comp, uint32_t w[16])
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
        // Two PRMTs per output word; the selectors repeat every 12 bytes (= 3 words).
        const uint32_t selT = comp == 0 ? 0x0630u : (comp == 1 ? 0x0741u : 0x0052u);
        const uint32_t selU = comp == 0 ? 0x5210u : (comp == 1 ? 0x6210u : 0x7410u);
#pragma unroll
        for (int r = 0; r < 8; r++) {
            uint32_t s[6];
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
            w[2 * r] = __byte_perm(__byte_perm(s[0], s[1], selT), s[2], selU);
            w[2 * r + 1] = __byte_perm(__byte_perm(s[3], s[4], selT), s[5], selU);
        }
    }
}

}  // namespace sdct
