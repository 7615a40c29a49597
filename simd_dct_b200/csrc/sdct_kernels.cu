// sdct_kernels.cu - sm_100a kernels of the 8x8 block-DCT hot path and their launchers.
//
// Work decomposition (all kernels): the blocks of one coefficient set (one frame, one
// component) form a linear sequence n = (band*eyes + eye)*blocks_per_row + bx, which is also
// the order of every output layout (A: n*64, B: (n/8)*512, C: byte n of each of 64 planes).
// A CTA owns SDCT_TB consecutive blocks of that sequence, one 8x8 block per thread, both
// butterfly passes in registers; results are staged in shared memory in the final layout and
// leave the SM as full-line coalesced stores.
#include "sdct_kernels.h"
#include "sdct_math.cuh"

namespace sdct {

constexpr int TB = SDCT_TB;  // threads (= blocks) per CTA

struct BlockPos {
    bool valid;
    uint32_t n_local;   // index inside the processed window
    uint64_t n_out;     // index inside the coefficient set
    const uint8_t *px;  // top-left pixel (component 0) of the block
    uint8_t *coef;      // start of the coefficient set of (frame, comp)
    uint32_t comp;
    uint32_t n0;        // first n_local of this CTA
    uint32_t count;     // valid blocks in this CTA
};

SDCT_DEV BlockPos locate(const KParams &p)
{
    BlockPos b;
    const uint32_t chunk = blockIdx.x % p.chunks;
    const uint32_t fc = blockIdx.x / p.chunks;
    b.comp = fc % p.comps;
    const uint32_t frame = fc / p.comps;
    const uint32_t total = p.rows * p.bpr;
    b.n0 = chunk * TB;
    b.count = min((uint32_t)TB, total - b.n0);
    b.n_local = b.n0 + threadIdx.x;
    b.valid = threadIdx.x < b.count;
    const uint32_t nl = b.valid ? b.n_local : b.n0;  // keep addresses in range
    const uint32_t row = nl / p.bpr + p.row0;
    const uint32_t bx = nl - (nl / p.bpr) * p.bpr;
    const uint32_t band = row / p.eyes, eye = row - band * p.eyes;
    b.n_out = (uint64_t)p.row0 * p.bpr + nl;
    b.px = p.pixels + (uint64_t)frame * p.pixel_frame_stride + (uint64_t)eye * p.eye_stride +
           (uint64_t)band * 8 * p.pitch + (uint64_t)bx * 8 * p.comps;
    b.coef = p.coefs + (uint64_t)frame * p.coef_frame_stride + (uint64_t)b.comp * p.coef_comp_stride;
    return b;
}

// ---- pixel block -> 16 words (row r = words 2r, 2r+1), component extraction fused ----------
template <int COMPS, bool VEC>
SDCT_DEV void load_block_bytes(const uint8_t *px, uint64_t pitch, uint32_t comp, uint32_t w[16])
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

// ---- forward transform of one block: bytes -> 16 packed coefficient words in layout order --
template <int LAYOUT, int MODE, bool X3>
SDCT_DEV void forward_block(const uint32_t w[16], const float *tab, uint32_t o[16])
{
    float b[64];
#pragma unroll
    for (int i = 0; i < 64; i++) {
        const float x = byte_to_float(w[i >> 2], i & 3);
        b[i] = (MODE == SDCT_K_EXACT && LAYOUT != SDCT_K_LAYOUT_B) ? div255(x) : x;
    }
    if (MODE == SDCT_K_FAST) {
#pragma unroll
        for (int r = 0; r < 8; r++) dct8_fast<1>(b + 8 * r);
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_fast<8>(b + c);
    } else if (LAYOUT == SDCT_K_LAYOUT_A) {  // rows, then columns (:343-358)
#pragma unroll
        for (int r = 0; r < 8; r++) dct8_scalar<1>(b + 8 * r);
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_scalar<8>(b + c);
    } else if (LAYOUT == SDCT_K_LAYOUT_C) {  // columns, then rows (:224-241)
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_scalar<8>(b + c);
#pragma unroll
        for (int r = 0; r < 8; r++) dct8_scalar<1>(b + 8 * r);
    } else {                                 // layout B: rows, then columns (:2158-2215)
#pragma unroll
        for (int r = 0; r < 8; r++) dct8_paired<1, X3>(b + 8 * r);
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_paired<8, X3>(b + c);
    }
    // b[u*8+v] = coefficient (u vertical, v horizontal).  Output byte i: A -> (u,v) = (i%8, i/8)
#pragma unroll
    for (int j = 0; j < 16; j++) {
        uint32_t q[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = 4 * j + k;
            const int src = (LAYOUT == SDCT_K_LAYOUT_A) ? (i & 7) * 8 + (i >> 3) : i;
            if (MODE == SDCT_K_FAST) q[k] = quant_fast(b[src], tab[i]);
            else if (LAYOUT == SDCT_K_LAYOUT_B) q[k] = quant_avx(b[src], tab[i]);
            else q[k] = quant_scalar(b[src], tab[i]);
        }
        o[j] = pack4(q[0], q[1], q[2], q[3]);
    }
}

// ---- staging: thread's 16 words -> shared memory image of the CTA's output -----------------
template <int LAYOUT>
SDCT_DEV void stage_coefs(uint8_t *sm, const uint32_t o[16])
{
    const uint32_t t = threadIdx.x;
    if (LAYOUT == SDCT_K_LAYOUT_A) {
        uint4 *d = reinterpret_cast<uint4 *>(sm + t * 64);
#pragma unroll
        for (int j = 0; j < 4; j++) d[j ^ ((t >> 1) & 3)] = make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
    } else if (LAYOUT == SDCT_K_LAYOUT_B) {
        uint8_t *d = sm + (t >> 3) * 512 + (t & 7);
#pragma unroll
        for (int i = 0; i < 64; i++) d[i * 8] = (uint8_t)(o[i >> 2] >> (8 * (i & 3)));
    } else {
#pragma unroll
        for (int i = 0; i < 64; i++) sm[i * TB + t] = (uint8_t)(o[i >> 2] >> (8 * (i & 3)));
    }
}

template <int LAYOUT>
SDCT_DEV void flush_coefs(const uint8_t *sm, const KParams &p, const BlockPos &b)
{
    const uint32_t t = threadIdx.x;
    const uint64_t n_first = (uint64_t)p.row0 * p.bpr + b.n0;
    if (LAYOUT != SDCT_K_LAYOUT_C) {
        uint8_t *dst = b.coef + n_first * 64;
        const uint32_t chunks16 = b.count * 4;
        if (p.coef_aligned16) {
            for (uint32_t k = t; k < chunks16; k += TB) {
                uint32_t src = k;
                if (LAYOUT == SDCT_K_LAYOUT_A) src = (k & ~3u) | ((k & 3) ^ ((k >> 3) & 3));
                reinterpret_cast<uint4 *>(dst)[k] = reinterpret_cast<const uint4 *>(sm)[src];
            }
        } else {
            for (uint32_t k = t; k < chunks16 * 16; k += TB) {
                uint32_t c = k >> 4;
                if (LAYOUT == SDCT_K_LAYOUT_A) c = (c & ~3u) | ((c & 3) ^ ((c >> 3) & 3));
                dst[k] = sm[c * 16 + (k & 15)];
            }
        }
    } else {
        const uint32_t warp = t >> 5, lane = t & 31;
        for (uint32_t i = warp; i < 64; i += TB / 32) {
            uint8_t *dst = b.coef + (uint64_t)i * p.coef_plane_stride + n_first;
            if (p.coef_aligned4 && (b.count & 3) == 0) {
                for (uint32_t k = lane; k < b.count / 4; k += 32)
                    reinterpret_cast<uint32_t *>(dst)[k] = reinterpret_cast<const uint32_t *>(sm + i * TB)[k];
            } else {
                for (uint32_t k = lane; k < b.count; k += 32) dst[k] = sm[i * TB + k];
            }
        }
    }
}

template <int LAYOUT, int MODE, bool X3, int COMPS, bool VEC>
__global__ void __launch_bounds__(SDCT_TB) forward_kernel(const __grid_constant__ KParams p)
{
    __shared__ __align__(16) uint8_t sm[TB * 64];
    const BlockPos b = locate(p);
    uint32_t w[16], o[16];
    load_block_bytes<COMPS, VEC>(b.px, p.pitch, b.comp, w);
    forward_block<LAYOUT, MODE, X3>(w, p.tab, o);
    stage_coefs<LAYOUT>(sm, o);
    __syncthreads();
    flush_coefs<LAYOUT>(sm, p, b);
}

// ---- inverse ------------------------------------------------------------------------------
// coefficient bytes of this thread's block -> 16 words in natural order (word j = i 4j..4j+3)
template <int LAYOUT>
SDCT_DEV void load_coefs(uint8_t *sm, const KParams &p, const BlockPos &b, uint32_t o[16])
{
    const uint32_t t = threadIdx.x;
    const uint64_t n_first = (uint64_t)p.row0 * p.bpr + b.n0;
    if (LAYOUT != SDCT_K_LAYOUT_C) {
        const uint8_t *src = b.coef + n_first * 64;
        const uint32_t chunks16 = b.count * 4;
        if (p.coef_aligned16) {
            for (uint32_t k = t; k < chunks16; k += TB)
                reinterpret_cast<uint4 *>(sm)[k] = __ldg(reinterpret_cast<const uint4 *>(src) + k);
        } else {
            for (uint32_t k = t; k < chunks16 * 16; k += TB) sm[k] = __ldg(src + k);
        }
    } else {
        const uint32_t warp = t >> 5, lane = t & 31;
        for (uint32_t i = warp; i < 64; i += TB / 32) {
            const uint8_t *src = b.coef + (uint64_t)i * p.coef_plane_stride + n_first;
            if (p.coef_aligned4 && (b.count & 3) == 0) {
                for (uint32_t k = lane; k < b.count / 4; k += 32)
                    reinterpret_cast<uint32_t *>(sm + i * TB)[k] = __ldg(reinterpret_cast<const uint32_t *>(src) + k);
            } else {
                for (uint32_t k = lane; k < b.count; k += 32) sm[i * TB + k] = __ldg(src + k);
            }
        }
    }
    __syncthreads();
    if (LAYOUT == SDCT_K_LAYOUT_A) {
        const uint32_t *s = reinterpret_cast<const uint32_t *>(sm + t * 64);
#pragma unroll
        for (int j = 0; j < 16; j++) o[j] = s[j];
    } else if (LAYOUT == SDCT_K_LAYOUT_B) {
        const uint8_t *s = sm + (t >> 3) * 512 + (t & 7);
#pragma unroll
        for (int j = 0; j < 16; j++)
            o[j] = pack4(s[(4 * j) * 8], s[(4 * j + 1) * 8], s[(4 * j + 2) * 8], s[(4 * j + 3) * 8]);
    } else {
#pragma unroll
        for (int j = 0; j < 16; j++)
            o[j] = pack4(sm[(4 * j) * TB + t], sm[(4 * j + 1) * TB + t], sm[(4 * j + 2) * TB + t], sm[(4 * j + 3) * TB + t]);
    }
}

// 16 coefficient words (layout order) -> 16 pixel words (row r = words 2r, 2r+1)
template <int LAYOUT>
SDCT_DEV void inverse_block(const uint32_t o[16], const float *tab, uint32_t w[16])
{
    float b[64];
#pragma unroll
    for (int i = 0; i < 64; i++) {
        const int dst = (LAYOUT == SDCT_K_LAYOUT_A) ? (i & 7) * 8 + (i >> 3) : i;
        // (q - 127) exactly: byte into the mantissa of 2^23, subtract 2^23 + 127
        const float q = __uint_as_float(__byte_perm(o[i >> 2], 0x4B000000u, 0x7650 + (i & 3))) - 8388735.0f;
        b[dst] = fmul(q, tab[i]);
    }
#pragma unroll
    for (int c = 0; c < 8; c++) idct8<8, false>(b + c);
#pragma unroll
    for (int r = 0; r < 8; r++) idct8<1, true>(b + 8 * r);
#pragma unroll
    for (int j = 0; j < 16; j++)
        w[j] = pack4(unit_to_u8(b[4 * j]), unit_to_u8(b[4 * j + 1]), unit_to_u8(b[4 * j + 2]), unit_to_u8(b[4 * j + 3]));
}

template <int COMPS, bool VEC>
SDCT_DEV void store_block_bytes(uint8_t *px, uint64_t pitch, uint32_t comp, const uint32_t w[16])
{
#pragma unroll
    for (int r = 0; r < 8; r++) {
        if (COMPS == 1 && VEC) {
            *reinterpret_cast<uint2 *>(px + r * pitch) = make_uint2(w[2 * r], w[2 * r + 1]);
        } else {
#pragma unroll
            for (int c = 0; c < 8; c++)
                px[r * pitch + c * COMPS + comp] = (uint8_t)(w[2 * r + (c >> 2)] >> (8 * (c & 3)));
        }
    }
}

template <int LAYOUT, int COMPS, bool VEC>
__global__ void __launch_bounds__(SDCT_TB) inverse_kernel(const __grid_constant__ KParams p)
{
    __shared__ __align__(16) uint8_t sm[TB * 64];
    const BlockPos b = locate(p);
    uint32_t o[16], w[16];
    load_coefs<LAYOUT>(sm, p, b, o);
    inverse_block<LAYOUT>(o, p.tab, w);
    if (b.valid) store_block_bytes<COMPS, VEC>(const_cast<uint8_t *>(b.px), p.pitch, b.comp, w);
}

// fused forward -> quantise -> dequantise -> inverse: coefficients stay in registers
template <int LAYOUT, int MODE, bool X3, int COMPS, bool VEC>
__global__ void __launch_bounds__(SDCT_TB) roundtrip_kernel(const __grid_constant__ KParams p)
{
    const BlockPos b = locate(p);
    uint32_t w[16], o[16];
    load_block_bytes<COMPS, VEC>(b.px, p.pitch, b.comp, w);
    forward_block<LAYOUT, MODE, X3>(w, p.tab, o);
    inverse_block<LAYOUT>(o, p.tab2, w);
    if (b.valid) {
        uint8_t *dst = p.pixels_out + (b.px - p.pixels);
        store_block_bytes<COMPS, VEC>(dst, p.pitch, b.comp, w);
    }
}

}  // namespace sdct

// ---- launchers ------------------------------------------------------------------------------
using namespace sdct;

template <int LAYOUT, int MODE, bool X3>
static cudaError_t launch_fwd2(int op, const KParams &p, unsigned grid, cudaStream_t s)
{
    const bool vec = p.pixels_aligned8 != 0;
#define SDCT_GO(K, C, V)                                                     \
    do {                                                                     \
        K<LAYOUT, MODE, X3, C, V><<<grid, SDCT_TB, 0, s>>>(p);               \
        return cudaGetLastError();                                           \
    } while (0)
    if (op == SDCT_K_OP_FORWARD) {
        if (p.comps == 1) { if (vec) SDCT_GO(forward_kernel, 1, true); else SDCT_GO(forward_kernel, 1, false); }
        else              { if (vec) SDCT_GO(forward_kernel, 3, true); else SDCT_GO(forward_kernel, 3, false); }
    } else {
        if (p.comps == 1) { if (vec) SDCT_GO(roundtrip_kernel, 1, true); else SDCT_GO(roundtrip_kernel, 1, false); }
        else              { if (vec) SDCT_GO(roundtrip_kernel, 3, true); else SDCT_GO(roundtrip_kernel, 3, false); }
    }
#undef SDCT_GO
}

template <int LAYOUT>
static cudaError_t launch_inv2(const KParams &p, unsigned grid, cudaStream_t s)
{
    const bool vec = p.pixels_aligned8 != 0;
    if (p.comps == 1) {
        if (vec) inverse_kernel<LAYOUT, 1, true><<<grid, SDCT_TB, 0, s>>>(p);
        else inverse_kernel<LAYOUT, 1, false><<<grid, SDCT_TB, 0, s>>>(p);
    } else {
        inverse_kernel<LAYOUT, 3, false><<<grid, SDCT_TB, 0, s>>>(p);
    }
    return cudaGetLastError();
}

extern "C" cudaError_t sdct_k_launch(int op, int layout, int mode, int x3, const KParams *pp, cudaStream_t s)
{
    const KParams &p = *pp;
    const unsigned long long grid64 = (unsigned long long)p.chunks * p.comps * p.frames;
    if (grid64 == 0) return cudaSuccess;
    if (grid64 > 0x7fffffffull) return cudaErrorInvalidConfiguration;
    const unsigned grid = (unsigned)grid64;
    if (op == SDCT_K_OP_INVERSE) {
        switch (layout) {
        case SDCT_K_LAYOUT_A: return launch_inv2<SDCT_K_LAYOUT_A>(p, grid, s);
        case SDCT_K_LAYOUT_B: return launch_inv2<SDCT_K_LAYOUT_B>(p, grid, s);
        default: return launch_inv2<SDCT_K_LAYOUT_C>(p, grid, s);
        }
    }
    if (mode == SDCT_K_FAST) {
        switch (layout) {
        case SDCT_K_LAYOUT_A: return launch_fwd2<SDCT_K_LAYOUT_A, SDCT_K_FAST, false>(op, p, grid, s);
        case SDCT_K_LAYOUT_B: return launch_fwd2<SDCT_K_LAYOUT_B, SDCT_K_FAST, false>(op, p, grid, s);
        default: return launch_fwd2<SDCT_K_LAYOUT_C, SDCT_K_FAST, false>(op, p, grid, s);
        }
    }
    switch (layout) {
    case SDCT_K_LAYOUT_A: return launch_fwd2<SDCT_K_LAYOUT_A, SDCT_K_EXACT, false>(op, p, grid, s);
    case SDCT_K_LAYOUT_B:
        return x3 ? launch_fwd2<SDCT_K_LAYOUT_B, SDCT_K_EXACT, true>(op, p, grid, s)
                  : launch_fwd2<SDCT_K_LAYOUT_B, SDCT_K_EXACT, false>(op, p, grid, s);
    default: return launch_fwd2<SDCT_K_LAYOUT_C, SDCT_K_EXACT, false>(op, p, grid, s);
    }
}
