// sdct_kernels.cu - sm_100a kernels of the 8x8 block-DCT hot path and their launchers.
//
// Work decomposition (all kernels): the blocks of one coefficient set (one frame, one component)
// form a linear sequence n = (band*eyes + eye)*blocks_per_row + bx, which is also the order of
// every output layout (A: n*64, B: (n/8)*512, C: byte n of each of 64 planes).  A CTA owns
// SDCT_CTA_BLOCKS = 256 consecutive blocks of that sequence; each of its 128 threads owns two of
// them (P, Q) and runs both butterfly passes of both blocks in registers as packed-fp32 pairs
// (sdct_math.cuh).  Quantised levels leave the registers through the saturating pack, are staged in
// shared memory as the byte image of the CTA's output, and reach HBM as 16-byte coalesced stores;
// pixels are read with 16-byte loads (one per row for the two adjacent blocks of a thread).
//   layouts B, C: P, Q = blocks 2t, 2t+1 (adjacent: bytes (P,Q) of one coefficient are adjacent too)
//   layout  A   : P, Q = blocks t, t+128 (a block's 64 bytes are contiguous; bank-friendly slots)
#include "sdct_kernels.h"
#include "sdct_math.cuh"
#include "sdct_io.cuh"

#include <atomic>

// This file is compiled once per SDCT_PART (simd_dct_b200/build.py) so that the template
// instantiations build in parallel: part 0 = dispatcher + self-test kernel, parts 1..8 = one
// (operation, layout / mode) family each.
#ifndef SDCT_CONV_MAGIC
#define SDCT_CONV_MAGIC 0
#endif
#ifndef SDCT_CONV_FOLD
#define SDCT_CONV_FOLD 1
#endif
#ifndef SDCT_PART
#error "compile with -DSDCT_PART=<0..8>"
#endif

namespace sdct {

constexpr int TB = SDCT_TB;              // threads per CTA
constexpr int CB = SDCT_CTA_BLOCKS;      // 8x8 blocks per CTA

// staging image of one CTA's coefficients
constexpr int B_GROUP_PITCH = 528;       // layout B: 512 B group + 16 B pad (conflict-free STS.U16)
constexpr int A_SLOT = 80;               // layout A: 64 B block + 16 B pad (conflict-free STS.128)
constexpr int C_PITCH = CB;              // layout C: 64 rows of CB bytes
constexpr int STAGE_BYTES = 32 * B_GROUP_PITCH > CB * A_SLOT ? 32 * B_GROUP_PITCH : CB * A_SLOT;

constexpr int IN_STAGE_BYTES = 8 * TB * 16;   // prefetched pixels of one tile: 8 rows x 128 threads x (P 8 B | Q 8 B)

// this thread's two blocks of a tile: P, Q = slots 2t, 2t+1 (layouts B, C) or t, t+128 (layout A)
struct BlockPos {
    const uint8_t *px[2];   // top-left pixel (component 0)
    bool valid[2];
    uint32_t comp;
    uint32_t color;         // interleaved pixels: convert to YCbCr on load
};

template <int LAYOUT>
SDCT_DEV BlockPos locate(const KParams &p, const Tile &t)
{
    BlockPos b;
    b.comp = t.comp;
    b.color = p.color;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const uint32_t slot = (LAYOUT == SDCT_K_LAYOUT_A) ? threadIdx.x + k * TB : 2 * threadIdx.x + k;
        b.valid[k] = slot < t.count;
        const uint32_t nl = t.n0 + (b.valid[k] ? slot : 0);  // keep addresses in range
        const uint32_t r = fastdiv(nl, p.div_bpr);
        const uint32_t bx = nl - r * p.bpr;
        const uint32_t row = r + p.row0;
        const uint32_t band = p.eyes == 2 ? row >> 1 : row, eye = p.eyes == 2 ? row & 1 : 0;
        b.px[k] = t.pix + (uint64_t)eye * p.eye_stride + (uint64_t)band * 8 * p.pitch + (uint64_t)bx * 8 * p.comps;
    }
    return b;
}

// ---- asynchronous pixel prefetch (cp.async): each thread copies the rows of ITS two blocks into
// ITS 16-byte slots of the stage, so no barrier is needed between the copy and the read ---------
SDCT_DEV void cp_async16(void *smem, const void *gmem)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
SDCT_DEV void cp_async8(void *smem, const void *gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
SDCT_DEV void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> SDCT_DEV void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int VEC>
SDCT_DEV void prefetch_pair(uint8_t *stage, const BlockPos &b, uint64_t pitch)
{
    uint8_t *slot = stage + threadIdx.x * 16;
#pragma unroll
    for (int r = 0; r < 8; r++) {
        if (VEC == 2) {
            cp_async16(slot + r * (TB * 16), b.px[0] + r * pitch);
        } else {
            cp_async8(slot + r * (TB * 16), b.px[0] + r * pitch);
            cp_async8(slot + r * (TB * 16) + 8, b.px[1] + r * pitch);
        }
    }
}

SDCT_DEV void read_prefetched(const uint8_t *stage, uint32_t wp[16], uint32_t wq[16])
{
    const uint8_t *slot = stage + threadIdx.x * 16;
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const uint4 v = *reinterpret_cast<const uint4 *>(slot + r * (TB * 16));
        wp[2 * r] = v.x; wp[2 * r + 1] = v.y; wq[2 * r] = v.z; wq[2 * r + 1] = v.w;
    }
}

template <int COMPS, int VEC>
SDCT_DEV void load_pair(const BlockPos &b, uint64_t pitch, uint32_t wp[16], uint32_t wq[16])
{
    if (COMPS == 1 && VEC == 2) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(b.px[0] + r * pitch));
            wp[2 * r] = v.x; wp[2 * r + 1] = v.y; wq[2 * r] = v.z; wq[2 * r + 1] = v.w;
        }
    } else {
        load_block<COMPS, (VEC ? 1 : 0)>(b.px[0], pitch, b.comp, wp, b.color);
        load_block<COMPS, (VEC ? 1 : 0)>(b.px[1], pitch, b.comp, wq, b.color);
    }
}

// ---- forward transform of the pair: pixel words -> 64 level bit patterns, b[u*8+v] ---------------
// (u = vertical, v = horizontal frequency).  TABIDX maps the natural index to the layout's.
template <int LAYOUT>
SDCT_DEV constexpr int tab_index(int nat) { return LAYOUT == SDCT_K_LAYOUT_A ? (nat & 7) * 8 + (nat >> 3) : nat; }

// does this (layout, mode) produce complemented levels (255 - q)?
template <int LAYOUT, int MODE>
SDCT_DEV constexpr bool complemented() { return MODE == SDCT_K_FAST || LAYOUT == SDCT_K_LAYOUT_B; }

template <int LAYOUT, int MODE, bool X3>
SDCT_DEV void forward_pair(const uint32_t wp[16], const uint32_t wq[16], const float2 *tab, F2 b[64])
{
    // FAST (every layout) and EXACT layout B transform the raw pixel values, rows first: the conversion is folded
    // into the first butterfly stage of the row pass (stage1_pixels, sdct_math.cuh) - no I2F, no extra FMA-pipe op
    constexpr bool FOLD = SDCT_CONV_FOLD && (MODE == SDCT_K_FAST || LAYOUT == SDCT_K_LAYOUT_B);
    if (FOLD) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const Stage1 a = stage1_pixels(wp[2 * r], wp[2 * r + 1], wq[2 * r], wq[2 * r + 1]);
            if (MODE == SDCT_K_FAST) dct8_fast_tail<1>(a, b + 8 * r);
            else dct8_paired_tail<1, X3 ? 1 : 0>(a, b + 8 * r);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 64; i++) {
#if SDCT_CONV_MAGIC
            // 2^23 + byte assembled by PRMT (ALU pipe), minus 2^23 (exact) on the FMA pipe: no XU conversion
            const uint32_t sel = 0x7540u | (uint32_t)(i & 3);
            const F2 m = make_float2(__uint_as_float(__byte_perm(wp[i >> 2], 0x4B000000u, sel)), __uint_as_float(__byte_perm(wq[i >> 2], 0x4B000000u, sel)));
            const F2 x = sub2(m, splat(8388608.f));
#else
            const F2 x = make_float2(byte_to_float(wp[i >> 2], i & 3), byte_to_float(wq[i >> 2], i & 3));
#endif
            // (layout C with the quirk flag = the reference's SSE encoders: x * (1.f / 255), :955-962 - a product, so the first
            //  pass below takes the protected adds)
            b[i] = (MODE == SDCT_K_EXACT && LAYOUT == SDCT_K_LAYOUT_C && X3) ? mulx<PFP>(x, SDCT_R255_HI)
                 : (MODE == SDCT_K_EXACT && LAYOUT != SDCT_K_LAYOUT_B) ? div255(x) : x;
        }
    }
    if (MODE == SDCT_K_FAST) {
        if (!FOLD) {
#pragma unroll
            for (int r = 0; r < 8; r++) dct8_fast<1>(b + 8 * r);
        }
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_fast<8>(b + c);
    } else if (LAYOUT == SDCT_K_LAYOUT_A) {  // rows, then columns (:343-358)
#pragma unroll
        for (int r = 0; r < 8; r++) dct8_scalar<1, false>(b + 8 * r);
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_scalar<8, true>(b + c);
    } else if (LAYOUT == SDCT_K_LAYOUT_C && X3) {  // SSE encoders (:964-1007): columns, then rows, with inplace_dct8_sse41
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_paired<8, true, 2>(b + c);
#pragma unroll
        for (int r = 0; r < 8; r++) dct8_paired<1, true, 2>(b + 8 * r);
    } else if (LAYOUT == SDCT_K_LAYOUT_C) {  // columns, then rows (:224-241)
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_scalar<8, false>(b + c);
#pragma unroll
        for (int r = 0; r < 8; r++) dct8_scalar<1, true>(b + 8 * r);
    } else {                                 // layout B: rows, then columns (:2158-2215)
        if (!FOLD) {
#pragma unroll
            for (int r = 0; r < 8; r++) dct8_paired<1, false, X3 ? 1 : 0>(b + 8 * r);
        }
#pragma unroll
        for (int c = 0; c < 8; c++) dct8_paired<8, true, X3 ? 1 : 0>(b + c);
    }
#pragma unroll
    for (int i = 0; i < 64; i++) {
        const F2 t = tab[tab_index<LAYOUT>(i)];
        b[i] = (MODE == SDCT_K_EXACT && LAYOUT == SDCT_K_LAYOUT_C && X3) ? quant_sse(b[i], t)
             : (MODE == SDCT_K_EXACT && LAYOUT != SDCT_K_LAYOUT_B) ? quant_scalar(b[i], t) : quant_rne_c(b[i], t);
    }
}

// ---- staging: level bit patterns -> shared-memory byte image of the CTA's output -----------------
template <int LAYOUT>
SDCT_DEV void stage_levels(uint8_t *sm, const F2 b[64])
{
    const uint32_t t = threadIdx.x;
    if (LAYOUT == SDCT_K_LAYOUT_B) {
        // blocks 2t, 2t+1 of group t/4: coefficient i -> bytes [i*8 + 2(t%4), +1]
        uint16_t *d = reinterpret_cast<uint16_t *>(sm + (t >> 2) * B_GROUP_PITCH + 2 * (t & 3));
#pragma unroll
        for (int i = 0; i < 64; i++) d[i * 4] = (uint16_t)pack_pq(b[i]);
    } else if (LAYOUT == SDCT_K_LAYOUT_C) {
        uint16_t *d = reinterpret_cast<uint16_t *>(sm + 2 * t);
#pragma unroll
        for (int i = 0; i < 64; i++) d[i * (C_PITCH / 2)] = (uint16_t)pack_pq(b[i]);
    } else {
        // byte i of a block = coefficient (u, v) = (i%8, i/8): 16-byte chunk j holds v = 2j, 2j+1
#pragma unroll
        for (int k = 0; k < 2; k++) {
            uint4 *d = reinterpret_cast<uint4 *>(sm + (t + k * TB) * A_SLOT);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                uint32_t w[4];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const int v = 2 * j + (q >> 1), u0 = 4 * (q & 1);
#define SDCT_L(u) (k ? b[(u) * 8 + v].y : b[(u) * 8 + v].x)
                    w[q] = pack4_sat(SDCT_L(u0), SDCT_L(u0 + 1), SDCT_L(u0 + 2), SDCT_L(u0 + 3));
#undef SDCT_L
                }
                d[j] = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
    }
}

SDCT_DEV uint4 xor4(uint4 v, uint32_t m) { return make_uint4(v.x ^ m, v.y ^ m, v.z ^ m, v.w ^ m); }

// shared-memory image -> global memory, coalesced; COMPL undoes the complemented levels
template <int LAYOUT, bool COMPL>
SDCT_DEV void flush_coefs(const uint8_t *sm, const KParams &p, const Tile &b)
{
    const uint32_t t = threadIdx.x;
    const uint64_t n_first = (uint64_t)p.row0 * p.bpr + b.n0;
    const uint32_t m32 = COMPL ? 0xffffffffu : 0u;
    if (LAYOUT != SDCT_K_LAYOUT_C) {
        uint8_t *dst = b.coef + n_first * 64;
        const uint32_t chunks16 = b.count * 4;
        auto src_off = [](uint32_t k) -> uint32_t {  // 16-byte chunk k of the output image
            return LAYOUT == SDCT_K_LAYOUT_A ? (k >> 2) * A_SLOT + (k & 3) * 16 : (k >> 5) * B_GROUP_PITCH + (k & 31) * 16;
        };
        if (p.coef_aligned16) {
#pragma unroll 4
            for (uint32_t k = t; k < chunks16; k += TB)
                reinterpret_cast<uint4 *>(dst)[k] = xor4(*reinterpret_cast<const uint4 *>(sm + src_off(k)), m32);
        } else {
            for (uint32_t k = t; k < chunks16 * 16; k += TB) dst[k] = sm[src_off(k >> 4) + (k & 15)] ^ (uint8_t)m32;
        }
    } else {
        if (p.coef_aligned16 && (b.count & 15) == 0) {
            // 16 chunks per plane row: a warp writes two 256-byte row segments per instruction
            const uint32_t per_row = b.count >> 4;
            for (uint32_t k = t; k < 64 * per_row; k += TB) {
                const uint32_t i = k / per_row, c = k - i * per_row;
                uint8_t *dst = b.coef + (uint64_t)i * p.coef_plane_stride + n_first;
                reinterpret_cast<uint4 *>(dst)[c] = xor4(*reinterpret_cast<const uint4 *>(sm + i * C_PITCH + c * 16), m32);
            }
        } else {
            const uint32_t warp = t >> 5, lane = t & 31;
            for (uint32_t i = warp; i < 64; i += TB / 32) {
                uint8_t *dst = b.coef + (uint64_t)i * p.coef_plane_stride + n_first;
                if (p.coef_aligned4 && (b.count & 3) == 0) {
                    for (uint32_t k = lane; k < b.count / 4; k += 32)
                        reinterpret_cast<uint32_t *>(dst)[k] = reinterpret_cast<const uint32_t *>(sm + i * C_PITCH)[k] ^ m32;
                } else {
                    for (uint32_t k = lane; k < b.count; k += 32) dst[k] = sm[i * C_PITCH + k] ^ (uint8_t)m32;
                }
            }
        }
    }
}

// 16-byte-aligned output: the bytes a warp stages are whole 16-byte chunks of the output that no other warp touches
// (B: its 64 blocks = 8 groups of 512 bytes; A: blocks 32w.. and TB + 32w.., 64 bytes each; C: 64 bytes of each of the
// 64 plane rows - block counts are multiples of 16 there), so the warp flushes its own part after a __syncwarp() and
// the CTA needs no barrier at all.  The flush is ONE basic block: eight unconditional 16-byte shared-memory loads
// (thread-constant base + immediate offsets), then eight PREDICATED 16-byte stores - no branch around a chunk, so the
// loads are all in flight together and ptxas can sink the stores into the arithmetic that follows (the caller
// flushes tile k - 1 at the start of tile k's row pass, when most data registers are still free).
SDCT_DEV void st_global_v4_if(void *ptr, uint4 v, bool pred)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %5, 0;\n\t@p st.global.v4.u32 [%0], {%1, %2, %3, %4};\n\t}"
                 ::"l"(ptr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"((uint32_t)pred) : "memory");
}

struct FlushJob {       // the tile whose staged image is still to be written (count == 0: none)
    uint8_t *coef;      // start of the coefficient set
    uint64_t n_first;   // first block of the tile in the set's block sequence
    uint32_t count;
};

template <int LAYOUT, bool COMPL>
SDCT_DEV void flush_coefs_warp(const uint8_t *sm, const KParams &p, const FlushJob &f)
{
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t m32 = COMPL ? 0xffffffffu : 0u;
    uint4 v[8];
    if (LAYOUT == SDCT_K_LAYOUT_C) {   // chunk j: plane row 8j + lane/4, bytes 64 warp + 16 (lane % 4) ..
        const uint32_t c = 64u * warp + 16u * (lane & 3);
        const uint8_t *src = sm + (lane >> 2) * C_PITCH + c;
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = *reinterpret_cast<const uint4 *>(src + j * 8 * C_PITCH);
        uint8_t *dst = f.coef + (uint64_t)(lane >> 2) * p.coef_plane_stride + f.n_first + c;
        const uint64_t step = 8 * p.coef_plane_stride;
#pragma unroll
        for (int j = 0; j < 8; j++, dst += step) st_global_v4_if(dst, xor4(v[j], m32), c < f.count);
        return;
    }
    const uint32_t chunks16 = f.count * 4;
    if (LAYOUT == SDCT_K_LAYOUT_B) {   // chunk k = 256 warp + 32 j + lane: group 8 warp + j, piece lane
        const uint8_t *src = sm + 8u * warp * B_GROUP_PITCH + lane * 16u;
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = *reinterpret_cast<const uint4 *>(src + j * B_GROUP_PITCH);
        const uint32_t k0 = 256u * warp + lane;
        uint8_t *dst = f.coef + f.n_first * 64 + 16ull * k0;
#pragma unroll
        for (int j = 0; j < 8; j++) st_global_v4_if(dst + j * 512, xor4(v[j], m32), k0 + 32u * j < chunks16);
    } else {                           // chunk k = (j / 4) 4 TB + 128 warp + 32 (j % 4) + lane: block k / 4, piece k % 4
        const uint8_t *src = sm + (32u * warp + (lane >> 2)) * A_SLOT + (lane & 3) * 16u;
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = *reinterpret_cast<const uint4 *>(src + ((j >> 2) * TB + 8 * (j & 3)) * A_SLOT);
        const uint32_t k0 = 128u * warp + lane;
        uint8_t *dst = f.coef + f.n_first * 64 + 16ull * k0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const uint32_t dk = (j >> 2) * 4u * TB + 32u * (j & 3);
            st_global_v4_if(dst + 16ull * dk, xor4(v[j], m32), k0 + dk < chunks16);
        }
    }
}

// can every warp move its own part of the coefficient image?  (uniform per launch)
template <int LAYOUT>
SDCT_DEV bool warp_local_coefs(const KParams &p)
{
    return p.coef_aligned16 && (LAYOUT != SDCT_K_LAYOUT_C || ((p.rows * p.bpr) & 15) == 0);
}

// Persistent CTAs: tile = blockIdx.x, + gridDim.x, ...  The pixels of tile k+1 are prefetched with
// cp.async while tile k is transformed; the output staging image is double-buffered so that one
// barrier per tile suffices (the image written in iteration k is flushed before barrier k+1, and
// written again only after it).
template <int COMPS, int VEC> __host__ __device__ constexpr bool prefetchable() { return COMPS == 1 && VEC >= 1; }

// WL (warp-local coefficients, uniform per launch) is a template parameter so that the deferred flush and the row pass
// it hides behind are one basic block.
template <int LAYOUT, int MODE, bool X3, int COMPS, int VEC, bool WL>
SDCT_DEV void forward_tiles(const KParams &p, uint8_t *smem)
{
    constexpr bool PF = prefetchable<COMPS, VEC>();
    constexpr bool COMPL = complemented<LAYOUT, MODE>();
    uint8_t *stage_in = smem + 2 * STAGE_BYTES;
    TileWalk w;   // the tile after the current one (the one whose pixels are prefetched)
    w.init(p, 0, 1);
    if (w.tile >= p.tiles) return;
    Tile t = tile_at<COMPS, CB>(p, w);
    if (PF) {
        prefetch_pair<VEC>(stage_in, locate<LAYOUT>(p, t), p.pitch);
        cp_async_commit();
    }
    w.next(p);
    FlushJob pending = {nullptr, 0, 0};   // WL: the previous tile's image, flushed inside this tile's row pass
    for (uint32_t it = 0;; it++, w.next(p)) {
        const bool more = w.tile < p.tiles;
        Tile tn = t;
        if (more) tn = tile_at<COMPS, CB>(p, w);
        uint32_t wp[16], wq[16];
        if (PF) {
            if (more) prefetch_pair<VEC>(stage_in + ((it + 1) & 1) * IN_STAGE_BYTES, locate<LAYOUT>(p, tn), p.pitch);
            cp_async_commit();
            cp_async_wait<1>();
            read_prefetched(stage_in + (it & 1) * IN_STAGE_BYTES, wp, wq);
        } else {
            load_pair<COMPS, VEC>(locate<LAYOUT>(p, t), p.pitch, wp, wq);
        }
        if (WL) flush_coefs_warp<LAYOUT, COMPL>(smem + ((it + 1) & 1) * STAGE_BYTES, p, pending);
        F2 v[64];
        forward_pair<LAYOUT, MODE, X3>(wp, wq, p.tab, v);
        uint8_t *sm = smem + (it & 1) * STAGE_BYTES;
        stage_levels<LAYOUT>(sm, v);
        if (WL) {
            __syncwarp();
            pending.coef = t.coef; pending.n_first = (uint64_t)p.row0 * p.bpr + t.n0; pending.count = t.count;
            if (!more) { flush_coefs_warp<LAYOUT, COMPL>(sm, p, pending); break; }
        } else {
            __syncthreads();
            flush_coefs<LAYOUT, COMPL>(sm, p, t);
            if (!more) break;
        }
        t = tn;
    }
}



// profiling builds only (-DSDCT_TRACE_CTA=1): globaltimer at entry and exit of every warp of the first 1024 CTAs, and the SM
#ifndef SDCT_TRACE_CTA
#define SDCT_TRACE_CTA 0
#endif
__device__ unsigned long long g_warp_times[SDCT_TRACE_CTA ? 1024 * 4 * 3 : 1];
SDCT_DEV void warp_stamp(int which)
{
#if SDCT_TRACE_CTA
    if ((threadIdx.x & 31) == 0 && blockIdx.x < 1024) {
        unsigned long long t;
        unsigned sm;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
        unsigned long long *q = g_warp_times + (blockIdx.x * 4 + (threadIdx.x >> 5)) * 3;
        q[which] = t;
        q[2] = sm;
    }
#else
    (void)which;
#endif
}

template <int LAYOUT, int MODE, bool X3, int COMPS, int VEC>
__global__ void __launch_bounds__(SDCT_TB, 3) forward_kernel(const __grid_constant__ KParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    warp_stamp(0);
    if (warp_local_coefs<LAYOUT>(p)) forward_tiles<LAYOUT, MODE, X3, COMPS, VEC, true>(p, smem);
    else forward_tiles<LAYOUT, MODE, X3, COMPS, VEC, false>(p, smem);
    warp_stamp(1);
}

// ---- inverse ---------------------------------------------------------------------------------------
// global coefficient bytes of this CTA -> shared-memory image (same image as the forward staging)
template <int LAYOUT>
SDCT_DEV void fetch_coefs(uint8_t *sm, const KParams &p, const Tile &b)
{
    const uint32_t t = threadIdx.x;
    const uint64_t n_first = (uint64_t)p.row0 * p.bpr + b.n0;
    if (LAYOUT != SDCT_K_LAYOUT_C) {
        const uint8_t *src = b.coef + n_first * 64;
        const uint32_t chunks16 = b.count * 4;
        auto dst_off = [](uint32_t k) -> uint32_t {
            return LAYOUT == SDCT_K_LAYOUT_A ? (k >> 2) * A_SLOT + (k & 3) * 16 : (k >> 5) * B_GROUP_PITCH + (k & 31) * 16;
        };
        if (p.coef_aligned16) {
#pragma unroll 4
            for (uint32_t k = t; k < chunks16; k += TB)
                *reinterpret_cast<uint4 *>(sm + dst_off(k)) = __ldg(reinterpret_cast<const uint4 *>(src) + k);
        } else {
            for (uint32_t k = t; k < chunks16 * 16; k += TB) sm[dst_off(k >> 4) + (k & 15)] = __ldg(src + k);
        }
    } else {
        if (p.coef_aligned16 && (b.count & 15) == 0) {
            const uint32_t per_row = b.count >> 4;
            for (uint32_t k = t; k < 64 * per_row; k += TB) {
                const uint32_t i = k / per_row, c = k - i * per_row;
                const uint8_t *src = b.coef + (uint64_t)i * p.coef_plane_stride + n_first;
                *reinterpret_cast<uint4 *>(sm + i * C_PITCH + c * 16) = __ldg(reinterpret_cast<const uint4 *>(src) + c);
            }
        } else {
            const uint32_t warp = t >> 5, lane = t & 31;
            for (uint32_t i = warp; i < 64; i += TB / 32) {
                const uint8_t *src = b.coef + (uint64_t)i * p.coef_plane_stride + n_first;
                for (uint32_t k = lane; k < b.count; k += 32) sm[i * C_PITCH + k] = __ldg(src + k);
            }
        }
    }
}

// the warp's own 16-byte chunks of a tile's coefficient image (layouts A, B), asynchronously
template <int LAYOUT>
SDCT_DEV void prefetch_coefs_warp(uint8_t *sm, const KParams &p, const Tile &b)
{
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint64_t n_first = (uint64_t)p.row0 * p.bpr + b.n0;
    if (LAYOUT == SDCT_K_LAYOUT_C) {
#pragma unroll
        for (uint32_t j = 0; j < 8; j++) {
            const uint32_t k = 32u * j + lane, i = k >> 2, c = 64u * warp + 16u * (k & 3);
            if (c < b.count) cp_async16(sm + i * C_PITCH + c, b.coef + (uint64_t)i * p.coef_plane_stride + n_first + c);
        }
        return;
    }
    const uint8_t *src = b.coef + n_first * 64;
    const uint32_t chunks16 = b.count * 4;
#pragma unroll
    for (uint32_t j = 0; j < 8; j++) {
        uint32_t k, off;
        if (LAYOUT == SDCT_K_LAYOUT_B) {
            k = 256u * warp + 32u * j + lane;
            off = (k >> 5) * B_GROUP_PITCH + (k & 31) * 16;
        } else {
            k = (j >> 2) * 512u + 128u * warp + 32u * (j & 3) + lane;
            off = (k >> 2) * A_SLOT + (k & 3) * 16;
        }
        if (k < chunks16) cp_async16(sm + off, src + 16ull * k);
    }
}

// shared-memory image -> the pair's levels as floats, natural index b[u*8+v]
template <int LAYOUT>
SDCT_DEV void unstage_levels(const uint8_t *sm, F2 b[64])
{
    const uint32_t t = threadIdx.x;
    if (LAYOUT == SDCT_K_LAYOUT_A) {
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const uint4 *s = reinterpret_cast<const uint4 *>(sm + (t + k * TB) * A_SLOT);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint4 q = s[j];
                const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int e = 0; e < 16; e++) {
                    const int i = 16 * j + e, u = i & 7, v = i >> 3;
                    const float f = byte_to_float(w[e >> 2], e & 3);
                    if (k) b[u * 8 + v].y = f; else b[u * 8 + v].x = f;
                }
            }
        }
    } else {
        const uint16_t *s = LAYOUT == SDCT_K_LAYOUT_B
                                ? reinterpret_cast<const uint16_t *>(sm + (t >> 2) * B_GROUP_PITCH + 2 * (t & 3))
                                : reinterpret_cast<const uint16_t *>(sm + 2 * t);
        constexpr int step = LAYOUT == SDCT_K_LAYOUT_B ? 4 : C_PITCH / 2;
#pragma unroll
        for (int i = 0; i < 64; i++) {
            const uint32_t h = s[i * step];
            b[i] = make_float2(byte_to_float(h, 0), byte_to_float(h, 1));
        }
    }
}

// levels (as floats, natural index) -> pixel bit patterns, b[y*8+x]
template <int LAYOUT>
SDCT_DEV void inverse_pair(const float2 *tab, const float2 *off, F2 b[64])
{
#pragma unroll
    for (int i = 0; i < 64; i++) b[i] = fma2(b[i], tab[tab_index<LAYOUT>(i)], off[tab_index<LAYOUT>(i)]);
#pragma unroll
    for (int c = 0; c < 8; c++) idct8<8>(b + c);
#pragma unroll
    for (int r = 0; r < 8; r++) idct8<1>(b + 8 * r);
#pragma unroll
    for (int i = 0; i < 64; i++) b[i] = pixel_bits(b[i]);
}

template <int COMPS, int VEC>
SDCT_DEV void store_block(uint8_t *px, uint64_t pitch, uint32_t comp, const uint32_t w[16])
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

// pixel bit patterns of the pair -> global memory
template <int COMPS, int VEC>
SDCT_DEV void store_pair(const BlockPos &b, const uint8_t *in_base, uint8_t *out_base, uint64_t pitch, const F2 v[64])
{
    uint32_t wp[16], wq[16];
#pragma unroll
    for (int j = 0; j < 16; j++) {
        wp[j] = pack4_sat(v[4 * j].x, v[4 * j + 1].x, v[4 * j + 2].x, v[4 * j + 3].x);
        wq[j] = pack4_sat(v[4 * j].y, v[4 * j + 1].y, v[4 * j + 2].y, v[4 * j + 3].y);
    }
    uint8_t *dp = out_base + (b.px[0] - in_base), *dq = out_base + (b.px[1] - in_base);
    if (COMPS == 1 && VEC == 2) {
        if (b.valid[0]) {  // VEC == 2 implies an even count: P valid <=> Q valid
#pragma unroll
            for (int r = 0; r < 8; r++)
                *reinterpret_cast<uint4 *>(dp + r * pitch) = make_uint4(wp[2 * r], wp[2 * r + 1], wq[2 * r], wq[2 * r + 1]);
        }
    } else {
        if (b.valid[0]) store_block<COMPS, (VEC ? 1 : 0)>(dp, pitch, b.comp, wp);
        if (b.valid[1]) store_block<COMPS, (VEC ? 1 : 0)>(dq, pitch, b.comp, wq);
    }
}

template <int LAYOUT, int COMPS, int VEC>
__global__ void __launch_bounds__(SDCT_TB, 3) inverse_kernel(const __grid_constant__ KParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    TileWalk w;
    w.init(p, 0, 1);
    if (warp_local_coefs<LAYOUT>(p)) {
        // 16-byte-aligned coefficients: the 16-byte chunks a warp's blocks need are whole
        // chunks of the input that no other warp reads (see flush_coefs_warp), so each warp prefetches its own
        // part of the NEXT tile with cp.async while it transforms the current one - no CTA barrier
        if (w.tile >= p.tiles) return;
        Tile t = tile_at<COMPS, CB>(p, w);
        prefetch_coefs_warp<LAYOUT>(smem, p, t);
        cp_async_commit();
        w.next(p);
        for (uint32_t it = 0;; it++, w.next(p)) {
            const bool more = w.tile < p.tiles;
            Tile tn = t;
            if (more) tn = tile_at<COMPS, CB>(p, w);
            __syncwarp();   // every lane has read the image the next copies overwrite
            if (more) prefetch_coefs_warp<LAYOUT>(smem + ((it + 1) & 1) * STAGE_BYTES, p, tn);
            cp_async_commit();
            cp_async_wait<1>();
            __syncwarp();
            F2 v[64];
            unstage_levels<LAYOUT>(smem + (it & 1) * STAGE_BYTES, v);
            inverse_pair<LAYOUT>(p.tab, p.off, v);
            store_pair<COMPS, VEC>(locate<LAYOUT>(p, t), p.pixels, const_cast<uint8_t *>(p.pixels), p.pitch, v);
            if (!more) break;
            t = tn;
        }
        return;
    }
    for (uint32_t it = 0; w.tile < p.tiles; it++, w.next(p)) {
        const Tile t = tile_at<COMPS, CB>(p, w);
        uint8_t *sm = smem + (it & 1) * STAGE_BYTES;
        fetch_coefs<LAYOUT>(sm, p, t);
        __syncthreads();
        F2 v[64];
        unstage_levels<LAYOUT>(sm, v);
        inverse_pair<LAYOUT>(p.tab, p.off, v);
        store_pair<COMPS, VEC>(locate<LAYOUT>(p, t), p.pixels, const_cast<uint8_t *>(p.pixels), p.pitch, v);
    }
}

// fused forward -> quantise -> dequantise -> inverse: coefficients never leave the registers
template <int LAYOUT, int MODE, bool X3, int COMPS, int VEC>
__global__ void __launch_bounds__(SDCT_TB, 3) roundtrip_kernel(const __grid_constant__ KParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    constexpr bool PF = prefetchable<COMPS, VEC>();
    uint8_t *stage_in = smem;
    TileWalk w;   // the tile after the current one
    w.init(p, 0, 1);
    if (w.tile >= p.tiles) return;
    Tile t = tile_at<COMPS, CB>(p, w);
    if (PF) {
        prefetch_pair<VEC>(stage_in, locate<LAYOUT>(p, t), p.pitch);
        cp_async_commit();
    }
    w.next(p);
    for (uint32_t it = 0;; it++, w.next(p)) {
        const bool more = w.tile < p.tiles;
        Tile tn = t;
        if (more) tn = tile_at<COMPS, CB>(p, w);
        uint32_t wp[16], wq[16];
        if (PF) {
            if (more) prefetch_pair<VEC>(stage_in + ((it + 1) & 1) * IN_STAGE_BYTES, locate<LAYOUT>(p, tn), p.pitch);
            cp_async_commit();
            cp_async_wait<1>();
            read_prefetched(stage_in + (it & 1) * IN_STAGE_BYTES, wp, wq);
        } else {
            load_pair<COMPS, VEC>(locate<LAYOUT>(p, t), p.pitch, wp, wq);
        }
        F2 v[64];
        forward_pair<LAYOUT, MODE, X3>(wp, wq, p.tab, v);
        // bit patterns -> saturated levels -> floats (the level a decoder would read back)
#pragma unroll
        for (int i = 0; i < 64; i++) {
            uint32_t h = pack_pq(v[i]);
            if (complemented<LAYOUT, MODE>()) h = ~h;
            v[i] = make_float2(byte_to_float(h, 0), byte_to_float(h, 1));
        }
        inverse_pair<LAYOUT>(p.tab2, p.off, v);
        store_pair<COMPS, VEC>(locate<LAYOUT>(p, t), p.pixels, p.pixels_out, p.pitch, v);
        if (!more) break;
        t = tn;
    }
}


// ---- interleaved 3-component pixels on the OUTPUT side (inverse, round trip) ---------------------------------------
// A CTA decodes the three components of the same 256 blocks one after the other, stages their pixel rows planar in
// shared memory ([component][pixel row][block slot][8 bytes]) and then writes interleaved rows: a thread takes the 16
// pixels of two adjacent blocks of one row (three 16-byte reads), interleaves them with 24 PRMTs into 48 contiguous
// bytes and stores three 16-byte vectors, so that every output byte crosses HBM once, in full sectors.  Needs 16-byte
// aligned pixel rows and an even number of blocks per row (a slot pair never straddles a block-row wrap); other
// geometries keep the byte-store path of store_block.
constexpr int PIX_IMG = CB * 64;   // 16 KiB per component

template <int LAYOUT>
SDCT_DEV void stage_pixels(uint8_t *img, const F2 v[64])
{
    const uint32_t t = threadIdx.x;
#pragma unroll
    for (int r = 0; r < 8; r++) {
        const uint32_t p0 = pack4_sat(v[8 * r].x, v[8 * r + 1].x, v[8 * r + 2].x, v[8 * r + 3].x), p1 = pack4_sat(v[8 * r + 4].x, v[8 * r + 5].x, v[8 * r + 6].x, v[8 * r + 7].x);
        const uint32_t q0 = pack4_sat(v[8 * r].y, v[8 * r + 1].y, v[8 * r + 2].y, v[8 * r + 3].y), q1 = pack4_sat(v[8 * r + 4].y, v[8 * r + 5].y, v[8 * r + 6].y, v[8 * r + 7].y);
        if (LAYOUT == SDCT_K_LAYOUT_A) {   // slots t, t + 128
            *reinterpret_cast<uint2 *>(img + r * (CB * 8) + t * 8) = make_uint2(p0, p1);
            *reinterpret_cast<uint2 *>(img + r * (CB * 8) + (t + TB) * 8) = make_uint2(q0, q1);
        } else {                           // slots 2t, 2t + 1
            *reinterpret_cast<uint4 *>(img + r * (CB * 8) + t * 16) = make_uint4(p0, p1, q0, q1);
        }
    }
}

// planar images of the three components -> interleaved pixel rows of the tile's blocks in `out_frame`
SDCT_DEV void flush_rgb(const uint8_t *img3, const KParams &p, const Tile &t, uint8_t *out_frame)
{
    const uint32_t k = threadIdx.x, slot = 2 * k;
    if (slot >= t.count) return;
    const uint32_t nl = t.n0 + slot;
    const uint32_t r0 = fastdiv(nl, p.div_bpr), bx = nl - r0 * p.bpr, row = r0 + p.row0;
    const uint32_t band = p.eyes == 2 ? row >> 1 : row, eye = p.eyes == 2 ? row & 1 : 0;
    uint8_t *dst = out_frame + (uint64_t)eye * p.eye_stride + (uint64_t)band * 8 * p.pitch + (uint64_t)bx * 24;
#pragma unroll 2
    for (int r = 0; r < 8; r++, dst += p.pitch) {
        const uint4 R = *reinterpret_cast<const uint4 *>(img3 + r * (CB * 8) + k * 16);
        const uint4 G = *reinterpret_cast<const uint4 *>(img3 + PIX_IMG + r * (CB * 8) + k * 16);
        const uint4 B = *reinterpret_cast<const uint4 *>(img3 + 2 * PIX_IMG + r * (CB * 8) + k * 16);
        const uint32_t rr[4] = {R.x, R.y, R.z, R.w}, gg[4] = {G.x, G.y, G.z, G.w}, bb[4] = {B.x, B.y, B.z, B.w};
        uint32_t o[12];
#pragma unroll
        for (int q = 0; q < 4; q++) {   // four pixels (R, G, B words q) -> 12 bytes
            if (p.color) { ycc_words_to_rgb(rr[q], gg[q], bb[q], o + 3 * q); continue; }
            o[3 * q] = __byte_perm(__byte_perm(rr[q], gg[q], 0x1040), bb[q], 0x3410);       // R0 G0 B0 R1
            o[3 * q + 1] = __byte_perm(__byte_perm(gg[q], rr[q], 0x2601), bb[q], 0x3250);   // G1 B1 R2 G2
            o[3 * q + 2] = __byte_perm(__byte_perm(bb[q], rr[q], 0x3072), gg[q], 0x3710);   // B2 R3 G3 B3
        }
        uint4 *d = reinterpret_cast<uint4 *>(dst);
        d[0] = make_uint4(o[0], o[1], o[2], o[3]);
        d[1] = make_uint4(o[4], o[5], o[6], o[7]);
        d[2] = make_uint4(o[8], o[9], o[10], o[11]);
    }
}

// can the output side take the staged interleave?  (uniform per launch; the launcher decides)
static inline bool rgb_out_fusable(const KParams &p) { return p.comps == 3 && p.pixel_aligned16 && (p.bpr & 1) == 0; }

// p.tiles counts PIXEL tiles here (chunks * frames); the three components are an inner loop
template <int LAYOUT>
__global__ void __launch_bounds__(SDCT_TB, 2) inverse_rgb_kernel(const __grid_constant__ KParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t *pix = smem + 2 * STAGE_BYTES;
    TileWalk w;
    w.init(p, 0, 1);
    for (uint32_t it = 0; w.tile < p.tiles; w.next(p)) {
        const Tile t = tile_at<1, CB>(p, w);
#pragma unroll 1
        for (uint32_t c = 0; c < 3; c++, it++) {
            Tile tc = t;
            tc.coef += (uint64_t)c * p.coef_comp_stride;
            uint8_t *sm = smem + (it & 1) * STAGE_BYTES;
            fetch_coefs<LAYOUT>(sm, p, tc);
            __syncthreads();
            F2 v[64];
            unstage_levels<LAYOUT>(sm, v);
            inverse_pair<LAYOUT>(p.tab, p.off, v);
            stage_pixels<LAYOUT>(pix + c * PIX_IMG, v);
        }
        __syncthreads();
        flush_rgb(pix, p, t, const_cast<uint8_t *>(t.pix));
        __syncthreads();
    }
}

template <int LAYOUT, int MODE, bool X3>
__global__ void __launch_bounds__(SDCT_TB, 3) roundtrip_rgb_kernel(const __grid_constant__ KParams p)
{
    extern __shared__ __align__(16) uint8_t pix[];
    TileWalk w;
    w.init(p, 0, 1);
    for (; w.tile < p.tiles; w.next(p)) {
        const Tile t = tile_at<1, CB>(p, w);
        BlockPos b = locate<LAYOUT>(p, t);
#pragma unroll 1
        for (uint32_t c = 0; c < 3; c++) {
            b.comp = c;
            uint32_t wp[16], wq[16];
            load_pair<3, 1>(b, p.pitch, wp, wq);
            F2 v[64];
            forward_pair<LAYOUT, MODE, X3>(wp, wq, p.tab, v);
#pragma unroll
            for (int i = 0; i < 64; i++) {
                uint32_t h = pack_pq(v[i]);
                if (complemented<LAYOUT, MODE>()) h = ~h;
                v[i] = make_float2(byte_to_float(h, 0), byte_to_float(h, 1));
            }
            inverse_pair<LAYOUT>(p.tab2, p.off, v);
            stage_pixels<LAYOUT>(pix + c * PIX_IMG, v);
        }
        __syncthreads();
        flush_rgb(pix, p, t, p.pixels_out + (t.pix - p.pixels));
        __syncthreads();
    }
}

#if SDCT_PART == 0
// ---- self-test: the 8-point kernels on raw floats (bitwise oracle comparison in tests/) ----------
// data: n vectors of 8 floats, transformed in place.  kind: 0 scalar, 1 paired, 2 paired+X3,
// 3 fast, 4 idct, 5 = x/255 of element value (div255), 6 = x * (1.f/255), +8 = "second pass" variant (scalar first stage),
// 14 = the SSE form of the paired graph (X1 sign)
__global__ void dct8_selftest_kernel(float *data, uint32_t n, int kind)
{
    const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (i >= n) return;
    const uint32_t j = min(i + 1, n - 1);
    F2 v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = make_float2(data[i * 8 + k], data[j * 8 + k]);
    switch (kind) {
    case 0: dct8_scalar<1, false>(v); break;
    case 8: dct8_scalar<1, true>(v); break;
    case 1: dct8_paired<1, false, 0>(v); break;
    case 9: dct8_paired<1, true, 0>(v); break;
    case 2: dct8_paired<1, false, 1>(v); break;
    case 10: dct8_paired<1, true, 1>(v); break;
    case 14: dct8_paired<1, true, 2>(v); break;   // the SSE form (X1 sign), as both passes of the layout-C quirk use it
    case 6:                                        // x * (1.f / 255) (the SSE encoders' conversion)
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = mulx<PFP>(v[k], SDCT_R255_HI);
        break;
    case 3: dct8_fast<1>(v); break;
    case 4: idct8<1>(v); break;
    case 5:
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = div255(v[k]);
        break;
    default: break;
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        data[i * 8 + k] = v[k].x;
        if (j != i) data[j * 8 + k] = v[k].y;
    }
}

#endif  // SDCT_PART == 0

}  // namespace sdct

// ---- launchers ------------------------------------------------------------------------------------
using namespace sdct;

// cudaFuncSetAttribute belongs to the current device's context, not to the process: one flag bit per device
static inline bool attr_pending(std::atomic<uint64_t> &done, uint64_t *bit)
{
    int dev = 0;
    cudaGetDevice(&dev);
    *bit = 1ull << (dev & 63);
    return !(done.load(std::memory_order_acquire) & *bit);
}

typedef cudaError_t (*part_fn)(int op, int x3, const KParams &p, unsigned grid, cudaStream_t s);
#define SDCT_DECL(name) cudaError_t name(int op, int x3, const KParams &p, unsigned grid, cudaStream_t s)
SDCT_DECL(sdct_part_exact_a); SDCT_DECL(sdct_part_exact_b); SDCT_DECL(sdct_part_exact_c);
SDCT_DECL(sdct_part_fast_a); SDCT_DECL(sdct_part_fast_b); SDCT_DECL(sdct_part_fast_c);
SDCT_DECL(sdct_part_inverse);

template <int LAYOUT, int MODE, bool X3>
static cudaError_t launch_fwd2(int op, const KParams &p, unsigned grid, cudaStream_t s)
{
#define SDCT_GO(K, C, V)                                                                              \
    do {                                                                                              \
        const size_t smem = (op == SDCT_K_OP_FORWARD ? 2 * STAGE_BYTES : 0) +                         \
                            (prefetchable<C, V>() ? 2 * IN_STAGE_BYTES : 0);                          \
        static std::atomic<uint64_t> attr_done{0};                                                    \
        uint64_t bit;                                                                                 \
        if (attr_pending(attr_done, &bit)) {                                                          \
            cudaError_t e = cudaFuncSetAttribute(K<LAYOUT, MODE, X3, C, V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
            if (e != cudaSuccess) return e;                                                           \
            attr_done.fetch_or(bit, std::memory_order_release);                                       \
        }                                                                                             \
        K<LAYOUT, MODE, X3, C, V><<<grid, SDCT_TB, smem, s>>>(p);                                     \
        return cudaGetLastError();                                                                    \
    } while (0)
#define SDCT_PICK(K)                                                                                   \
    do {                                                                                               \
        if (p.comps == 1) {                                                                            \
            if (p.pixel_vec == 2 && LAYOUT != SDCT_K_LAYOUT_A) SDCT_GO(K, 1, 2);                      \
            else if (p.pixel_vec >= 1) SDCT_GO(K, 1, 1);                                               \
            else SDCT_GO(K, 1, 0);                                                                     \
        } else {                                                                                       \
            if (p.pixel_vec >= 1) SDCT_GO(K, 3, 1); else SDCT_GO(K, 3, 0);                             \
        }                                                                                              \
    } while (0)
    if (op == SDCT_K_OP_ROUNDTRIP && rgb_out_fusable(p)) {   // staged interleave on the output side; tiles = pixel tiles
        KParams q = p;
        q.tiles = p.tiles / 3;
        const unsigned resident3 = 3u * (p.sm_count ? p.sm_count : 148u);
        static std::atomic<uint64_t> attr_done{0};
        uint64_t bit;
        if (attr_pending(attr_done, &bit)) {
            cudaError_t e = cudaFuncSetAttribute(roundtrip_rgb_kernel<LAYOUT, MODE, X3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * PIX_IMG);
            if (e != cudaSuccess) return e;
            attr_done.fetch_or(bit, std::memory_order_release);
        }
        roundtrip_rgb_kernel<LAYOUT, MODE, X3><<<(q.tiles < resident3 ? q.tiles : resident3), SDCT_TB, 3 * PIX_IMG, s>>>(q);
        return cudaGetLastError();
    }
    if (op == SDCT_K_OP_FORWARD) SDCT_PICK(forward_kernel);
    else SDCT_PICK(roundtrip_kernel);
#undef SDCT_PICK
#undef SDCT_GO
}

template <int LAYOUT>
static cudaError_t launch_inv2(const KParams &p, unsigned grid, cudaStream_t s)
{
    const size_t smem = 2 * STAGE_BYTES;
#define SDCT_INV(C, V)                                                                                                   \
    do {                                                                                                                 \
        static std::atomic<uint64_t> attr_done{0};                                                                       \
        uint64_t bit;                                                                                                    \
        if (attr_pending(attr_done, &bit)) {                                                                             \
            cudaError_t e = cudaFuncSetAttribute(inverse_kernel<LAYOUT, C, V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
            if (e != cudaSuccess) return e;                                                                              \
            attr_done.fetch_or(bit, std::memory_order_release);                                                          \
        }                                                                                                                \
        inverse_kernel<LAYOUT, C, V><<<grid, SDCT_TB, smem, s>>>(p);                                                     \
        return cudaGetLastError();                                                                                       \
    } while (0)
    if (rgb_out_fusable(p)) {   // staged interleave on the output side; tiles = pixel tiles
        KParams q = p;
        q.tiles = p.tiles / 3;
        const unsigned resident2 = 2u * (p.sm_count ? p.sm_count : 148u);
        const size_t smem_rgb = 2 * STAGE_BYTES + 3 * PIX_IMG;
        static std::atomic<uint64_t> attr_done{0};
        uint64_t bit;
        if (attr_pending(attr_done, &bit)) {
            cudaError_t e = cudaFuncSetAttribute(inverse_rgb_kernel<LAYOUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_rgb);
            if (e != cudaSuccess) return e;
            attr_done.fetch_or(bit, std::memory_order_release);
        }
        inverse_rgb_kernel<LAYOUT><<<(q.tiles < resident2 ? q.tiles : resident2), SDCT_TB, smem_rgb, s>>>(q);
        return cudaGetLastError();
    }
    if (p.comps == 1) {
        if (p.pixel_vec == 2 && LAYOUT != SDCT_K_LAYOUT_A) SDCT_INV(1, 2);
        else if (p.pixel_vec >= 1) SDCT_INV(1, 1);
        else SDCT_INV(1, 0);
    } else {
        SDCT_INV(3, 0);
    }
#undef SDCT_INV
}

#if SDCT_PART == 1
SDCT_DECL(sdct_part_exact_a) { (void)x3; return launch_fwd2<SDCT_K_LAYOUT_A, SDCT_K_EXACT, false>(op, p, grid, s); }
#elif SDCT_PART == 2
#if SDCT_TRACE_CTA
extern "C" int sdct_k_warp_times(unsigned long long *host) { return (int)cudaMemcpyFromSymbol(host, g_warp_times, sizeof(g_warp_times)); }
#endif
SDCT_DECL(sdct_part_exact_b)
{
    return x3 ? launch_fwd2<SDCT_K_LAYOUT_B, SDCT_K_EXACT, true>(op, p, grid, s)
              : launch_fwd2<SDCT_K_LAYOUT_B, SDCT_K_EXACT, false>(op, p, grid, s);
}
#elif SDCT_PART == 3
SDCT_DECL(sdct_part_exact_c)   // x3 here: the layout's reference-SIMD quirk = the SSE stereo encoders' arithmetic (SDCT_QUIRK_SSE)
{
    return x3 ? launch_fwd2<SDCT_K_LAYOUT_C, SDCT_K_EXACT, true>(op, p, grid, s)
              : launch_fwd2<SDCT_K_LAYOUT_C, SDCT_K_EXACT, false>(op, p, grid, s);
}
#elif SDCT_PART == 4
SDCT_DECL(sdct_part_fast_a) { (void)x3; return launch_fwd2<SDCT_K_LAYOUT_A, SDCT_K_FAST, false>(op, p, grid, s); }
#elif SDCT_PART == 5
SDCT_DECL(sdct_part_fast_b) { (void)x3; return launch_fwd2<SDCT_K_LAYOUT_B, SDCT_K_FAST, false>(op, p, grid, s); }
#elif SDCT_PART == 6
SDCT_DECL(sdct_part_fast_c) { (void)x3; return launch_fwd2<SDCT_K_LAYOUT_C, SDCT_K_FAST, false>(op, p, grid, s); }
#elif SDCT_PART == 7
SDCT_DECL(sdct_part_inverse)
{
    (void)op; (void)x3;
    switch (p.layout) {
    case SDCT_K_LAYOUT_A: return launch_inv2<SDCT_K_LAYOUT_A>(p, grid, s);
    case SDCT_K_LAYOUT_B: return launch_inv2<SDCT_K_LAYOUT_B>(p, grid, s);
    default: return launch_inv2<SDCT_K_LAYOUT_C>(p, grid, s);
    }
}
#elif SDCT_PART == 0
extern "C" cudaError_t sdct_k_launch(int op, int layout, int mode, int x3, const KParams *pp, cudaStream_t s)
{
    KParams p = *pp;
    p.layout = (uint32_t)layout;
    const unsigned long long tiles = (unsigned long long)p.chunks * p.comps * p.frames;
    if (tiles == 0) return cudaSuccess;
    if (tiles > 0xffff0000ull) return cudaErrorInvalidConfiguration;
    p.tiles = (uint32_t)tiles;
    p.div_chunks = make_fastdiv(p.chunks);
    p.div_bpr = make_fastdiv(p.bpr);
    // persistent CTAs: three per SM (168 registers/thread), each walking tiles blockIdx.x, +grid, ...
    const unsigned resident = 3u * (p.sm_count ? p.sm_count : 148u);
    const unsigned grid = (unsigned)(tiles < resident ? tiles : resident);
    if (op == SDCT_K_OP_INVERSE) return sdct_part_inverse(op, x3, p, grid, s);
    static const part_fn exact[3] = {sdct_part_exact_a, sdct_part_exact_b, sdct_part_exact_c};
    static const part_fn fast[3] = {sdct_part_fast_a, sdct_part_fast_b, sdct_part_fast_c};
    if (layout < 0 || layout > 2) return cudaErrorInvalidValue;
    return (mode == SDCT_K_FAST ? fast : exact)[layout](op, x3, p, grid, s);
}

extern "C" cudaError_t sdct_k_selftest(float *d_data, uint32_t n, int kind, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    const unsigned threads = 128, pairs = (n + 1) / 2;
    dct8_selftest_kernel<<<(pairs + threads - 1) / threads, threads, 0, s>>>(d_data, n, kind);
    return cudaGetLastError();
}
#endif
