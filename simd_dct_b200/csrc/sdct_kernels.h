// sdct_kernels.h - launch interface between the C-ABI host layer (sdct_api.cpp) and the
// sm_100a kernels (sdct_kernels.cu).  Internal; the public surface is include/sdct.h.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#define SDCT_TB 128          // threads per CTA
#define SDCT_CTA_BLOCKS 256  // 8x8 blocks per CTA (two per thread)
#define SDCT_TC_TILE 128     // tensor-core kernels: 8x8 blocks per tile (one per TMEM lane)

#define SDCT_K_LAYOUT_A 0
#define SDCT_K_LAYOUT_B 1
#define SDCT_K_LAYOUT_C 2
#define SDCT_K_EXACT 0
#define SDCT_K_FAST 1
#define SDCT_K_TENSOR 2
#define SDCT_K_OP_FORWARD 0
#define SDCT_K_OP_INVERSE 1
#define SDCT_K_OP_ROUNDTRIP 2

struct KFastDiv { uint32_t m, sh1, sh2; };   // n / d = (t + ((n - t) >> sh1)) >> sh2, t = umulhi(n, m)

struct KParams {
    const uint8_t *pixels;    // forward/roundtrip: input; inverse: output
    uint8_t *coefs;           // forward: output; inverse: input (unused by roundtrip)
    uint8_t *pixels_out;      // roundtrip output
    uint64_t pitch, eye_stride, pixel_frame_stride;
    uint64_t coef_frame_stride, coef_comp_stride, coef_plane_stride;
    uint32_t bpr;             // 8x8 blocks per pixel row
    uint32_t eyes;
    uint32_t row0, rows;      // block-row window: row = band*eyes + eye
    uint32_t comps;
    uint32_t frames;
    uint32_t layout;          // SDCT_K_LAYOUT_* (filled in by sdct_k_launch)
    uint32_t chunks;          // tiles per (frame, component)
    uint32_t tiles;           // chunks * comps * frames (filled in by sdct_k_launch)
    uint32_t sm_count;        // SMs of the device (persistent grid size)
    uint32_t pixel_vec;       // 0 = byte access, 1 = 8-byte rows, 2 = 16-byte rows for an adjacent block pair
    uint32_t coef_aligned16, coef_aligned4;
    uint32_t color;           // interleaved pixels: 1 = RGB <-> YCbCr fused into the load / the store (SDCT_COLOR_YCBCR)
    uint32_t pixel_aligned16; // pixel base pointers and strides are multiples of 16 (whatever the component count)
    // scale tables, each value duplicated for the two packed lanes (P, Q)
    float2 tab[64];           // forward: quantiser scale; inverse: dequantiser scale
    float2 tab2[64];          // roundtrip: dequantiser scale
    float2 off[64];           // inverse / roundtrip: -127 * dequantiser scale
    KFastDiv div_chunks, div_bpr;  // tensor-core kernels
    uint32_t tc_ablate;       // profiling only ($SDCT_TC_ABLATE): bit0 no MMA, 1 no A write, 2 no epilogue, 3 no stores, 4 no loads
    const void *tc_matrix;    // tensor-core kernels: the 16 KiB B operand (fp16 digits of 2^22 C(x)C, UMMA canonical layout)
    const void *tc_matrix2;   // fused tensor-core round trip: the dequantising inverse operand E (tc_matrix = the forward operand)
    // tensor-core kernels, TMA pixel path: CUtensorMap over the pixels as u32[frame][eye][row][W / 4]
    // (filled in by sdct_k_launch_tc; must stay 64-byte aligned inside the kernel parameter block)
    alignas(64) unsigned char tc_tensor_map[128];
    // layout C coefficients as u8[frame][component][plane 0..63][block], box 128 blocks x 64 planes, 128-byte swizzle:
    // the tensor-core inverse loads its tiles through it, the tensor-core forward stores them
    alignas(64) unsigned char tc_coef_map[128];
};

extern "C" cudaError_t sdct_k_launch(int op, int layout, int mode, int x3, const KParams *p, cudaStream_t s);

// switches of the tensor-core launcher (the context reads $SDCT_TC_* once, in sdct_create)
struct SdctTcOptions {
    bool no_tma = false;         // $SDCT_TC_NO_TMA: thread-copy paths instead of tensor maps
    bool no_rgb_fuse = false;    // $SDCT_TC_NO_RGB_FUSE
    bool no_tma_store = false;   // $SDCT_TC_NO_TMA_STORE
    bool no_rt_fuse = false;     // $SDCT_TC_NO_RT_FUSE: SDCT_TENSOR round trip always as two kernels
    uint32_t ablate = 0;         // $SDCT_TC_ABLATE (profiling builds only)
};
// encoded CUtensorMaps, keyed by (base pointer, dimensions, strides): cuTensorMapEncodeTiled runs once per
// distinct buffer geometry instead of once or twice per launch
struct SdctTcMapCache {
    struct Entry { const void *base; uint64_t dims[4], strides[3]; uint32_t kind; bool ok; uint64_t used; unsigned char map[128]; };
    static constexpr int N = 16;
    Entry e[N];
    int n = 0;
    uint64_t clock = 0, hits = 0, misses = 0;
};
// tensor-core form (sdct_tc.cu); p->tab holds the folded scales in D-column order.  `maps` may be null (no caching);
// the caller serialises access to it.
extern "C" cudaError_t sdct_k_launch_tc(int op, int layout, const KParams *p, const SdctTcOptions *opt, SdctTcMapCache *maps, cudaStream_t s);
// fused tensor-core round trip (single planes with the TMA pixel path): p->tab = forward scales in POSITIONAL column order,
// p->off = the inverse's pixel offsets, p->tab2[0] = its 2^-S, p->tc_matrix / tc_matrix2 = forward (positional order) /
// inverse operand.  cudaErrorNotSupported when the geometry has no TMA path (the caller then runs the two kernels).
extern "C" cudaError_t sdct_k_launch_tc_roundtrip(int layout, const KParams *p, const SdctTcOptions *opt, SdctTcMapCache *maps, cudaStream_t s);
// D column j of the tensor-core kernels -> coefficient position in the layout's own index order
static inline int sdct_tc_column_position(int layout, int j)
{
    return layout == SDCT_K_LAYOUT_A ? 16 * ((j & 7) >> 1) + 2 * (j >> 3) + (j & 1) : j;
}
// tensor-core inverse: D column j -> pixel y*8+x of the block, two forms of the pixel epilogue:
//   rows = true  (the fused round trip): column j IS pixel j.  A consumer thread reads its own TMEM lane
//     (tcgen05.ld.32x32b: thread = block), i.e. whole pixel rows of one block; a warp's store of a row covers 256 contiguous bytes.
//   rows = false (the inverse kernel): a thread receives columns 8k + 2q + e of two lanes (16x256b loads); chunk k / 4 is then
//     the pixel row 2q + k/4 and (k % 4, e) its eight pixels.
// (measured both ways on both kernels, 64 x 1080p: inverse 58 us chunk form / 61-63 us row form, RGB inverse 205 / 245 us;
//  fused round trip 94 us chunk form / 83 us row form)
static inline int sdct_tc_inverse_column_pixel(int j, bool rows = false)
{
    if (rows) return j;
    const int k = j >> 3, q = (j & 7) >> 1, e = j & 1;
    return (2 * q + (k >> 2)) * 8 + 2 * (k & 3) + e;
}
// per-plane statistics of layout-C coefficient sets (sdct_stats.cu): d_stats = frames * comps * 64 records of four
// 64-bit counters {zero levels, zero runs, sum |level - 127|, OR |level - 127|} over the job's band window
extern "C" cudaError_t sdct_k_plane_stats(const KParams *p, void *d_stats, cudaStream_t s);
// the 8-point kernels on n raw 8-float vectors (device memory, in place); see dct8_selftest_kernel
extern "C" cudaError_t sdct_k_selftest(float *d_data, uint32_t n, int kind, cudaStream_t s);
