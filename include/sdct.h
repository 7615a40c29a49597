/*
 * sdct.h - C ABI of the B200 (sm_100a) 8x8 block-DCT library (libsdct.so).
 *
 * This is the drop-in boundary behind the reference's public header: plain pointers and
 * sizes, no C++ or torch types.  Each entry point cites the reference interface it replaces
 * (paths are into rainerzufalldererste/simd_dct).  The three C++-linkage functions of
 * src/simd_dct.h:29-31 are exported by the same library (simd_dct_b200/csrc/simd_dct_shim.cpp,
 * declared in include/simd_dct.h) and forward to the sdct_ref_* functions below.
 *
 * There is no CPU fallback: every compute entry point returns SDCT_NO_DEVICE / SDCT_CUDA_ERROR
 * (and sets sdct_last_error()) when no usable sm_100 device is present.
 */
#ifndef SDCT_H
#define SDCT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SDCT_ABI_VERSION 1

/* status: values 0..2 are the reference's simdDctResult (src/simd_dct.h:22-27) */
enum {
    SDCT_OK = 0,
    SDCT_INVALID_PARAMETER = 1,
    SDCT_NOT_SUPPORTED = 2,
    SDCT_CUDA_ERROR = 3,
    SDCT_NO_DEVICE = 4
};

/* coefficient layouts */
enum {
    SDCT_LAYOUT_A = 0, /* block-contiguous, 64 B per block, byte [v*8+u]  - simdDCT_EncodeQuantizeBuffer (src/simd_dct.cpp:300-395) */
    SDCT_LAYOUT_B = 1, /* 8-block groups, 512 B, byte [i*8+blk]           - simdDCT_EncodeQuantize32ReorderBuffer (:2064-2262) */
    SDCT_LAYOUT_C = 2  /* 64 coefficient planes, eyes interleaved per band - simdDCT_EncodeQuantizeReorderStereoBuffer (:177-298) */
};

/* flags */
enum {
    SDCT_EXACT = 0,         /* the reference's fp32 operation order; byte-identical to its CPU output */
    SDCT_FAST = 1,          /* FMA-contracted arithmetic defined in oracle/sdct_oracle.c (<= 1 LSB from SDCT_EXACT) */
    SDCT_QUIRK_AVX_X3 = 2,  /* layout B: reproduce the reference's X3 sign (src/simd_dct.cpp:2181); see DESIGN.md */
    SDCT_QUIRK_SSE = 16,    /* layout C: the bytes of the reference's SSE4.1 / SSSE3 / SSE2 stereo encoders (src/simd_dct.cpp:896-1103,
                               what its dispatcher runs on every x86-64 CPU, :78-83) instead of its scalar path's: x * (1.f / 255),
                               inplace_dct8_sse41 with its X1 sign (:788), cvtps_epi32(c * q + 127.f).  Forward and round trip, width % 16 == 0 */
    SDCT_COLOR_YCBCR = 8,   /* interleaved 3-component pixels: RGB -> YCbCr (JFIF / BT.601 full range, 8-bit fixed point, exact integer
                               arithmetic defined in oracle/sdct_oracle.c) fused into the forward load, YCbCr -> RGB into the inverse
                               store; the coefficient sets of a frame are then Y, Cb, Cr.  Combines with every arithmetic mode.
                               The inverse side needs 16-byte-aligned pixel rows and an even number of blocks per row. */
    SDCT_TENSOR = 4         /* tensor-core (tcgen05) integer form of the true DCT, defined in oracle/sdct_oracle.c mode TENSOR.
                               Forward: every geometry.  Inverse: one or three components, 16-byte-aligned pixel rows and
                               coefficients, an even number of blocks per row (SDCT_NOT_SUPPORTED otherwise).  Round trip: one fused kernel
                               for single planes with rows of >= 128 blocks, the two kernels back to back otherwise (same bytes).  Excludes SDCT_FAST and SDCT_QUIRK_AVX_X3.
                               Distance from SDCT_EXACT (measured on every workload of tests/test_gpu_mismatch_rates.py, recorded in
                               profiles/r02_mismatch_rates.json): SDCT_FAST and SDCT_TENSOR levels differ by at most 1 in at most
                               3e-4 of the bytes (asserted bound 1e-3); the SDCT_TENSOR inverse differs from the fp32 inverse by
                               at most 1 in at most 8e-4 of the pixels, and both are within 1 of a float64 inverse in at most
                               7e-4 (asserted bound 2e-3). */
};

typedef struct sdct_context sdct_context;

/*
 * One batched job over device-resident buffers.
 *
 * Pixel side:  frame f, eye e, component k, pixel (y, x) lives at
 *     pixels[f*pixel_frame_stride + e*eye_stride + y*pitch + x*components + k]
 * Coefficient side: the coefficient set of (frame f, component k) starts at
 *     coefs[f*coef_frame_stride + k*coef_comp_stride]
 * and has width*height*eyes bytes in the chosen layout (layout C: plane i at + i*coef_plane_stride).
 * Only bands [band_begin, band_end) (a band = 8 pixel rows) are processed; band_end = 0 means
 * height/8.  Outputs of unprocessed bands are left untouched, as in the reference's
 * startY/endY window (src/simd_dct.cpp:268-282).
 */
typedef struct sdct_job {
    int32_t layout;
    int32_t flags;
    int32_t components;        /* 1 = planar plane, 3 = interleaved 3-component pixels */
    int32_t eyes;              /* layout C: 1 or 2; other layouts: 1 */
    uint64_t width, height;    /* one plane of one eye, in pixels; width%8==0 (B: %64), height%8==0 */
    uint64_t pitch;            /* bytes between pixel rows */
    uint64_t eye_stride;       /* bytes between the eyes' pixel planes (layout C, eyes == 2) */
    uint64_t frames;
    uint64_t pixel_frame_stride;
    uint64_t coef_frame_stride;
    uint64_t coef_comp_stride;
    uint64_t coef_plane_stride;
    uint64_t band_begin, band_end;
} sdct_job;

/* fills strides for a dense batch: frames x [eyes x] height x width x components pixels */
void sdct_job_dense(sdct_job *job, int layout, int flags, int components, int eyes,
                    uint64_t width, uint64_t height, uint64_t frames);

/* context = one device + its streams, staging buffers and launch counters
 * (replaces the caller-side _DetectCPUFeatures() contract, src/simd_platform.h:131).  The $SDCT_* environment
 * switches (SDCT_HOST_CHUNK_MB, SDCT_HOST_SLOTS, SDCT_HOST_NO_RAMP, SDCT_HOST_ENGINE_STREAMS, SDCT_TC_NO_TMA, ...) are read here, once. */
int sdct_create(int device, sdct_context **out);
void sdct_destroy(sdct_context *ctx);
const char *sdct_last_error(void);
int sdct_abi_version(void);
int sdct_device_count(void);
int sdct_sm_count(const sdct_context *ctx);
/* number of kernel launches issued through this context so far */
uint64_t sdct_kernel_launches(const sdct_context *ctx);

/* device-resident operations; `stream` is a cudaStream_t (NULL = the context's own non-blocking stream;
 * pass cudaStreamLegacy, i.e. (void *)1, for the legacy default stream).
 * lut = the caller's 64-entry quantisation table in host memory (pQuantizeLUT of
 * src/simd_dct.h:29-31); the scale tables are derived from it exactly as
 * src/simd_dct.cpp:194-212 / :2236-2239 do. */
int sdct_forward_device(sdct_context *ctx, const sdct_job *job, const void *d_pixels, const float *lut,
                        void *d_coefs, void *stream);
int sdct_inverse_device(sdct_context *ctx, const sdct_job *job, const void *d_coefs, const float *lut,
                        void *d_pixels, void *stream);
/* fused forward -> quantise -> dequantise -> inverse; coefficients never leave the SM */
int sdct_roundtrip_device(sdct_context *ctx, const sdct_job *job, const void *d_pixels_in, const float *lut,
                          void *d_pixels_out, void *stream);

/* The same round trip as TWO kernels (forward -> quantised levels in scratch memory -> inverse) for callers that cannot allocate
 * device memory at call time (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
 * reference.  The levels live in a file-scope __device__ array of
 * SDCT_NOALLOC_MAX_COEF_BYTES (32 MiB, plus 16 KiB for the SDCT_TENSOR inverse operand) that every GPU which loads the library
 * holds until process exit; the call path contains no cudaMalloc / cudaFree / memory-pool routine.  Same arguments, same flags
 * and the same output bytes as sdct_roundtrip_device.  Jobs with frames * width * height * eyes * components above the maximum
 * are rejected with SDCT_NOT_SUPPORTED (sdct_last_error() names both sizes).  Every call on a GPU shares the array (it keeps
 * the previous call's contents; nothing in it needs clearing, the forward kernel writes every byte the inverse reads): calls
 * whose kernels overlap on the device - different streams, threads or contexts - race and must be serialised by the caller. */
#define SDCT_NOALLOC_MAX_COEF_BYTES (32u << 20)
int sdct_roundtrip_noalloc_device(sdct_context *ctx, const sdct_job *job, const void *d_pixels_in, const float *lut,
                                  void *d_pixels_out, void *stream);

/* The step after the path (layout C only: each of its 64 planes is one homogeneous stream, src/simd_dct.cpp:1061-1069):
 * per coefficient plane of every (frame, component) set, over the job's band window, how many levels are zero (byte 127,
 * the quantiser's bias), in how many runs they come (in block order), the sum of |level - 127| and the bitwise OR of all |level - 127| (its
 * top set bit = the magnitude bit width) - what an entropy coder needs to choose a code per plane.  d_stats: frames * components * 64 records, device memory. */
typedef struct sdct_plane_stat {
    uint64_t zeros, zero_runs, sum_abs, or_abs;
} sdct_plane_stat;
int sdct_plane_stats_device(sdct_context *ctx, const sdct_job *job, const void *d_coefs, sdct_plane_stat *d_stats, void *stream);

/* host-buffer operations: chunked H2D -> kernel -> D2H pipeline ($SDCT_HOST_SLOTS chunks in flight, default 4, each a
 * dense 16-byte-aligned sub-job on the device); pinned caller memory is DMA'd directly (one stream per engine: uploads,
 * kernels and downloads each in their own queue, ordered by events; $SDCT_HOST_ENGINE_STREAMS=0 selects one stream per
 * slot), pageable memory goes through pinned staging.  Only the bytes of the job's window are written - row padding and unprocessed bands stay untouched.
 * Thread-safe: concurrent calls on one context (e.g. disjoint band windows of one image, the reference's threading
 * model, src/simd_dct.cpp:1073-1087) run on separate pipelines and overlap. */
int sdct_forward_host(sdct_context *ctx, const sdct_job *job, const void *h_pixels, const float *lut,
                      void *h_coefs);
int sdct_inverse_host(sdct_context *ctx, const sdct_job *job, const void *h_coefs, const float *lut,
                      void *h_pixels);
/* the same over several devices in ONE call: contiguous frame ranges (frames >= n_ctx) or band ranges of the job's
 * window go to the contexts' devices, one host thread and pipeline each; no inter-device exchange */
int sdct_forward_host_multi(sdct_context *const *ctxs, int n_ctx, const sdct_job *job, const void *h_pixels,
                            const float *lut, void *h_coefs);
int sdct_inverse_host_multi(sdct_context *const *ctxs, int n_ctx, const sdct_job *job, const void *h_coefs,
                            const float *lut, void *h_pixels);

/* test hook: runs one of the device 8-point kernels on n vectors of 8 floats (host memory, in place)
 * so that tests can compare the GPU arithmetic bit for bit with the CPU restatement.
 * kind: 0 forward/scalar order (src/simd_dct.cpp:148-171), 1 forward/AVX2 order (:2160-2183),
 * 2 = 1 with the reference's X3 sign (:2181), 3 forward/FAST, 4 inverse, 5 element-wise x/255.f;
 * kinds 8, 9, 10 are 0, 1, 2 in their second-pass form (first butterfly stage issued as scalar adds) */
int sdct_selftest_dct8(sdct_context *ctx, int kind, float *h_data, size_t n);

/* test hook (no device needed): the integer matrix K[c*64+p] = round(2^22 C[u][y] C[v][x]) of the
 * SDCT_TENSOR form, as the library builds it; tests compare it with the oracle's restatement */
void sdct_tensor_matrix(int32_t *k4096);

/* synchronise the context's streams */
int sdct_synchronize(sdct_context *ctx);

/* pinned host memory helpers for callers that want zero-copy staging */
void *sdct_host_alloc(size_t bytes);
void sdct_host_free(void *p);

/*
 * The reference's call semantics on HOST buffers, used by the simd_dct.h shim:
 * sizeY is the stereo (double) height, only rows below sizeY/2 are visited, [startY, endY] is
 * the inclusive window of src/simd_dct.cpp:268-282 / :375-387 / :2243-2255, return value is
 * the reference's simdDctResult.  They run on lazily created process-wide contexts: device $SDCT_DEVICE (default 0),
 * or - $SDCT_DEVICES = "all" | "0,1,..." - the call's bands sharded over several devices (sdct_forward_host_multi).
 * Calls from several threads overlap (see sdct_forward_host).
 */
int sdct_ref_encode_quantize_buffer(const uint8_t *from, uint8_t *to, const float *lut,
                                    size_t size_x, size_t size_y, size_t start_y, size_t end_y);        /* src/simd_dct.h:29 */
int sdct_ref_encode_quantize_reorder_stereo_buffer(const uint8_t *from, uint8_t *to, const float *lut,
                                                   size_t size_x, size_t size_y, size_t start_y, size_t end_y); /* :30 */
int sdct_ref_encode_quantize32_reorder_buffer(const uint8_t *from, uint8_t *to, const float *lut,
                                              size_t size_x, size_t size_y, size_t start_y, size_t end_y);      /* :31 */
/* flags applied by the three functions above (default SDCT_EXACT | SDCT_QUIRK_AVX_X3 = reference bytes) */
void sdct_ref_set_flags(int flags);
int sdct_ref_get_flags(void);
/* number of devices the three functions above shard a call over (creates the process-wide contexts; 0 on failure) */
int sdct_ref_device_count(void);

/* contiguous shard [begin, end) of `total` units for `rank` of `world` (frames or bands; no collective) */
void sdct_shard_range(uint64_t total, int rank, int world, uint64_t *begin, uint64_t *end);

#ifdef __cplusplus
}
#endif
#endif /* SDCT_H */
