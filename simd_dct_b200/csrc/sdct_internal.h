// sdct_internal.h - shared between sdct_api.cpp and sdct_host.cpp (not installed).
#pragma once
#include "../../include/sdct.h"
#include "sdct_kernels.h"

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <memory>
#include <mutex>
#include <vector>

int sdct_fail(int code, const char *fmt, ...);
#define fail sdct_fail

#define CU_TRY(expr)                                                                              \
    do {                                                                                          \
        cudaError_t e_ = (expr);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail(SDCT_CUDA_ERROR, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_),  \
                        __FILE__, __LINE__);                                                      \
    } while (0)

// makes `device` current for a scope and restores the caller's current device afterwards: the library never changes
// the calling thread's current device behind its back (a caller that drives several contexts keeps its own)
struct DeviceScope {
    int prev = -1, dev;
    cudaError_t err;
    explicit DeviceScope(int device) : dev(device)
    {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
    }
    ~DeviceScope()
    {
        if (prev >= 0 && prev != dev) cudaSetDevice(prev);
    }
    DeviceScope(const DeviceScope &) = delete;
    DeviceScope &operator=(const DeviceScope &) = delete;
};

constexpr int kSlots = 8;  // slots of one host pipeline; the depth in use is SdctEnv::host_slots <= kSlots

// one in-flight chunk of the host path: its stream, the device buffers of the chunk's dense sub-job and the
// pinned staging buffers used when the caller's memory is pageable
struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;                    // the slot's chunk has left the device (its download is complete)
    cudaEvent_t ev_in = nullptr, ev_k = nullptr;   // engine-stream mode: upload complete, kernel complete
    uint8_t *d_pix = nullptr, *d_coef = nullptr;
    size_t d_pix_cap = 0, d_coef_cap = 0;
    uint8_t *h_in = nullptr, *h_out = nullptr;
    size_t h_in_cap = 0, h_out_cap = 0;
};

// A host-buffer call owns one Pipeline from the context's pool for its duration, so that N caller threads
// (the reference's concurrency model: disjoint startY/endY windows, src/simd_dct.cpp:1073-1087) overlap.
struct Pipeline {
    Slot slots[kSlots];
    bool busy = false;
};

// $SDCT_* switches, read ONCE per context in sdct_create (no getenv on any launch path)
struct SdctEnv {
    uint64_t host_chunk_bytes = 16ull << 20;
    int host_slots = 4;
    bool host_no_ramp = false;
    bool host_engine_streams = true;    // $SDCT_HOST_ENGINE_STREAMS=0: one stream per slot instead of one per engine (sdct_host.cpp)
    bool host_trace = false;   // $SDCT_HOST_TRACE: per-chunk timeline of a host call on stderr (profiling)
    SdctTcOptions tc;
};

struct sdct_context {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    SdctEnv env;
    std::atomic<uint64_t> launches{0};
    std::mutex pipes_mu;   // guards the pool only; a call runs on its own pipeline without the lock
    std::vector<std::unique_ptr<Pipeline>> pipes;
    void *d_tc_matrix[3] = {nullptr, nullptr, nullptr};  // tensor-core B operand: [0] layout-A column order, [1] natural order, [2] layout A positional
    // tensor-core inverse operands depend on the LUT: a small cache keyed by (layout, LUT).  An entry is uploaded
    // synchronously before its key becomes visible and is only overwritten after a device synchronise.
    struct InvOperand { void *d = nullptr; int layout = -1; bool rows = false; float lut[64]; float off_cols[64]; int S = 0; uint64_t used = 0; };
    InvOperand inv_cache[8];
    uint64_t inv_clock = 0;
    std::mutex inv_mu;
    // stream-ordered scratch memory (SDCT_TENSOR round trip: the coefficients between the two kernels): an own pool that
    // keeps its memory (the default pool hands freed memory back to the driver at every synchronisation point)
    cudaMemPool_t scratch_pool = nullptr;
    SdctTcMapCache tc_maps;   // encoded CUtensorMaps keyed by (pointer, geometry); guarded by its own mutex
    std::mutex tc_maps_mu;
};

struct Geometry {
    uint32_t bpr, bands_total, band_begin, band_end, eyes, comps;
};
int sdct_validate(const sdct_job *j, Geometry *g);
Pipeline *sdct_acquire_pipeline(sdct_context *ctx);
void sdct_release_pipeline(sdct_context *ctx, Pipeline *p);
void sdct_destroy_pipeline(Pipeline *p);

// sdct_noalloc.cu: the allocation-free two-kernel round trip keeps its scratch in file-scope __device__ arrays; this is the
// per-device record of their ADDRESSES (cudaGetSymbolAddress, once per device) and of what the operand array holds
struct SdctStaticScratch {
    uint8_t *coefs = nullptr;      // SDCT_NOALLOC_MAX_COEF_BYTES: the levels between the forward and the inverse kernel
    void *inv_operand = nullptr;   // 16 KiB: the tensor-core inverse's LUT-dependent operand (SDCT_TENSOR jobs)
    int layout = -1;               // key of inv_operand's current contents: (layout, lut); -1 = nothing valid
    float lut[64];
    float off_cols[64];
    int S = 0;
    std::mutex mu;                 // guards the key; the arrays themselves are shared by every call (overlapping calls race)
};
// sdct_api.cpp: forward kernel -> ss->coefs -> inverse kernel on `stream`; calls no allocating routine
int sdct_roundtrip_two_pass_static(sdct_context *ctx, const sdct_job *j, const void *d_in, const float *lut, void *d_out,
                                   void *stream, SdctStaticScratch *ss, uint64_t max_coef_bytes);

// tensor-core form: K[c][p] = round(2^22 * C[u][y] * C[v][x]), c = u*8+v, p = y*8+x (sdct_tc_matrix.cpp)
void sdct_tc_integer_matrix(int32_t K[64 * 64]);
// the 16 KiB B operand for `layout`: fp16 digits of K in the UMMA canonical K-major layout
void sdct_tc_operand(int layout, uint16_t out[128 * 64], bool positional = false);   // positional: column c = coefficient position c
// the LUT-dependent 16 KiB B operand of the tensor-core inverse; returns the scaling exponent S, fills the
// per-column offsets -127 * sum_i E[p][i]
int sdct_tc_inverse_operand(int layout, const float tab[64], uint16_t out[128 * 64], float off_cols[64], bool rows = false);
