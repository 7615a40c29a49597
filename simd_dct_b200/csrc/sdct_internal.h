// sdct_internal.h - shared between sdct_api.cpp and sdct_host.cpp (not installed).
#pragma once
#include "../../include/sdct.h"
#include "sdct_kernels.h"

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <mutex>

int sdct_fail(int code, const char *fmt, ...);
#define fail sdct_fail

#define CU_TRY(expr)                                                                              \
    do {                                                                                          \
        cudaError_t e_ = (expr);                                                                  \
        if (e_ != cudaSuccess)                                                                    \
            return fail(SDCT_CUDA_ERROR, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_),  \
                        __FILE__, __LINE__);                                                      \
    } while (0)

constexpr int kSlots = 4;  // pipeline depth of the host path

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
    uint8_t *d_pix = nullptr, *d_coef = nullptr;
    size_t d_pix_cap = 0, d_coef_cap = 0;
    uint8_t *h_in = nullptr, *h_out = nullptr;  // pinned staging for pageable callers
    size_t h_in_cap = 0, h_out_cap = 0;
};

struct sdct_context {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    Slot slots[kSlots];
    std::atomic<uint64_t> launches{0};
    std::mutex host_mu;  // the host pipeline owns the slots
};

struct Geometry {
    uint32_t bpr, bands_total, band_begin, band_end, eyes, comps;
};
int sdct_validate(const sdct_job *j, Geometry *g);
