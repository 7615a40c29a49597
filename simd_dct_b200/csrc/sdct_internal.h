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

constexpr int kSlots = 8;  // slots of the host path; the pipeline depth in use is host_depth() <= kSlots

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
    void *d_tc_matrix[2] = {nullptr, nullptr};  // tensor-core B operand: [0] layout-A column order, [1] natural order
    // tensor-core inverse operands depend on the LUT: a small cache keyed by (layout, LUT)
    struct InvOperand { void *d = nullptr; int layout = -1; float lut[64]; float off_cols[64]; int S = 0; };
    InvOperand inv_cache[4];
    unsigned inv_next = 0;
    std::mutex inv_mu;
};

struct Geometry {
    uint32_t bpr, bands_total, band_begin, band_end, eyes, comps;
};
int sdct_validate(const sdct_job *j, Geometry *g);

// tensor-core form: K[c][p] = round(2^22 * C[u][y] * C[v][x]), c = u*8+v, p = y*8+x (sdct_tc_matrix.cpp)
void sdct_tc_integer_matrix(int32_t K[64 * 64]);
// the 16 KiB B operand for `layout`: fp16 digits of K in the UMMA canonical K-major layout
void sdct_tc_operand(int layout, uint16_t out[128 * 64]);
// the LUT-dependent 16 KiB B operand of the tensor-core inverse; returns the scaling exponent S, fills the
// per-column offsets -127 * sum_i E[p][i]
int sdct_tc_inverse_operand(int layout, const float tab[64], uint16_t out[128 * 64], float off_cols[64]);
