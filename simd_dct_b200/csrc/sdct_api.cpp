// sdct_api.cpp - host side of the C ABI declared in include/sdct.h: contexts, job validation,
// scale-table construction, the chunked host<->device pipeline and the reference-call wrappers.
// No compute happens here and there is no CPU fallback: every path ends in sdct_k_launch().
#include "sdct_internal.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

namespace {
thread_local char g_err[512] = "";
}

int sdct_fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int sdct_validate(const sdct_job *j, Geometry *g)
{
    if (!j) return fail(SDCT_INVALID_PARAMETER, "job is null");
    if (j->layout < SDCT_LAYOUT_A || j->layout > SDCT_LAYOUT_C) return fail(SDCT_INVALID_PARAMETER, "unknown layout %d", j->layout);
    if (j->components != 1 && j->components != 3) return fail(SDCT_NOT_SUPPORTED, "components must be 1 or 3 (got %d)", j->components);
    const int eyes = j->layout == SDCT_LAYOUT_C ? j->eyes : 1;
    if (eyes != 1 && eyes != 2) return fail(SDCT_INVALID_PARAMETER, "eyes must be 1 or 2 (got %d)", j->eyes);
    // src/simd_dct.cpp:76 / :98 (multiple of 8) and :118 (layout B: multiple of 64)
    if ((j->width & 7) || (j->height & 7)) return fail(SDCT_NOT_SUPPORTED, "width and height must be multiples of 8");
    if (j->layout == SDCT_LAYOUT_B && (j->width & 63)) return fail(SDCT_NOT_SUPPORTED, "layout B needs width %% 64 == 0");
    if (j->width / 8 > 0x7fffffffull || j->height / 8 * eyes > 0x7fffffffull || j->frames > 0x7fffffffull)
        return fail(SDCT_NOT_SUPPORTED, "image too large");
    if ((j->width / 8) * (j->height / 8) * eyes > 0xffffffffull) return fail(SDCT_NOT_SUPPORTED, "more than 2^32 blocks per plane");
    if (j->pitch < j->width * (uint64_t)j->components) return fail(SDCT_INVALID_PARAMETER, "pitch smaller than a pixel row");
    g->bpr = (uint32_t)(j->width / 8);
    g->bands_total = (uint32_t)(j->height / 8);
    g->band_begin = (uint32_t)std::min<uint64_t>(j->band_begin, g->bands_total);
    g->band_end = j->band_end == 0 ? g->bands_total : (uint32_t)std::min<uint64_t>(j->band_end, g->bands_total);
    if (g->band_end < g->band_begin) g->band_end = g->band_begin;
    g->eyes = (uint32_t)eyes;
    g->comps = (uint32_t)j->components;
    return SDCT_OK;
}

namespace {

// scale tables, in float exactly as the reference builds them (EXACT) or as oracle/sdct_oracle.c
// defines them (FAST, inverse); every value is duplicated for the two packed lanes of the kernels
// tensor-core form: tab[j] = quantiser scale of D column j with 2^-22 (the matrix scaling) folded in
int build_tc_tables(const sdct_job *j, const float *lut, float2 tab[64])
{
    if (!lut) return fail(SDCT_INVALID_PARAMETER, "quantisation LUT is null");
    const bool lb = j->layout == SDCT_LAYOUT_B;
    for (int c = 0; c < 64; c++) {
        const int pos = sdct_tc_column_position(j->layout, c);
        const float d = lut[pos] * 0.95f;                                   // vr, src/simd_dct.cpp:191
        const float f = std::ldexp(lb ? 255.0f / d : 1.f / d, -22);         // :2236-2239 / :194-212 (x255 and /255 cancel)
        // |t| <= 255 * 64 * 2^20 < 2^34
        if (!std::isfinite(f) || std::fabs(f) * 17179869184.f >= 2147483648.f)
            return fail(SDCT_INVALID_PARAMETER, "LUT entry %d (%g) gives an unusable scale", pos, (double)lut[pos]);
        tab[c] = make_float2(f, f);
    }
    return SDCT_OK;
}

int build_tables(const sdct_job *j, const float *lut, int op, float2 tab[64], float2 tab2[64], float2 off[64])
{
    if (!lut) return fail(SDCT_INVALID_PARAMETER, "quantisation LUT is null");
    const bool fast = (j->flags & SDCT_FAST) != 0;
    const bool lb = j->layout == SDCT_LAYOUT_B;
    const bool sse = j->layout == SDCT_LAYOUT_C && (j->flags & SDCT_QUIRK_SSE) && op != SDCT_K_OP_INVERSE;
    for (int i = 0; i < 64; i++) {
        const float d = lut[i] * 0.95f;                               // vr, src/simd_dct.cpp:191
        float f, cmax = 0.f;
        if (fast) { f = lb ? 0.125f * (255.0f / d) : 0.125f * (1.f / d); cmax = 16384.f; }
        else if (lb) { f = 255.0f / d; cmax = 2048.f; }               // :2236-2239
        else if (sse) { f = 255.0f / d; cmax = 16.f; }                // :908-923 (coefficients of pixels / 255: |c| <= 8)
        else f = 1.f / d;                                             // :194-212
        const float inv = lb ? (d / 255.f) * 0.125f : d * 0.125f;     // -> pixel units after the unnormalised passes
        const float o = -(127.f * inv);
        if (!std::isfinite(f) || !std::isfinite(inv))
            return fail(SDCT_INVALID_PARAMETER, "LUT entry %d (%g) gives a non-finite scale", i, (double)lut[i]);
        // the round-to-nearest quantisers are exact while |coefficient * scale| < 2^31 (beyond that the
        // reference's cvtps_epi32 returns INT_MIN, src/simd_dct.cpp:2224, which is not reproduced)
        if (op != SDCT_K_OP_INVERSE && cmax != 0.f && std::fabs(f) * cmax >= 2147483648.f)
            return fail(SDCT_INVALID_PARAMETER, "LUT entry %d (%g) is too small: |coefficient * scale| could reach 2^31", i, (double)lut[i]);
        if (op == SDCT_K_OP_INVERSE) { tab[i] = make_float2(inv, inv); tab2[i] = make_float2(0.f, 0.f); }
        else { tab[i] = make_float2(f, f); tab2[i] = make_float2(inv, inv); }
        off[i] = make_float2(o, o);
    }
    return SDCT_OK;
}

bool aligned(const void *p, uint64_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

// the $SDCT_* switches of a context, read once (sdct_create)
void read_env(SdctEnv *e)
{
    auto num = [](const char *name, long lo, long hi, long dflt) {
        const char *v = getenv(name);
        if (!v) return dflt;
        const long n = atol(v);
        return n >= lo && n <= hi ? n : dflt;
    };
    e->host_chunk_bytes = (uint64_t)num("SDCT_HOST_CHUNK_MB", 1, 4096, 16) << 20;
    e->host_slots = (int)num("SDCT_HOST_SLOTS", 1, kSlots, 4);
    e->host_no_ramp = getenv("SDCT_HOST_NO_RAMP") != nullptr;
    e->host_trace = getenv("SDCT_HOST_TRACE") != nullptr;
    e->host_engine_streams = num("SDCT_HOST_ENGINE_STREAMS", 0, 1, 1) != 0;
    e->tc.no_tma = getenv("SDCT_TC_NO_TMA") != nullptr;
    e->tc.no_rgb_fuse = getenv("SDCT_TC_NO_RGB_FUSE") != nullptr;
    e->tc.no_tma_store = getenv("SDCT_TC_NO_TMA_STORE") != nullptr;
    e->tc.no_rt_fuse = getenv("SDCT_TC_NO_RT_FUSE") != nullptr;
    e->tc.ablate = (uint32_t)num("SDCT_TC_ABLATE", 0, 0x7fffffff, 0);
}

// The LUT-dependent operand of the tensor-core inverse (cached per (layout, LUT)).  A miss uploads it SYNCHRONOUSLY (the
// key becomes visible only after the copy has completed, so a hit on any stream may launch without an event); before an
// entry's device buffer is overwritten the device is synchronised (kernels in flight on other streams may read it).
struct InvOperandRef { void *d; int S; float off_cols[64]; };
int inverse_operand(sdct_context *ctx, int layout, const float *lut, const float scale[64], InvOperandRef *out, bool rows = false)
{
    std::lock_guard<std::mutex> lock(ctx->inv_mu);
    sdct_context::InvOperand *op_entry = nullptr;
    for (auto &e : ctx->inv_cache)
        if (e.d && e.layout == layout && e.rows == rows && memcmp(e.lut, lut, sizeof(e.lut)) == 0) op_entry = &e;
    if (!op_entry) {
        sdct_context::InvOperand *victim = &ctx->inv_cache[0];
        for (auto &e : ctx->inv_cache) {
            if (e.layout < 0) { victim = &e; break; }
            if (e.used < victim->used) victim = &e;
        }
        std::vector<uint16_t> operand(128 * 64);
        float off_cols[64];
        const int S = sdct_tc_inverse_operand(layout, scale, operand.data(), off_cols, rows);
        if (victim->layout >= 0) CU_TRY(cudaDeviceSynchronize());
        victim->layout = -1;   // invalid while its buffer is rewritten
        if (!victim->d) CU_TRY(cudaMalloc(&victim->d, operand.size() * sizeof(uint16_t)));
        CU_TRY(cudaMemcpy(victim->d, operand.data(), operand.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
        victim->S = S;
        memcpy(victim->off_cols, off_cols, sizeof(off_cols));
        memcpy(victim->lut, lut, sizeof(victim->lut));
        victim->rows = rows;
        victim->layout = layout;
        op_entry = victim;
    }
    op_entry->used = ++ctx->inv_clock;
    out->d = op_entry->d;
    out->S = op_entry->S;
    memcpy(out->off_cols, op_entry->off_cols, sizeof(out->off_cols));
    return SDCT_OK;
}

// The same operand in the allocation-free path's file-scope array (sdct_noalloc.cu) instead of a cache entry: uploaded on the
// call's stream when (layout, LUT) differ from what the array holds; no allocation of device memory on any path.
int static_inverse_operand(SdctStaticScratch *ss, int layout, const float *lut, const float scale[64], InvOperandRef *out, cudaStream_t st)
{
    std::lock_guard<std::mutex> lock(ss->mu);
    if (ss->layout != layout || memcmp(ss->lut, lut, sizeof(ss->lut)) != 0) {
        std::vector<uint16_t> operand(128 * 64);
        ss->layout = -1;
        ss->S = sdct_tc_inverse_operand(layout, scale, operand.data(), ss->off_cols, false);
        // pageable source: the call returns once the bytes are staged, so `operand` may go out of scope
        CU_TRY(cudaMemcpyAsync(ss->inv_operand, operand.data(), operand.size() * sizeof(uint16_t), cudaMemcpyHostToDevice, st));
        memcpy(ss->lut, lut, sizeof(ss->lut));
        ss->layout = layout;
    }
    out->d = ss->inv_operand;
    out->S = ss->S;
    memcpy(out->off_cols, ss->off_cols, sizeof(out->off_cols));
    return SDCT_OK;
}

int run_device(sdct_context *ctx, int op, const sdct_job *j, const uint8_t *pixels, uint8_t *coefs,
               uint8_t *pixels_out, const float *lut, cudaStream_t stream, SdctStaticScratch *static_operand = nullptr)
{
    Geometry g;
    int rc = sdct_validate(j, &g);
    if (rc) return rc;
    if (!ctx) return fail(SDCT_INVALID_PARAMETER, "context is null");
    if (!pixels || (op != SDCT_K_OP_ROUNDTRIP && !coefs) || (op == SDCT_K_OP_ROUNDTRIP && !pixels_out))
        return fail(SDCT_INVALID_PARAMETER, "null buffer");
    KParams p;
    memset(&p, 0, sizeof(p));
    const bool tensor = (j->flags & SDCT_TENSOR) != 0;
    if (tensor && op == SDCT_K_OP_ROUNDTRIP) {
        // single planes with a TMA pixel path: the fused kernel below; everything else: the forward and the inverse kernel back to
        // back on the caller's stream, the coefficients in a stream-ordered scratch allocation with dense 16-byte-aligned strides
        if (!ctx) return fail(SDCT_INVALID_PARAMETER, "context is null");
        if (!pixels || !pixels_out) return fail(SDCT_INVALID_PARAMETER, "null buffer");
        sdct_job cj = *j;
        cj.coef_comp_stride = j->width * j->height * (uint64_t)g.eyes;
        cj.coef_plane_stride = cj.coef_comp_stride / 64;
        cj.coef_frame_stride = cj.coef_comp_stride * g.comps;
        const size_t bytes = (size_t)(j->frames * cj.coef_frame_stride);
        if (bytes == 0) return SDCT_OK;
        DeviceScope dev(ctx->device);
        CU_TRY(dev.err);
        cudaStream_t st = stream ? stream : ctx->stream;
        if (g.comps == 1 && !(j->flags & (SDCT_FAST | SDCT_QUIRK_AVX_X3 | SDCT_COLOR_YCBCR)) && lut) {
            // single planes: ONE kernel that keeps the levels in tensor memory between the two products (sdct_tc.cu:
            // roundtrip_tc_kernel) where the geometry has a TMA pixel path; same bytes as the two kernels
            KParams q;
            memset(&q, 0, sizeof(q));
            float2 inv_tab[64], unused[64];
            if ((rc = build_tables(j, lut, SDCT_K_OP_INVERSE, inv_tab, unused, unused))) return rc;
            const bool lb = j->layout == SDCT_LAYOUT_B;
            bool ok = true;
            for (int c = 0; c < 64 && ok; c++) {   // forward scales in POSITIONAL column order (D1's columns are E's K index)
                const float d = lut[c] * 0.95f;
                const float f = std::ldexp(lb ? 255.0f / d : 1.f / d, -22);
                ok = std::isfinite(f) && std::fabs(f) * 17179869184.f < 2147483648.f;
                q.tab[c] = make_float2(f, f);
            }
            InvOperandRef ref;
            float scale[64];
            for (int i = 0; i < 64; i++) scale[i] = inv_tab[i].x;
            if (ok && inverse_operand(ctx, j->layout, lut, scale, &ref, /*rows*/ true) == SDCT_OK) {
                for (int c = 0; c < 64; c++) q.off[c] = make_float2(ref.off_cols[c], ref.off_cols[c]);
                const float unscale = std::ldexp(1.f, -ref.S);
                q.tab2[0] = make_float2(unscale, unscale);
                q.tc_matrix = ctx->d_tc_matrix[j->layout == SDCT_LAYOUT_A ? 2 : 1];
                q.tc_matrix2 = ref.d;
                q.pixels = pixels; q.pixels_out = pixels_out;
                q.pitch = j->pitch; q.eye_stride = j->eye_stride; q.pixel_frame_stride = j->pixel_frame_stride;
                q.bpr = g.bpr; q.eyes = g.eyes; q.comps = 1; q.frames = (uint32_t)j->frames;
                q.sm_count = (uint32_t)ctx->sm_count;
                q.row0 = g.band_begin * g.eyes;
                q.rows = (g.band_end - g.band_begin) * g.eyes;
                const uint64_t pstr = j->pitch | j->eye_stride | j->pixel_frame_stride;
                q.pixel_vec = (aligned(pixels, 16) && aligned(pixels_out, 16) && (pstr & 15) == 0 && (g.bpr & 1) == 0 && !(j->pitch >> 32)) ? 2 : 0;
                cudaError_t e;
                {
                    std::lock_guard<std::mutex> lock(ctx->tc_maps_mu);
                    e = sdct_k_launch_tc_roundtrip(j->layout, &q, &ctx->env.tc, &ctx->tc_maps, st);
                }
                if (e == cudaSuccess) {
                    ctx->launches.fetch_add(1, std::memory_order_relaxed);
                    return SDCT_OK;
                }
                if (e != cudaErrorNotSupported) CU_TRY(e);
                cudaGetLastError();
            }
        }
        void *scratch = nullptr;
        CU_TRY(cudaMallocFromPoolAsync(&scratch, bytes, ctx->scratch_pool, st));
        rc = run_device(ctx, SDCT_K_OP_FORWARD, &cj, pixels, (uint8_t *)scratch, nullptr, lut, st);
        if (rc == SDCT_OK) rc = run_device(ctx, SDCT_K_OP_INVERSE, &cj, pixels_out, (uint8_t *)scratch, nullptr, lut, st);
        cudaFreeAsync(scratch, st);
        return rc;
    }
    if (tensor && (j->flags & (SDCT_FAST | SDCT_QUIRK_AVX_X3 | SDCT_QUIRK_SSE))) return fail(SDCT_INVALID_PARAMETER, "SDCT_TENSOR excludes SDCT_FAST and the SDCT_QUIRK_* flags");
    if (j->flags & SDCT_QUIRK_SSE) {   // the reference's SSE stereo encoders (src/simd_dct.cpp:896-1103): layout C, reference arithmetic only
        if (j->layout != SDCT_LAYOUT_C || (j->flags & SDCT_FAST)) return fail(SDCT_INVALID_PARAMETER, "SDCT_QUIRK_SSE applies to layout C with SDCT_EXACT");
        if (j->width & 15) return fail(SDCT_NOT_SUPPORTED, "SDCT_QUIRK_SSE needs width %% 16 == 0 (the reference's SSE encoders take two blocks at a time)");
    }
    const bool tensor_inv = tensor && op == SDCT_K_OP_INVERSE;
    rc = (tensor && !tensor_inv) ? build_tc_tables(j, lut, p.tab) : build_tables(j, lut, op, p.tab, p.tab2, p.off);
    if (rc) return rc;
    p.pixels = pixels; p.coefs = coefs; p.pixels_out = pixels_out;
    p.pitch = j->pitch; p.eye_stride = j->eye_stride; p.pixel_frame_stride = j->pixel_frame_stride;
    p.coef_frame_stride = j->coef_frame_stride; p.coef_comp_stride = j->coef_comp_stride;
    p.coef_plane_stride = j->coef_plane_stride;
    p.bpr = g.bpr; p.eyes = g.eyes; p.comps = g.comps; p.frames = (uint32_t)j->frames;
    p.sm_count = (uint32_t)ctx->sm_count;
    p.row0 = g.band_begin * g.eyes;
    p.rows = (g.band_end - g.band_begin) * g.eyes;
    const uint64_t blocks = (uint64_t)p.rows * p.bpr;
    p.chunks = (uint32_t)((blocks + SDCT_CTA_BLOCKS - 1) / SDCT_CTA_BLOCKS);
    const uint64_t pstr = j->pitch | j->eye_stride | j->pixel_frame_stride;
    auto pix_aligned = [&](uint64_t a) {
        return aligned(pixels, a) && (pstr & (a - 1)) == 0 && (op != SDCT_K_OP_ROUNDTRIP || aligned(pixels_out, a));
    };
    p.pixel_vec = (pix_aligned(16) && g.comps == 1 && (g.bpr & 1) == 0) ? 2 : (pix_aligned(8) ? 1 : 0);
    p.pixel_aligned16 = pix_aligned(16);
    if (j->flags & SDCT_COLOR_YCBCR) {
        if (g.comps != 3) return fail(SDCT_INVALID_PARAMETER, "SDCT_COLOR_YCBCR needs components = 3 (interleaved RGB)");
        if (op != SDCT_K_OP_FORWARD && !(p.pixel_aligned16 && (g.bpr & 1) == 0))
            return fail(SDCT_NOT_SUPPORTED, "SDCT_COLOR_YCBCR on the output side needs 16-byte-aligned pixel rows and an even number of blocks per row");
        p.color = 1;
    }
    const uint64_t cstr = j->coef_frame_stride | j->coef_comp_stride;
    p.coef_aligned16 = coefs && aligned(coefs, 16) && (cstr & 15) == 0;
    p.coef_aligned4 = coefs && aligned(coefs, 4) && ((cstr | j->coef_plane_stride) & 3) == 0;
    if (j->layout == SDCT_LAYOUT_C)  // 16-byte plane segments: every plane and the window start must be aligned too
        p.coef_aligned16 = p.coef_aligned16 && (j->coef_plane_stride & 15) == 0 && (((uint64_t)p.row0 * p.bpr) & 15) == 0;
    if (blocks == 0 || j->frames == 0) return SDCT_OK;
    DeviceScope dev(ctx->device);
    CU_TRY(dev.err);
    const int mode = (j->flags & SDCT_FAST) ? SDCT_K_FAST : SDCT_K_EXACT;
    // the layout's reference-SIMD quirk: layout B the AVX2 path's X3 sign, layout C the SSE encoders' arithmetic
    const int x3 = (j->layout == SDCT_LAYOUT_C ? (j->flags & SDCT_QUIRK_SSE) : (j->flags & SDCT_QUIRK_AVX_X3)) ? 1 : 0;
    if (tensor_inv) {
        // the tensor-core inverse moves whole 16-byte pieces on both sides; other geometries have no kernel (the fp32
        // inverse is a different arithmetic definition, so it is not substituted silently)
        const bool pix_ok = g.comps == 1 ? p.pixel_vec == 2 : (p.pixel_aligned16 && (g.bpr & 1) == 0);
        if (!pix_ok || !p.coef_aligned16 || (j->layout == SDCT_LAYOUT_C && (blocks & 15)) || (j->pitch >> 32))
            return fail(SDCT_NOT_SUPPORTED, "SDCT_TENSOR inverse needs 16-byte-aligned pixel rows and coefficients and "
                                            "an even number of blocks per row (layout C: a multiple of 16 blocks)");
        cudaStream_t st = stream ? stream : ctx->stream;
        {
            InvOperandRef ref;
            float scale[64];
            for (int i = 0; i < 64; i++) scale[i] = p.tab[i].x;
            if ((rc = static_operand ? static_inverse_operand(static_operand, j->layout, lut, scale, &ref, st)
                                     : inverse_operand(ctx, j->layout, lut, scale, &ref))) return rc;
            p.tc_matrix = ref.d;
            for (int c = 0; c < 64; c++) p.tab[c] = make_float2(ref.off_cols[c], ref.off_cols[c]);
            const float unscale = std::ldexp(1.f, -ref.S);
            p.tab2[0] = make_float2(unscale, unscale);
        }
        {
            std::lock_guard<std::mutex> lock(ctx->tc_maps_mu);
            const cudaError_t e = sdct_k_launch_tc(op, j->layout, &p, &ctx->env.tc, &ctx->tc_maps, st);
            if (e == cudaErrorNotSupported) return fail(SDCT_NOT_SUPPORTED, "SDCT_TENSOR inverse: this geometry does not fit a tensor map");
            CU_TRY(e);
        }
    } else if (tensor) {
        p.tc_matrix = ctx->d_tc_matrix[j->layout == SDCT_LAYOUT_A ? 0 : 1];
        std::lock_guard<std::mutex> lock(ctx->tc_maps_mu);
        CU_TRY(sdct_k_launch_tc(op, j->layout, &p, &ctx->env.tc, &ctx->tc_maps, stream ? stream : ctx->stream));
    } else {
        CU_TRY(sdct_k_launch(op, j->layout, mode, x3, &p, stream ? stream : ctx->stream));
    }
    ctx->launches.fetch_add(1, std::memory_order_relaxed);
    return SDCT_OK;
}

}  // namespace

// The two-kernel round trip with the levels in caller-chosen memory (sdct_noalloc.cu passes its file-scope array): what the
// SDCT_TENSOR branch of run_device does with a pool allocation, for every arithmetic, without one.
int sdct_roundtrip_two_pass_static(sdct_context *ctx, const sdct_job *j, const void *d_in, const float *lut, void *d_out,
                                   void *stream, SdctStaticScratch *ss, uint64_t max_coef_bytes)
{
    Geometry g;
    int rc = sdct_validate(j, &g);
    if (rc) return rc;
    if (!ctx) return fail(SDCT_INVALID_PARAMETER, "context is null");
    if (!d_in || !d_out) return fail(SDCT_INVALID_PARAMETER, "null buffer");
    sdct_job cj = *j;   // dense, 16-byte-aligned coefficient strides (the array is 256-byte aligned)
    cj.coef_comp_stride = j->width * j->height * (uint64_t)g.eyes;
    cj.coef_plane_stride = cj.coef_comp_stride / 64;
    cj.coef_frame_stride = cj.coef_comp_stride * g.comps;
    if (j->frames != 0 && cj.coef_frame_stride > max_coef_bytes / j->frames)
        return fail(SDCT_NOT_SUPPORTED, "job has %llu x %llu coefficient bytes; the allocation-free entry point holds at most %llu "
                                        "(SDCT_NOALLOC_MAX_COEF_BYTES) - use sdct_roundtrip_device",
                    (unsigned long long)j->frames, (unsigned long long)cj.coef_frame_stride, (unsigned long long)max_coef_bytes);
    if (j->frames * cj.coef_frame_stride == 0) return SDCT_OK;
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    rc = run_device(ctx, SDCT_K_OP_FORWARD, &cj, (const uint8_t *)d_in, ss->coefs, nullptr, lut, st);
    if (rc == SDCT_OK) rc = run_device(ctx, SDCT_K_OP_INVERSE, &cj, (const uint8_t *)d_out, ss->coefs, nullptr, lut, st, ss);
    return rc;
}

extern "C" {

void sdct_job_dense(sdct_job *job, int layout, int flags, int components, int eyes, uint64_t width,
                    uint64_t height, uint64_t frames)
{
    memset(job, 0, sizeof(*job));
    if (layout != SDCT_LAYOUT_C) eyes = 1;
    job->layout = layout; job->flags = flags; job->components = components; job->eyes = eyes;
    job->width = width; job->height = height; job->frames = frames;
    job->pitch = width * components;
    job->eye_stride = job->pitch * height;
    job->pixel_frame_stride = job->eye_stride * eyes;
    job->coef_plane_stride = width * height * eyes / 64;
    job->coef_comp_stride = width * height * eyes;
    job->coef_frame_stride = job->coef_comp_stride * components;
}

const char *sdct_last_error(void) { return g_err; }
int sdct_abi_version(void) { return SDCT_ABI_VERSION; }

int sdct_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int sdct_create(int device, sdct_context **out)
{
    if (!out) return fail(SDCT_INVALID_PARAMETER, "out is null");
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(SDCT_NO_DEVICE, "no CUDA device visible - this library has no CPU fallback");
    }
    if (device < 0 || device >= n) return fail(SDCT_INVALID_PARAMETER, "device %d out of range (0..%d)", device, n - 1);
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(SDCT_NO_DEVICE, "device %d is sm_%d%d; kernels are built for sm_100a only", device, prop.major, prop.minor);
    DeviceScope dev(device);
    CU_TRY(dev.err);
    std::unique_ptr<sdct_context, void (*)(sdct_context *)> c(new sdct_context(), sdct_destroy);   // released on every early return
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    read_env(&c->env);
    CU_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        CU_TRY(cudaMemPoolCreate(&c->scratch_pool, &props));
        uint64_t keep = ~0ull;
        CU_TRY(cudaMemPoolSetAttribute(c->scratch_pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    {   // tensor-core B operands (16 KiB each)
        std::vector<uint16_t> op(128 * 64);
        const int layouts[3] = {SDCT_K_LAYOUT_A, SDCT_K_LAYOUT_B, SDCT_K_LAYOUT_A};
        for (int i = 0; i < 3; i++) {   // [2]: layout A in positional column order (the fused round trip)
            sdct_tc_operand(layouts[i], op.data(), i == 2);
            CU_TRY(cudaMalloc(&c->d_tc_matrix[i], op.size() * sizeof(uint16_t)));
            CU_TRY(cudaMemcpy(c->d_tc_matrix[i], op.data(), op.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
        }
    }
    *out = c.release();
    return SDCT_OK;
}

void sdct_destroy(sdct_context *c)
{
    if (!c) return;
    DeviceScope dev(c->device);
    for (auto &p : c->pipes) sdct_destroy_pipeline(p.get());
    c->pipes.clear();
    if (c->stream) { cudaStreamSynchronize(c->stream); cudaStreamDestroy(c->stream); }
    if (c->scratch_pool) cudaMemPoolDestroy(c->scratch_pool);
    cudaFree(c->d_tc_matrix[0]); cudaFree(c->d_tc_matrix[1]); cudaFree(c->d_tc_matrix[2]);
    for (auto &e : c->inv_cache) cudaFree(e.d);
    delete c;
}

int sdct_sm_count(const sdct_context *c) { return c ? c->sm_count : 0; }
uint64_t sdct_kernel_launches(const sdct_context *c) { return c ? c->launches.load() : 0; }

int sdct_synchronize(sdct_context *c)
{
    if (!c) return fail(SDCT_INVALID_PARAMETER, "context is null");
    DeviceScope dev(c->device);
    CU_TRY(dev.err);
    CU_TRY(cudaStreamSynchronize(c->stream));
    std::lock_guard<std::mutex> lock(c->pipes_mu);
    for (auto &p : c->pipes)
        for (Slot &s : p->slots)
            if (s.stream) CU_TRY(cudaStreamSynchronize(s.stream));
    return SDCT_OK;
}

int sdct_forward_device(sdct_context *c, const sdct_job *j, const void *d_pixels, const float *lut, void *d_coefs, void *stream)
{
    return run_device(c, SDCT_K_OP_FORWARD, j, (const uint8_t *)d_pixels, (uint8_t *)d_coefs, nullptr, lut, (cudaStream_t)stream);
}

int sdct_inverse_device(sdct_context *c, const sdct_job *j, const void *d_coefs, const float *lut, void *d_pixels, void *stream)
{
    return run_device(c, SDCT_K_OP_INVERSE, j, (const uint8_t *)d_pixels, (uint8_t *)const_cast<void *>(d_coefs), nullptr, lut, (cudaStream_t)stream);
}

int sdct_roundtrip_device(sdct_context *c, const sdct_job *j, const void *d_in, const float *lut, void *d_out, void *stream)
{
    return run_device(c, SDCT_K_OP_ROUNDTRIP, j, (const uint8_t *)d_in, nullptr, (uint8_t *)d_out, lut, (cudaStream_t)stream);
}

int sdct_plane_stats_device(sdct_context *c, const sdct_job *j, const void *d_coefs, sdct_plane_stat *d_stats, void *stream)
{
    Geometry g;
    int rc = sdct_validate(j, &g);
    if (rc) return rc;
    if (!c) return fail(SDCT_INVALID_PARAMETER, "context is null");
    if (!d_coefs || !d_stats) return fail(SDCT_INVALID_PARAMETER, "null buffer");
    if (j->layout != SDCT_LAYOUT_C) return fail(SDCT_NOT_SUPPORTED, "plane statistics are defined for layout C (64 coefficient planes)");
    KParams p;
    memset(&p, 0, sizeof(p));
    p.coefs = (uint8_t *)const_cast<void *>(d_coefs);
    p.coef_frame_stride = j->coef_frame_stride; p.coef_comp_stride = j->coef_comp_stride; p.coef_plane_stride = j->coef_plane_stride;
    p.bpr = g.bpr; p.comps = g.comps; p.frames = (uint32_t)j->frames; p.sm_count = (uint32_t)c->sm_count;
    p.row0 = g.band_begin * g.eyes; p.rows = (g.band_end - g.band_begin) * g.eyes;
    const uint64_t cstr = j->coef_frame_stride | j->coef_comp_stride | j->coef_plane_stride;
    p.coef_aligned16 = aligned(d_coefs, 16) && (cstr & 15) == 0 && (((uint64_t)p.row0 * p.bpr) & 15) == 0;
    DeviceScope dev(c->device);
    CU_TRY(dev.err);
    CU_TRY(sdct_k_plane_stats(&p, d_stats, stream ? (cudaStream_t)stream : c->stream));
    c->launches.fetch_add(1, std::memory_order_relaxed);
    return SDCT_OK;
}

void sdct_tensor_matrix(int32_t *k4096) { if (k4096) sdct_tc_integer_matrix(k4096); }

int sdct_selftest_dct8(sdct_context *c, int kind, float *h_data, size_t n)
{
    if (!c || !h_data) return fail(SDCT_INVALID_PARAMETER, "null argument");
    if (n == 0) return SDCT_OK;
    if (n > 0x7fffffffu / 8) return fail(SDCT_NOT_SUPPORTED, "too many vectors");
    DeviceScope dev(c->device);
    CU_TRY(dev.err);
    float *d = nullptr;
    CU_TRY(cudaMalloc((void **)&d, n * 8 * sizeof(float)));
    cudaError_t e = cudaMemcpyAsync(d, h_data, n * 8 * sizeof(float), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = sdct_k_selftest(d, (uint32_t)n, kind, c->stream);
    if (e == cudaSuccess) c->launches.fetch_add(1, std::memory_order_relaxed);
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_data, d, n * 8 * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if (e != cudaSuccess) return fail(SDCT_CUDA_ERROR, "selftest failed: %s", cudaGetErrorString(e));
    return SDCT_OK;
}

void *sdct_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        fail(SDCT_CUDA_ERROR, "cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}

void sdct_host_free(void *p) { if (p) cudaFreeHost(p); }

void sdct_shard_range(uint64_t total, int rank, int world, uint64_t *begin, uint64_t *end)
{
    if (world < 1) world = 1;
    if (rank < 0) rank = 0;
    if (rank >= world) rank = world - 1;
    const uint64_t q = total / (uint64_t)world, r = total % (uint64_t)world;
    const uint64_t b = q * (uint64_t)rank + std::min<uint64_t>((uint64_t)rank, r);
    if (begin) *begin = b;
    if (end) *end = b + q + ((uint64_t)rank < r ? 1 : 0);
}

}  // extern "C"
