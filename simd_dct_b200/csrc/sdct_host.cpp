// sdct_host.cpp - host-buffer entry points: the chunked H2D -> kernel -> D2H pipeline behind
// sdct_forward_host / sdct_inverse_host, and the reference-call wrappers that the simd_dct.h
// shim binds (src/simd_dct.h:29-31).  Copies and kernels of different chunks overlap on
// kSlots streams; pageable caller memory goes through pinned staging, pinned memory
// (sdct_host_alloc / cudaHostRegister) is DMA'd directly.
#include "sdct_internal.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace {

struct Span { uint64_t off, len; };

void merge_spans(std::vector<Span> &v)
{
    std::sort(v.begin(), v.end(), [](const Span &a, const Span &b) { return a.off < b.off; });
    size_t o = 0;
    for (size_t i = 0; i < v.size(); i++) {
        if (v[i].len == 0) continue;
        if (o && v[o - 1].off + v[o - 1].len >= v[i].off) {
            v[o - 1].len = std::max(v[o - 1].off + v[o - 1].len, v[i].off + v[i].len) - v[o - 1].off;
        } else {
            v[o++] = v[i];
        }
    }
    v.resize(o);
}

struct Chunk {
    uint64_t f0, f1;      // frames
    uint32_t b0, b1;      // bands
    std::vector<Span> pix, coef;
};

void chunk_spans(const sdct_job &j, const Geometry &g, Chunk &c)
{
    const uint64_t row_bytes = j.width * (uint64_t)g.comps;
    const uint64_t nb = (uint64_t)(c.b1 - c.b0);
    for (uint64_t f = c.f0; f < c.f1; f++) {
        for (uint32_t e = 0; e < g.eyes; e++)
            c.pix.push_back({f * j.pixel_frame_stride + e * j.eye_stride + (uint64_t)c.b0 * 8 * j.pitch,
                             nb ? (nb * 8 - 1) * j.pitch + row_bytes : 0});
        const uint64_t n0 = (uint64_t)c.b0 * g.eyes * g.bpr, n1 = (uint64_t)c.b1 * g.eyes * g.bpr;
        for (uint32_t k = 0; k < g.comps; k++) {
            const uint64_t base = f * j.coef_frame_stride + k * j.coef_comp_stride;
            if (j.layout == SDCT_LAYOUT_C)
                for (int i = 0; i < 64; i++) c.coef.push_back({base + i * j.coef_plane_stride + n0, n1 - n0});
            else
                c.coef.push_back({base + n0 * 64, (n1 - n0) * 64});
        }
    }
    merge_spans(c.pix);
    merge_spans(c.coef);
}

uint64_t span_bytes(const std::vector<Span> &v) { uint64_t s = 0; for (const Span &x : v) s += x.len; return s; }

bool is_pinned(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

int grow_device(uint8_t **p, size_t *cap, size_t need)
{
    if (*cap >= need) return SDCT_OK;
    if (*p) CU_TRY(cudaFree(*p));
    *p = nullptr; *cap = 0;
    CU_TRY(cudaMalloc((void **)p, need));
    *cap = need;
    return SDCT_OK;
}

int grow_pinned(uint8_t **p, size_t *cap, size_t need)
{
    if (*cap >= need) return SDCT_OK;
    if (*p) CU_TRY(cudaFreeHost(*p));
    *p = nullptr; *cap = 0;
    CU_TRY(cudaHostAlloc((void **)p, need, cudaHostAllocDefault));
    *cap = need;
    return SDCT_OK;
}

struct Pending {  // a staged D2H that still has to be scattered into pageable caller memory
    bool active = false;
    std::vector<Span> spans;
    uint8_t *dst = nullptr;
};

int drain(Slot &s, Pending &p)
{
    if (!p.active) return SDCT_OK;
    CU_TRY(cudaEventSynchronize(s.done));
    uint64_t o = 0;
    for (const Span &x : p.spans) { memcpy(p.dst + x.off, s.h_out + o, x.len); o += x.len; }
    p.active = false;
    return SDCT_OK;
}

int host_run(sdct_context *ctx, int op, const sdct_job *job, const uint8_t *h_pixels, uint8_t *h_coefs,
             const float *lut)
{
    if (!ctx) return fail(SDCT_INVALID_PARAMETER, "context is null");
    Geometry g;
    int rc = sdct_validate(job, &g);
    if (rc) return rc;
    if (!h_pixels || !h_coefs) return fail(SDCT_INVALID_PARAMETER, "null buffer");
    if (!lut) return fail(SDCT_INVALID_PARAMETER, "quantisation LUT is null");
    if (job->frames == 0 || g.band_end == g.band_begin) return SDCT_OK;
    std::lock_guard<std::mutex> lock(ctx->host_mu);
    CU_TRY(cudaSetDevice(ctx->device));

    // ---- chunking: by frames for batches, by band ranges inside a frame otherwise ----------
    const uint64_t band_bytes = (uint64_t)g.eyes * 8 * job->width * g.comps;
    const uint64_t frame_bytes = band_bytes * (g.band_end - g.band_begin);
    const char *env = getenv("SDCT_HOST_CHUNK_MB");
    const uint64_t target = (env && atoi(env) > 0 ? (uint64_t)atoi(env) : 16) << 20;
    // The first chunk's upload and the last chunk's download overlap with nothing, so the chunk sizes ramp:
    // per/8, per/4, per/2, per, ..., per, per/2, per/4, per/8 units (frames or bands) when the job is long enough.
    auto ramp = [](uint64_t total, uint64_t per) {
        std::vector<uint64_t> sizes;
        const uint64_t edge[3] = {std::max<uint64_t>(1, per / 8), std::max<uint64_t>(1, per / 4), std::max<uint64_t>(1, per / 2)};
        const uint64_t edges = edge[0] + edge[1] + edge[2];
        if (getenv("SDCT_HOST_NO_RAMP") || total < 2 * edges + per) {
            for (uint64_t u = 0; u < total; u += per) sizes.push_back(std::min(per, total - u));
            return sizes;
        }
        for (int i = 0; i < 3; i++) sizes.push_back(edge[i]);
        for (uint64_t left = total - 2 * edges; left > 0;) {
            const uint64_t n = std::min(per, left);
            sizes.push_back(n);
            left -= n;
        }
        for (int i = 2; i >= 0; i--) sizes.push_back(edge[i]);
        return sizes;
    };
    std::vector<Chunk> chunks;
    if (job->frames > 1 && frame_bytes <= 2 * target) {
        const uint64_t per = std::max<uint64_t>(1, target / std::max<uint64_t>(frame_bytes, 1));
        uint64_t f = 0;
        for (uint64_t n : ramp(job->frames, per)) {
            Chunk c; c.f0 = f; c.f1 = f + n; c.b0 = g.band_begin; c.b1 = g.band_end;
            chunks.push_back(std::move(c));
            f += n;
        }
    } else {
        const uint64_t per = std::max<uint64_t>(1, target / std::max<uint64_t>(band_bytes, 1));
        const std::vector<uint64_t> sizes = ramp(g.band_end - g.band_begin, per);
        for (uint64_t f = 0; f < job->frames; f++) {
            uint32_t bnd = g.band_begin;
            for (uint64_t n : sizes) {
                Chunk c; c.f0 = f; c.f1 = f + 1; c.b0 = bnd; c.b1 = bnd + (uint32_t)n;
                chunks.push_back(std::move(c));
                bnd += (uint32_t)n;
            }
        }
    }

    const bool fwd = op == SDCT_K_OP_FORWARD;
    const uint8_t *h_in = fwd ? h_pixels : h_coefs;
    uint8_t *h_out = fwd ? h_coefs : const_cast<uint8_t *>(h_pixels);
    const bool in_pinned = is_pinned(h_in), out_pinned = is_pinned(h_out);
    // pipeline depth: chunks in flight ($SDCT_HOST_SLOTS for experiments)
    const char *senv = getenv("SDCT_HOST_SLOTS");
    const int depth = senv && atoi(senv) >= 1 && atoi(senv) <= kSlots ? atoi(senv) : 4;
    Pending pending[kSlots];

    for (size_t ci = 0; ci < chunks.size(); ci++) {
        Chunk &c = chunks[ci];
        chunk_spans(*job, g, c);
        Slot &s = ctx->slots[ci % depth];
        Pending &pd = pending[ci % depth];
        rc = drain(s, pd);
        if (rc) return rc;
        if (ci >= (size_t)depth) CU_TRY(cudaEventSynchronize(s.done));  // slot buffers free again

        const uint64_t plo = c.pix.front().off, phi = c.pix.back().off + c.pix.back().len;
        const uint64_t clo = c.coef.front().off, chi = c.coef.back().off + c.coef.back().len;
        // keep the vector-path alignment of the caller's layout: round the window start down
        const uint64_t plo_a = plo & ~(uint64_t)15, clo_a = clo & ~(uint64_t)15;
        if ((rc = grow_device(&s.d_pix, &s.d_pix_cap, phi - plo_a))) return rc;
        if ((rc = grow_device(&s.d_coef, &s.d_coef_cap, chi - clo_a))) return rc;
        const std::vector<Span> &sin = fwd ? c.pix : c.coef, &sout = fwd ? c.coef : c.pix;
        uint8_t *d_in = fwd ? s.d_pix : s.d_coef, *d_out = fwd ? s.d_coef : s.d_pix;
        const uint64_t in_lo = fwd ? plo_a : clo_a, out_lo = fwd ? clo_a : plo_a;

        // H2D
        if (in_pinned) {
            for (const Span &x : sin)
                CU_TRY(cudaMemcpyAsync(d_in + (x.off - in_lo), h_in + x.off, x.len, cudaMemcpyHostToDevice, s.stream));
        } else {
            if ((rc = grow_pinned(&s.h_in, &s.h_in_cap, span_bytes(sin)))) return rc;
            uint64_t o = 0;
            for (const Span &x : sin) {
                memcpy(s.h_in + o, h_in + x.off, x.len);
                CU_TRY(cudaMemcpyAsync(d_in + (x.off - in_lo), s.h_in + o, x.len, cudaMemcpyHostToDevice, s.stream));
                o += x.len;
            }
        }
        // kernel on this chunk's window of the caller's layout
        sdct_job sub = *job;
        sub.frames = c.f1 - c.f0;
        sub.band_begin = c.b0; sub.band_end = c.b1;
        const uint8_t *kp = s.d_pix - plo_a + c.f0 * job->pixel_frame_stride;
        uint8_t *kc = s.d_coef - clo_a + c.f0 * job->coef_frame_stride;
        rc = fwd ? sdct_forward_device(ctx, &sub, kp, lut, kc, s.stream)
                 : sdct_inverse_device(ctx, &sub, kc, lut, const_cast<uint8_t *>(kp), s.stream);
        if (rc) return rc;
        // D2H
        if (out_pinned) {
            for (const Span &x : sout)
                CU_TRY(cudaMemcpyAsync(h_out + x.off, d_out + (x.off - out_lo), x.len, cudaMemcpyDeviceToHost, s.stream));
        } else {
            if ((rc = grow_pinned(&s.h_out, &s.h_out_cap, span_bytes(sout)))) return rc;
            uint64_t o = 0;
            for (const Span &x : sout) {
                CU_TRY(cudaMemcpyAsync(s.h_out + o, d_out + (x.off - out_lo), x.len, cudaMemcpyDeviceToHost, s.stream));
                o += x.len;
            }
            pd.active = true; pd.spans = sout; pd.dst = h_out;
        }
        CU_TRY(cudaEventRecord(s.done, s.stream));
    }
    for (int i = 0; i < kSlots; i++) {
        if ((rc = drain(ctx->slots[i], pending[i]))) return rc;
        CU_TRY(cudaStreamSynchronize(ctx->slots[i].stream));
    }
    return SDCT_OK;
}

// ---- process-wide context for the reference-style calls -----------------------------------
std::mutex g_mu;
sdct_context *g_ctx = nullptr;
std::atomic<int> g_ref_flags{SDCT_EXACT | SDCT_QUIRK_AVX_X3};

int global_context(sdct_context **out)
{
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_ctx) {
        const char *env = getenv("SDCT_DEVICE");
        int rc = sdct_create(env ? atoi(env) : 0, &g_ctx);
        if (rc) return rc;
    }
    *out = g_ctx;
    return SDCT_OK;
}

// shared body of the three reference-style calls.  win_scale: the window test compares
// win_scale*y with startY/endY (2 everywhere, except the scalar layout-A loop, :377-387)
int ref_call(int layout, const uint8_t *from, uint8_t *to, const float *lut, size_t sx, size_t sy,
             size_t start_y, size_t end_y, size_t align_x, size_t win_scale)
{
    if (!from || !to) return fail(SDCT_INVALID_PARAMETER, "null buffer");               // :75 :97 :117
    if ((sx & (align_x - 1)) || (sy & 7)) return fail(SDCT_NOT_SUPPORTED, "unsupported size"); // :76 :98 :118
    if (!lut) return fail(SDCT_INVALID_PARAMETER, "quantisation LUT is null (the reference would dereference it)");
    const size_t half = sy / 2;                       // only rows below sizeY/2 are visited
    const size_t bands = (half + 7) / 8;              // `for (y = 0; y < sizeY/2; y += 8)`
    if (layout == SDCT_LAYOUT_C && (sy & 15))
        return fail(SDCT_NOT_SUPPORTED, "stereo layout needs sizeY %% 16 == 0 (the reference reads out of bounds otherwise)");
    if (sx == 0 || bands == 0) return SDCT_OK;
    // band k (y = 8k) runs iff start_y <= win_scale*8k <= end_y
    const size_t step = 8 * win_scale;
    const size_t b0 = (start_y + step - 1) / step;
    const size_t b1 = std::min(bands, end_y / step + 1);
    if (b0 >= b1) return SDCT_OK;
    sdct_context *ctx;
    int rc = global_context(&ctx);
    if (rc) return rc;
    sdct_job j;
    const int eyes = layout == SDCT_LAYOUT_C ? 2 : 1;
    sdct_job_dense(&j, layout, g_ref_flags.load(), 1, eyes, sx, bands * 8, 1);
    if (layout == SDCT_LAYOUT_C) {
        j.eye_stride = sx * sy / 2;                   // :292
        j.coef_plane_stride = sx * sy / 64;           // :257
    }
    j.band_begin = b0; j.band_end = b1;
    return host_run(ctx, SDCT_K_OP_FORWARD, &j, from, to, lut);
}

}  // namespace

extern "C" {

int sdct_forward_host(sdct_context *c, const sdct_job *j, const void *h_pixels, const float *lut, void *h_coefs)
{
    return host_run(c, SDCT_K_OP_FORWARD, j, (const uint8_t *)h_pixels, (uint8_t *)h_coefs, lut);
}

int sdct_inverse_host(sdct_context *c, const sdct_job *j, const void *h_coefs, const float *lut, void *h_pixels)
{
    return host_run(c, SDCT_K_OP_INVERSE, j, (const uint8_t *)h_pixels, (uint8_t *)const_cast<void *>(h_coefs), lut);
}

int sdct_ref_encode_quantize_buffer(const uint8_t *from, uint8_t *to, const float *lut, size_t sx, size_t sy,
                                    size_t start_y, size_t end_y)
{
    return ref_call(SDCT_LAYOUT_A, from, to, lut, sx, sy, start_y, end_y, 8, 1);
}

int sdct_ref_encode_quantize_reorder_stereo_buffer(const uint8_t *from, uint8_t *to, const float *lut, size_t sx,
                                                   size_t sy, size_t start_y, size_t end_y)
{
    return ref_call(SDCT_LAYOUT_C, from, to, lut, sx, sy, start_y, end_y, 8, 2);
}

int sdct_ref_encode_quantize32_reorder_buffer(const uint8_t *from, uint8_t *to, const float *lut, size_t sx,
                                              size_t sy, size_t start_y, size_t end_y)
{
    return ref_call(SDCT_LAYOUT_B, from, to, lut, sx, sy, start_y, end_y, 64, 2);
}

void sdct_ref_set_flags(int flags) { g_ref_flags.store(flags); }
int sdct_ref_get_flags(void) { return g_ref_flags.load(); }

}  // extern "C"
