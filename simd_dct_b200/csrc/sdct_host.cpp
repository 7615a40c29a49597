// sdct_host.cpp - host-buffer entry points: the chunked H2D -> kernel -> D2H pipeline behind
// sdct_forward_host / sdct_inverse_host (one device) and sdct_*_host_multi (several devices, one
// call), and the reference-call wrappers that the simd_dct.h shim binds (src/simd_dct.h:29-31).
//
// A job is cut into chunks (frame ranges of a batch, band ranges of a large frame).  Every chunk is a
// DENSE SUB-JOB on the device: its window of the caller's image is copied into a compact,
// 16-byte-aligned device layout (row pitch rounded up to 16, only the window's rows, only the
// window's coefficient bytes), so the kernels always run their aligned vector / TMA paths whatever
// the caller's pitch and alignment, a small window over large frames costs no device memory, and
// bytes outside the window - row padding included - are never written back.  Copies are linear
// where source and destination are contiguous and 2-D (cudaMemcpy2DAsync: pixel rows with a padded
// pitch; the 64 plane segments of a layout-C window) otherwise.  Copies and kernels of different
// chunks overlap on the slots' streams; pageable caller memory goes through pinned staging that
// has the device layout (one linear transfer per chunk), pinned memory is DMA'd directly.
//
// Concurrency: a call owns one Pipeline of the context's pool for its duration, so caller threads
// with disjoint windows (the reference's threading model, src/simd_dct.cpp:1073-1087) overlap; the
// multi-device form runs one such pipeline per device on its own host thread.
#include "sdct_internal.h"

#include <chrono>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Copy {   // width x rows bytes between host (h_off, h_pitch) and device (d_off, d_pitch); rows == 1: linear
    uint64_t h_off, d_off, width, rows, h_pitch, d_pitch;
};

void normalise(std::vector<Copy> &v)
{
    for (Copy &c : v)
        if (c.rows > 1 && c.h_pitch == c.width && c.d_pitch == c.width) { c.width *= c.rows; c.rows = 1; }
    std::sort(v.begin(), v.end(), [](const Copy &a, const Copy &b) { return a.h_off < b.h_off; });
    size_t o = 0;
    for (size_t i = 0; i < v.size(); i++) {
        if (v[i].width == 0 || v[i].rows == 0) continue;
        if (o && v[o - 1].rows == 1 && v[i].rows == 1 && v[o - 1].h_off + v[o - 1].width == v[i].h_off &&
            v[o - 1].d_off + v[o - 1].width == v[i].d_off)
            v[o - 1].width += v[i].width;
        else
            v[o++] = v[i];
    }
    v.resize(o);
}

uint64_t align16(uint64_t x) { return (x + 15) & ~(uint64_t)15; }

struct Chunk {
    uint64_t f0, f1;      // frames
    uint32_t b0, b1;      // bands
};

struct Plan {             // a chunk as a dense sub-job on the device + its copies
    sdct_job dev;
    std::vector<Copy> pix, coef;
    uint64_t pix_bytes = 0, coef_bytes = 0;
};

Plan plan_chunk(const sdct_job &j, const Geometry &g, const Chunk &c)
{
    Plan p;
    const uint64_t nb = c.b1 - c.b0, frames = c.f1 - c.f0;
    const uint64_t row_bytes = j.width * (uint64_t)g.comps, dpitch = align16(row_bytes);
    const uint64_t n = nb * g.eyes * g.bpr;            // blocks of one coefficient set in the window
    sdct_job &d = p.dev;
    d = j;
    d.eyes = (int32_t)g.eyes;
    d.height = nb * 8;
    d.pitch = dpitch;
    d.eye_stride = dpitch * nb * 8;
    d.frames = frames;
    d.pixel_frame_stride = d.eye_stride * g.eyes;
    d.band_begin = 0; d.band_end = nb;
    d.coef_plane_stride = j.layout == SDCT_LAYOUT_C ? align16(n) : n;
    d.coef_comp_stride = j.layout == SDCT_LAYOUT_C ? 64 * d.coef_plane_stride : n * 64;
    d.coef_frame_stride = d.coef_comp_stride * g.comps;
    p.pix_bytes = d.pixel_frame_stride * frames;
    p.coef_bytes = d.coef_frame_stride * frames;
    const uint64_t n0 = (uint64_t)c.b0 * g.eyes * g.bpr;
    for (uint64_t f = 0; f < frames; f++) {
        for (uint32_t e = 0; e < g.eyes; e++)
            p.pix.push_back({(c.f0 + f) * j.pixel_frame_stride + e * j.eye_stride + (uint64_t)c.b0 * 8 * j.pitch,
                             f * d.pixel_frame_stride + e * d.eye_stride, row_bytes, nb * 8, j.pitch, dpitch});
        for (uint32_t k = 0; k < g.comps; k++) {
            const uint64_t hb = (c.f0 + f) * j.coef_frame_stride + k * j.coef_comp_stride;
            const uint64_t db = f * d.coef_frame_stride + k * d.coef_comp_stride;
            if (j.layout == SDCT_LAYOUT_C) p.coef.push_back({hb + n0, db, n, 64, j.coef_plane_stride, d.coef_plane_stride});
            else p.coef.push_back({hb + n0 * 64, db, n * 64, 1, 0, 0});
        }
    }
    normalise(p.pix);
    normalise(p.coef);
    return p;
}

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

// host <-> device copies of one chunk side, straight between the caller's (pinned) memory and the device buffer
int copy_direct(const std::vector<Copy> &v, uint8_t *host, uint8_t *dev, bool to_device, cudaStream_t st)
{
    for (const Copy &c : v) {
        uint8_t *h = host + c.h_off, *d = dev + c.d_off;
        if (c.rows == 1) {
            CU_TRY(to_device ? cudaMemcpyAsync(d, h, c.width, cudaMemcpyHostToDevice, st)
                             : cudaMemcpyAsync(h, d, c.width, cudaMemcpyDeviceToHost, st));
        } else {
            CU_TRY(to_device ? cudaMemcpy2DAsync(d, c.d_pitch, h, c.h_pitch, c.width, c.rows, cudaMemcpyHostToDevice, st)
                             : cudaMemcpy2DAsync(h, c.h_pitch, d, c.d_pitch, c.width, c.rows, cudaMemcpyDeviceToHost, st));
        }
    }
    return SDCT_OK;
}

// pageable caller memory <-> pinned staging that has the device layout
void stage_rows(const std::vector<Copy> &v, uint8_t *host, uint8_t *stage, bool to_stage)
{
    for (const Copy &c : v)
        for (uint64_t r = 0; r < c.rows; r++) {
            uint8_t *h = host + c.h_off + r * c.h_pitch, *s = stage + c.d_off + r * c.d_pitch;
            if (to_stage) memcpy(s, h, c.width); else memcpy(h, s, c.width);
        }
}

struct Pending {  // a staged D2H that still has to be scattered into pageable caller memory
    bool active = false;
    std::vector<Copy> copies;
    uint8_t *dst = nullptr;
};

int drain(Slot &s, Pending &p)
{
    if (!p.active) return SDCT_OK;
    CU_TRY(cudaEventSynchronize(s.done));
    stage_rows(p.copies, p.dst, s.h_out, false);
    p.active = false;
    return SDCT_OK;
}

// chunk sizes in units (frames or bands).  The first chunk's upload and the last chunk's download overlap with
// nothing, so the sizes ramp: per/8, per/4, per/2, per, ..., per, per/2, per/4, per/8 when the job is long enough.
std::vector<uint64_t> ramp(uint64_t total, uint64_t per, bool no_ramp)
{
    std::vector<uint64_t> sizes;
    const uint64_t edge[3] = {std::max<uint64_t>(1, per / 8), std::max<uint64_t>(1, per / 4), std::max<uint64_t>(1, per / 2)};
    const uint64_t edges = edge[0] + edge[1] + edge[2];
    if (no_ramp || total < 2 * edges + per) {
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
}

int run_chunks(sdct_context *ctx, Pipeline *pipe, int op, const sdct_job *job, const Geometry &g, const uint8_t *h_pixels,
               uint8_t *h_coefs, const float *lut)
{
    DeviceScope dev(ctx->device);
    CU_TRY(dev.err);
    int rc;
    // ---- chunking: by frames for batches, by band ranges inside a frame otherwise ----------
    const uint64_t band_bytes = (uint64_t)g.eyes * 8 * job->width * g.comps;
    const uint64_t frame_bytes = band_bytes * (g.band_end - g.band_begin);
    const uint64_t target = ctx->env.host_chunk_bytes;
    std::vector<Chunk> chunks;
    if (job->frames > 1 && frame_bytes <= 2 * target) {
        const uint64_t per = std::max<uint64_t>(1, target / std::max<uint64_t>(frame_bytes, 1));
        uint64_t f = 0;
        for (uint64_t n : ramp(job->frames, per, ctx->env.host_no_ramp)) {
            chunks.push_back({f, f + n, g.band_begin, g.band_end});
            f += n;
        }
    } else {
        const uint64_t per = std::max<uint64_t>(1, target / std::max<uint64_t>(band_bytes, 1));
        const std::vector<uint64_t> sizes = ramp(g.band_end - g.band_begin, per, ctx->env.host_no_ramp);
        for (uint64_t f = 0; f < job->frames; f++) {
            uint32_t bnd = g.band_begin;
            for (uint64_t n : sizes) {
                chunks.push_back({f, f + 1, bnd, bnd + (uint32_t)n});
                bnd += (uint32_t)n;
            }
        }
    }

    const bool fwd = op == SDCT_K_OP_FORWARD;
    uint8_t *h_in = const_cast<uint8_t *>(fwd ? h_pixels : h_coefs);
    uint8_t *h_out = fwd ? h_coefs : const_cast<uint8_t *>(h_pixels);
    const bool in_pinned = is_pinned(h_in), out_pinned = is_pinned(h_out);
    const int depth = ctx->env.host_slots;
    Pending pending[kSlots];
    // Two ways to put the chunks on the device.  Slot streams (pageable memory, where the host has to stage and scatter
    // between the copies anyway; $SDCT_HOST_ENGINE_STREAMS=0): chunk i's copy in, kernel and copy out run on the stream of
    // slot i % depth.  Engine streams (pinned caller memory, the default): ONE stream per engine - every upload on
    // slots[0].stream, every kernel on slots[1].stream, every download on slots[2].stream, chunk-to-chunk and buffer-reuse
    // order carried by events, the host never blocks before the final synchronise - so that each copy engine sees its own
    // copies back to back in one queue.  Measured on 64 x 1080p (133 MB each way, one B200, PCIe 5 x16): 3.13 ms against
    // 3.22 ms with slot streams (raw concurrent copies of the same bytes with no dependency between them: 2.68-2.75 ms).
    // Tried and not kept: two stream triples alternating (3.19), chunk copies issued in 4 MB pieces (3.38), stream
    // priorities for either engine (no change), chunks of 8 / 12 / 24 / 32 MB (3.22 / 3.12 / 3.19 / 3.30), no ramp (3.13).
    const bool engines = ctx->env.host_engine_streams && in_pinned && out_pinned;
    static_assert(kSlots >= 3, "engine streams are taken from the slots' streams");
    const cudaStream_t st_in = pipe->slots[0].stream, st_k = pipe->slots[1].stream, st_out = pipe->slots[2].stream;
    // $SDCT_HOST_TRACE (profiling only): four events per chunk and the host's submit times, printed after the call
    const bool trace = ctx->env.host_trace;
    std::vector<cudaEvent_t> tev;
    std::vector<double> t_submit;
    const auto t_begin = std::chrono::steady_clock::now();
    auto stamp = [&](cudaStream_t st) {
        if (!trace) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, st);
        tev.push_back(e);
    };

    // the slots' device buffers at the size of the call's largest chunk BEFORE anything is in flight: growing one inside the
    // loop would free a buffer that the slot's previous chunk may still use (cudaFree waits for the whole device)
    {
        size_t max_pix = 0, max_coef = 0;
        for (const Chunk &c : chunks) {
            const Plan pl = plan_chunk(*job, g, c);
            max_pix = std::max<size_t>(max_pix, pl.pix_bytes);
            max_coef = std::max<size_t>(max_coef, pl.coef_bytes);
        }
        for (size_t i = 0; i < std::min<size_t>((size_t)depth, chunks.size()); i++) {
            if ((rc = grow_device(&pipe->slots[i].d_pix, &pipe->slots[i].d_pix_cap, max_pix))) return rc;
            if ((rc = grow_device(&pipe->slots[i].d_coef, &pipe->slots[i].d_coef_cap, max_coef))) return rc;
        }
    }

    for (size_t ci = 0; ci < chunks.size(); ci++) {
        const Plan plan = plan_chunk(*job, g, chunks[ci]);
        Slot &s = pipe->slots[ci % depth];
        Pending &pd = pending[ci % depth];
        if (engines) {   // slot buffers free again once chunk ci - depth has left the device: ordered ON the device, the host runs ahead
            if (ci >= (size_t)depth) CU_TRY(cudaStreamWaitEvent(st_in, s.done, 0));
        } else {
            if ((rc = drain(s, pd))) return rc;
            if (ci >= (size_t)depth) CU_TRY(cudaEventSynchronize(s.done));  // slot buffers free again
        }
        const cudaStream_t sin = engines ? st_in : s.stream, sk = engines ? st_k : s.stream, sout = engines ? st_out : s.stream;
        if (trace) t_submit.push_back(std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_begin).count());
        stamp(sin);
        if ((rc = grow_device(&s.d_pix, &s.d_pix_cap, plan.pix_bytes))) return rc;
        if ((rc = grow_device(&s.d_coef, &s.d_coef_cap, plan.coef_bytes))) return rc;
        const std::vector<Copy> &cin = fwd ? plan.pix : plan.coef, &cout = fwd ? plan.coef : plan.pix;
        uint8_t *d_in = fwd ? s.d_pix : s.d_coef, *d_out = fwd ? s.d_coef : s.d_pix;
        const uint64_t in_bytes = fwd ? plan.pix_bytes : plan.coef_bytes, out_bytes = fwd ? plan.coef_bytes : plan.pix_bytes;

        // H2D
        if (in_pinned) {
            if ((rc = copy_direct(cin, h_in, d_in, true, sin))) return rc;
        } else {
            if ((rc = grow_pinned(&s.h_in, &s.h_in_cap, in_bytes))) return rc;
            stage_rows(cin, h_in, s.h_in, true);
            CU_TRY(cudaMemcpyAsync(d_in, s.h_in, in_bytes, cudaMemcpyHostToDevice, sin));
        }
        stamp(sin);
        if (engines) {
            CU_TRY(cudaEventRecord(s.ev_in, sin));
            CU_TRY(cudaStreamWaitEvent(sk, s.ev_in, 0));
        }
        // kernel: the chunk as a dense job on the slot's buffers
        rc = fwd ? sdct_forward_device(ctx, &plan.dev, s.d_pix, lut, s.d_coef, sk)
                 : sdct_inverse_device(ctx, &plan.dev, s.d_coef, lut, s.d_pix, sk);
        if (rc) return rc;
        stamp(sk);
        if (engines) {
            CU_TRY(cudaEventRecord(s.ev_k, sk));
            CU_TRY(cudaStreamWaitEvent(sout, s.ev_k, 0));
        }
        // D2H
        if (out_pinned) {
            if ((rc = copy_direct(cout, h_out, d_out, false, sout))) return rc;
        } else {
            if ((rc = grow_pinned(&s.h_out, &s.h_out_cap, out_bytes))) return rc;
            CU_TRY(cudaMemcpyAsync(s.h_out, d_out, out_bytes, cudaMemcpyDeviceToHost, sout));
            pd.active = true; pd.copies = cout; pd.dst = h_out;
        }
        stamp(sout);
        CU_TRY(cudaEventRecord(s.done, sout));
    }
    for (int i = 0; i < kSlots; i++) {
        if ((rc = drain(pipe->slots[i], pending[i]))) return rc;
        if (pipe->slots[i].stream) CU_TRY(cudaStreamSynchronize(pipe->slots[i].stream));
    }
    if (trace) {
        const double t_end = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_begin).count();
        fprintf(stderr, "sdct host trace: %zu chunks, %d slots, call %.0f us (times in us: host submit | device h2d begin, h2d end, kernel end, d2h end; device clock starts at chunk 0's h2d begin)\n",
                chunks.size(), depth, t_end);
        for (size_t ci = 0; ci < chunks.size(); ci++) {
            float t[4];
            for (int k = 0; k < 4; k++) cudaEventElapsedTime(&t[k], tev[0], tev[4 * ci + k]);
            fprintf(stderr, "  chunk %2zu frames %llu-%llu bands %u-%u | %8.0f | %8.0f %8.0f %8.0f %8.0f\n", ci, (unsigned long long)chunks[ci].f0,
                    (unsigned long long)chunks[ci].f1, chunks[ci].b0, chunks[ci].b1, t_submit[ci], t[0] * 1e3, t[1] * 1e3, t[2] * 1e3, t[3] * 1e3);
        }
        for (cudaEvent_t e : tev) cudaEventDestroy(e);
    }
    return SDCT_OK;
}

int host_run(sdct_context *ctx, int op, const sdct_job *job, const uint8_t *h_pixels, uint8_t *h_coefs, const float *lut)
{
    if (!ctx) return fail(SDCT_INVALID_PARAMETER, "context is null");
    Geometry g;
    int rc = sdct_validate(job, &g);
    if (rc) return rc;
    if (!h_pixels || !h_coefs) return fail(SDCT_INVALID_PARAMETER, "null buffer");
    if (!lut) return fail(SDCT_INVALID_PARAMETER, "quantisation LUT is null");
    if (job->frames == 0 || g.band_end == g.band_begin) return SDCT_OK;
    Pipeline *pipe = sdct_acquire_pipeline(ctx);
    if (!pipe) return SDCT_CUDA_ERROR;   // message set by sdct_acquire_pipeline
    rc = run_chunks(ctx, pipe, op, job, g, h_pixels, h_coefs, lut);
    if (rc) {   // leave no work of a failed call in flight on the pipeline's streams
        for (Slot &s : pipe->slots)
            if (s.stream) cudaStreamSynchronize(s.stream);
        cudaGetLastError();
    }
    sdct_release_pipeline(ctx, pipe);
    return rc;
}

// one call over several devices: contiguous frame ranges (batches) or band ranges (one large frame) per device,
// each on its own host thread and pipeline; no inter-device exchange (blocks are independent)
int host_run_multi(sdct_context *const *ctxs, int n, int op, const sdct_job *job, const uint8_t *h_pixels, uint8_t *h_coefs,
                   const float *lut)
{
    if (!ctxs || n < 1) return fail(SDCT_INVALID_PARAMETER, "no contexts");
    for (int i = 0; i < n; i++)
        if (!ctxs[i]) return fail(SDCT_INVALID_PARAMETER, "context %d is null", i);
    if (n == 1) return host_run(ctxs[0], op, job, h_pixels, h_coefs, lut);
    Geometry g;
    int rc = sdct_validate(job, &g);
    if (rc) return rc;
    if (!h_pixels || !h_coefs) return fail(SDCT_INVALID_PARAMETER, "null buffer");
    if (!lut) return fail(SDCT_INVALID_PARAMETER, "quantisation LUT is null");
    const bool by_frames = job->frames >= (uint64_t)n;
    const uint64_t total = by_frames ? job->frames : (uint64_t)(g.band_end - g.band_begin);
    std::vector<int> rcs(n, SDCT_OK);
    std::vector<std::string> msgs(n);
    std::vector<std::thread> threads;
    for (int i = 0; i < n; i++) {
        uint64_t b, e;
        sdct_shard_range(total, i, n, &b, &e);
        if (e == b) continue;
        threads.emplace_back([=, &rcs, &msgs]() {
            sdct_job sub = *job;
            const uint8_t *px = h_pixels;
            uint8_t *cf = h_coefs;
            if (by_frames) {
                sub.frames = e - b;
                px += b * job->pixel_frame_stride;
                cf += b * job->coef_frame_stride;
            } else {   // base pointers stay those of the whole image: a band's output position does not depend on the window
                sub.band_begin = g.band_begin + b;
                sub.band_end = g.band_begin + e;
            }
            rcs[i] = host_run(ctxs[i], op, &sub, px, cf, lut);
            if (rcs[i]) msgs[i] = sdct_last_error();
        });
    }
    for (std::thread &t : threads) t.join();
    for (int i = 0; i < n; i++)
        if (rcs[i]) return fail(rcs[i], "device %d: %s", ctxs[i]->device, msgs[i].c_str());
    return SDCT_OK;
}

// ---- process-wide contexts for the reference-style calls -----------------------------------
// $SDCT_DEVICES = "all" or a comma-separated list: the call is sharded over those devices (one process, no
// collective); otherwise the single device $SDCT_DEVICE (default 0).
std::mutex g_mu;
std::vector<sdct_context *> g_ctxs;
std::atomic<int> g_ref_flags{SDCT_EXACT | SDCT_QUIRK_AVX_X3};

int global_contexts(std::vector<sdct_context *> *out)
{
    std::lock_guard<std::mutex> lock(g_mu);
    if (g_ctxs.empty()) {
        std::vector<int> devs;
        const char *list = getenv("SDCT_DEVICES");
        if (list && *list) {
            if (!strcmp(list, "all")) {
                const int n = sdct_device_count();
                for (int i = 0; i < n; i++) devs.push_back(i);
            } else {
                for (const char *p = list; *p;) {
                    char *end;
                    const long d = strtol(p, &end, 10);
                    if (end == p) break;
                    devs.push_back((int)d);
                    p = *end == ',' ? end + 1 : end;
                }
            }
        }
        if (devs.empty()) {
            const char *env = getenv("SDCT_DEVICE");
            devs.push_back(env ? atoi(env) : 0);
        }
        std::vector<sdct_context *> made;
        for (int d : devs) {
            sdct_context *c = nullptr;
            const int rc = sdct_create(d, &c);
            if (rc) {
                for (sdct_context *m : made) sdct_destroy(m);
                return rc;
            }
            made.push_back(c);
        }
        g_ctxs = made;
    }
    *out = g_ctxs;
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
    std::vector<sdct_context *> ctxs;
    int rc = global_contexts(&ctxs);
    if (rc) return rc;
    sdct_job j;
    const int eyes = layout == SDCT_LAYOUT_C ? 2 : 1;
    int flags = g_ref_flags.load();   // each quirk belongs to one entry point: the X3 sign to layout B, the SSE arithmetic to layout C
    if (layout != SDCT_LAYOUT_B) flags &= ~SDCT_QUIRK_AVX_X3;
    if (layout != SDCT_LAYOUT_C) flags &= ~SDCT_QUIRK_SSE;
    sdct_job_dense(&j, layout, flags, 1, eyes, sx, bands * 8, 1);
    if (layout == SDCT_LAYOUT_C) {
        j.eye_stride = sx * sy / 2;                   // :292
        j.coef_plane_stride = sx * sy / 64;           // :257
    }
    j.band_begin = b0; j.band_end = b1;
    return host_run_multi(ctxs.data(), (int)ctxs.size(), SDCT_K_OP_FORWARD, &j, from, to, lut);
}

}  // namespace

// ---- pipeline pool ----------------------------------------------------------------------------
Pipeline *sdct_acquire_pipeline(sdct_context *ctx)
{
    {
        std::lock_guard<std::mutex> lock(ctx->pipes_mu);
        for (auto &p : ctx->pipes)
            if (!p->busy) { p->busy = true; return p.get(); }
    }
    DeviceScope dev(ctx->device);
    if (dev.err != cudaSuccess) { fail(SDCT_CUDA_ERROR, "cudaSetDevice(%d) failed", ctx->device); return nullptr; }
    std::unique_ptr<Pipeline> p(new Pipeline());
    p->busy = true;
    for (Slot &s : p->slots) {
        cudaError_t e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s.ev_in, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s.ev_k, cudaEventDisableTiming);
        if (e != cudaSuccess) {
            fail(SDCT_CUDA_ERROR, "creating a host pipeline failed: %s", cudaGetErrorString(e));
            sdct_destroy_pipeline(p.get());
            return nullptr;
        }
    }
    std::lock_guard<std::mutex> lock(ctx->pipes_mu);
    ctx->pipes.push_back(std::move(p));
    return ctx->pipes.back().get();
}

void sdct_release_pipeline(sdct_context *ctx, Pipeline *p)
{
    std::lock_guard<std::mutex> lock(ctx->pipes_mu);
    p->busy = false;
}

void sdct_destroy_pipeline(Pipeline *p)
{
    for (Slot &s : p->slots) {
        if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); s.stream = nullptr; }
        if (s.done) { cudaEventDestroy(s.done); s.done = nullptr; }
        if (s.ev_in) { cudaEventDestroy(s.ev_in); s.ev_in = nullptr; }
        if (s.ev_k) { cudaEventDestroy(s.ev_k); s.ev_k = nullptr; }
        cudaFree(s.d_pix); cudaFree(s.d_coef);
        cudaFreeHost(s.h_in); cudaFreeHost(s.h_out);
        s.d_pix = s.d_coef = s.h_in = s.h_out = nullptr;
        s.d_pix_cap = s.d_coef_cap = s.h_in_cap = s.h_out_cap = 0;
    }
}

extern "C" {

int sdct_forward_host(sdct_context *c, const sdct_job *j, const void *h_pixels, const float *lut, void *h_coefs)
{
    return host_run(c, SDCT_K_OP_FORWARD, j, (const uint8_t *)h_pixels, (uint8_t *)h_coefs, lut);
}

int sdct_inverse_host(sdct_context *c, const sdct_job *j, const void *h_coefs, const float *lut, void *h_pixels)
{
    return host_run(c, SDCT_K_OP_INVERSE, j, (const uint8_t *)h_pixels, (uint8_t *)const_cast<void *>(h_coefs), lut);
}

int sdct_forward_host_multi(sdct_context *const *ctxs, int n, const sdct_job *j, const void *h_pixels, const float *lut, void *h_coefs)
{
    return host_run_multi(ctxs, n, SDCT_K_OP_FORWARD, j, (const uint8_t *)h_pixels, (uint8_t *)h_coefs, lut);
}

int sdct_inverse_host_multi(sdct_context *const *ctxs, int n, const sdct_job *j, const void *h_coefs, const float *lut, void *h_pixels)
{
    return host_run_multi(ctxs, n, SDCT_K_OP_INVERSE, j, (const uint8_t *)h_pixels, (uint8_t *)const_cast<void *>(h_coefs), lut);
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
int sdct_ref_device_count(void)
{
    std::vector<sdct_context *> ctxs;
    return global_contexts(&ctxs) == SDCT_OK ? (int)ctxs.size() : 0;
}

}  // extern "C"
