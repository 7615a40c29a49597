#!/usr/bin/env python
"""bench.py - throughput of the 8x8 block-DCT hot path on B200 (contract: see DESIGN.md "Measurement").

Workload (BASELINE.json configs[1]): 64 frames of 1920x1080 single-plane u8, forward DCT +
quantise + reorder into layout B (the reference's fastest path, simdDCT_EncodeQuantize32ReorderBuffer),
SDCT_EXACT | SDCT_QUIRK_AVX_X3 = byte-identical to the reference's AVX2 output.
A step = one pass over the whole batch.  With N GPUs every rank runs the same batch size on its own
frames (weak scaling, no collective on the data path); value = blocks of all ranks / max-over-ranks time.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "8x8 blocks/s, forward DCT+quantise+reorder (layout B), 64x1080p luma"
UNIT = "blocks/s"
BYTES_PER_BLOCK = 128  # 64 B u8 read + 64 B u8 written


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def synth_frames(frames, h, w, seed):
    """uniform random u8 (worst case for the quantiser) - synthetic, seeded"""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, (frames, h, w), dtype=np.uint8)


class ClockSampler:
    """SM clock / throttle reasons sampled through NVML while the GPU runs the timed work
    (B200_PROFILING.md "clocks line").  NVML is polled from a thread (about 1 kHz) because the
    timed region of this workload lasts milliseconds, far below nvidia-smi's sampling period."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, torch_index):
        self.samples = []       # (t, sm_mhz, reasons_mask, power_w)
        self.marks = {}
        self.h = None
        self.err = None
        try:
            import pynvml
            import torch
            self.nv = pynvml
            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(torch_index).uuid)
            if not uuid.startswith("GPU-"):
                uuid = "GPU-" + uuid
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(torch_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)
        self._stop = threading.Event()
        self.thread = None

    def _reasons(self):
        nv = self.nv
        for name in ("nvmlDeviceGetCurrentClocksEventReasons", "nvmlDeviceGetCurrentClocksThrottleReasons"):
            fn = getattr(nv, name, None)
            if fn is not None:
                try:
                    return int(fn(self.h))
                except Exception:  # noqa: BLE001
                    continue
        return 0

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                except Exception:  # noqa: BLE001
                    pw = None
                self.samples.append((time.perf_counter(), mhz, self._reasons(), pw))
            except Exception as e:  # noqa: BLE001
                self.err = repr(e)
                return
            time.sleep(0.0005)

    def start(self):
        if self.h is not None:
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()

    def mark(self, name):
        self.marks[name] = time.perf_counter()

    def stop(self, t0=None, t1=None):
        if self.thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: %s" % self.err]}
        self._stop.set()
        self.thread.join(timeout=2)
        sel = [x for x in self.samples if (t0 is None or x[0] >= t0) and (t1 is None or x[0] <= t1)]
        mask = 0
        for x in sel:
            mask |= x[2]
        pw = [x[3] for x in sel if x[3] is not None]
        return {"sm_mhz": statistics.median(x[1] for x in sel) if sel else None, "sm_min_mhz": min((x[1] for x in sel), default=None),
                "sm_max_mhz": self.max_mhz, "reasons": sorted(v for k, v in self.REASONS.items() if mask & k),
                "samples": len(sel), "power_w_max": max(pw) if pw else None}


# ---- CPU reference arm --------------------------------------------------------------------------
def cpu_reference_runner(w, h, layout="B"):
    """returns (kind, isa, fn(frames_u8[F,H,W], lut, out[F,H*W], threads)) timing the reference's own
    CPU implementation of the path (oracle/_ref, the unmodified reference built with its own
    flags) or, if that did not travel, the oracle port.  layout B: the AVX-512VL / AVX2 path
    (src/simd_dct.cpp:1869 / :2064); layout A: the default dispatch (SSE4.1, :1540); layout C: the SSE4.1
    stereo path (:896), a frame passed as two eyes of H/2 rows."""
    import oraclelib as ol
    from concurrent.futures import ThreadPoolExecutor

    isa = None
    if ol.have_reference("fast"):
        ref = ol.Reference("fast")
        flags = ol.cpu_flags()
        if layout == "B":
            isa = "avx512vl" if {"avx512vl", "avx512f"} <= flags else ("avx2" if "avx2" in flags else None)
        else:
            isa = "sse41" if "sse4_1" in flags else "scalar"
        if isa:
            ref.set_isa(isa)
            fn = ref.fn[layout]
            kind = "reference"

            def one(src, dst, lut):
                # layouts A, B: a full H-row plane = sizeY 2H (only rows < sizeY/2 are visited); layout C: sizeY = H
                sy = h - h % 16 if layout == "C" else 2 * h   # (two eyes of whole bands: sizeY % 16 == 0)
                fn(src.ctypes.data_as(ol._u8p), dst.ctypes.data_as(ol._u8p), lut.ctypes.data_as(ol._f32p), w, sy, 0, sy)
    if isa is None:
        o = ol.Oracle()
        kind, isa = "port", "scalar-c"
        lay = {"A": ol.LAYOUT_A, "B": ol.LAYOUT_B, "C": ol.LAYOUT_C}[layout]
        eyes = 2 if layout == "C" else 1

        def one(src, dst, lut):
            he = (h - h % 16) // 2 if eyes == 2 else h
            o.lib.sdct_oracle_forward(lay, ol.MODE_EXACT, ol.QUIRK_AVX_X3, src.ctypes.data_as(ol._u8p), w, 1, w, he, eyes,
                                      he * w, lut.ctypes.data_as(ol._f32p), dst.ctypes.data_as(ol._u8p), eyes * he * w // 64)

    def run(frames, lut, out, threads):
        n = frames.shape[0]

        def work(t):
            for f in range(t, n, threads):
                one(frames[f], out[f], lut)
        if threads == 1:
            work(0)
            return
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(work, range(threads)))
    return kind, isa, run


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def time_cpu(w, h, frames_np, lut, reps, layout="B", threads=None, budget_s=8.0):
    """best of `reps` passes over `frames_np` (bounded by budget_s) -> blocks/s"""
    kind, isa, run = cpu_reference_runner(w, h, layout)
    threads = threads or host_threads()
    out = np.empty((frames_np.shape[0], h * w + 4096), dtype=np.uint8)   # + slack: the SSE layout-A stores overrun
    run(frames_np, lut, out, threads)  # warm
    best = float("inf")
    t_start = time.perf_counter()
    done = 0
    for _ in range(reps):
        t0 = time.perf_counter()
        run(frames_np, lut, out, threads)
        best = min(best, time.perf_counter() - t0)
        done += 1
        if time.perf_counter() - t_start > budget_s:
            break
    hh = h - h % 16 if layout == "C" else h
    blocks = frames_np.shape[0] * (hh // 8) * (w // 8)
    return {"value": blocks / best, "unit": UNIT, "cores": threads, "kind": kind, "gbps": blocks / best * BYTES_PER_BLOCK / 1e9,
            "sample": f"{frames_np.shape[0]} frames {w}x{h}, layout {layout}, {isa}, best of {done}, {threads} thread(s)"}, out[:, :h * w]


def strict_reference_bytes(frames_np, lut, w, h):
    """the bytes the reference writes for these frames: the STRICT build of the unmodified reference, AVX2 path
    (the parity target of layout B; AVX-512VL under -ffast-math differs from it in ~1e-5 of the bytes);
    the oracle port when the host has no AVX2 or oracle/_ref did not travel.  -> (bytes[F, H*W], source)"""
    import oraclelib as ol
    out = np.empty((frames_np.shape[0], h * w), dtype=np.uint8)
    if ol.have_reference("strict") and "avx2" in ol.cpu_flags():
        ref = ol.Reference("strict")
        ref.set_isa("avx2")
        for f in range(frames_np.shape[0]):
            rc, o = ref.call("B", frames_np[f], lut, w, 2 * h, 0, 2 * h)
            assert rc == 0
            out[f] = o[:h * w]
        return out, "reference (strict build, avx2 path)"
    o = ol.Oracle()
    for f in range(frames_np.shape[0]):
        out[f] = o.forward(ol.LAYOUT_B, frames_np[f], lut, quirks=ol.QUIRK_AVX_X3)
    return out, "oracle port (reference not runnable on this host)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=64)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--preheat", type=float, default=0.4, help="seconds of untimed back-to-back steps before the timed region (clock ramp)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs)")
    ap.add_argument("--no-variants", action="store_true", help="skip the secondary (non-headline) measurements")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    F, H, W = args.frames, args.height, args.width
    blocks_per_rank = F * (H // 8) * (W // 8)
    import oraclelib as ol
    lut = ol.lut(1.0)  # the reference CLI's default table (src/main.cpp:179-189)
    config = {"workload": f"configs[1]: {F} frames {W}x{H} single-plane u8 per GPU, forward DCT+quantise, layout B "
                          "(simdDCT_EncodeQuantize32ReorderBuffer), default LUT, exact+X3 quirk = reference bytes",
              "frames_per_gpu": F, "width": W, "height": H, "layout": "B", "arithmetic": "exact",
              "l2": "input (132.7 MB) + output (132.7 MB) per step exceed the 126 MB L2; no flush needed",
              "sharding": f"frames, {world} rank(s), no collective"}

    # ---------------- reference arm: the reference's CPU implementation on the host cores ----------
    if args.impl == "reference":
        if rank != 0:
            return 0
        frames_np = synth_frames(F, H, W, seed=1234)
        kind, isa, run = cpu_reference_runner(W, H)
        threads = host_threads()
        out = np.empty((F, H * W), dtype=np.uint8)
        for _ in range(args.warmup):
            run(frames_np, lut, out, threads)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            run(frames_np, lut, out, threads)
        dt = (time.perf_counter() - t0) / args.steps
        v = blocks_per_rank / dt
        line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": kind,
                                 "sample": f"each step = the full {F}-frame batch, {isa}, {threads} threads"},
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gbps": v * BYTES_PER_BLOCK / 1e9}
        print(json.dumps(line), flush=True)
        return 0

    # ---------------- our arm ------------------------------------------------------------------------
    import torch
    import simd_dct_b200 as sd

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (no CPU fallback exists)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        # NCCL prints its version banner on stdout when $NCCL_DEBUG asks for it; stdout carries ONE JSON line, so the
        # communicator is brought up (init + first collective) with fd 1 pointing at stderr
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    os.environ["SDCT_DEVICE"] = str(local_rank)  # device of the simd_dct.h shim's process-wide context
    ctx = sd.Context(local_rank)
    frames_np = synth_frames(F, H, W, seed=1234 + rank)
    d_in = torch.from_numpy(frames_np).cuda()
    flags = sd.EXACT | sd.QUIRK_AVX_X3
    job = sd.dense_job(sd.LAYOUT_B, W, H, frames=F, flags=flags)
    d_out = torch.empty(F * H * W, dtype=torch.uint8, device="cuda")

    def timed(fn, steps, warmup, preheat_s=0.0):
        """warm-up (>= `warmup` steps, plus `preheat_s` seconds of the same step so that the SM
        clock has ramped), then exactly `steps` steps between two CUDA events recorded on the
        stream the kernels are launched on (torch's current stream - Context passes it through)."""
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        t_end = time.perf_counter() + preheat_s
        while time.perf_counter() < t_end:
            for _ in range(50):
                fn()
            torch.cuda.synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = ctx.kernel_launches
        w0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        w1 = time.perf_counter()
        ms = e0.elapsed_time(e1)
        barrier()
        return max_over_ranks(ms) / steps, ctx.kernel_launches - n0, (w0, w1)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    t_heat = time.perf_counter()
    ms_step, launches, (w0, w1) = timed(lambda: ctx.forward(d_in, lut, job, out=d_out), args.steps, args.warmup, preheat_s=args.preheat)
    headline_out = d_out.clone()   # the bytes the timed launches wrote (d_out is reused by the variants below)
    clocks = None
    if rank == 0:
        # the same kernel runs back to back from t_heat + warm-up to w1: the clock samples of that
        # window are "under load"; the timed region itself is its last `steps` launches
        clocks = sampler.stop(t_heat + 0.1, w1)
        clocks["window"] = "0.3 s of back-to-back launches of the timed kernel ending with the timed region (%.2f ms)" % ((w1 - w0) * 1e3)
    value = world * blocks_per_rank / (ms_step * 1e-3)

    # roofline of the dominant (only) kernel: algorithmic bytes per launch / average launch duration
    peak, peak_src = peaks()
    achieved = blocks_per_rank * BYTES_PER_BLOCK / (ms_step * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("forward_B_exact_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "kernel": "sdct::forward_kernel<layout B, exact, X3>",
                "algorithmic_bytes_per_launch": blocks_per_rank * BYTES_PER_BLOCK}

    def rate(ms, nblocks, bytes_per_block=BYTES_PER_BLOCK, ranks=world):
        gbps = nblocks * bytes_per_block / (ms * 1e-3) / 1e9
        return {"us_per_launch": ms * 1e3, "blocks_per_s": ranks * nblocks / (ms * 1e-3), "gbps_per_gpu": gbps, "frac_of_hbm_peak": gbps / peak}

    # ---- parity of the timed bytes, checked in this run (rank 0; the first PF frames of the batch) ----------------------
    # EXACT: against the bytes of the unmodified reference (strict build, AVX2 path).  FAST / TENSOR are this repo's own
    # arithmetic: their distance from the EXACT bytes with the same X3 setting is reported as a rate, the
    # bound is <= 1 level.
    parity = None
    if rank == 0:
        PF = min(2, F)
        want, src = strict_reference_bytes(frames_np[:PF], lut, W, H)
        got = headline_out.view(F, -1)[:PF].cpu().numpy()
        parity = {"frames_checked": PF, "against": src, "exact_bytes_identical": bool(np.array_equal(got, want)),
                  "exact_bytes_differing": int(np.count_nonzero(got != want))}
        truth = ctx.forward(d_in[:PF], lut, sd.dense_job(sd.LAYOUT_B, W, H, frames=PF, flags=sd.EXACT)).cpu().numpy().astype(np.int16)
        for name, fl in (("fast", sd.FAST), ("tensor", sd.TENSOR)):
            lv = ctx.forward(d_in[:PF], lut, sd.dense_job(sd.LAYOUT_B, W, H, frames=PF, flags=fl)).cpu().numpy().astype(np.int16)
            d = np.abs(lv - truth)
            parity[name + "_mismatch_rate"] = float(np.mean(d != 0))
            parity[name + "_max_abs"] = int(d.max())
        parity["note"] = ("fast/tensor rates are vs EXACT without the X3 quirk (the true DCT) on this workload: uniform random pixels under "
                          "the default LUT, where layout B's quantiser scale reaches 2 684 per unit coefficient")

    # ---- e2e: HOST buffers through the reference-facing call (the simd_dct.h symbol exported by libsdct.so), H2D +
    # kernel + D2H inside the timed region.  The batch is handed over the way a caller of the reference would hand it
    # over: the F frames stacked into one plane of F*H rows (blocks are independent, so the bytes equal the per-frame
    # results), sizeY = 2*F*H because the reference visits only rows below sizeY/2 (src/simd_dct.cpp:2243).
    e2e = None
    if not args.no_e2e:
        h_in = torch.from_numpy(frames_np).pin_memory()
        h_out = torch.empty(F * H * W, dtype=torch.uint8).pin_memory()
        e2e_steps = max(3, min(args.steps, 10))
        nbytes = F * H * W

        def wall(fn, steps, warm=3):
            for _ in range(warm):
                fn()
            barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                fn()
            dt = max_over_ranks((time.perf_counter() - t0) / steps)
            barrier()
            return dt

        def shim_step(fn, sy):
            def step():
                rc = fn(h_in, h_out, lut, W, sy, 0, sy)
                assert rc == sd.sdr_Success, sd.last_error()
            return step
        dt = wall(shim_step(sd.simdDCT_EncodeQuantize32ReorderBuffer, 2 * F * H), e2e_steps)
        assert torch.equal(h_out.to(headline_out.device), headline_out), "host pipeline and device path disagree"
        # the platform's ceiling for this traffic: the same bytes as raw concurrent pinned copies (all ranks at once)
        s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
        d_a, d_b = torch.empty_like(d_in), torch.empty_like(d_out)

        def raw_copies():
            with torch.cuda.stream(s_up):
                d_a.copy_(h_in, non_blocking=True)
            with torch.cuda.stream(s_dn):
                h_out.copy_(d_b, non_blocking=True)
            s_up.synchronize()
            s_dn.synchronize()
        dt_raw = wall(raw_copies, e2e_steps)
        del d_a, d_b
        e2e = {"value": world * blocks_per_rank / dt, "unit": UNIT, "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": nbytes, "ms_per_step": dt * 1e3, "steps": e2e_steps,
               "gbps_each_way": nbytes / dt / 1e9,
               "copy_ceiling": {"ms_per_step": dt_raw * 1e3, "gbps_each_way_per_gpu": nbytes / dt_raw / 1e9, "aggregate_gbps_each_way": world * nbytes / dt_raw / 1e9,
                                "how": "the step's bytes as one pinned H2D and one pinned D2H copy on two streams, all ranks concurrently, wall clock, max over ranks"},
               "frac_of_copy_ceiling": dt_raw / dt,
               "api": "simdDCT_EncodeQuantize32ReorderBuffer (simd_dct.h shim of libsdct.so) on pinned host buffers; "
                      "synchronous call = chunked H2D / kernel / D2H pipeline, wall clock"}
        if not args.no_variants:
            # the other two entry points on the same bytes (layout C: the batch as two eyes of F*H/2 rows)
            other = {}
            dta = wall(shim_step(sd.simdDCT_EncodeQuantizeBuffer, 2 * F * H), max(3, e2e_steps // 2), warm=2)
            other["layout_A"] = {"ms_per_step": dta * 1e3, "blocks_per_s": world * blocks_per_rank / dta, "gbps_each_way": nbytes / dta / 1e9}
            if (F * H) % 16 == 0:
                dtc = wall(shim_step(sd.simdDCT_EncodeQuantizeReorderStereoBuffer, F * H), max(3, e2e_steps // 2), warm=2)
                other["layout_C"] = {"ms_per_step": dtc * 1e3, "blocks_per_s": world * blocks_per_rank / dtc, "gbps_each_way": nbytes / dtc / 1e9}
            # pageable caller memory (what an unmodified caller's malloc'd buffers are): staged through pinned memory
            p_in, p_out = frames_np, np.empty(F * H * W, dtype=np.uint8)

            def pageable():
                assert sd.simdDCT_EncodeQuantize32ReorderBuffer(p_in, p_out, lut, W, 2 * F * H, 0, 2 * F * H) == sd.sdr_Success
            dtp = wall(pageable, 3, warm=1)
            other["layout_B_pageable"] = {"ms_per_step": dtp * 1e3, "blocks_per_s": world * blocks_per_rank / dtp, "gbps_each_way": nbytes / dtp / 1e9}
            e2e["other_entry_points"] = other
            # one process driving every visible GPU through ONE host call (rank 0, the other ranks idle at the barrier)
            ndev = torch.cuda.device_count()
            if ndev > 1:
                if rank == 0:
                    ctxs = [sd.Context(d) for d in range(ndev)]
                    mj = sd.dense_job(sd.LAYOUT_B, W, H, frames=F, flags=flags)

                    def multi():
                        sd.forward_host_multi(ctxs, h_in, lut, mj, out=h_out)
                    for _ in range(2):
                        multi()
                    t0 = time.perf_counter()
                    for _ in range(5):
                        multi()
                    dtm = (time.perf_counter() - t0) / 5
                    e2e["single_process_all_gpus"] = {"gpus": ndev, "ms_per_step": dtm * 1e3, "blocks_per_s": blocks_per_rank / dtm,
                                                      "gbps_each_way": nbytes / dtm / 1e9,
                                                      "api": "sdct_forward_host_multi: one call, frames sharded over all devices, one host thread each"}
                    torch.cuda.set_device(local_rank)
                    assert torch.equal(h_out.to(headline_out.device), headline_out)
                    for c in ctxs:
                        c.close()
                    torch.cuda.set_device(local_rank)
                barrier()
        del h_in, h_out

    # ---- secondary measurements, keyed by BASELINE.json config (informative; `value` stays configs[1] EXACT) --------------
    variants = {}
    if not args.no_variants:
        vs = max(10, args.steps // 2)

        def tm(fn, steps=vs):
            ms, _, _ = timed(fn, steps, 3)
            return ms
        L = (("A", sd.LAYOUT_A), ("B", sd.LAYOUT_B), ("C", sd.LAYOUT_C))

        def exact_flags(layout):
            return sd.EXACT | (sd.QUIRK_AVX_X3 if layout == sd.LAYOUT_B else 0)

        # configs[1]: the headline batch, every layout / arithmetic / direction
        c1 = {}
        d_back = torch.empty_like(d_in)
        for lname, layout in L:
            j = sd.dense_job(layout, W, H, frames=F, flags=exact_flags(layout))
            jf = sd.dense_job(layout, W, H, frames=F, flags=sd.FAST)
            jt = sd.dense_job(layout, W, H, frames=F, flags=sd.TENSOR)   # tcgen05 form (own exact-integer definition, not the reference's fp32 rounding)
            c1[f"forward_{lname}_exact"] = rate(tm(lambda: ctx.forward(d_in, lut, j, out=d_out)), blocks_per_rank)
            c1[f"forward_{lname}_fast"] = rate(tm(lambda: ctx.forward(d_in, lut, jf, out=d_out)), blocks_per_rank)
            c1[f"forward_{lname}_tensor"] = rate(tm(lambda: ctx.forward(d_in, lut, jt, out=d_out)), blocks_per_rank)
            ctx.forward(d_in, lut, j, out=d_out)
            c1[f"inverse_{lname}"] = rate(tm(lambda: ctx.inverse(d_out, lut, j, out=d_back)), blocks_per_rank)
            c1[f"inverse_{lname}_tensor"] = rate(tm(lambda: ctx.inverse(d_out, lut, jt, out=d_back)), blocks_per_rank)
            c1[f"roundtrip_fused_{lname}_exact"] = rate(tm(lambda: ctx.roundtrip(d_in, lut, j, out=d_back)), blocks_per_rank)
            try:
                c1[f"roundtrip_fused_{lname}_tensor"] = rate(tm(lambda: ctx.roundtrip(d_in, lut, jt, out=d_back)), blocks_per_rank)
            except sd.SdctError as e:
                c1[f"roundtrip_fused_{lname}_tensor"] = {"unsupported": str(e)}
        variants["configs[1] 64x1080p luma, per GPU (weak)"] = c1
        del d_back

        # configs[2]: 256 frames 3840x2160x3 interleaved RGB -> 3 planar coefficient sets, per GPU (weak scaling)
        F2, H2, W2 = 256, 2160, 3840
        nb2 = F2 * 3 * (H2 // 8) * (W2 // 8)
        gen = torch.Generator(device="cuda").manual_seed(99 + rank)
        rgb = torch.randint(0, 256, (F2, H2, W2, 3), dtype=torch.uint8, device="cuda", generator=gen)
        coef2 = torch.empty(F2 * H2 * W2 * 3, dtype=torch.uint8, device="cuda")
        c2 = {}
        for lname, layout in L:
            for mname, fl in (("exact", exact_flags(layout)), ("tensor", sd.TENSOR)):
                if lname == "A" and mname == "exact":
                    continue
                j = sd.dense_job(layout, W2, H2, frames=F2, components=3, flags=fl)
                c2[f"forward_rgb_{lname}_{mname}"] = rate(tm(lambda: ctx.forward(rgb, lut, j, out=coef2), steps=5), nb2)
        # the colour front-end: RGB -> YCbCr fused into the same load
        jy = sd.dense_job(sd.LAYOUT_B, W2, H2, frames=F2, components=3, flags=sd.TENSOR | sd.COLOR_YCBCR)
        c2["forward_rgb_to_ycbcr_B_tensor"] = rate(tm(lambda: ctx.forward(rgb, lut, jy, out=coef2), steps=5), nb2)
        variants["configs[2] 256x4K interleaved RGB -> planar + forward, per GPU (weak)"] = c2
        # ... and the decode of those coefficient sets back to interleaved pixels (fp32; TENSOR where supported)
        back2 = torch.empty_like(rgb)
        for lname, layout in L[1:]:
            for mname, fl in (("fp32", exact_flags(layout)), ("tensor", sd.TENSOR)):
                j = sd.dense_job(layout, W2, H2, frames=F2, components=3, flags=fl)
                try:
                    c2[f"inverse_rgb_{lname}_{mname}"] = rate(tm(lambda: ctx.inverse(coef2, lut, j, out=back2), steps=5), nb2)
                except sd.SdctError as e:
                    c2[f"inverse_rgb_{lname}_{mname}"] = {"unsupported": str(e)}
        # the step after the path: per-plane statistics of the layout-C sets, one streaming read (64 B / block)
        jc = sd.dense_job(sd.LAYOUT_C, W2, H2, frames=F2, components=3, flags=sd.TENSOR)
        ctx.forward(rgb, lut, jc, out=coef2)
        c2["plane_stats_C (64 B read per block)"] = rate(tm(lambda: ctx.plane_stats(coef2, jc), steps=5), nb2, 64)
        del rgb, coef2, back2

        # configs[3]: inverse-only decode of 4K coefficient planes, 64 frames in total sharded by frame over the ranks
        # (strong scaling: the job is fixed, every rank decodes its contiguous frame range; no collective)
        from simd_dct_b200 import sharding
        F3 = 64
        c3 = {}
        for lname, layout in L:
            whole = sd.dense_job(layout, W2, H2, frames=F3, flags=exact_flags(layout))
            sub, _, _ = sharding.shard_frames(whole, rank, world)
            mine = int(sub.frames)
            px = torch.randint(0, 256, (max(mine, 1), H2, W2), dtype=torch.uint8, device="cuda", generator=gen)
            cf = ctx.forward(px, lut, sub) if mine else torch.empty(0, dtype=torch.uint8, device="cuda")
            for mname, fl in (("fp32", exact_flags(layout)), ("tensor", sd.TENSOR)):
                jj = sd.dense_job(layout, W2, H2, frames=mine, flags=fl)
                ms = tm(lambda: ctx.inverse(cf, lut, jj, out=px) if mine else None)
                r = rate(ms, F3 * (H2 // 8) * (W2 // 8) / world, ranks=world)
                r["frames_this_rank"] = mine
                c3[f"inverse_{lname}_{mname}"] = r
            del px, cf
        variants[f"configs[3] inverse-only, 64x4K planes sharded by frame over {world} GPU(s) (strong)"] = c3

        # configs[4]: 32768x32768x3 interleaved image, forward + inverse round trip, band ranges sharded over the ranks
        # (strong scaling; every rank holds only its own band range: 4096 / world bands)
        W5 = H5 = 32768
        bands = H5 // 8
        b0, b1 = sd.shard_range(bands, rank, world)
        rows = (b1 - b0) * 8
        nb5 = 3 * (b1 - b0) * (W5 // 8)
        img = torch.randint(0, 256, (max(rows, 8), W5, 3), dtype=torch.uint8, device="cuda", generator=gen)
        out5 = torch.empty_like(img)
        c5 = {}
        lut5 = ol.lut(255.0 * 4)   # a LUT under which layout B's round trip is meaningful (the default one saturates)
        tot5 = 3 * bands * (W5 // 8)
        for lname, layout in L[1:2]:
            je = sd.dense_job(layout, W5, rows, components=3, flags=sd.EXACT)
            jt = sd.dense_job(layout, W5, rows, components=3, flags=sd.TENSOR)
            r = rate(tm(lambda: ctx.roundtrip(img, lut5, je, out=out5), steps=5), tot5 / world, ranks=world)
            c5[f"roundtrip_fused_{lname}_exact"] = r
            try:
                c5[f"roundtrip_fused_{lname}_tensor"] = rate(tm(lambda: ctx.roundtrip(img, lut5, jt, out=out5), steps=5), tot5 / world, ranks=world)
            except sd.SdctError as e:
                c5[f"roundtrip_fused_{lname}_tensor"] = {"unsupported": str(e)}
            cf5 = torch.empty(rows * W5 * 3, dtype=torch.uint8, device="cuda")

            def two(j):
                ctx.forward(img, lut5, j, out=cf5)
                ctx.inverse(cf5, lut5, j, out=out5)
            c5[f"forward_then_inverse_{lname}_exact (256 B/block)"] = rate(tm(lambda: two(je), steps=5), tot5 / world, 2 * BYTES_PER_BLOCK, ranks=world)
            try:
                c5[f"forward_then_inverse_{lname}_tensor (256 B/block)"] = rate(tm(lambda: two(jt), steps=5), tot5 / world, 2 * BYTES_PER_BLOCK, ranks=world)
            except sd.SdctError as e:
                c5[f"forward_then_inverse_{lname}_tensor (256 B/block)"] = {"unsupported": str(e)}
            del cf5
        c5["bands_this_rank"] = b1 - b0
        del img, out5
        variants[f"configs[4] 32768x32768x3 round trip, bands sharded over {world} GPU(s) (strong)"] = c5

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": config, "gbps": value * BYTES_PER_BLOCK / 1e9, "roofline": roofline,
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "parity": parity, "variants": variants}
    if variants:
        c1 = variants["configs[1] 64x1080p luma, per GPU (weak)"]
        pick = lambda k: {kk: c1[k][kk] for kk in ("us_per_launch", "gbps_per_gpu", "frac_of_hbm_peak")} if "us_per_launch" in c1.get(k, {}) else None  # noqa: E731
        line["roofline_inverse"] = {"kernel": "inverse_kernel<B> (fp32) / inverse_tc_kernel<B> (tensor)", "fp32": pick("inverse_B"), "tensor": pick("inverse_B_tensor"),
                                    "note": "parity unpinned: the reference has no inverse"}
        line["roofline_roundtrip"] = {"kernel": "roundtrip_kernel<B> (fp32, fused) / roundtrip_tc_kernel<B> (tensor cores, fused: levels stay in shared / tensor memory)", "fp32": pick("roundtrip_fused_B_exact"),
                                      "tensor": pick("roundtrip_fused_B_tensor"), "algorithmic_bytes_per_block": BYTES_PER_BLOCK}
    if rank == 0:
        # the reference's CPU implementation beside it, on this box's host cores (bounded samples; the other ranks wait)
        n_cpu = F if world == 1 else min(F, 16)
        cpu, _ = time_cpu(W, H, frames_np[:n_cpu], lut, reps=3)
        cpu["single_thread"], _ = time_cpu(W, H, frames_np[:4], lut, reps=5, threads=1, budget_s=4.0)   # the reference's own method: best of N, one thread (src/main.cpp:510-530)
        if world == 1:
            cpu["layout_A"], _ = time_cpu(W, H, frames_np[:16], lut, reps=2, layout="A", budget_s=5.0)
            cpu["layout_C"], _ = time_cpu(W, H, frames_np[:16], lut, reps=2, layout="C", budget_s=5.0)
            cpu["layout_C_single_thread"], _ = time_cpu(W, H, frames_np[:2], lut, reps=3, layout="C", threads=1, budget_s=4.0)
        line["cpu_baseline"] = cpu
    barrier()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
