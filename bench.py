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
def cpu_reference_runner(w, h):
    """returns (kind, isa, fn(frames_u8[F,H,W], lut, out[F,H*W], threads)) timing the reference's own
    CPU implementation of the path (oracle/_ref, the unmodified reference built with its own
    flags) or, if that did not travel, the oracle port."""
    import ctypes as C
    import oraclelib as ol
    from concurrent.futures import ThreadPoolExecutor

    if ol.have_reference("fast"):
        ref = ol.Reference("fast")
        flags = ol.cpu_flags()
        isa = "avx512vl" if {"avx512vl", "avx512f"} <= flags else ("avx2" if "avx2" in flags else None)
        if isa:
            ref.set_isa(isa)
            fn = ref.fn["B"]
            kind = "reference"

            def one(src, dst, lut):
                # full H-row plane through the reference: sizeY = 2H (only rows < sizeY/2 are visited)
                fn(src.ctypes.data_as(ol._u8p), dst.ctypes.data_as(ol._u8p), lut.ctypes.data_as(ol._f32p), w, 2 * h, 0, 2 * h)
        else:
            isa = None
    else:
        isa = None
    if isa is None:
        o = ol.Oracle()
        kind, isa = "port", "scalar-c"

        def one(src, dst, lut):
            o.lib.sdct_oracle_forward(ol.LAYOUT_B, ol.MODE_EXACT, ol.QUIRK_AVX_X3, src.ctypes.data_as(ol._u8p), w, 1, w, h, 1, 0,
                                      lut.ctypes.data_as(ol._f32p), dst.ctypes.data_as(ol._u8p), 0)

    def run(frames, lut, out, threads):
        n = frames.shape[0]

        def work(t):
            for f in range(t, n, threads):
                one(frames[f], out[f], lut)
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(work, range(threads)))
    return kind, isa, run


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def time_cpu(w, h, frames_np, lut, reps):
    kind, isa, run = cpu_reference_runner(w, h)
    threads = host_threads()
    out = np.empty((frames_np.shape[0], h * w), dtype=np.uint8)
    run(frames_np, lut, out, threads)  # warm
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        run(frames_np, lut, out, threads)
        best = min(best, time.perf_counter() - t0)
    blocks = frames_np.shape[0] * (h // 8) * (w // 8)
    return {"value": blocks / best, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"{frames_np.shape[0]} frames {w}x{h}, layout B, {isa}, best of {reps}, {threads} threads"}, out


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

    # e2e: HOST buffers through the reference-facing call (the simd_dct.h symbol exported by
    # libsdct.so), H2D + kernel + D2H inside the timed region.  The batch is handed over the way a
    # caller of the reference would hand it over: the F frames stacked into one plane of F*H rows
    # (blocks are independent, so the bytes equal the per-frame results), sizeY = 2*F*H because the
    # reference visits only rows below sizeY/2 (src/simd_dct.cpp:2243).
    e2e = None
    if not args.no_e2e:
        h_in = torch.from_numpy(frames_np).pin_memory()
        h_out = torch.empty(F * H * W, dtype=torch.uint8).pin_memory()
        e2e_steps = max(3, min(args.steps, 10))

        def e2e_step():
            rc = sd.simdDCT_EncodeQuantize32ReorderBuffer(h_in, h_out, lut, W, 2 * F * H, 0, 2 * F * H)
            assert rc == sd.sdr_Success, sd.last_error()
        for _ in range(3):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        dt = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
        barrier()
        e2e = {"value": world * blocks_per_rank / dt, "unit": UNIT, "h2d_bytes_per_step": F * H * W,
               "d2h_bytes_per_step": F * H * W, "ms_per_step": dt * 1e3, "steps": e2e_steps,
               "gbps_each_way": F * H * W / dt / 1e9,
               "api": "simdDCT_EncodeQuantize32ReorderBuffer (simd_dct.h shim of libsdct.so) on pinned host buffers; "
                      "synchronous call = chunked H2D / kernel / D2H pipeline, wall clock"}
        assert torch.equal(h_out.cuda(), d_out), "host pipeline and device path disagree"

    # secondary measurements (same batch; informative, not the headline)
    variants = {}
    if not args.no_variants:
        vs = max(10, args.steps // 2)

        def add(name, fn, bytes_per_block=BYTES_PER_BLOCK, nblocks=blocks_per_rank):
            ms, _, _ = timed(fn, vs, 3)
            variants[name] = {"blocks_per_s": world * nblocks / (ms * 1e-3), "ms_per_step": ms,
                              "gbps_per_gpu": nblocks * bytes_per_block / (ms * 1e-3) / 1e9,
                              "frac_of_hbm_peak": nblocks * bytes_per_block / (ms * 1e-3) / 1e9 / peak}
        for lname, layout in (("A", sd.LAYOUT_A), ("B", sd.LAYOUT_B), ("C", sd.LAYOUT_C)):
            fl = sd.EXACT | (sd.QUIRK_AVX_X3 if layout == sd.LAYOUT_B else 0)
            j = sd.dense_job(layout, W, H, frames=F, flags=fl)
            if layout != sd.LAYOUT_B:
                add(f"forward_{lname}_exact", lambda j=j: ctx.forward(d_in, lut, j, out=d_out))
            jf = sd.dense_job(layout, W, H, frames=F, flags=sd.FAST)
            add(f"forward_{lname}_fast", lambda jf=jf: ctx.forward(d_in, lut, jf, out=d_out))
            jt = sd.dense_job(layout, W, H, frames=F, flags=sd.TENSOR)   # tcgen05 form (own exact-integer definition, not the reference's fp32 rounding)
            add(f"forward_{lname}_tensor", lambda jt=jt: ctx.forward(d_in, lut, jt, out=d_out))
            ctx.forward(d_in, lut, j, out=d_out)
            d_back = torch.empty_like(d_in)
            add(f"inverse_{lname}", lambda j=j, d_back=d_back: ctx.inverse(d_out, lut, j, out=d_back))
            add(f"inverse_{lname}_tensor", lambda jt=jt, d_back=d_back: ctx.inverse(d_out, lut, jt, out=d_back))
            add(f"roundtrip_fused_{lname}_exact", lambda j=j, d_back=d_back: ctx.roundtrip(d_in, lut, j, out=d_back))

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": config, "gbps": value * BYTES_PER_BLOCK / 1e9, "roofline": roofline,
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "variants": variants}
    if rank == 0 and world == 1:
        cpu, cpu_out = time_cpu(W, H, frames_np, lut, reps=3)
        line["cpu_baseline"] = cpu
        # the timed GPU output is the reference's output (when the reference itself is the baseline)
        if cpu["kind"] == "reference" and "avx2" in cpu["sample"]:
            line["cpu_baseline"]["bytes_identical_to_gpu"] = bool(np.array_equal(cpu_out.reshape(-1), d_out.cpu().numpy()))
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
