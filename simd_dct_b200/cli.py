"""Benchmark / inspection CLI on the GPU, with the reference CLI's workflow (src/main.cpp): raw 8-bit grayscale file in,
each selected encoder timed `--runs` times through the simd_dct.h entry points (host buffers -> libsdct.so), the
reference's MiB/s report (best, mean, mean -+ sigma; per FILE size like the reference, src/main.cpp:34-80) and an
optional `--to` dump of the last output.

    python -m simd_dct_b200.cli <raw_grayscale_image_file> <resolutionX> <resolutionY>
        --to <file>  --quality <n>  --runs <n>  --mode enc-quant | enc-quant32 | enc-quant-stereo   (as src/main.cpp:84-102)
        --device <n | all | 0,1,..>   GPU(s) the call is sharded over ($SDCT_DEVICE / $SDCT_DEVICES)
        --flags exact | fast | tensor arithmetic (exact = the reference's bytes; enc-quant32 then carries the X3 quirk)
        --no-quirk                    exact arithmetic without the reference's X3 sign (the true DCT)
        --sse-quirk                   enc-quant-stereo with the bytes of the reference's SSE stereo encoders (what its dispatcher
                                      runs on an x86-64 CPU, src/simd_dct.cpp:78-83) instead of its scalar path's; X % 16 == 0
        --inverse                     decode the coefficients back to pixels after the (last) encode; --to then dumps
                                      the pixels (the reference has no inverse: this repo's definition)
        --resident                    also time the kernel alone on device-resident buffers (CUDA events)

What the reference's --cpu-core / --max-simd select (a CPU core, an instruction set) has no meaning here and is
accepted and ignored.  clk/byte is replaced by blocks/s and GB/s (64 B read + 64 B written per 8x8 block)."""
from __future__ import annotations

import argparse
import math
import os
import sys
import time

import numpy as np

MODES = {  # name -> (title as the reference prints it, layout name, shim function name)
    "enc-quant": ("Encode Quantize", "A", "simdDCT_EncodeQuantizeBuffer"),
    "enc-quant32": ("Encode Quantize 32 Reorder", "B", "simdDCT_EncodeQuantize32ReorderBuffer"),
    "enc-quant-stereo": ("EncQuantReordStereo", "C", "simdDCT_EncodeQuantizeReorderStereoBuffer"),
}


def perf_line(ns, file_size, blocks):
    """the reference's report (src/main.cpp:34-80) with the clk/byte columns replaced by blocks/s and GB/s"""
    mib = file_size / (1024.0 * 1024.0)
    best, mean = min(ns), sum(ns) / len(ns)
    sd = math.sqrt(sum((x - mean) ** 2 for x in ns) / (len(ns) - 1)) if len(ns) > 1 else 0.0
    rate = lambda t: mib / (t * 1e-9)  # noqa: E731
    lo = rate(mean + sd)
    hi = rate(mean - sd) if mean > sd else float("inf")
    return ("| %8.2f MiB/s | %8.2f MiB/s (%8.2f ~ %8.2f) | %9.3e blocks/s | %7.2f GB/s"
            % (rate(best), rate(mean), lo, hi, blocks / (best * 1e-9), blocks * 128 / best))


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="python -m simd_dct_b200.cli", description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("file")
    ap.add_argument("resolutionX", type=int)
    ap.add_argument("resolutionY", type=int)
    ap.add_argument("--to", dest="out_file")
    ap.add_argument("--quality", type=int)
    ap.add_argument("--runs", type=int, default=128)
    ap.add_argument("--mode", action="append", choices=sorted(MODES))
    ap.add_argument("--cpu-core")
    ap.add_argument("--max-simd")
    ap.add_argument("--device", default=None)
    ap.add_argument("--flags", default="exact", choices=["exact", "fast", "tensor"])
    ap.add_argument("--no-quirk", action="store_true")
    ap.add_argument("--sse-quirk", action="store_true")
    ap.add_argument("--inverse", action="store_true")
    ap.add_argument("--resident", action="store_true")
    a = ap.parse_args(argv)
    if a.resolutionX <= 0 or a.resolutionY <= 0:
        print("Invalid Resolution Specified. Aborting.")
        return 1
    if not 1 <= a.runs <= 1024:   # src/main.cpp:22
        print("Invalid Parameter.")
        return 1
    if a.device is not None:      # read when the process-wide contexts are created, i.e. at the first call
        os.environ.pop("SDCT_DEVICES", None)
        if a.device == "all" or "," in a.device:
            os.environ["SDCT_DEVICES"] = a.device
        else:
            os.environ["SDCT_DEVICE"] = a.device
    try:
        data = np.fromfile(a.file, dtype=np.uint8)
    except OSError:
        print("Failed to read file.")
        return 1
    if data.size == 0:
        print("Invalid File size / failed to read file.")
        return 1

    import simd_dct_b200 as sd
    from simd_dct_b200 import _lib
    _lib.load()   # fails loudly when libsdct.so has not been built: there is no CPU path behind this tool
    lut = sd.quality_lut(1.0).copy()
    if a.quality is not None:     # src/main.cpp:214-217: every entry is multiplied by the quality level
        lut = (lut * np.float32(a.quality)).astype(np.float32)
    flags = {"exact": sd.EXACT, "fast": sd.FAST, "tensor": sd.TENSOR}[a.flags]
    x, y = a.resolutionX, a.resolutionY
    if data.size < x * y:
        print("File smaller than resolutionX * resolutionY. Aborting.")
        return 1
    file_size = data.size
    out = np.zeros(file_size, dtype=np.uint8)   # (the reference mallocs it and never clears it)
    half_bands = (y // 2 + 7) // 8
    modes = a.mode or list(MODES)               # src/main.cpp:446-447: none selected = all
    last = None
    print("device(s): %s, arithmetic: %s%s" % (a.device or os.environ.get("SDCT_DEVICES") or os.environ.get("SDCT_DEVICE", "0"), a.flags,
                                               "" if a.flags != "exact" else (" (true DCT)" if a.no_quirk else " (reference bytes)")))
    for m in ("enc-quant", "enc-quant32", "enc-quant-stereo"):
        if m not in modes:
            continue
        title, lname, fname = MODES[m]
        fn = getattr(sd, fname)
        layout = {"A": sd.LAYOUT_A, "B": sd.LAYOUT_B, "C": sd.LAYOUT_C}[lname]
        mode_flags = flags | (sd.QUIRK_AVX_X3 if (a.flags == "exact" and layout == sd.LAYOUT_B and not a.no_quirk) else 0)
        mode_flags |= sd.QUIRK_SSE if (a.flags == "exact" and layout == sd.LAYOUT_C and a.sse_quirk) else 0
        sd.set_reference_flags(mode_flags)
        blocks = (x // 8) * half_bands * (2 if lname == "C" else 1)
        ns = []
        for run in range(a.runs):
            t0 = time.perf_counter_ns()
            rc = fn(data, out, lut, x, y, 0, y)
            dt = time.perf_counter_ns() - t0
            ns.append(dt)
            sys.stdout.write("\r%3d: %5.2f MiB/s" % (run + 1, (file_size / (1024.0 * 1024.0)) / (dt * 1e-9)))
            if rc != sd.sdr_Success:
                print("Encode failed with error code 0x%X (%s). Aborting." % (rc, sd.last_error()))
                return 1
        print("\r%-30s %s" % (title, perf_line(ns, file_size, blocks)))
        last = (layout, lname, blocks)
        if a.resident:
            import torch
            ctx = sd.Context(int(os.environ.get("SDCT_DEVICE", "0")))
            eyes = 2 if lname == "C" else 1
            job = sd.dense_job(layout, x, half_bands * 8, eyes=eyes, flags=mode_flags)
            if lname == "C":
                job.eye_stride, job.coef_plane_stride = x * y // 2, x * y // 64
            with torch.cuda.device(ctx.device):
                d_in = torch.from_numpy(data[:x * y]).cuda()
                d_out = torch.zeros(x * y, dtype=torch.uint8, device="cuda")
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.runs + 1)]
                ctx.forward(d_in, lut, job, out=d_out)
                ev[0].record()
                for run in range(a.runs):
                    ctx.forward(d_in, lut, job, out=d_out)
                    ev[run + 1].record()
                torch.cuda.synchronize()
                kns = [ev[i].elapsed_time(ev[i + 1]) * 1e6 for i in range(a.runs)]
            print("%-30s %s" % ("  kernel only (resident)", perf_line(kns, file_size, blocks)))
            ctx.close()
    if a.inverse and last is not None:
        layout, lname, blocks = last
        ctx = sd.Context(int(os.environ.get("SDCT_DEVICE", "0")))
        eyes = 2 if lname == "C" else 1
        job = sd.dense_job(layout, x, half_bands * 8, eyes=eyes, flags=sd.TENSOR if a.flags == "tensor" else sd.EXACT)
        if lname == "C":
            job.eye_stride, job.coef_plane_stride = x * y // 2, x * y // 64
        pix = np.zeros(file_size, dtype=np.uint8)
        ns = []
        for run in range(a.runs):
            t0 = time.perf_counter_ns()
            ctx.inverse_host(out, lut, job, out=pix)
            ns.append(time.perf_counter_ns() - t0)
        print("%-30s %s" % ("Decode (" + lname + ")", perf_line(ns, file_size, blocks)))
        ctx.close()
        out = pix
    if a.out_file:
        try:
            out[:x * y].tofile(a.out_file)    # src/main.cpp:604: sizeX * sizeY bytes
        except OSError:
            print("Failed to open destination file. Aborting.\n")
            return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
