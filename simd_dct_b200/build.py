"""Builds simd_dct_b200/libsdct.so (CUDA kernels for sm_100a + the C ABI + the simd_dct.h shim)
with nvcc, in-tree.  nvcc cross-compiles without a GPU; the .so travels to the GPU box."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsdct.so")
KERNEL_PARTS = 8  # sdct_kernels.cu is compiled once per -DSDCT_PART (parallel template instantiation)
SOURCES = ["sdct_kernels.cu", "sdct_tc.cu", "sdct_stats.cu", "sdct_noalloc.cu", "sdct_tc_matrix.cpp", "sdct_api.cpp", "sdct_host.cpp", "simd_dct_shim.cpp"]
HEADERS = ["sdct_kernels.h", "sdct_math.cuh", "sdct_io.cuh", "sdct_internal.h", "../../include/sdct.h", "../../include/simd_dct.h"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-Wall", "-x", "cu"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found - cannot build libsdct.so")


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False, defs=None, out: str = LIB, only_kernels: bool = False) -> str:
    """defs / out: experiment builds (tools/build_variants.py) - extra -D switches, written to another .so that
    $SDCT_LIB selects at load time; the product build is the default call."""
    if defs is None and out == LIB and not force and not _stale():
        return LIB
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build" if out == LIB else os.path.join("build", os.path.splitext(os.path.basename(out))[0]))
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    units = [(src, None) for src in SOURCES if src != "sdct_kernels.cu"] + [("sdct_kernels.cu", k) for k in range(KERNEL_PARTS)]
    for src, part in units:
        stem = os.path.splitext(src)[0] + ("" if part is None else ".part%d" % part)
        obj = os.path.join(objdir, stem + ".o")
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ([] if part is None else ["-DSDCT_PART=%d" % part])
        if os.environ.get("SDCT_NVCC_DEFS"):         # experiments: extra -D switches
            cmd += ["-D" + d for d in os.environ["SDCT_NVCC_DEFS"].split(",")]
        cmd += ["-D" + d for d in (defs or [])]
        if os.environ.get("SDCT_TC_TRACE") == "1":   # profiling build: role stamps + ablation switches in sdct_tc.cu
            cmd.append("-DSDCT_TC_TRACE=1")
        cmd += ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, p in procs:
        log, _ = p.communicate()
        if verbose and log:
            print(log)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), log))
    cmd = [nvcc, "-shared", "-o", out] + objs
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
