#!/usr/bin/env python
"""Raw host<->device copy ceiling of the box: N processes (one per GPU), each moving the bench step's bytes
(132.7 MB pinned H2D and 132.7 MB pinned D2H, concurrently on two streams), all at once.

    python tools/pcie_probe.py --gpus N [--out profiles/r02_copy_ceiling.json] [--affinity]

Prints / writes per-GPU and aggregate GB/s each way for h2d only, d2h only and both, plus what `nvidia-smi topo`
and the CPU say about the host side.  --affinity pins process r to its share of the host cores (cores are dealt out
evenly); the default leaves the scheduler alone, as torchrun does.  This is the ceiling bench.py's e2e leg is
compared with (e2e.frac_of_copy_ceiling); it contains no kernel."""
import argparse
import json
import multiprocessing as mp
import os
import subprocess
import time

N_BYTES = 132710400


def worker(rank, world, affinity, chunk_mb, barrier, q):
    import torch
    if affinity:
        cores = sorted(os.sched_getaffinity(0))
        per = max(1, len(cores) // world)
        os.sched_setaffinity(0, set(cores[rank * per:(rank + 1) * per]) or set(cores))
    torch.cuda.set_device(rank)
    h_in = torch.empty(N_BYTES, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(N_BYTES, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(N_BYTES, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(N_BYTES, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    chunk = chunk_mb << 20 if chunk_mb else N_BYTES

    def run(h2d, d2h, reps):
        torch.cuda.synchronize()
        barrier.wait()
        t0 = time.perf_counter()
        for _ in range(reps):
            for off in range(0, N_BYTES, chunk):
                e = min(N_BYTES, off + chunk)
                if h2d:
                    with torch.cuda.stream(s1):
                        d_in[off:e].copy_(h_in[off:e], non_blocking=True)
                if d2h:
                    with torch.cuda.stream(s2):
                        h_out[off:e].copy_(d_out[off:e], non_blocking=True)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
        barrier.wait()
        return dt
    res = {}
    for name, a, b in (("h2d", 1, 0), ("d2h", 0, 1), ("both", 1, 1)):
        run(a, b, 2)
        res[name] = run(a, b, 10)
    q.put((rank, res))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--out", default=None)
    ap.add_argument("--affinity", action="store_true")
    ap.add_argument("--chunk-mb", type=int, default=0, help="issue the copies in chunks of this size (0 = one copy per direction)")
    a = ap.parse_args()
    mp.set_start_method("spawn")
    barrier, q = mp.Barrier(a.gpus), mp.Queue()
    procs = [mp.Process(target=worker, args=(r, a.gpus, a.affinity, a.chunk_mb, barrier, q)) for r in range(a.gpus)]
    for p in procs:
        p.start()
    got = dict(q.get() for _ in procs)
    for p in procs:
        p.join()
    out = {"gpus": a.gpus, "bytes_each_way_per_gpu": N_BYTES, "affinity": a.affinity, "chunk_mb": a.chunk_mb, "modes": {}}
    for name in ("h2d", "d2h", "both"):
        slow = max(got[r][name] for r in got)
        out["modes"][name] = {"per_gpu_gbps_each_way": [round(N_BYTES / got[r][name] / 1e9, 2) for r in sorted(got)],
                              "aggregate_gbps_each_way": round(a.gpus * N_BYTES / slow / 1e9, 2), "slowest_ms": round(slow * 1e3, 3)}
        print(name, out["modes"][name], flush=True)
    try:
        out["host"] = {"cpus": len(os.sched_getaffinity(0)),
                       "lscpu": subprocess.run("lscpu | grep -E 'Model name|^CPU\\(s\\)|NUMA|Socket'", shell=True, capture_output=True, text=True).stdout.strip().splitlines(),
                       "topo": subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout.strip().splitlines()[:14]}
    except Exception as e:  # noqa: BLE001
        out["host"] = {"error": repr(e)}
    if a.out:
        prev = {}
        if os.path.exists(a.out):
            try:
                prev = json.load(open(a.out))
            except Exception:  # noqa: BLE001
                prev = {}
        prev[f"gpus={a.gpus},affinity={int(a.affinity)},chunk_mb={a.chunk_mb}"] = out
        json.dump(prev, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
