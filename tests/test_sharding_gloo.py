"""The N > 1 path on CPU: two gloo ranks shard a batch by frames and one image by bands exactly as
bench.py / a multi-GPU caller does, run their share (with the CPU oracle standing in for the
kernels, which need a GPU), and the gathered bytes equal the single-process result."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, layout, q):
    import oraclelib as ol
    import simd_dct_b200 as sd
    from simd_dct_b200 import sharding
    from conftest import make_image
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert sharding.env_rank() == (rank, rank, world)
        o = ol.Oracle()
        lut = ol.lut(8)
        f, h, w = 5, 32, 128
        frames = np.stack([make_image("smooth", h, w, seed=s) for s in range(f)])
        # --- batch sharded by frames -------------------------------------------------------
        job = sd.dense_job(layout, w, h, frames=f)
        sub, poff, coff = sharding.shard_frames(job, rank, world)
        assert poff % (h * w) == 0 and coff % (h * w) == 0
        f0 = poff // (h * w)
        mine = [o.forward(layout, frames[f0 + k], lut, quirks=ol.QUIRK_AVX_X3) for k in range(int(sub.frames))]
        mine = np.concatenate(mine) if mine else np.zeros(0, np.uint8)
        parts = [None] * world
        dist.all_gather_object(parts, (coff, mine))
        full = np.zeros(f * h * w, np.uint8)
        for off, p in parts:
            full[off:off + p.size] = p
        want = np.concatenate([o.forward(layout, frames[k], lut, quirks=ol.QUIRK_AVX_X3) for k in range(f)])
        ok_frames = bool(np.array_equal(full, want))
        # --- one image sharded by bands ----------------------------------------------------
        img = make_image("rand", 72, 128, seed=9)   # 9 bands over 2 ranks: 5 + 4
        job = sd.dense_job(layout, 128, 72)
        sub = sharding.shard_bands(job, rank, world)
        b0, b1 = int(sub.band_begin), int(sub.band_end)
        whole = o.forward(layout, img, lut, quirks=ol.QUIRK_AVX_X3)
        piece = o.forward(layout, img[b0 * 8:b1 * 8], lut, quirks=ol.QUIRK_AVX_X3)
        spans = [None] * world
        dist.all_gather_object(spans, (b0, b1, piece))
        out = np.zeros_like(whole)
        bpr = 128 // 8
        for (s0, s1, p) in spans:
            n0, n1 = s0 * bpr, s1 * bpr
            if layout == sd.LAYOUT_C:
                ps, pl = whole.size // 64, (n1 - n0)
                for i in range(64):
                    out[i * ps + n0:i * ps + n1] = p[i * pl:(i + 1) * pl]
            else:
                out[n0 * 64:n1 * 64] = p
        ok_bands = bool(np.array_equal(out, whole)) and sorted((s[0], s[1]) for s in spans) == [(0, 5), (5, 9)]
        t = sharding.max_over_ranks(1.0 + rank, dist)
        q.put((rank, ok_frames, ok_bands, t))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("layout", [0, 1, 2])
def test_two_rank_sharding_matches_single_process(layout):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, layout, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_frames, ok_bands, t in res:
        assert ok_frames and ok_bands, (rank, ok_frames, ok_bands)
        assert t == 2.0  # max over ranks of (1 + rank)
