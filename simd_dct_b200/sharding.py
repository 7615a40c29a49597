"""Multi-GPU partitioning of the hot path: one process per GPU, every rank works on its own frames
(batch configs) or its own band range (one huge image); blocks are independent, so there is no
data-path collective.  torch.distributed is used only for rendezvous,
barriers and the max-over-ranks timing reduction."""
from __future__ import annotations

import copy
import os
from typing import Tuple

from .api import Job, shard_range


def env_rank() -> Tuple[int, int, int]:
    """(rank, local_rank, world_size) as torchrun exports them"""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def shard_frames(job: Job, rank: int, world: int):
    """-> (sub_job, pixel_byte_offset, coef_byte_offset) for this rank's contiguous frame range"""
    b, e = shard_range(int(job.frames), rank, world)
    sub = copy.copy(job)
    sub.frames = e - b
    return sub, b * int(job.pixel_frame_stride), b * int(job.coef_frame_stride)


def shard_bands(job: Job, rank: int, world: int):
    """-> sub_job whose [band_begin, band_end) is this rank's share of the job's band window; base
    pointers stay those of the whole image (every layout places a band's output independently of
    which bands are processed, as the reference's startY/endY window does, src/simd_dct.cpp:268-282)"""
    total = int(job.height) // 8
    lo = int(job.band_begin)
    hi = int(job.band_end) if int(job.band_end) else total
    b, e = shard_range(hi - lo, rank, world)
    sub = copy.copy(job)
    sub.band_begin, sub.band_end = lo + b, lo + e
    if e == b:          # empty share: band_end = 0 would mean "to the end"
        sub.frames = 0
    return sub


def max_over_ranks(value: float, dist=None, device=None) -> float:
    """the slowest rank's time (multi-GPU numbers are the max over ranks, never a wall clock)"""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([value], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
