"""Clip sharding across the GPUs of one box.

Clips never interact on the inference path (frames of a clip are coupled by temporal attention, pixels
of a frame by spatial attention, but the clip index is a pure batch dimension, reference
src/pipelines/pipeline_stable_video_diffusion_efficient_interp.py:542), so multi-GPU is one process
per GPU, clip i -> rank i mod world, weights replicated and NO collective on the data path.  The only
communication is the timing reduction of a benchmark (max over ranks) and, in training, DDP's own
gradient all-reduce.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_clips(n_clips: int, rank: int, world: int) -> list[int]:
    """Indices of the clips rank `rank` owns (round-robin, like one process per CUDA_VISIBLE_DEVICES
    in reference src/test_base.sh:17)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, n_clips, world))


def max_over_ranks(value: float, device=None) -> float:
    """The job-level time of a step is the slowest rank's (device-timed) value."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def sum_over_ranks(value: float, device=None) -> float:
    """Units processed by the whole job (clips, clip-frames) = sum of the per-rank counts."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t[0])
