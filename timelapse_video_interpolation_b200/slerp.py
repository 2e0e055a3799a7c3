"""First/last-frame latent slerp.

The reference file `frame_slerp_interp.py` is 0 bytes (only unused flags survive:
src/train_decouples.py:1566-1567, src/test.py:648-649), so the arithmetic is specified here and in
oracle/slerp_ref.py and implemented by csrc/slerp.cu:

    c = <v0, v1> / (|v0| |v1|)        over all elements of a clip, accumulated in fp64
    |c| > dot_threshold (0.9995) or a zero vector  ->  out_i = (1 - t_i) v0 + t_i v1
    otherwise  theta = acos(c),  out_i = sin((1 - t_i) theta)/sin(theta) v0 + sin(t_i theta)/sin(theta) v1
    t_i = i / (T - 1), i = 0..T-1;  coefficients are computed in fp64, rounded to fp32, and each
    output element is fl(fl(a*v0) + fl(b*v1)) (two roundings of the products, no FMA).
"""
from __future__ import annotations

import torch

from . import ops

DOT_THRESHOLD = 0.9995


def slerp(v0: torch.Tensor, v1: torch.Tensor, num_frames: int, dot_threshold: float = DOT_THRESHOLD) -> torch.Tensor:
    """v0, v1: [clips, ...] fp32 CUDA latents of the first / last frame -> [clips, num_frames, ...]."""
    if num_frames < 2:
        raise ValueError("num_frames must be >= 2")
    return ops.slerp_f32(v0, v1, num_frames, dot_threshold)
