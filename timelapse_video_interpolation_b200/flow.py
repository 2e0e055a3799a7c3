"""Optical-flow path: drop-in `compute_optical_flow_farneback`, `cosine_similarity` and the backward-warp consistency check.

`cosine_similarity(flow_a, flow_b)` has the name, arity and return type of the reference function
(dataset_construction/step4_flow_check.py:25-34): two (H, W, 2) float32 numpy arrays in OpenCV's HWC
layout -> np.float32 scalar.  CUDA tensors (optionally batched) are accepted too and then a CUDA
tensor is returned without a host sync.

`backward_warp_check` is the north-star extension (the reference file has no warp);
its spec is F.grid_sample(bilinear, zeros, align_corners=True) + the usual forward/backward
consistency inequality, see oracle/flow_ref.py.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops

OCC_ALPHA = 0.01
OCC_BETA = 0.5


def compute_optical_flow_farneback(img1, img2):
    """Drop-in for the reference function of the same name (dataset_construction/step4_flow_check.py:13-22): two BGR uint8
    images (H, W, 3) -> flow (H, W, 2) float32, the computation of cv2.cvtColor(BGR2GRAY) + cv2.calcOpticalFlowFarneback(
    0.5, 3, 15, 3, 5, 1.2, 0) on the GPU.  numpy in -> numpy out; CUDA uint8 tensors in -> CUDA tensor out (no host sync)."""
    if isinstance(img1, np.ndarray):
        a = torch.from_numpy(np.ascontiguousarray(img1, dtype=np.uint8)).cuda(non_blocking=True)
        b = torch.from_numpy(np.ascontiguousarray(img2, dtype=np.uint8)).cuda(non_blocking=True)
        return ops.farneback(a, b).cpu().numpy()
    return ops.farneback(img1, img2)


def check_clip_flows(frames, threshold: float = 0.5):
    """The per-video body of dataset_construction/step4_flow_check.py:55-81 without leaving the device: Farneback flow of
    every consecutive frame pair, cosine similarity of every consecutive flow pair, outlier flag `sim < threshold` (:80).
    frames: [N, H, W, 3] BGR uint8 (numpy or CUDA tensor), N >= 3 (:50).  Returns (flows [N-1, H, W, 2] fp32 CUDA,
    sims [N-2] fp32 CUDA, flagged [N-2] bool CUDA); nothing synchronises the host."""
    t = torch.from_numpy(np.ascontiguousarray(frames, dtype=np.uint8)).cuda(non_blocking=True) if isinstance(frames, np.ndarray) else frames
    if t.dim() != 4 or t.shape[0] < 3 or t.shape[-1] != 3 or t.dtype != torch.uint8:
        raise ValueError("frames must be [N >= 3, H, W, 3] uint8 (the reference skips shorter clips, :50-51)")
    N, H, W, _ = t.shape
    flows = ops.farneback_clip(t)   # every frame's pyramid is built once (half the per-pair preparation work)
    sims = ops.flow_cosine_f32(flows[:-1], flows[1:], hwc=True)
    return flows, sims, sims < threshold


def cosine_similarity(flow_a, flow_b):
    if isinstance(flow_a, np.ndarray):
        a = torch.from_numpy(np.ascontiguousarray(flow_a, dtype=np.float32))
        b = torch.from_numpy(np.ascontiguousarray(flow_b, dtype=np.float32))
        if a.shape != b.shape or a.numel() % 2 != 0:
            raise ValueError("flow fields must have the same shape and 2 components per pixel")
        # the reference reshapes to (-1, 2): any shape whose flattening is pixel-major works
        a = a.reshape(1, 1, -1, 2).cuda(non_blocking=True)
        b = b.reshape(1, 1, -1, 2).cuda(non_blocking=True)
        return np.float32(ops.flow_cosine_f32(a, b, hwc=True).cpu().numpy()[0])
    a, b = flow_a, flow_b
    single = a.dim() == 3
    if single:
        a, b = a.unsqueeze(0), b.unsqueeze(0)
    out = ops.flow_cosine_f32(a, b, hwc=True)
    return out[0] if single else out


def cosine_similarity_chw(flow_a: torch.Tensor, flow_b: torch.Tensor) -> torch.Tensor:
    """Same reduction for planar [B, 2, H, W] flows (the layout the warp kernel uses)."""
    return ops.flow_cosine_f32(flow_a, flow_b, hwc=False)


def backward_warp_check(img2: torch.Tensor, f_fwd: torch.Tensor, f_bwd: torch.Tensor, alpha: float = OCC_ALPHA,
                        beta: float = OCC_BETA):
    """Returns (warped [B,C,H,W] fp32, occluded [B,H,W] bool, valid [B,H,W] bool)."""
    warped, mask = ops.backward_warp_f32(img2, f_fwd, f_bwd, alpha, beta)
    return warped, (mask & 1).bool(), (mask & 2).bool()
