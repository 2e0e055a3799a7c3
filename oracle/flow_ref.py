"""CPU restatement of the optical-flow checks (test infrastructure).

cosine_similarity: dataset_construction/step4_flow_check.py:25-34, restated with the same numpy calls
(pinned against the reference's own function through tests/golden/flow_cosine_*.npz).
backward_warp_check: north-star extension, PARITY UNPINNED by the reference (it has no warp,
SURVEY.md D2); spec = SURVEY.md §8.1 row A7, built on torch.nn.functional.grid_sample on CPU.
"""
from __future__ import annotations

import numpy as np
import numpy.linalg as LA
import torch
import torch.nn.functional as F


def cosine_similarity(flow_a: np.ndarray, flow_b: np.ndarray):
    vec_a = flow_a.reshape(-1, 2)                       # :26
    vec_b = flow_b.reshape(-1, 2)                       # :27
    dot_product = np.sum(vec_a * vec_b, axis=1)         # :29
    norm_a = LA.norm(vec_a, axis=1) + 1e-6              # :30
    norm_b = LA.norm(vec_b, axis=1) + 1e-6              # :31
    cos_sim = dot_product / (norm_a * norm_b)           # :33
    return np.mean(cos_sim)                             # :34


def _grid(f_fwd: torch.Tensor) -> torch.Tensor:
    B, _, H, W = f_fwd.shape
    ys, xs = torch.meshgrid(torch.arange(H, dtype=f_fwd.dtype), torch.arange(W, dtype=f_fwd.dtype), indexing="ij")
    px = xs[None] + f_fwd[:, 0]
    py = ys[None] + f_fwd[:, 1]
    gx = 2.0 * px / max(W - 1, 1) - 1.0
    gy = 2.0 * py / max(H - 1, 1) - 1.0
    return torch.stack([gx, gy], dim=-1), px, py


def backward_warp_check(img2: torch.Tensor, f_fwd: torch.Tensor, f_bwd: torch.Tensor, alpha: float = 0.01,
                        beta: float = 0.5, dtype=torch.float64):
    """Returns (warped, occluded, valid, margin): margin = lhs - rhs of the occlusion inequality, so
    tests can ignore pixels whose decision is within rounding of the threshold."""
    img2, f_fwd, f_bwd = (t.detach().to("cpu", dtype) for t in (img2, f_fwd, f_bwd))
    B, _, H, W = f_fwd.shape
    grid, px, py = _grid(f_fwd)
    kw = dict(mode="bilinear", padding_mode="zeros", align_corners=True)
    warped = F.grid_sample(img2, grid, **kw)
    bt = F.grid_sample(f_bwd, grid, **kw)
    sx, sy = f_fwd[:, 0] + bt[:, 0], f_fwd[:, 1] + bt[:, 1]
    lhs = sx * sx + sy * sy
    rhs = alpha * (f_fwd[:, 0] ** 2 + f_fwd[:, 1] ** 2 + bt[:, 0] ** 2 + bt[:, 1] ** 2) + beta
    occ = lhs > rhs
    valid = (px >= 0) & (px <= W - 1) & (py >= 0) & (py <= H - 1)
    return warped, occ, valid, lhs - rhs
