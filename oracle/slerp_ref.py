"""CPU restatement of the latent slerp spec (test infrastructure).  PARITY UNPINNED: the reference's
frame_slerp_interp.py is a 0-byte file (unused flags at src/train_decouples.py:1566-1567,
src/test.py:648-649), so this file and SURVEY.md §8.1 row A1 ARE the specification.

The algorithm is the one every Stable-Diffusion latent-interpolation script uses — the public `slerp(t, v0, v1,
DOT_THRESHOLD=0.9995)` gist by Andrej Karpathy / Xander Steenbrugge ("stablediffusionwalk.py", 2022; copied into the
diffusers community pipelines `examples/community/interpolate_stable_diffusion.py` and `unclip_*_interpolation.py`):
    dot = sum(v0 * v1 / (norm(v0) * norm(v1)))
    if abs(dot) > DOT_THRESHOLD:  v2 = (1 - t) * v0 + t * v1                       # nearly colinear: lerp
    else:  theta_0 = arccos(dot); theta_t = theta_0 * t
           v2 = sin(theta_0 - theta_t) / sin(theta_0) * v0 + sin(theta_t) / sin(theta_0) * v1
with the whole latent of a clip as one vector and t = i / (T - 1).  That named algorithm — threshold, lerp fallback, the
two sine coefficients — is what is restated below (coefficients in fp64, the blend in fp32 like the fp32 torch original).
"""
from __future__ import annotations

import numpy as np

DOT_THRESHOLD = 0.9995


def slerp_coefficients(v0: np.ndarray, v1: np.ndarray, T: int, dot_threshold: float = DOT_THRESHOLD):
    """fp64 blend coefficients (a_i, b_i), i = 0..T-1 for one clip; rounded to fp32 by the caller."""
    x0 = v0.astype(np.float64).ravel()
    x1 = v1.astype(np.float64).ravel()
    dot, n0, n1 = float(x0 @ x1), float(x0 @ x0), float(x1 @ x1)
    t = np.arange(T, dtype=np.float64) / (T - 1)
    den = np.sqrt(n0) * np.sqrt(n1)
    a, b = 1.0 - t, t.copy()
    if den > 0.0:
        c = min(1.0, max(-1.0, dot / den))
        if abs(c) <= float(np.float32(dot_threshold)):
            theta = np.arccos(c)
            s = np.sin(theta)
            a = np.sin((1.0 - t) * theta) / s
            b = np.sin(t * theta) / s
    return a, b


def slerp_ref(v0: np.ndarray, v1: np.ndarray, T: int, dot_threshold: float = DOT_THRESHOLD) -> np.ndarray:
    """v0, v1: [clips, ...] fp32 -> [clips, T, ...] fp32; out = fl32(fl32(a*v0) + fl32(b*v1))."""
    v0 = np.asarray(v0, dtype=np.float32)
    v1 = np.asarray(v1, dtype=np.float32)
    out = np.empty((v0.shape[0], T) + v0.shape[1:], dtype=np.float32)
    for c in range(v0.shape[0]):
        a, b = slerp_coefficients(v0[c], v1[c], T, dot_threshold)
        a32, b32 = a.astype(np.float32), b.astype(np.float32)
        for i in range(T):
            out[c, i] = (a32[i] * v0[c]).astype(np.float32) + (b32[i] * v1[c]).astype(np.float32)
    return out
