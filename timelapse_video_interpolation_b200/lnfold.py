"""LayerNorm folded into the GEMM that consumes it (SURVEY.md §8.6 N1 "LN prologue").

Every LayerNorm of the two transformer blocks feeds exactly one projection (reference src/models/attention.py:255+:276
norm1 -> to_q/k/v, :300+:311 norm2 -> to_q, :321+:343 norm3 -> ff.net[0], :441-446 norm_in -> ff_in.net[0], :451, :457,
:462).  With mean / rstd the row statistics of x,

    LN(x) W^T + b  =  rstd * (x Wg^T  -  mean * colsum)  +  (W beta + b),      Wg = gamma (.) W,  colsum_n = sum_k Wg[n, k]

so the projection can read the RAW rows: `ops.layernorm_stats` makes one read-only pass for (mean, rstd), and
`ops.gemm_bf16_tn_ln` applies the affine map to the fp32 accumulator in its epilogue (before GEGLU / bias / residual).
The normalised activations never reach HBM: one 2·M·C-byte write and one launch less per LayerNorm.
`colsum` is summed from the bf16-ROUNDED Wg, so the `mean` term cancels the mean part of the accumulator exactly.

No-grad path only (the tape path keeps the LayerNorm Function, which has a backward).

MEASURED (B200, L0: M = 258 048 rows, C = 320; tools/prof_lnfold.py, profiles/r2_lnfold.txt) and therefore OFF by default
(`TVI_LN_FOLD=1` turns it on): the statistics pass costs 0.039 ms against 0.072 ms for the full LayerNorm, but the GEMM
is bound by the latency of its epilogue chain, and the extra FMA + shared-memory read per element cost
more than the LayerNorm saved — QKV 0.166 -> 0.206 ms, ff_in + GEGLU 0.455 -> 0.555 ms; the L0 module forward is
unchanged within noise (7.42 vs 7.40 ms).  What does pay is skipping the LayerNorm in front of a one-key cross-attention
(blocks._cross_attn), which needs no arithmetic at all.
"""
from __future__ import annotations

import os
import weakref

import torch

from . import caches, ops

__all__ = ["enabled", "supported", "folded"]

_cache: weakref.WeakKeyDictionary = weakref.WeakKeyDictionary()


def enabled() -> bool:
    return os.environ.get("TVI_LN_FOLD", "0") == "1"


def supported(norm, x: torch.Tensor) -> bool:
    """The statistics kernel is built for the SVD widths; anything else takes the stand-alone LayerNorm."""
    return (enabled() and isinstance(getattr(norm, "normalized_shape", None), (tuple, list, torch.Size))
            and tuple(norm.normalized_shape) == (x.shape[-1],) and ops.layernorm_stats_supported(x.shape[-1]))


def _sig(t):
    return None if t is None else (t.data_ptr(), t._version, t.dtype)


def folded(owner, key, w_bf16: torch.Tensor, bias, norm):
    """(Wg bf16 [N, K], colsum fp32 [N], bias' bf16 [N]) for the projection `w_bf16` [N, K] (already in the row order the
    GEMM wants, e.g. fused q|k|v or GEGLU-interleaved; `bias` in the same order or None) behind LayerNorm `norm`.
    Cached per (owner module, key); rebuilt when the weight tensor object, a parameter version or the cache generation
    changes (caches.py)."""
    gamma, beta = getattr(norm, "weight", None), getattr(norm, "bias", None)
    sig = (caches.generation(), _sig(w_bf16), _sig(bias), _sig(gamma), _sig(beta))
    per = _cache.get(owner)
    if per is None:
        per = _cache.setdefault(owner, {})
    hit = per.get(key)
    if hit is not None and hit[0] == sig and hit[1] is w_bf16:
        return hit[2]
    with torch.no_grad():
        wf = w_bf16.detach().float()
        wg = (wf * gamma.detach().float()[None, :] if gamma is not None else wf).to(torch.bfloat16).contiguous()
        colsum = wg.float().sum(dim=1).contiguous()
        cb = torch.zeros(wf.shape[0], dtype=torch.float32, device=wf.device)
        if beta is not None:
            cb = wf @ beta.detach().float()
        if bias is not None:
            cb = cb + bias.detach().float()
        out = (wg, colsum, cb.to(torch.bfloat16).contiguous())
    per[key] = (sig, w_bf16, out)      # holds w_bf16: `is` above cannot match a recycled address
    return out
