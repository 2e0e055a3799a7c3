"""Kernel sequencing of one attention call (forward), shared by the processor and the benchmarks.

`attention_forward` is what `TviAttnProcessor.__call__` runs after validating the module: it only
orders libtvi launches and owns the intermediate buffers.  Backward (training config 5) is not
built yet: a call that needs gradients raises instead of silently detaching.
"""
from __future__ import annotations

import torch

from . import _lib, ops


def _to_bf16(t):
    return t if t is None or t.dtype == torch.bfloat16 else t.detach().to(torch.bfloat16)


def self_attention(x: torch.Tensor, w_qkv: torch.Tensor, w_o: torch.Tensor, b_o, heads: int, scale: float) -> torch.Tensor:
    """x [B', S, C] bf16 -> [B', S, C].  Reference: AttnProcessor2_0.__call__ with
    encoder_hidden_states=None (src/models/attention_processor.py:1223-1261)."""
    Bp, S, C = x.shape
    inner = w_qkv.shape[0] // 3
    qkv = ops.gemm_bf16_tn(x.view(Bp * S, C), w_qkv)  # [B'*S, 3*inner], heads stay packed along the row
    q, k, v = qkv[:, :inner], qkv[:, inner:2 * inner], qkv[:, 2 * inner:]
    if S <= 64:
        o = ops.attn_temporal(q, k, v, Bp, heads, S, S, scale)
    else:
        o = ops.attn_spatial(q, k, v, Bp, heads, S, S, scale)
    y = ops.gemm_bf16_tn(o, w_o, b_o)
    return y.view(Bp, S, -1)


def cross_attention(x: torch.Tensor, e: torch.Tensor, w_q: torch.Tensor, w_kv: torch.Tensor, w_o: torch.Tensor, b_o,
                    heads: int, scale: float, bias=None, stats=None) -> torch.Tensor:
    """x [B', S, C], e [B', L, Ckv] -> [B', S, C].  Reference: same processor with
    encoder_hidden_states given (keys/values from the context, :1225-1231)."""
    Bp, S, C = x.shape
    L = e.shape[1]
    if e.shape[0] != Bp:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "encoder_hidden_states batch does not match hidden_states")
    if L > 8:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"cross-attention with {L} context tokens (libtvi builds 1..8)")
    inner = w_kv.shape[0] // 2
    kv = ops.gemm_bf16_tn(e.view(Bp * L, e.shape[2]), w_kv)  # [B'*L, 2*inner]
    k, v = kv[:, :inner], kv[:, inner:]
    if L == 1 and stats is None:
        # softmax over one key == 1: every query row of a group gets to_out(v_group); project the
        # B' value rows once and broadcast them (Q is never needed)
        yv = ops.gemm_bf16_tn(v, w_o, b_o)  # [B', C]
        y = ops.attn_fewkeys(None, yv, yv, Bp * S, S, heads, 1, 1.0)  # L == 1: pure row broadcast
        return y.view(Bp, S, -1)
    q = ops.gemm_bf16_tn(x.view(Bp * S, C), w_q)
    o = ops.attn_fewkeys(q, k, v, Bp * S, S, heads, L, scale, Tq=S if (bias is not None or stats is not None) else 1,
                         bias=bias, stats=stats)
    y = ops.gemm_bf16_tn(o, w_o, b_o)
    return y.view(Bp, S, -1)


def attention_forward(proc, attn, x, e, heads: int, scale: float, wo, bo, bias, stats) -> torch.Tensor:
    needs_grad = torch.is_grad_enabled() and (
        x.requires_grad or (e is not None and e.requires_grad) or wo.requires_grad
        or attn.to_q.weight.requires_grad or attn.to_k.weight.requires_grad or attn.to_v.weight.requires_grad)
    if needs_grad:
        raise NotImplementedError(
            "libtvi attention backward is not built yet: call under torch.no_grad() / inference_mode() "
            "(training config 5 is tracked in DESIGN.md)")
    w_o = _to_bf16(wo).contiguous()
    b_o = _to_bf16(bo).contiguous() if bo is not None else None
    if e is None:
        return self_attention(x, proc._fused(attn, ("to_q", "to_k", "to_v")), w_o, b_o, heads, scale)
    return cross_attention(x, e, proc._fused(attn, ("to_q",)), proc._fused(attn, ("to_k", "to_v")), w_o, b_o, heads,
                           scale, bias, stats)
