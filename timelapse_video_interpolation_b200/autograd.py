"""Kernel sequencing of one attention call, forward and backward, shared by the processor and the benchmarks.

`attention_forward` is what `TviAttnProcessor.__call__` runs after validating the module: it only
orders libtvi launches and owns the intermediate buffers.  When gradients are required (training
config 5: src/train_decouples.py back-propagates through every Attention module and updates the
`temporal_transformer_block*` parameters) the same launches run inside `_AttentionFn`, whose backward
is again nothing but libtvi kernels: projection dgrad / wgrad / bias column sums and the three
attention-core backward kernels (include/tvi.h).  torch autograd is only the tape.
"""
from __future__ import annotations

import torch
from torch.autograd.function import once_differentiable

from . import _lib, caches, lnfold, ops


def _to_bf16(t):
    return t if t is None or t.dtype == torch.bfloat16 else t.detach().to(torch.bfloat16)


def _self_core(q, k, v, Bp, S, heads, scale, lse=None):
    if S <= 64:
        return ops.attn_temporal(q, k, v, Bp, heads, S, S, scale)
    return ops.attn_spatial(q, k, v, Bp, heads, S, S, scale, lse=lse)


def _out_proj(o, w_o, b_o, residual):
    if residual is None:
        return ops.gemm_bf16_tn(o, w_o, b_o)
    return ops.gemm_bf16_tn_ex(o, w_o, b_o, residual=residual.reshape(o.shape[0], residual.shape[-1]))


def _proj(x2, w, ln):
    """x2 @ w^T, or LayerNorm(x2) @ w^T through the folded GEMM when ln = (owner module, key, norm) (lnfold.py)."""
    if ln is None:
        return ops.gemm_bf16_tn(x2, w)
    owner, key, norm = ln
    wg, colsum, cb = lnfold.folded(owner, key, w, None, norm)
    return ops.gemm_bf16_tn_ln(x2, wg, cb, colsum, ops.layernorm_stats(x2, float(norm.eps)))


def self_attention(x: torch.Tensor, w_qkv: torch.Tensor, w_o: torch.Tensor, b_o, heads: int, scale: float,
                   residual=None, frames=None, ln=None) -> torch.Tensor:
    """x [B', S, C] bf16 -> [B', S, C].  Reference: AttnProcessor2_0.__call__ with
    encoder_hidden_states=None (src/models/attention_processor.py:1223-1261).
    `frames` = T: x is [B*T, S, C] in the SPATIAL token order and attention runs over the T tokens of every (b, s)
    — the temporal block's self-attention without the permutes of src/models/attention.py:436-438 / :474-476."""
    Bp, S, C = x.shape
    inner = w_qkv.shape[0] // 3
    qkv = _proj(x.view(Bp * S, C), w_qkv, ln)  # [B'*S, 3*inner], heads stay packed along the row
    q, k, v = qkv[:, :inner], qkv[:, inner:2 * inner], qkv[:, 2 * inner:]
    if frames is not None:
        o = ops.attn_temporal(q, k, v, Bp // frames * S, heads, frames, frames, scale, seq_inner=S)
    else:
        o = _self_core(q, k, v, Bp, S, heads, scale)
    y = _out_proj(o, w_o, b_o, residual)
    return y.view(Bp, S, y.shape[-1])


def cross_attention(x: torch.Tensor, e: torch.Tensor, w_q: torch.Tensor, w_kv: torch.Tensor, w_o: torch.Tensor, b_o,
                    heads: int, scale: float, bias=None, stats=None, residual=None, frames=None, ln=None) -> torch.Tensor:
    """x [B', S, C], e [B', L, Ckv] -> [B', S, C].  Reference: same processor with
    encoder_hidden_states given (keys/values from the context, :1225-1231).
    `frames` = T: x is [B*T, S, C] in the spatial token order, e is the temporal block's [B*S, L, Ckv] context; row
    (b, t, s) attends to the L keys of group (b, s) (one key: it receives that group's projected value row)."""
    Bp, S, C = x.shape
    L = e.shape[1]
    inner = w_kv.shape[0] // 2
    if frames is not None:
        if e.shape[0] != Bp // frames * S:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "encoder_hidden_states batch does not match hidden_states (expects B*S rows)")
        if L > 8:
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"cross-attention with {L} context tokens (libtvi builds 1..8)")
        kv = ops.gemm_bf16_tn(e.view(-1, e.shape[2]), w_kv)                              # [B*S*L, 2*inner]
        if L == 1 and stats is None:
            yv = ops.gemm_bf16_tn(kv[:, inner:], w_o, b_o)  # [B*S, C]
            res2 = residual.reshape(Bp * S, residual.shape[-1]) if residual is not None else None
            return ops.bcast_rows_add(yv, res2, frames, group_inner=S, nrows=Bp * S).view(Bp, S, yv.shape[-1])
        # more keys, the ramp bias or the statistics: the few-keys core reads the T query rows of group (b, s) with a
        # row stride of S (bias / statistics rows are the frame index t)
        q = _proj(x.view(Bp * S, C), w_q, ln)
        o = ops.attn_fewkeys(q, kv[:, :inner], kv[:, inner:], Bp * S, frames, heads, L, scale,
                             Tq=frames if (bias is not None or stats is not None) else 1, bias=bias, stats=stats, group_inner=S)
        y = _out_proj(o, w_o, b_o, residual)
        return y.view(Bp, S, y.shape[-1])
    _check_context(x, e)
    kv = ops.gemm_bf16_tn(e.view(Bp * L, e.shape[2]), w_kv)  # [B'*L, 2*inner]
    k, v = kv[:, :inner], kv[:, inner:]
    if L == 1 and stats is None:
        # softmax over one key == 1: every query row of a group gets to_out(v_group); project the
        # B' value rows once and broadcast them (Q is never needed)
        yv = ops.gemm_bf16_tn(v, w_o, b_o)  # [B', C]
        if residual is not None:
            return ops.bcast_rows_add(yv, residual.reshape(Bp * S, residual.shape[-1]), S).view(Bp, S, yv.shape[-1])
        y = ops.attn_fewkeys(None, yv, yv, Bp * S, S, heads, 1, 1.0)  # L == 1: pure row broadcast
        return y.view(Bp, S, y.shape[-1])
    q = _proj(x.view(Bp * S, C), w_q, ln)
    o = ops.attn_fewkeys(q, k, v, Bp * S, S, heads, L, scale, Tq=S if (bias is not None or stats is not None) else 1,
                         bias=bias, stats=stats)
    y = _out_proj(o, w_o, b_o, residual)
    return y.view(Bp, S, y.shape[-1])


def _check_context(x, e):
    if e.shape[0] != x.shape[0]:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "encoder_hidden_states batch does not match hidden_states")
    if e.shape[1] > 8:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"cross-attention with {e.shape[1]} context tokens (libtvi builds 1..8)")


class _AttentionFn(torch.autograd.Function):
    """One attention call on the autograd tape.  Inputs that may receive gradients: x, e and the six
    parameter tensors (SURVEY.md §8.2 row B "Autograd"); `cfg` carries the derived bf16 fused weights and
    the non-differentiable options (ramp bias: a constant; statistics: the reference detaches them through
    `.item()`, src/train_decouples.py:460-471, :1027-1029)."""

    @staticmethod
    def forward(ctx, x, e, wq, wk, wv, wo, bo, cfg, residual=None):
        # `residual` (shape of the output): added in the out-projection epilogue (the blocks' `attn_output + hidden_states`);
        # its gradient is dy itself
        res2 = None if residual is None else residual.detach().contiguous().view(-1, residual.shape[-1])
        heads, scale = cfg["heads"], cfg["scale"]
        w_o, b_o = cfg["w_o"], cfg["b_o"]
        Bp, S, C = x.shape
        # frames = T: x is [B*T, S, C] in the SPATIAL token order and attention runs over the T tokens of every (b, s) (the
        # temporal block without its permute copies, forward and backward): the cores address the interleaved sequences
        frames = cfg.get("frames")
        ctx.cfg = cfg
        ctx.shape = (Bp, S, C)
        ctx.param_dtypes = tuple(None if t is None else t.dtype for t in (wq, wk, wv, wo, bo))
        if e is None:
            w_qkv = cfg["w_qkv"]
            inner = w_qkv.shape[0] // 3
            x2 = x.view(Bp * S, C)
            qkv = ops.gemm_bf16_tn(x2, w_qkv)
            q, k, v = qkv[:, :inner], qkv[:, inner:2 * inner], qkv[:, 2 * inner:]
            lse = None
            if frames is not None:
                o = ops.attn_temporal(q, k, v, Bp // frames * S, heads, frames, frames, scale, seq_inner=S)
            else:
                lse = torch.empty(Bp * S, heads, dtype=torch.float32, device=x.device) if S > 64 else None
                o = _self_core(q, k, v, Bp, S, heads, scale, lse=lse)
            y = _out_proj(o, w_o, b_o, res2)
            ctx.mode = "self"
            ctx.save_for_backward(x2, qkv, o, lse, w_qkv, w_o)
            return y.view(Bp, S, y.shape[-1])
        if frames is None:
            _check_context(x, e)
        elif e.shape[0] != Bp // frames * S or e.shape[1] > 8:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "encoder_hidden_states must be [B*S, L <= 8, C] with frames=")
        L = e.shape[1]
        ctx.e_shape = tuple(e.shape)
        # rows per context group and how the groups are laid out: contiguous [group, S] or interleaved [B, T, S]
        rpg, ginner = (S, 1) if frames is None else (frames, S)
        w_q, w_kv = cfg["w_q"], cfg["w_kv"]
        inner = w_kv.shape[0] // 2
        e2 = e.reshape(-1, e.shape[2])
        kv = ops.gemm_bf16_tn(e2, w_kv)
        k, v = kv[:, :inner], kv[:, inner:]
        bias, stats = cfg.get("bias"), cfg.get("stats")
        if L == 1 and stats is None:
            yv = ops.gemm_bf16_tn(v, w_o, b_o)
            if res2 is not None or frames is not None:
                y = ops.bcast_rows_add(yv, res2, rpg, group_inner=ginner, nrows=Bp * S)
            else:
                y = ops.attn_fewkeys(None, yv, yv, Bp * S, S, heads, 1, 1.0)
            ctx.mode = "cross1"
            ctx.save_for_backward(e2, kv, w_kv, w_o)
            return y.view(Bp, S, y.shape[-1])
        x2 = x.view(Bp * S, C)
        q = ops.gemm_bf16_tn(x2, w_q)
        o = ops.attn_fewkeys(q, k, v, Bp * S, rpg, heads, L, scale, Tq=rpg if (bias is not None or stats is not None) else 1,
                             bias=bias, stats=stats, group_inner=ginner)
        y = _out_proj(o, w_o, b_o, res2)
        ctx.mode = "cross"
        ctx.save_for_backward(x2, e2, q, kv, o, w_q, w_kv, w_o)
        return y.view(Bp, S, y.shape[-1])

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        cfg = ctx.cfg
        heads, scale = cfg["heads"], cfg["scale"]
        Bp, S, C = ctx.shape
        need_x, need_e, need_wq, need_wk, need_wv, need_wo, need_bo = ctx.needs_input_grad[:7]
        need_res = len(ctx.needs_input_grad) > 8 and ctx.needs_input_grad[8]
        dt_wq, dt_wk, dt_wv, dt_wo, dt_bo = ctx.param_dtypes
        dy2 = dy.contiguous().view(Bp * S, dy.shape[-1])
        if dy2.dtype != torch.bfloat16:
            dy2 = dy2.to(torch.bfloat16)
        dx = de = dwq = dwk = dwv = dwo = dbo = None

        def cast(t, dt):
            return t if t is None or t.dtype == dt else t.to(dt)

        if ctx.mode == "self":
            x2, qkv, o, lse, w_qkv, w_o = ctx.saved_tensors
            inner = w_qkv.shape[0] // 3
            if need_wo:
                dwo = ops.gemm_bf16_wgrad(dy2, o)
            if need_bo:
                dbo = ops.colsum_bf16(dy2).view(-1)
            do = ops.gemm_bf16_dgrad(dy2, w_o)
            q, k, v = qkv[:, :inner], qkv[:, inner:2 * inner], qkv[:, 2 * inner:]
            dqkv = torch.empty_like(qkv)  # dq | dk | dv written in place as column slices
            dq, dk, dv = dqkv[:, :inner], dqkv[:, inner:2 * inner], dqkv[:, 2 * inner:]
            frames = cfg.get("frames")
            if frames is not None:
                ops.attn_temporal_bwd(q, k, v, do, Bp // frames * S, heads, frames, frames, scale, dq=dq, dk=dk, dv=dv, seq_inner=S)
            elif S <= 64:
                ops.attn_temporal_bwd(q, k, v, do, Bp, heads, S, S, scale, dq=dq, dk=dk, dv=dv)
            else:
                ops.attn_spatial_bwd(q, k, v, o, do, lse, Bp, heads, S, S, scale, dq=dq, dk=dk, dv=dv)
            if need_x:
                dx = ops.gemm_bf16_dgrad(dqkv, w_qkv).view(Bp, S, C)
            if need_wq and need_wk and need_wv:
                dw = ops.gemm_bf16_wgrad(dqkv, x2)
                dwq, dwk, dwv = dw[:inner], dw[inner:2 * inner], dw[2 * inner:]
            else:
                dwq = ops.gemm_bf16_wgrad(dq, x2) if need_wq else None
                dwk = ops.gemm_bf16_wgrad(dk, x2) if need_wk else None
                dwv = ops.gemm_bf16_wgrad(dv, x2) if need_wv else None
        elif ctx.mode == "cross1":
            e2, kv, w_kv, w_o = ctx.saved_tensors
            inner = w_kv.shape[0] // 2
            v = kv[:, inner:]
            frames = cfg.get("frames")
            # [groups, C]: every query row of a group received the same row
            dyv = ops.colsum_bf16(dy2, S) if frames is None else ops.colsum_bf16(dy2, frames, group_inner=S)
            if need_wo:
                dwo = ops.gemm_bf16_wgrad(dyv, v)
            if need_bo:
                dbo = ops.colsum_bf16(dyv).view(-1)
            if need_e or need_wv:
                dv = ops.gemm_bf16_dgrad(dyv, w_o)
                if need_e:
                    de = ops.gemm_bf16_dgrad(dv, w_kv[inner:]).view(ctx.e_shape)
                if need_wv:
                    dwv = ops.gemm_bf16_wgrad(dv, e2)
            # softmax over one key is constant: nothing flows to the query path or to the key projection.  Behind a fused
            # LayerNorm + residual entry (cfg["dx_none_ok"]) "nothing" is None: that Function still runs (the residual's
            # gradient reaches it) and passes dres through without touching a zero tensor of the activations' size
            if need_x and not cfg.get("dx_none_ok"):
                dx = torch.zeros(Bp, S, C, dtype=torch.bfloat16, device=dy.device)
            if need_wq:
                dwq = torch.zeros(cfg["w_q"].shape, dtype=torch.bfloat16, device=dy.device)
            if need_wk:
                dwk = torch.zeros(inner, e2.shape[1], dtype=torch.bfloat16, device=dy.device)
        else:
            x2, e2, q, kv, o, w_q, w_kv, w_o = ctx.saved_tensors
            inner = w_kv.shape[0] // 2
            L = ctx.e_shape[1]
            frames = cfg.get("frames")
            rpg, ginner = (S, 1) if frames is None else (frames, S)
            k, v = kv[:, :inner], kv[:, inner:]
            if need_wo:
                dwo = ops.gemm_bf16_wgrad(dy2, o)
            if need_bo:
                dbo = ops.colsum_bf16(dy2).view(-1)
            do = ops.gemm_bf16_dgrad(dy2, w_o)
            bias = cfg.get("bias")
            dkv = torch.empty_like(kv)
            dq, _, _ = ops.attn_fewkeys_bwd(q, k, v, do, Bp * S, rpg, heads, L, scale, Tq=rpg if bias is not None else 1, bias=bias,
                                            dk=dkv[:, :inner], dv=dkv[:, inner:], group_inner=ginner)
            if need_x:
                dx = ops.gemm_bf16_dgrad(dq, w_q).view(Bp, S, C)
            if need_wq:
                dwq = ops.gemm_bf16_wgrad(dq, x2)
            if need_e:
                de = ops.gemm_bf16_dgrad(dkv, w_kv).view(ctx.e_shape)
            if need_wk and need_wv:
                dw = ops.gemm_bf16_wgrad(dkv, e2)
                dwk, dwv = dw[:inner], dw[inner:]
            else:
                dwk = ops.gemm_bf16_wgrad(dkv[:, :inner], e2) if need_wk else None
                dwv = ops.gemm_bf16_wgrad(dkv[:, inner:], e2) if need_wv else None
        return (dx, de, cast(dwq, dt_wq), cast(dwk, dt_wk), cast(dwv, dt_wv), cast(dwo, dt_wo), cast(dbo, dt_bo), None,
                dy if need_res else None)


def attention_forward(proc, attn, x, e, heads: int, scale: float, wo, bo, bias, stats, residual=None, frames=None,
                      pre_norm=None) -> torch.Tensor:
    """`pre_norm` (an nn.LayerNorm): the call computes attention(pre_norm(x), ...); on the no-grad path the LayerNorm is
    folded into the query-side projection (lnfold.py), elsewhere it runs as its own kernel first."""
    wq, wk, wv = attn.to_q.weight, attn.to_k.weight, attn.to_v.weight
    needs_grad = torch.is_grad_enabled() and any(
        t is not None and t.requires_grad for t in (x, e, wq, wk, wv, wo, bo) + (
            tuple(pre_norm.parameters()) if pre_norm is not None else ()))
    caches.note_call(needs_grad)
    w_o = _to_bf16(wo).contiguous()
    b_o = _to_bf16(bo).contiguous() if bo is not None else None
    if not needs_grad and not (residual is not None and torch.is_grad_enabled() and residual.requires_grad):
        ln = None
        if pre_norm is not None:
            if lnfold.supported(pre_norm, x):
                ln = (attn, "to_q" if e is not None else "qkv", pre_norm)
            else:
                from .feedforward import layer_norm
                x = layer_norm(pre_norm, x)
        if e is None:
            return self_attention(x, proc._fused(attn, ("to_q", "to_k", "to_v")), w_o, b_o, heads, scale, residual, frames, ln)
        return cross_attention(x, e, proc._fused(attn, ("to_q",)), proc._fused(attn, ("to_k", "to_v")), w_o, b_o, heads,
                               scale, bias, stats, residual, frames, ln)
    # training: the residual rides in the out-projection epilogue (its gradient is dy itself); when it is the LayerNorm's own
    # input — the blocks' `x + attn(norm(x))` — the LayerNorm backward kernel adds that gradient to dx in the same pass
    cfg = {"heads": heads, "scale": scale, "w_o": w_o, "b_o": b_o, "frames": frames}
    if pre_norm is not None:
        from .feedforward import layer_norm, layer_norm_residual
        if residual is x:
            x, residual = layer_norm_residual(pre_norm, x)
            cfg["dx_none_ok"] = True
        else:
            x = layer_norm(pre_norm, x)
    if residual is not None and residual.shape != x.shape[:-1] + (w_o.shape[0],):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "residual must have the shape of the output")
    if e is None:
        cfg["w_qkv"] = proc._fused(attn, ("to_q", "to_k", "to_v"))
    else:
        cfg["w_q"] = proc._fused(attn, ("to_q",))
        cfg["w_kv"] = proc._fused(attn, ("to_k", "to_v"))
        cfg["bias"] = bias
        cfg["stats"] = stats
    return _AttentionFn.apply(x, e, wq, wk, wv, wo, bo, cfg, residual)
