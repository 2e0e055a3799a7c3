"""`TransformerSpatioTemporalModel.forward` on libtvi (reference src/models/transformer_temporal.py:253-360, the module
the author marks "Most Important Block"; SURVEY.md §8.6 rows N1 "AlphaBlender" and N2 "layout glue").

Same signature, same results; what changes is where the bytes go:
  * `self.norm` + `permute(0, 2, 3, 1).reshape` (:300-304)  -> one GroupNorm kernel pair that writes token-major
  * `proj_in` / `proj_out` (:306, :353)                      -> the tcgen05 GEMM
  * `time_proj` + `time_pos_embed` (:308-315)                -> input-independent (frame index only): computed once per
                                                               (weights, B, T) in fp32 and cached
  * `hidden_states + emb` (:333)                             -> row-broadcast add kernel
  * the two blocks (:329-347)                                -> blocks.py forwards (the temporal one without permutes)
  * `time_mixer` (:349-351)                                  -> blend kernel, one fp32 alpha per frame
  * `reshape.permute(0, 3, 1, 2).contiguous() + residual`    -> one transpose-add kernel
`install_spatio_temporal(model)` rebinds every module with this attribute layout.

Training (src/train.py back-propagates through this forward, the spatial block under `torch.utils.checkpoint(...,
use_reentrant=False)` at :321-327): when a gradient is needed the call goes on the autograd tape — proj_in / proj_out, the
blocks' LayerNorms, attention and FeedForward through their libtvi autograd Functions (forward AND backward kernels),
the glue (GroupNorm + NCHW<->token transposes, frame-embedding add, AlphaBlender, module residual) through the same
forward kernels as inference plus their backward kernels (layout_ops.cu: `tvi_groupnorm_tokens_bwd_bf16`,
`tvi_nchw_to_tokens_bf16`, `tvi_blend_rows_bwd_bf16`, column sums for the embedding).
"""
from __future__ import annotations

import sys
import types
import weakref
from dataclasses import dataclass
from typing import Optional

import torch
import torch.utils.checkpoint

from . import _lib, caches, ops
from .blocks import (_training, basic_transformer_block_forward, install_blocks,
                     temporal_basic_transformer_block_forward)
from .feedforward import _bf16 as _bf16_param
from .feedforward import linear as _linear

__all__ = ["transformer_spatio_temporal_forward", "install_spatio_temporal", "uninstall_spatio_temporal"]


@dataclass
class TransformerTemporalModelOutput:
    """Stand-in for the reference's output dataclass (:28-38) when its module is not importable."""
    sample: torch.Tensor


def _frame_embedding(model, batch_size: int, num_frames: int, device) -> torch.Tensor:
    """emb[b*T + t] = time_pos_embed(time_proj(t)) (:308-315): depends on the frame index and the weights only."""
    tp = model.time_pos_embed
    ws = (tp.linear_1.weight, tp.linear_1.bias, tp.linear_2.weight, tp.linear_2.bias)
    sig = (caches.generation(), batch_size, num_frames, str(device)) + tuple((w.data_ptr(), w._version) for w in ws)
    hit = getattr(model, "_tvi_frame_emb", None)
    if hit is not None and hit[0] == sig:
        return hit[1]
    with torch.no_grad():
        idx = torch.arange(num_frames, device=device).repeat(batch_size, 1).reshape(-1)
        t_emb = model.time_proj(idx).to(torch.bfloat16)          # the reference casts to hidden_states.dtype (:313)
        emb = tp(t_emb.to(tp.linear_1.weight.dtype)).to(torch.bfloat16).contiguous()
    model._tvi_frame_emb = (sig, emb)
    return emb


def _alpha(mixer, image_only_indicator: torch.Tensor) -> torch.Tensor:
    """One fp32 alpha per frame: AlphaBlender.get_alpha for ndims == 3, with the reference's cast to the activation
    dtype (`.to(x_spatial.dtype)`) applied so that the blend weights are the bf16-rounded ones."""
    strategy = getattr(mixer, "merge_strategy", "learned_with_images")
    # the same indicator tensor comes back for every module of every denoising step: cache on (tensor, version)
    sig = (caches.generation(), image_only_indicator.data_ptr(), image_only_indicator._version, tuple(image_only_indicator.shape),
           mixer.mix_factor.data_ptr(), mixer.mix_factor._version)
    if torch.cuda.is_current_stream_capturing():  # a CUDA graph must recompute alpha from the static indicator buffer
        return _alpha_uncached(mixer, image_only_indicator, strategy)
    hit = getattr(mixer, "_tvi_alpha", None)
    if hit is not None and hit[0] == sig and hit[1]() is image_only_indicator:
        return hit[2]
    a = _alpha_uncached(mixer, image_only_indicator, strategy)
    mixer._tvi_alpha = (sig, weakref.ref(image_only_indicator), a)
    return a


def _alpha_uncached(mixer, image_only_indicator: torch.Tensor, strategy: str) -> torch.Tensor:
    ind = image_only_indicator.reshape(-1)
    mf = mixer.mix_factor.detach().float().reshape(-1)[:1]
    if strategy == "fixed":
        a = mf.expand(ind.numel())
    elif strategy == "learned":
        a = torch.sigmoid(mf).expand(ind.numel())
    elif strategy == "learned_with_images":
        a = torch.where(ind.bool(), torch.ones(1, device=ind.device), torch.sigmoid(mf).to(ind.device))
    else:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"AlphaBlender merge_strategy {strategy!r}")
    a = a.to(torch.bfloat16).float()
    if getattr(mixer, "switch_spatial_to_temporal_mix", False):
        a = 1.0 - a
    return a.contiguous()


def transformer_spatio_temporal_forward(model, hidden_states: torch.Tensor, encoder_hidden_states=None,
                                        image_only_indicator: Optional[torch.Tensor] = None, return_dict: bool = True):
    """Same signature as TransformerSpatioTemporalModel.forward (:253-259)."""
    if hidden_states.dim() != 4:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "hidden_states must be [batch*frames, channels, height, width]")
    if image_only_indicator is None:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "image_only_indicator [batch, frames] is required (it carries num_frames)")
    if not hidden_states.is_cuda:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "hidden_states is not on a CUDA device (there is no CPU fallback)")
    if hidden_states.dtype != torch.bfloat16:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"hidden_states is {hidden_states.dtype}; the libtvi path is bf16")
    BF, C, H, W = hidden_states.shape
    T = image_only_indicator.shape[-1]
    B = BF // T
    S = H * W
    if B * T != BF:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "batch*frames is not a multiple of image_only_indicator.shape[-1]")
    if isinstance(encoder_hidden_states, tuple):                                     # :274-279
        spatial_ctx, temporal_ctx = encoder_hidden_states
    else:
        spatial_ctx = temporal_ctx = encoder_hidden_states
    # :281-296 — the first frame's context of every clip, broadcast over the H*W positions ([hw, batch] major, as written)
    tc = temporal_ctx[None, :].reshape(B, T, -1, temporal_ctx.shape[-1])[:, 0]
    time_context = tc[None, :].broadcast_to(S, B, tc.shape[-2], tc.shape[-1]).reshape(S * B, tc.shape[-2], tc.shape[-1])

    if _training(model, hidden_states):
        out = _forward_on_tape(model, hidden_states, spatial_ctx, time_context, image_only_indicator, B, T, H, W)
    else:
        out = _forward_fused(model, hidden_states, spatial_ctx, time_context, image_only_indicator, B, T, H, W)
    if not return_dict:
        return (out,)
    mod = sys.modules.get(type(model).__module__)
    cls = getattr(mod, "TransformerTemporalModelOutput", None) or TransformerTemporalModelOutput
    return cls(sample=out)


def _forward_fused(model, hidden_states, spatial_ctx, time_context, image_only_indicator, B, T, H, W):
    """No-grad path: fused GroupNorm / layout / blend kernels, cached frame embedding and alpha."""
    BF, S = B * T, H * W
    residual = hidden_states.contiguous()
    norm = model.norm
    x = ops.groupnorm_tokens(residual, _bf16_param(norm.weight), _bf16_param(norm.bias), int(norm.num_groups), float(norm.eps))
    x = ops.gemm_bf16_tn(x, _bf16_param(model.proj_in.weight), _bf16_param(model.proj_in.bias))   # [BF, S, inner]
    emb = _frame_embedding(model, B, T, hidden_states.device)                                     # [BF, C]
    if emb.shape[-1] != x.shape[-1]:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "in_channels != inner_dim: the frame embedding cannot be added (:333)")
    alpha = _alpha(model.time_mixer, image_only_indicator.to(hidden_states.device))
    for block, temporal_block in zip(model.transformer_blocks, model.temporal_transformer_blocks):
        x = basic_transformer_block_forward(block, x, encoder_hidden_states=spatial_ctx)           # :337-339
        x_mix = ops.bcast_rows_add(emb, x, S)                                                     # :341-342
        x_mix = temporal_basic_transformer_block_forward(temporal_block, x_mix.view(BF, S, x_mix.shape[-1]), num_frames=T,
                                                         encoder_hidden_states=time_context)      # :343-347
        x = ops.blend_rows(x, x_mix, alpha, S)                                                    # :349-351
    y = ops.gemm_bf16_tn(x, _bf16_param(model.proj_out.weight), _bf16_param(model.proj_out.bias))  # :353
    return ops.tokens_to_nchw_add(y.view(BF, S, y.shape[-1]), residual, H, W)                      # :354-357


def _cast_grad(g, like):
    return None if (g is None or like is None) else (g if g.dtype == like.dtype else g.to(like.dtype)).view(like.shape)


class _GroupNormTokensFn(torch.autograd.Function):
    """`self.norm(x).permute(0, 2, 3, 1).reshape(BF, H*W, C)` (:300-304) on the tape: groupnorm_tokens forward, the
    three-pass backward of layout_ops.cu (dx in NCHW, dgamma, dbeta; fixed-order sums)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, groups, eps):
        xc = x.detach().contiguous()
        y = ops.groupnorm_tokens(xc, _bf16_param(gamma), _bf16_param(beta), groups, eps)
        ctx.save_for_backward(xc, gamma, beta)
        ctx.groups, ctx.eps = groups, eps
        return y

    @staticmethod
    def backward(ctx, dn):
        xc, gamma, beta = ctx.saved_tensors
        need_x, need_g, need_b = ctx.needs_input_grad[:3]
        dx, dg, db = ops.groupnorm_tokens_bwd(xc, dn.contiguous(), _bf16_param(gamma), ctx.groups, ctx.eps, need_dx=need_x,
                                              need_affine=need_g or need_b)
        return dx, _cast_grad(dg, gamma) if need_g else None, _cast_grad(db, beta) if need_b else None, None, None


class _TokensToNchwAddFn(torch.autograd.Function):
    """`y.reshape(BF, H, W, C).permute(0, 3, 1, 2).contiguous() + residual` (:354-357): d tokens = transpose(d out)."""

    @staticmethod
    def forward(ctx, y, residual, H, W):
        return ops.tokens_to_nchw_add(y.detach().contiguous(), residual.detach().contiguous(), H, W)

    @staticmethod
    def backward(ctx, dout):
        dout = dout.contiguous()
        return (ops.nchw_to_tokens(dout) if ctx.needs_input_grad[0] else None), (dout if ctx.needs_input_grad[1] else None), None, None


class _BcastAddFn(torch.autograd.Function):
    """`hidden_states + emb` (:341-342), emb one row per frame: d emb = the column sums over the frame's H*W rows."""

    @staticmethod
    def forward(ctx, emb, x, S):
        ctx.S = S
        return ops.bcast_rows_add(emb.detach().contiguous(), x.detach().contiguous(), S).view(x.shape)

    @staticmethod
    def backward(ctx, d):
        d = d.contiguous()
        demb = ops.colsum_bf16(d.view(-1, d.shape[-1]), ctx.S) if ctx.needs_input_grad[0] else None
        return demb, (d if ctx.needs_input_grad[1] else None), None


class _BlendFn(torch.autograd.Function):
    """AlphaBlender.forward (:349-351) with one fp32 alpha per frame; d alpha = sum over the frame of dy (x_s - x_t)."""

    @staticmethod
    def forward(ctx, xs, xt, alpha, S):
        xs2, xt2, a = xs.detach().contiguous(), xt.detach().contiguous(), alpha.detach().float().contiguous()
        ctx.save_for_backward(xs2, xt2, a)
        ctx.S = S
        return ops.blend_rows(xs2, xt2, a, S)

    @staticmethod
    def backward(ctx, dy):
        xs2, xt2, a = ctx.saved_tensors
        dxs, dxt, da = ops.blend_rows_bwd(dy.contiguous().view(-1, dy.shape[-1]), xs2.view(-1, xs2.shape[-1]),
                                          xt2.view(-1, xt2.shape[-1]), a, ctx.S, ctx.needs_input_grad[2])
        return (dxs.view(xs2.shape) if ctx.needs_input_grad[0] else None), (dxt.view(xt2.shape) if ctx.needs_input_grad[1] else None), da, None


def _forward_on_tape(model, hidden_states, spatial_ctx, time_context, image_only_indicator, B, T, H, W):
    """Training path (see the module docstring): same order of operations as :300-357, every activation-sized step a
    libtvi kernel in both directions (the Functions above); torch ops only touch the [B*T, C] embedding and the [B*T]
    blend weights."""
    BF, S = B * T, H * W
    norm = model.norm
    x = _GroupNormTokensFn.apply(hidden_states, norm.weight, norm.bias, int(norm.num_groups), float(norm.eps))   # :300-304
    x = _linear(model.proj_in, x)                                                                 # :306
    idx = torch.arange(T, device=hidden_states.device).repeat(B, 1).reshape(-1)                   # :308-315
    tp = model.time_pos_embed
    t_emb = model.time_proj(idx).to(hidden_states.dtype)
    emb = tp(t_emb.to(tp.linear_1.weight.dtype)).to(hidden_states.dtype)                          # [BF, C]
    if emb.shape[-1] != x.shape[-1]:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "in_channels != inner_dim: the frame embedding cannot be added (:333)")
    # the reference casts alpha to the activation dtype (`.to(x_spatial.dtype)`): the blend weights are the bf16-rounded ones
    alpha = _alpha_on_tape(model.time_mixer, image_only_indicator.to(hidden_states.device)).reshape(-1)
    alpha = alpha.to(hidden_states.dtype).float()                                                 # [BF] fp32
    use_ckpt = bool(getattr(model, "gradient_checkpointing", False)) and model.training
    for block, temporal_block in zip(model.transformer_blocks, model.temporal_transformer_blocks):
        if use_ckpt:  # :321-327
            x = torch.utils.checkpoint.checkpoint(block, x, None, spatial_ctx, None, use_reentrant=False)
        else:
            x = block(x, encoder_hidden_states=spatial_ctx)                                        # :329-333
        x_mix = _BcastAddFn.apply(emb, x, S)                                                       # :341-342
        x_mix = temporal_block(x_mix, num_frames=T, encoder_hidden_states=time_context)            # :343-347
        x = _BlendFn.apply(x, x_mix, alpha, S)                                                     # :349-351 (AlphaBlender.forward)
    y = _linear(model.proj_out, x)                                                                 # :353
    return _TokensToNchwAddFn.apply(y.view(BF, S, y.shape[-1]), hidden_states, H, W)               # :354-357


def _alpha_on_tape(mixer, image_only_indicator: torch.Tensor) -> torch.Tensor:
    """AlphaBlender.get_alpha for ndims == 3 as differentiable torch ops (mix_factor may train): [B*T, 1, 1]."""
    strategy = getattr(mixer, "merge_strategy", "learned_with_images")
    ind = image_only_indicator.reshape(-1)
    mf = mixer.mix_factor.float().reshape(-1)[:1]
    if strategy == "fixed":
        a = mf.expand(ind.numel())
    elif strategy == "learned":
        a = torch.sigmoid(mf).expand(ind.numel())
    elif strategy == "learned_with_images":
        a = torch.where(ind.bool(), torch.ones(1, device=ind.device), torch.sigmoid(mf).to(ind.device))
    else:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"AlphaBlender merge_strategy {strategy!r}")
    if getattr(mixer, "switch_spatial_to_temporal_mix", False):
        a = 1.0 - a
    return a.reshape(-1, 1, 1)


def _is_spatio_temporal(m) -> bool:
    return all(hasattr(m, a) for a in ("norm", "proj_in", "transformer_blocks", "temporal_transformer_blocks", "time_pos_embed",
                                       "time_proj", "time_mixer", "proj_out"))


def install_spatio_temporal(model: torch.nn.Module, processor=None) -> int:
    """Rebinds `forward` of every TransformerSpatioTemporalModel-shaped module (and of its blocks); returns how many."""
    caches.invalidate_caches()
    n = 0
    for _, m in model.named_modules():
        if _is_spatio_temporal(m):
            install_blocks(m, processor)
            if not hasattr(m, "_tvi_saved_model_forward"):
                m._tvi_saved_model_forward = m.__dict__.get("forward")
            m.forward = types.MethodType(transformer_spatio_temporal_forward, m)
            n += 1
    return n


def uninstall_spatio_temporal(model: torch.nn.Module) -> int:
    from .blocks import uninstall_blocks
    caches.invalidate_caches()
    n = 0
    for _, m in model.named_modules():
        if hasattr(m, "_tvi_saved_model_forward"):
            if m._tvi_saved_model_forward is None:
                m.__dict__.pop("forward", None)
            else:
                m.forward = m._tvi_saved_model_forward
            del m._tvi_saved_model_forward
            uninstall_blocks(m)
            n += 1
    return n
