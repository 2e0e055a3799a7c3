"""Block-level forwards on libtvi ("needs a block-level swap").

Drop-in replacements for
  BasicTransformerBlock.forward          reference src/models/attention.py:234-347 (the `use_layer_norm` branch SVD uses)
  TemporalBasicTransformerBlock.forward  reference src/models/attention.py:423-478
Every arithmetic step is a libtvi kernel: LayerNorm -> attention (processor path, residual add fused into the
out-projection epilogue) -> LayerNorm -> cross-attention (+ residual) -> LayerNorm -> FeedForward (GEGLU fused into the
first GEMM, bias + residual into the second).  On the no-grad path the LayerNorm in front of a one-key
cross-attention — whose queries are never read — is skipped; every other LayerNorm is handed to the projection it feeds
as `pre_norm=`, which either runs it as its own kernel (default) or folds it into the GEMM (lnfold.py, `TVI_LN_FOLD=1`:
built and parity-tested, measured slower at the SVD shapes, see there).  The `[B*T, S, C] <-> [B*S, T, C]` re-layout of the temporal block
(:436-438, :474-476) is folded away (row N2): LayerNorm / FeedForward / projections are row-wise, so the block runs on
the spatial token order in place and only the attention cores address the T tokens of a (b, s) sequence with a
row stride of S (`tvi_attn_temporal_fwd_interleaved`, `tvi_bcast_rows_add_bf16(group_inner=S)` for one-key contexts,
`tvi_attn_fewkeys_fwd_interleaved` for 2..8 keys, the ramp bias and the logit statistics).

`install_blocks(model)` re-binds `forward` on every module with the block's attribute layout (duck-typed) and installs
the attention processor on its Attention modules.  Training (config 5) works through the same forwards: when a gradient is
required LayerNorm, attention and FeedForward go on the autograd tape through their libtvi Functions (forward and backward
kernels): the residual adds ride in the GEMM epilogues and their gradients in the LayerNorm backward, and the temporal
block stays in the spatial token order in both directions (interleaved forward AND backward cores).
"""
from __future__ import annotations

import types
from typing import Optional

import torch

from . import _lib, caches
from .feedforward import feed_forward
from .processor import TviAttnProcessor

__all__ = ["basic_transformer_block_forward", "temporal_basic_transformer_block_forward", "install_blocks", "uninstall_blocks"]


def _no_grad_only(block, x):
    """The MODULE-level forward (GroupNorm / layout kernels) has no backward; the block forwards do (see _training)."""
    if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in block.parameters())):
        raise NotImplementedError("the libtvi TransformerSpatioTemporalModel forward is inference-only (GroupNorm / layout "
                                  "kernels have no backward): call under torch.no_grad(), or install only the blocks for training")


def _training(block, x) -> bool:
    """True when the call must go on the autograd tape: LayerNorm, attention and FeedForward then run through their
    autograd Functions (libtvi forward AND backward kernels); residual adds and the temporal re-layout stay torch ops."""
    return torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in block.parameters()))


def _proc(attn):
    proc = attn.processor
    if not isinstance(proc, TviAttnProcessor):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "block forward: the Attention module does not carry a TviAttnProcessor "
                                                  "(call install_blocks / install first)")
    return proc


def _attn(attn, x, e, residual, frames=None, pre_norm=None):
    return _proc(attn)(attn, x, encoder_hidden_states=e, residual=residual, frames=frames, pre_norm=pre_norm)


def _cross_attn(block, norm, attn, x, e, frames=None):
    """norm -> cross-attention -> + x.  With ONE context token the softmax is 1 and the queries are never read
    (autograd.cross_attention), so on the no-grad path the LayerNorm in front of to_q is dead work: skip it (0.07 ms of
    HBM traffic per call at L0; SVD's image-embedding context has exactly one token)."""
    proc = _proc(attn)
    if (e is not None and e.shape[1] == 1 and proc.stats_sink is None and not proc.ramp_bias
            and not _training(block, x) and not (torch.is_grad_enabled() and e.requires_grad)):
        return proc(attn, x, encoder_hidden_states=e, residual=x, frames=frames)   # x only lends its shape
    return proc(attn, x, encoder_hidden_states=e, residual=x, frames=frames, pre_norm=norm)


def _check_basic(block):
    for flag in ("use_ada_layer_norm", "use_ada_layer_norm_zero", "use_ada_layer_norm_single"):
        if getattr(block, flag, False):
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"block forward: {flag} is set; SVD blocks use plain LayerNorm")
    if getattr(block, "pos_embed", None) is not None or getattr(block, "_chunk_size", None) is not None:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "block forward: pos_embed / chunked feed-forward are not supported")


def basic_transformer_block_forward(block, hidden_states: torch.Tensor, attention_mask=None,
                                    encoder_hidden_states: Optional[torch.Tensor] = None, encoder_attention_mask=None,
                                    timestep=None, cross_attention_kwargs=None, class_labels=None) -> torch.Tensor:
    """Same signature as BasicTransformerBlock.forward (:234-243)."""
    _check_basic(block)
    if attention_mask is not None or encoder_attention_mask is not None:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "block forward: attention masks are not supported (SVD never passes one)")
    if cross_attention_kwargs and any(k != "scale" for k in cross_attention_kwargs):
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "block forward: unsupported cross_attention_kwargs")
    x = hidden_states.contiguous()
    # 0. self-attention (:270-287): norm1 -> attn1 -> + hidden_states
    x = _attn(block.attn1, x, encoder_hidden_states if getattr(block, "only_cross_attention", False) else None, x,
              pre_norm=block.norm1)
    # 3. cross-attention (:296-317)
    if getattr(block, "attn2", None) is not None:
        x = _cross_attn(block, block.norm2, block.attn2, x, encoder_hidden_states)
    # 4. feed-forward (:321-343): norm3 -> ff -> + hidden_states
    return feed_forward(block.ff, x, residual=x, pre_norm=block.norm3)


def temporal_basic_transformer_block_forward(block, hidden_states: torch.Tensor, num_frames: int,
                                             encoder_hidden_states: Optional[torch.Tensor] = None,
                                             fold_layout: bool = True) -> torch.Tensor:
    """Same signature as TemporalBasicTransformerBlock.forward (:423-428); `fold_layout=False` keeps the reference's two
    permute copies (same results bit for bit — every kernel computes a row independently of its position)."""
    if getattr(block, "_chunk_size", None) is not None:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "block forward: chunked feed-forward is not supported")
    BF, S, C = hidden_states.shape
    B = BF // num_frames
    attn2 = getattr(block, "attn2", None)
    in_place = fold_layout and num_frames <= 64 and (
        attn2 is None or encoder_hidden_states is None or encoder_hidden_states.shape[1] <= 8)
    if in_place:
        # every op but the attention cores is row-wise: stay in the [B*T, S, C] token order, no permute copies
        x = hidden_states.contiguous()
        x = feed_forward(block.ff_in, x, residual=x if block.is_res else None, pre_norm=block.norm_in)   # :441-449
        x = _attn(block.attn1, x, None, x, frames=num_frames, pre_norm=block.norm1)                      # :451-453
        if attn2 is not None:
            x = _cross_attn(block, block.norm2, attn2, x, encoder_hidden_states, frames=num_frames)   # :456-460
        return feed_forward(block.ff, x, residual=x if block.is_res else None, pre_norm=block.norm3)     # :462-472
    # [B*T, S, C] -> [B*S, T, C] (:435-438)
    x = hidden_states.reshape(B, num_frames, S, C).permute(0, 2, 1, 3).reshape(B * S, num_frames, C)
    x = feed_forward(block.ff_in, x, residual=x if block.is_res else None, pre_norm=block.norm_in)   # :441-449
    x = _attn(block.attn1, x, None, x, pre_norm=block.norm1)                                         # :451-453
    if getattr(block, "attn2", None) is not None:
        x = _cross_attn(block, block.norm2, block.attn2, x, encoder_hidden_states)   # :456-460
    x = feed_forward(block.ff, x, residual=x if block.is_res else None, pre_norm=block.norm3)        # :462-472
    # [B*S, T, C] -> [B*T, S, C] (:474-476)
    return x.reshape(B, S, num_frames, x.shape[-1]).permute(0, 2, 1, 3).reshape(B * num_frames, S, x.shape[-1])


def _is_basic(m) -> bool:
    return all(hasattr(m, a) for a in ("norm1", "attn1", "norm3", "ff")) and not hasattr(m, "norm_in")


def _is_temporal(m) -> bool:
    return all(hasattr(m, a) for a in ("norm_in", "ff_in", "norm1", "attn1", "norm3", "ff", "is_res"))


def install_blocks(model: torch.nn.Module, processor: Optional[TviAttnProcessor] = None) -> int:
    """Re-binds the forward of every (Temporal)BasicTransformerBlock-shaped module; returns how many."""
    proc = processor or TviAttnProcessor()
    caches.invalidate_caches()
    n = 0
    for _, m in model.named_modules():
        if _is_temporal(m):
            fwd = temporal_basic_transformer_block_forward
        elif _is_basic(m):
            fwd = basic_transformer_block_forward
        else:
            continue
        for a in ("attn1", "attn2"):
            attn = getattr(m, a, None)
            if attn is not None and not isinstance(getattr(attn, "processor", None), TviAttnProcessor):
                if not hasattr(attn, "_tvi_saved_processor"):
                    attn._tvi_saved_processor = attn.processor
                attn.set_processor(proc) if hasattr(attn, "set_processor") else setattr(attn, "processor", proc)
        if not hasattr(m, "_tvi_saved_block_forward"):
            m._tvi_saved_block_forward = m.__dict__.get("forward")
        m.forward = types.MethodType(fwd, m)
        n += 1
    return n


def uninstall_blocks(model: torch.nn.Module) -> int:
    caches.invalidate_caches()
    n = 0
    for _, m in model.named_modules():
        if hasattr(m, "_tvi_saved_block_forward"):
            if m._tvi_saved_block_forward is None:
                m.__dict__.pop("forward", None)
            else:
                m.forward = m._tvi_saved_block_forward
            del m._tvi_saved_block_forward
            n += 1
    return n
