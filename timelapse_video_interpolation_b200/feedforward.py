"""FeedForward / GEGLU / LayerNorm of the transformer blocks on libtvi.

Drop-in for `FeedForward.forward(hidden_states, scale=1.0)` (reference src/models/attention.py:481-535; its
`net[0]` is upstream diffusers' `GEGLU`: `proj = nn.Linear(dim, 2*inner)`, `hidden, gate = proj(x).chunk(2, -1)`,
`hidden * F.gelu(gate)`; `net[2] = nn.Linear(inner, dim_out)`) and for the `nn.LayerNorm` calls in front of it
(:321, :441, :462).  Two tcgen05 GEMM launches per call:

    h = GEGLU(x)     one GEMM over the re-ordered [2*inner, dim] projection weight, bias + value * gelu(gate)
                     fused in the epilogue: the [tokens, 2*inner] intermediate never reaches HBM
    y = h W2^T + b2 (+ residual)   bias and the block's residual add fused in the epilogue

`install_feed_forward(model)` re-binds `forward` on every FeedForward-shaped module (duck-typed like the attention
plugin).  Inference only for now: a call that needs gradients raises (the backward of this path is not built).
"""
from __future__ import annotations

import types
import weakref
from typing import Optional

import torch

from . import _lib, ops

__all__ = ["feed_forward", "layer_norm", "install_feed_forward", "uninstall_feed_forward"]

_cache: weakref.WeakKeyDictionary = weakref.WeakKeyDictionary()  # weak: an id() key could be recycled by a later module


def _is_feed_forward(m) -> bool:
    net = getattr(m, "net", None)
    return (net is not None and len(net) >= 3 and hasattr(net[0], "proj") and hasattr(net[0].proj, "weight")
            and hasattr(net[2], "weight"))


def _geglu_weights(proj):
    """proj.weight [2*inner, dim] (value rows then gate rows, the `.chunk(2, -1)` order) -> rows interleaved in
    blocks of H = tvi_geglu_block(inner); cached per (data_ptr, version) so state_dict keys stay untouched."""
    w, b = proj.weight, proj.bias
    sig = (w.data_ptr(), w._version, w.dtype, None if b is None else (b.data_ptr(), b._version))
    hit = _cache.get(proj)
    if hit is not None and hit[0] == sig:
        return hit[1], hit[2]
    inner = w.shape[0] // 2
    H = ops.geglu_block(inner)
    if H == 0:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"GEGLU inner width {inner} is not a multiple of 32")
    with torch.no_grad():
        wb = w.detach().to(torch.bfloat16)
        wi = torch.stack([wb[:inner].view(inner // H, H, -1), wb[inner:].view(inner // H, H, -1)], dim=1)
        wi = wi.reshape(2 * inner, -1).contiguous()
        bi = None
        if b is not None:
            bb = b.detach().to(torch.bfloat16)
            bi = torch.stack([bb[:inner].view(inner // H, H), bb[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous()
    _cache[proj] = (sig, wi, bi)
    return wi, bi


def _bf16(t):
    return None if t is None else (t if t.dtype == torch.bfloat16 else t.detach().to(torch.bfloat16)).contiguous()


def feed_forward(ff, hidden_states: torch.Tensor, scale: float = 1.0, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`ff(hidden_states)` of a FeedForward module with GEGLU activation; `residual` (same shape as the output)
    is added in the second GEMM's epilogue (the blocks' `ff_output + hidden_states`)."""
    if not _is_feed_forward(ff):
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "feed_forward: expected a FeedForward module with a GEGLU projection")
    if getattr(ff.net[0].proj, "lora_layer", None) is not None or getattr(ff.net[2], "lora_layer", None) is not None:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "feed_forward: LoRA layers are not supported by the libtvi path")
    for m in ff.net:
        if isinstance(m, torch.nn.Dropout) and m.p != 0.0 and ff.training:
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "feed_forward: dropout with p > 0 is not supported")
    x = hidden_states
    if not x.is_cuda:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "feed_forward: hidden_states is not on a CUDA device (there is no CPU fallback)")
    if x.dtype != torch.bfloat16:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"feed_forward: hidden_states is {x.dtype}; the libtvi path is bf16")
    if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in ff.parameters())):
        raise NotImplementedError("libtvi FeedForward backward is not built: call under torch.no_grad() or keep the stock "
                                  "FeedForward for training (DESIGN.md §7)")
    x = x.contiguous()
    wi, bi = _geglu_weights(ff.net[0].proj)
    h = ops.gemm_bf16_tn_ex(x.view(-1, x.shape[-1]), wi, bi, geglu=True)
    out_lin = ff.net[2]
    r2 = None
    if residual is not None:
        r2 = residual.contiguous().view(-1, residual.shape[-1])
    y = ops.gemm_bf16_tn_ex(h, _bf16(out_lin.weight), _bf16(out_lin.bias), residual=r2)
    return y.view(*x.shape[:-1], -1)


def layer_norm(norm, x: torch.Tensor) -> torch.Tensor:
    """`norm(x)` for an nn.LayerNorm over the last dimension (fp32 statistics in the kernel)."""
    if tuple(norm.normalized_shape) != (x.shape[-1],):
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "layer_norm: only normalisation over the last dimension is built")
    if x.dtype != torch.bfloat16 or not x.is_cuda:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "layer_norm: expected a CUDA bfloat16 tensor")
    return ops.layernorm_bf16(x.contiguous(), _bf16(norm.weight), _bf16(norm.bias), norm.eps)


def install_feed_forward(model: torch.nn.Module) -> int:
    """Re-binds `forward` of every FeedForward-shaped module of `model` to the libtvi path; returns how many."""
    n = 0
    for _, m in model.named_modules():
        if not _is_feed_forward(m):
            continue
        if not hasattr(m, "_tvi_saved_ff_forward"):
            m._tvi_saved_ff_forward = m.__dict__.get("forward")

        def fwd(self, hidden_states, scale: float = 1.0, *a, **k):
            return feed_forward(self, hidden_states, scale)

        m.forward = types.MethodType(fwd, m)
        n += 1
    return n


def uninstall_feed_forward(model: torch.nn.Module) -> int:
    n = 0
    for _, m in model.named_modules():
        if hasattr(m, "_tvi_saved_ff_forward"):
            if m._tvi_saved_ff_forward is None:
                m.__dict__.pop("forward", None)
            else:
                m.forward = m._tvi_saved_ff_forward
            del m._tvi_saved_ff_forward
            n += 1
    return n
