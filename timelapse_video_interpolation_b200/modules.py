"""`Attention` — a diffusers-free parameter container with the reference module's interface.

diffusers is not installed on the build or GPU boxes, so tests and benchmarks need an
object the processor can be installed on.  This mirrors the SVD-relevant subset of
`Attention.__init__` / `set_processor` / `forward` (reference src/models/attention_processor.py:88-207,
:371-404, :496-528): same attribute names, same parameter shapes and state_dict keys
(`to_q.weight`, `to_k.weight`, `to_v.weight`, `to_out.0.weight`, `to_out.0.bias`), same dispatch to
`self.processor`.  It holds no attention arithmetic of its own.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn


class Attention(nn.Module):
    def __init__(self, query_dim: int, cross_attention_dim: Optional[int] = None, heads: int = 8, dim_head: int = 64,
                 dropout: float = 0.0, bias: bool = False, out_bias: bool = True, processor=None):
        super().__init__()
        self.inner_dim = dim_head * heads
        self.cross_attention_dim = cross_attention_dim if cross_attention_dim is not None else query_dim
        self.upcast_attention = False
        self.upcast_softmax = False
        self.rescale_output_factor = 1.0
        self.residual_connection = False
        self.dropout = dropout
        self.scale = dim_head ** -0.5            # reference :127
        self.heads = heads                       # reference :129
        self.sliceable_head_dim = heads
        self.added_kv_proj_dim = None
        self.only_cross_attention = False
        self.group_norm = None
        self.spatial_norm = None
        self.norm_cross = None
        self.to_q = nn.Linear(query_dim, self.inner_dim, bias=bias)                    # :181
        self.to_k = nn.Linear(self.cross_attention_dim, self.inner_dim, bias=bias)     # :185
        self.to_v = nn.Linear(self.cross_attention_dim, self.inner_dim, bias=bias)     # :186
        self.to_out = nn.ModuleList([nn.Linear(self.inner_dim, query_dim, bias=out_bias), nn.Dropout(dropout)])  # :195-197
        if processor is None:
            from .processor import TviAttnProcessor
            processor = TviAttnProcessor()
        self.set_processor(processor)

    def set_processor(self, processor, _remove_lora: bool = False) -> None:
        # reference :371-404: a processor that is not an nn.Module is simply assigned
        if hasattr(self, "processor") and isinstance(self.processor, nn.Module) and not isinstance(processor, nn.Module):
            self._modules.pop("processor")
        self.processor = processor

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **cross_attention_kwargs):
        # reference :522-528
        return self.processor(self, hidden_states, encoder_hidden_states=encoder_hidden_states,
                              attention_mask=attention_mask, **cross_attention_kwargs)


class GEGLU(nn.Module):
    """Parameter container of upstream diffusers' GEGLU (`proj = nn.Linear(dim_in, 2 * dim_out)`); the eager
    forward below is only used where no libtvi forward is installed (CPU-side tests of the host logic)."""

    def __init__(self, dim_in: int, dim_out: int):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)

    def forward(self, hidden_states, scale: float = 1.0):
        hidden_states, gate = self.proj(hidden_states).chunk(2, dim=-1)
        return hidden_states * torch.nn.functional.gelu(gate)


class FeedForward(nn.Module):
    """diffusers-free mirror of the reference `FeedForward` container (src/models/attention.py:481-535) for the
    `geglu` activation SVD uses: same sub-module layout and state_dict keys (`net.0.proj.*`, `net.2.*`)."""

    def __init__(self, dim: int, dim_out: Optional[int] = None, mult: int = 4, dropout: float = 0.0):
        super().__init__()
        inner_dim = int(dim * mult)
        dim_out = dim_out if dim_out is not None else dim
        self.net = nn.ModuleList([GEGLU(dim, inner_dim), nn.Dropout(dropout), nn.Linear(inner_dim, dim_out)])

    def forward(self, hidden_states, scale: float = 1.0):
        for module in self.net:  # reference :528-535
            hidden_states = module(hidden_states)
        return hidden_states


class BasicTransformerBlock(nn.Module):
    """Parameter container with the attribute layout of the reference's spatial block for SVD
    (src/models/attention.py:90-232 with norm_type="layer_norm", GEGLU feed-forward, attention_bias=False):
    norm1, attn1, norm2, attn2, norm3, ff.  The eager forward is the reference's order of operations."""

    def __init__(self, dim: int, heads: int, cross_attention_dim: Optional[int] = None, processor=None):
        super().__init__()
        self.use_layer_norm = True
        self.use_ada_layer_norm = self.use_ada_layer_norm_zero = self.use_ada_layer_norm_single = False
        self.only_cross_attention = False
        self.pos_embed = None
        self._chunk_size = None
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, None, heads=heads, processor=processor)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, cross_attention_dim, heads=heads, processor=processor)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)

    def forward(self, hidden_states, attention_mask=None, encoder_hidden_states=None, **kw):
        hidden_states = self.attn1(self.norm1(hidden_states)) + hidden_states
        hidden_states = self.attn2(self.norm2(hidden_states), encoder_hidden_states=encoder_hidden_states) + hidden_states
        return self.ff(self.norm3(hidden_states)) + hidden_states


class TemporalBasicTransformerBlock(nn.Module):
    """Attribute layout of the reference's temporal block (src/models/attention.py:350-478)."""

    def __init__(self, dim: int, time_mix_inner_dim: int, heads: int, cross_attention_dim: Optional[int] = None, processor=None):
        super().__init__()
        self.is_res = dim == time_mix_inner_dim
        self._chunk_size = None
        self.norm_in = nn.LayerNorm(dim)
        self.ff_in = FeedForward(dim, dim_out=time_mix_inner_dim)
        self.norm1 = nn.LayerNorm(time_mix_inner_dim)
        self.attn1 = Attention(time_mix_inner_dim, None, heads=heads, processor=processor)
        self.norm2 = nn.LayerNorm(time_mix_inner_dim)
        self.attn2 = Attention(time_mix_inner_dim, cross_attention_dim, heads=heads, processor=processor)
        self.norm3 = nn.LayerNorm(time_mix_inner_dim)
        self.ff = FeedForward(time_mix_inner_dim)
