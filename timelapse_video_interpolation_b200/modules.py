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

    def __init__(self, dim: int, heads: int, cross_attention_dim: Optional[int] = None, processor=None,
                 attention_head_dim: Optional[int] = None):
        super().__init__()
        dh = attention_head_dim if attention_head_dim is not None else (dim // heads if dim % heads == 0 and dim // heads in (64, 128) else 64)
        self.use_layer_norm = True
        self.use_ada_layer_norm = self.use_ada_layer_norm_zero = self.use_ada_layer_norm_single = False
        self.only_cross_attention = False
        self.pos_embed = None
        self._chunk_size = None
        self.norm1 = nn.LayerNorm(dim)
        self.attn1 = Attention(dim, None, heads=heads, dim_head=dh, processor=processor)
        self.norm2 = nn.LayerNorm(dim)
        self.attn2 = Attention(dim, cross_attention_dim, heads=heads, dim_head=dh, processor=processor)
        self.norm3 = nn.LayerNorm(dim)
        self.ff = FeedForward(dim)

    def forward(self, hidden_states, attention_mask=None, encoder_hidden_states=None, **kw):
        hidden_states = self.attn1(self.norm1(hidden_states)) + hidden_states
        hidden_states = self.attn2(self.norm2(hidden_states), encoder_hidden_states=encoder_hidden_states) + hidden_states
        return self.ff(self.norm3(hidden_states)) + hidden_states


class TemporalBasicTransformerBlock(nn.Module):
    """Attribute layout of the reference's temporal block (src/models/attention.py:350-478)."""

    def __init__(self, dim: int, time_mix_inner_dim: int, heads: int, cross_attention_dim: Optional[int] = None, processor=None,
                 attention_head_dim: Optional[int] = None):
        super().__init__()
        t = time_mix_inner_dim
        dh = attention_head_dim if attention_head_dim is not None else (t // heads if t % heads == 0 and t // heads in (64, 128) else 64)
        self.is_res = dim == time_mix_inner_dim
        self._chunk_size = None
        self.norm_in = nn.LayerNorm(dim)
        self.ff_in = FeedForward(dim, dim_out=time_mix_inner_dim)
        self.norm1 = nn.LayerNorm(time_mix_inner_dim)
        self.attn1 = Attention(time_mix_inner_dim, None, heads=heads, dim_head=dh, processor=processor)
        self.norm2 = nn.LayerNorm(time_mix_inner_dim)
        self.attn2 = Attention(time_mix_inner_dim, cross_attention_dim, heads=heads, dim_head=dh, processor=processor)
        self.norm3 = nn.LayerNorm(time_mix_inner_dim)
        self.ff = FeedForward(time_mix_inner_dim)


class Timesteps(nn.Module):
    """Sinusoidal frame-position embedding (upstream diffusers 0.24 `models/embeddings.py:Timesteps` /
    `get_timestep_embedding`, not vendored by the reference): exponent_i = -ln(10000) * i / (half - shift),
    emb = t * exp(exponent), [sin | cos] swapped to [cos | sin] when `flip_sin_to_cos`."""

    def __init__(self, num_channels: int, flip_sin_to_cos: bool, downscale_freq_shift: float):
        super().__init__()
        self.num_channels, self.flip_sin_to_cos, self.downscale_freq_shift = num_channels, flip_sin_to_cos, downscale_freq_shift

    def forward(self, timesteps):
        import math
        half = self.num_channels // 2
        exponent = -math.log(10000) * torch.arange(half, dtype=torch.float32, device=timesteps.device) / (half - self.downscale_freq_shift)
        emb = timesteps[:, None].float() * torch.exp(exponent)[None, :]
        emb = torch.cat([torch.sin(emb), torch.cos(emb)], dim=-1)
        if self.flip_sin_to_cos:
            emb = torch.cat([emb[:, half:], emb[:, :half]], dim=-1)
        if self.num_channels % 2 == 1:
            emb = torch.nn.functional.pad(emb, (0, 1, 0, 0))
        return emb


class TimestepEmbedding(nn.Module):
    """linear_1 -> SiLU -> linear_2 (upstream diffusers 0.24 `models/embeddings.py:TimestepEmbedding`)."""

    def __init__(self, in_channels: int, time_embed_dim: int, out_dim: Optional[int] = None):
        super().__init__()
        self.linear_1 = nn.Linear(in_channels, time_embed_dim)
        self.act = nn.SiLU()
        self.linear_2 = nn.Linear(time_embed_dim, out_dim if out_dim is not None else time_embed_dim)

    def forward(self, sample):
        return self.linear_2(self.act(self.linear_1(sample)))


class AlphaBlender(nn.Module):
    """upstream diffusers 0.24 `models/resnet.py:AlphaBlender`, `merge_strategy="learned_with_images"` (the one
    TransformerSpatioTemporalModel builds, src/models/transformer_temporal.py:243): alpha = 1 where
    image_only_indicator, sigmoid(mix_factor) elsewhere; out = alpha * x_spatial + (1 - alpha) * x_temporal."""

    def __init__(self, alpha: float, merge_strategy: str = "learned_with_images", switch_spatial_to_temporal_mix: bool = False):
        super().__init__()
        self.merge_strategy = merge_strategy
        self.switch_spatial_to_temporal_mix = switch_spatial_to_temporal_mix
        if merge_strategy == "fixed":
            self.register_buffer("mix_factor", torch.Tensor([alpha]))
        else:
            self.register_parameter("mix_factor", nn.Parameter(torch.Tensor([alpha])))

    def get_alpha(self, image_only_indicator, ndims: int):
        if self.merge_strategy == "fixed":
            alpha = self.mix_factor
        elif self.merge_strategy == "learned":
            alpha = torch.sigmoid(self.mix_factor)
        else:
            alpha = torch.where(image_only_indicator.bool(), torch.ones(1, 1, device=image_only_indicator.device),
                                torch.sigmoid(self.mix_factor)[..., None])
            alpha = alpha.reshape(-1)[:, None, None] if ndims == 3 else alpha[:, None, :, None, None]
        return alpha

    def forward(self, x_spatial, x_temporal, image_only_indicator=None):
        alpha = self.get_alpha(image_only_indicator, x_spatial.ndim).to(x_spatial.dtype)
        if self.switch_spatial_to_temporal_mix:
            alpha = 1.0 - alpha
        return alpha * x_spatial + (1.0 - alpha) * x_temporal


class TransformerSpatioTemporalModel(nn.Module):
    """Attribute layout and state_dict keys of the reference module (src/models/transformer_temporal.py:207-251)."""

    def __init__(self, num_attention_heads: int = 16, attention_head_dim: int = 88, in_channels: int = 320,
                 out_channels: Optional[int] = None, num_layers: int = 1, cross_attention_dim: Optional[int] = None):
        super().__init__()
        if attention_head_dim not in (64, 128):
            raise ValueError("libtvi builds d = 64 (every level of the SVD checkpoint) and d = 128 heads (the class default "
                             "num_attention_heads=(5, 10, 10, 20) at C = 1280)")
        inner_dim = num_attention_heads * attention_head_dim
        self.num_attention_heads, self.attention_head_dim = num_attention_heads, attention_head_dim
        self.inner_dim, self.in_channels = inner_dim, in_channels
        self.norm = nn.GroupNorm(num_groups=32, num_channels=in_channels, eps=1e-6)
        self.proj_in = nn.Linear(in_channels, inner_dim)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner_dim, num_attention_heads, cross_attention_dim, attention_head_dim=attention_head_dim)
             for _ in range(num_layers)])
        self.temporal_transformer_blocks = nn.ModuleList(
            [TemporalBasicTransformerBlock(inner_dim, inner_dim, num_attention_heads, cross_attention_dim,
                                           attention_head_dim=attention_head_dim) for _ in range(num_layers)])
        self.time_pos_embed = TimestepEmbedding(in_channels, in_channels * 4, out_dim=in_channels)
        self.time_proj = Timesteps(in_channels, True, 0)
        self.time_mixer = AlphaBlender(alpha=0.5, merge_strategy="learned_with_images")
        self.out_channels = in_channels if out_channels is None else out_channels
        self.proj_out = nn.Linear(inner_dim, in_channels)
        self.gradient_checkpointing = False

    def forward(self, hidden_states, encoder_hidden_states=None, image_only_indicator=None, return_dict: bool = True):
        raise RuntimeError("parameter container: call spatio_temporal.install_spatio_temporal(model) to bind the libtvi forward")
