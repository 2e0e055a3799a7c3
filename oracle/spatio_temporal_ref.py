"""CPU restatement of TransformerSpatioTemporalModel.forward (test infrastructure; never imported by the product).

  TransformerSpatioTemporalModel.forward   src/models/transformer_temporal.py:253-360
  Timesteps / TimestepEmbedding            upstream diffusers 0.24 models/embeddings.py (un-vendored; restated)
  AlphaBlender ("learned_with_images")     upstream diffusers 0.24 models/resnet.py      (un-vendored; restated)
Pinned by tests/golden/spatio_temporal.npz: outputs of the reference's OWN TransformerSpatioTemporalModel class over
those three restated leaves (tests/golden/make_golden.py:make_spatio_temporal); the leaves themselves are unpinned.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from .blocks_ref import basic_transformer_block_ref, temporal_basic_transformer_block_ref


def timesteps_ref(t: torch.Tensor, num_channels: int, flip_sin_to_cos: bool = True, shift: float = 0.0) -> torch.Tensor:
    half = num_channels // 2
    exponent = -math.log(10000) * torch.arange(half, dtype=torch.float32) / (half - shift)
    emb = t[:, None].float() * torch.exp(exponent)[None, :]
    emb = torch.cat([torch.sin(emb), torch.cos(emb)], dim=-1)
    if flip_sin_to_cos:
        emb = torch.cat([emb[:, half:], emb[:, :half]], dim=-1)
    return emb


def _p(t, dtype):
    return None if t is None else t.detach().to("cpu", dtype)


def alpha_ref(mixer, image_only_indicator: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    a = torch.where(image_only_indicator.to("cpu").bool(), torch.ones(1, 1), torch.sigmoid(_p(mixer.mix_factor, torch.float32))[..., None])
    return a.reshape(-1)[:, None, None].to(dtype)


def transformer_spatio_temporal_ref(model, hidden_states, encoder_hidden_states, image_only_indicator, dtype=torch.float32,
                                    alpha_dtype=None):
    """`alpha_dtype`: the reference casts alpha to the activation dtype (`.to(x_spatial.dtype)`); pass torch.bfloat16 to
    restate a bf16 run's blend weights while keeping the rest of the arithmetic in `dtype`."""
    x = hidden_states.detach().to("cpu", dtype)
    BF, C, H, W = x.shape
    T = image_only_indicator.shape[-1]
    B = BF // T
    if isinstance(encoder_hidden_states, tuple):
        sp, tm = encoder_hidden_states
    else:
        sp = tm = encoder_hidden_states
    sp, tm = sp.detach().to("cpu", dtype), tm.detach().to("cpu", dtype)
    tc = tm[None, :].reshape(B, T, -1, tm.shape[-1])[:, 0]                                               # :281-289
    tc = tc[None, :].broadcast_to(H * W, B, tc.shape[-2], tc.shape[-1]).reshape(H * W * B, tc.shape[-2], tc.shape[-1])  # :291-296
    residual = x
    x = F.group_norm(x, model.norm.num_groups, _p(model.norm.weight, dtype), _p(model.norm.bias, dtype), model.norm.eps)  # :300
    x = x.permute(0, 2, 3, 1).reshape(BF, H * W, C)                                                      # :302-304
    x = F.linear(x, _p(model.proj_in.weight, dtype), _p(model.proj_in.bias, dtype))                      # :306
    idx = torch.arange(T).repeat(B, 1).reshape(-1)                                                       # :308-310
    t_emb = timesteps_ref(idx, model.in_channels).to(dtype)                                              # :311-313
    tp = model.time_pos_embed
    emb = F.linear(F.silu(F.linear(t_emb, _p(tp.linear_1.weight, dtype), _p(tp.linear_1.bias, dtype))),
                   _p(tp.linear_2.weight, dtype), _p(tp.linear_2.bias, dtype))[:, None, :]                # :314-315
    alpha = alpha_ref(model.time_mixer, image_only_indicator, alpha_dtype or dtype).to(dtype)
    for block, tblock in zip(model.transformer_blocks, model.temporal_transformer_blocks):
        x = basic_transformer_block_ref(block, x, sp, dtype)                                             # :337-339
        xm = temporal_basic_transformer_block_ref(tblock, x + emb, T, tc, dtype)                         # :341-347
        x = alpha * x + (1.0 - alpha) * xm                                                               # :349-351
    x = F.linear(x, _p(model.proj_out.weight, dtype), _p(model.proj_out.bias, dtype))                    # :353
    x = x.reshape(BF, H, W, C).permute(0, 3, 1, 2).contiguous()                                          # :354
    return x + residual                                                                                  # :356
