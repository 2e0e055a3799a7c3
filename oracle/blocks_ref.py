"""CPU restatement of the two transformer-block forwards (test infrastructure), built from the attention and
feed-forward restatements:
  BasicTransformerBlock.forward          src/models/attention.py:234-347 (use_layer_norm branch; no gligen / ada norms)
  TemporalBasicTransformerBlock.forward  src/models/attention.py:423-478
Pinned by tests/golden/blocks.npz (outputs of the reference's own block classes, GEGLU leaf restated) in
tests/test_oracle_golden.py.
"""
from __future__ import annotations

import torch

from .attention_ref import attn_processor_ref
from .feedforward_ref import feed_forward_ref, layer_norm_ref


def basic_transformer_block_ref(block, hidden_states, encoder_hidden_states=None, dtype=torch.float32):
    x = hidden_states.detach().to("cpu", dtype)
    x = attn_processor_ref(block.attn1, layer_norm_ref(block.norm1, x, dtype), None, dtype) + x              # :270-287
    if getattr(block, "attn2", None) is not None:
        x = attn_processor_ref(block.attn2, layer_norm_ref(block.norm2, x, dtype), encoder_hidden_states, dtype) + x  # :296-317
    return feed_forward_ref(block.ff, layer_norm_ref(block.norm3, x, dtype), dtype=dtype) + x               # :321-343


def temporal_basic_transformer_block_ref(block, hidden_states, num_frames, encoder_hidden_states=None, dtype=torch.float32):
    x = hidden_states.detach().to("cpu", dtype)
    BF, S, C = x.shape
    B = BF // num_frames
    x = x[None, :].reshape(B, num_frames, S, C).permute(0, 2, 1, 3).reshape(B * S, num_frames, C)        # :435-438
    residual = x
    x = feed_forward_ref(block.ff_in, layer_norm_ref(block.norm_in, x, dtype), dtype=dtype)               # :441-446
    if block.is_res:
        x = x + residual                                                                                  # :448-449
    x = attn_processor_ref(block.attn1, layer_norm_ref(block.norm1, x, dtype), None, dtype) + x           # :451-453
    if getattr(block, "attn2", None) is not None:
        x = attn_processor_ref(block.attn2, layer_norm_ref(block.norm2, x, dtype), encoder_hidden_states, dtype) + x  # :456-460
    ff = feed_forward_ref(block.ff, layer_norm_ref(block.norm3, x, dtype), dtype=dtype)                   # :462-467
    x = ff + x if block.is_res else ff                                                                    # :469-472
    C2 = x.shape[-1]
    return x[None, :].reshape(B, S, num_frames, C2).permute(0, 2, 1, 3).reshape(B * num_frames, S, C2)    # :474-476
