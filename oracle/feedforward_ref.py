"""CPU restatement of the FeedForward / GEGLU / LayerNorm math of the transformer blocks (test infrastructure).

  FeedForward.forward                  src/models/attention.py:526-535 (iterates net: GEGLU, Dropout(p=0), Linear)
  GEGLU.forward                        upstream diffusers 0.24 `models/activations.py` (NOT in the reference tree, the
                                       `thirdparty/diffusers` submodule is empty): proj -> chunk(2, dim=-1) ->
                                       hidden_states * F.gelu(gate)  (erf form).  Parity for this LEAF is UNPINNED by
                                       the reference; tests/golden/blocks.npz pins everything around it by driving
                                       the reference's own FeedForward / block classes over a GEGLU with exactly this
                                       forward (tests/golden/make_golden.py:make_blocks).
  residual add                         src/models/attention.py:343, :449, :472 (`ff_output + hidden_states`)
  nn.LayerNorm                         :270, :300, :321, :441, :446, :462
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def feed_forward_ref(ff, hidden_states: torch.Tensor, residual=None, dtype=torch.float32) -> torch.Tensor:
    x = hidden_states.detach().to("cpu", dtype)
    proj, out = ff.net[0].proj, ff.net[2]
    h = F.linear(x, proj.weight.detach().to("cpu", dtype), None if proj.bias is None else proj.bias.detach().to("cpu", dtype))
    hidden, gate = h.chunk(2, dim=-1)
    h = hidden * F.gelu(gate)
    y = F.linear(h, out.weight.detach().to("cpu", dtype), None if out.bias is None else out.bias.detach().to("cpu", dtype))
    if residual is not None:
        y = y + residual.detach().to("cpu", dtype)
    return y


def layer_norm_ref(norm, x: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    w = None if norm.weight is None else norm.weight.detach().to("cpu", dtype)
    b = None if norm.bias is None else norm.bias.detach().to("cpu", dtype)
    return F.layer_norm(x.detach().to("cpu", dtype), tuple(norm.normalized_shape), w, b, norm.eps)
