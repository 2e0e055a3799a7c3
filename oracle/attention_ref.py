"""CPU restatement of the reference's attention processor math (test infrastructure).

Follows, operation by operation, the explicit eager path of the reference:
  Attention.forward -> processor dispatch      src/models/attention_processor.py:496-528
  head_to_batch_dim / batch_to_head_dim        :530-568
  get_attention_scores (baddbmm alpha=scale -> softmax -> cast)   :570-615
  AttnProcessor.__call__ order of operations   :694-766  (== AttnProcessor2_0 :1187-1263 up to
                                               SDPA's internal accumulation order)
and the monkey-patched temporal forward of src/train_decouples.py:377-495 (ramp bias :437-457,
per-frame mean-logit statistics :460-471).

`attn` is any object with to_q/to_k/to_v/to_out[0] (.weight, .bias), .heads, .scale — the reference's
own Attention, upstream diffusers' or the package's parameter container.  Math runs on CPU in
`dtype` (fp32 by default, fp64 for tie-breaking) on copies of the parameters.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def _w(lin, dtype):
    w = lin.weight.detach().to("cpu", dtype)
    b = lin.bias.detach().to("cpu", dtype) if getattr(lin, "bias", None) is not None else None
    return w, b


def head_to_batch_dim(t: torch.Tensor, heads: int) -> torch.Tensor:
    # :546-568  [B, S, h*d] -> [B*h, S, d]
    B, S, D = t.shape
    return t.reshape(B, S, heads, D // heads).permute(0, 2, 1, 3).reshape(B * heads, S, D // heads)


def batch_to_head_dim(t: torch.Tensor, heads: int) -> torch.Tensor:
    # :530-544  [B*h, S, d] -> [B, S, h*d]
    Bh, S, d = t.shape
    return t.reshape(Bh // heads, heads, S, d).permute(0, 2, 1, 3).reshape(Bh // heads, S, heads * d)


def get_attention_scores(query, key, scale: float, bias=None):
    # :598-615 with attention_mask None: baddbmm(beta=0, alpha=scale) -> softmax(dim=-1)
    scores = scale * torch.bmm(query, key.transpose(-1, -2))
    if bias is not None:
        scores = scores + bias
    return scores.softmax(dim=-1), scores


def attn_processor_ref(attn, hidden_states: torch.Tensor, encoder_hidden_states=None, dtype=torch.float32,
                       round_to=None) -> torch.Tensor:
    """AttnProcessor.__call__ (:699-766) for the SVD case (3-D input, no mask/norms/residual).

    round_to: optional dtype (torch.bfloat16) to round to after every nn.Linear and after the
    attention core, which is where the reference's bf16 model rounds.
    """
    def rnd(t):
        return t.to(round_to).to(dtype) if round_to is not None else t

    x = hidden_states.detach().to("cpu", dtype)
    e = x if encoder_hidden_states is None else encoder_hidden_states.detach().to("cpu", dtype)
    heads = int(attn.heads)
    wq, bq = _w(attn.to_q, dtype)
    wk, bk = _w(attn.to_k, dtype)
    wv, bv = _w(attn.to_v, dtype)
    wo, bo = _w(attn.to_out[0], dtype)
    query = rnd(F.linear(x, wq, bq))           # :729
    key = rnd(F.linear(e, wk, bk))             # :736
    value = rnd(F.linear(e, wv, bv))           # :737
    query = head_to_batch_dim(query, heads)    # :745-747
    key = head_to_batch_dim(key, heads)
    value = head_to_batch_dim(value, heads)
    probs, _ = get_attention_scores(query, key, float(attn.scale))  # :749
    out = rnd(torch.bmm(probs, value))         # :750
    out = batch_to_head_dim(out, heads)        # :751
    out = rnd(F.linear(out, wo, bo))           # :754 (dropout :756 is p=0; /rescale_output_factor == 1 :764)
    return out


def ca_temporal_forward_ref(attn, hidden_states, encoder_hidden_states=None, use_attentionmask=False,
                            dtype=torch.float32):
    """Patched temporal forward of src/train_decouples.py:377-495.

    Returns (output, stats) where stats is None for self-attention and otherwise a [T, 2] tensor:
    stats[f, 0] / stats[f, 1] = mean over (batch, heads, keys of the first / second half) of the
    pre-softmax logits (after the optional ramp bias) of query frame f — the values the reference
    pulls out with 2*T .item() calls (:460-471).
    """
    x = hidden_states.detach().to("cpu", dtype)
    do_cross = encoder_hidden_states is not None
    e = encoder_hidden_states.detach().to("cpu", dtype) if do_cross else x
    heads = int(attn.heads)
    wq, bq = _w(attn.to_q, dtype)
    wk, bk = _w(attn.to_k, dtype)
    wv, bv = _w(attn.to_v, dtype)
    wo, bo = _w(attn.to_out[0], dtype)
    B = e.shape[0]
    q = F.linear(x, wq, bq).view(B, -1, heads, wq.shape[0] // heads).transpose(1, 2)   # :397, :411
    k = F.linear(e, wk, bk).view(B, -1, heads, wq.shape[0] // heads).transpose(1, 2)   # :403, :416
    v = F.linear(e, wv, bv).view(B, -1, heads, wq.shape[0] // heads).transpose(1, 2)   # :404, :426
    L, S = q.size(-2), k.size(-2)
    w = q @ k.transpose(-2, -1) * float(attn.scale)                                    # :447 (scale = d^-0.5)
    stats = None
    if use_attentionmask and do_cross:
        assert S == 2, "Expecting only start and end frames as KV"                     # :439
        ws = torch.linspace(0, 1, steps=L, dtype=dtype)
        w = w + torch.stack([1 - ws, ws], dim=1)[None, None]                           # :440-457
    if do_cross:
        half = int(S / 2)
        stats = torch.stack([w[:, :, :, :half].mean(dim=(0, 1, 3)) if half > 0 else torch.full((L,), float("nan"), dtype=dtype),
                             w[:, :, :, half:].mean(dim=(0, 1, 3))], dim=1)           # :460-471
    p = torch.softmax(w, dim=-1)                                                       # :474
    o = (p @ v).transpose(1, 2).reshape(B, -1, wq.shape[0])                            # :476, :483
    return F.linear(o, wo, bo), stats                                                  # :486


def attn_processor_grads_ref(attn, hidden_states, encoder_hidden_states, grad_output, use_attentionmask=False,
                             dtype=torch.float32):
    """Gradients of the processor call by torch autograd on CPU — what the reference's training scripts
    compute (src/train_decouples.py back-propagates through AttnProcessor2_0 / the patched temporal
    forward with plain autograd; there is no custom backward in the reference).

    Same math as attn_processor_ref / ca_temporal_forward_ref (the optional (T,2) ramp bias of
    :437-457 is a constant).  Returns (output, grads) with grads keyed
    x, e, to_q, to_k, to_v, to_out_w, to_out_b; `e` is None for self-attention.
    """
    x = hidden_states.detach().to("cpu", dtype).requires_grad_(True)
    do_cross = encoder_hidden_states is not None
    e = encoder_hidden_states.detach().to("cpu", dtype).requires_grad_(True) if do_cross else None
    heads = int(attn.heads)
    ws = {n: getattr(attn, n).weight.detach().to("cpu", dtype).requires_grad_(True) for n in ("to_q", "to_k", "to_v")}
    wo = attn.to_out[0].weight.detach().to("cpu", dtype).requires_grad_(True)
    bo = attn.to_out[0].bias.detach().to("cpu", dtype).requires_grad_(True)
    src = e if do_cross else x
    query = head_to_batch_dim(F.linear(x, ws["to_q"]), heads)
    key = head_to_batch_dim(F.linear(src, ws["to_k"]), heads)
    value = head_to_batch_dim(F.linear(src, ws["to_v"]), heads)
    bias = None
    if use_attentionmask and do_cross:
        L = x.shape[1]
        w = torch.linspace(0, 1, steps=L, dtype=dtype)
        bias = torch.stack([1 - w, w], dim=1)[None]
    probs, _ = get_attention_scores(query, key, float(attn.scale), bias)
    out = F.linear(batch_to_head_dim(torch.bmm(probs, value), heads), wo, bo)
    out.backward(grad_output.detach().to("cpu", dtype))
    grads = {"x": x.grad, "e": e.grad if do_cross else None, "to_q": ws["to_q"].grad, "to_k": ws["to_k"].grad,
             "to_v": ws["to_v"].grad, "to_out_w": wo.grad, "to_out_b": bo.grad}
    return out.detach(), grads
