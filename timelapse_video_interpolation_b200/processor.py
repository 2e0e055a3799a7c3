"""`TviAttnProcessor` — the diffusers AttentionProcessor plugin backed by libtvi.

Drop-in for `AttnProcessor2_0` / `XFormersAttnProcessor` / `AttnProcessor`
(reference src/models/attention_processor.py:1178-1263, :1081-1175, :694-766): same call signature,
same result layout, installed with `attn.set_processor(p)` / `unet.set_attn_processor(p)` (reference
src/models/unet_spatio_temporal_condition.py:278-310) or with `install(model)` below, which also
re-owns the `forward` that src/train_decouples.py:375-516 monkey-patches onto `*temporal*` modules.

The processor is duck-typed on the `attn` module (two `Attention` class identities reach it,
reference src/models/transformer_temporal.py:22-24): it reads `heads`, `scale`, `to_q/to_k/to_v`,
`to_out[0]`, and raises on anything outside the SVD invariants — there is no fallback path.

Per call (all kernels from libtvi, on the current CUDA stream, no host sync):
  self-attention : fused-QKV tcgen05 GEMM -> attention core (temporal kernel if the sequence has
                   <= 64 tokens, flash spatial kernel otherwise) reading heads in place from the
                   packed QKV buffer -> out-projection GEMM with bias.
  cross-attention: fused-KV GEMM on the context rows; with one key the softmax is exactly 1, so
                   the out projection is applied to the B' value rows and broadcast; otherwise Q GEMM
                   -> few-keys core (optional (T,2) ramp bias + fused mean-logit statistics,
                   reference src/train_decouples.py:437-471) -> out GEMM.
"""
from __future__ import annotations

import types
import weakref
from typing import Optional

import torch

from . import _lib, caches, ops
from .autograd import attention_forward

__all__ = ["TviAttnProcessor", "install", "uninstall"]


def _linear_weight(lin, what: str) -> torch.Tensor:
    if getattr(lin, "lora_layer", None) is not None:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"{what}: LoRA layers are not supported by the libtvi path")
    return lin.weight


class TviAttnProcessor:
    """AttentionProcessor whose arithmetic runs in hand-written sm_100a kernels.

    Options (used by `install` for the train_decouples / test.py instrumentation mode):
      ramp_bias : add the reference's linspace (T,2) bias to the temporal cross-attention logits
                  (src/train_decouples.py:437-457; requires exactly 2 context tokens).
      stats_sink: callable(stats[T, L] fp32 CUDA tensor, attn) invoked after every temporal
                  cross-attention call with the per-frame mean pre-softmax logits
                  (src/train_decouples.py:460-471, but without the 2*T host syncs).
      cast_inputs: accept fp16/fp32 activations and weights by casting to bf16 and back (the kernels
                  compute in bf16 with fp32 accumulation only). Off by default.
    """

    def __init__(self, ramp_bias: bool = False, stats_sink=None, cast_inputs: bool = False):
        self.ramp_bias = ramp_bias
        self.stats_sink = stats_sink
        self.cast_inputs = cast_inputs
        # keyed by the module OBJECT (weakly): an id() key can be recycled by a later module whose parameters land on
        # the freed addresses, which would serve the dead module's fused weights
        self._wcache: weakref.WeakKeyDictionary = weakref.WeakKeyDictionary()
        self._bias_cache: dict = {}

    # ------------------------------------------------------------------ weight caches
    def _fused(self, attn, names: tuple[str, ...]) -> torch.Tensor:
        """Derived, never-registered fused weight ([sum(out), in], bf16); rebuilt when any source
        parameter changes (optimizer step bumps `_version`) or the cache generation advances (caches.py: `.data`
        writes such as EMA copy_to are invisible to `_version`), so state_dict keys are untouched."""
        ws = [_linear_weight(getattr(attn, n), n) for n in names]
        sig = (caches.generation(),) + tuple((w.data_ptr(), w._version, w.dtype, w.device) for w in ws)
        per_module = self._wcache.get(attn)
        if per_module is None:
            per_module = self._wcache.setdefault(attn, {})
        hit = per_module.get(names)
        if hit is not None and hit[0] == sig:
            return hit[1]
        with torch.no_grad():
            fused = torch.cat([w.detach().to(torch.bfloat16) for w in ws], dim=0).contiguous() if len(ws) > 1 \
                else ws[0].detach().to(torch.bfloat16).contiguous()
        per_module[names] = (sig, fused)
        return fused

    def _ramp(self, T: int, device) -> torch.Tensor:
        key = (T, device)
        b = self._bias_cache.get(key)
        if b is None:
            w = torch.linspace(0, 1, steps=T, device=device, dtype=torch.float32)
            b = torch.stack([1 - w, w], dim=1).contiguous()
            self._bias_cache[key] = b
        return b

    # ------------------------------------------------------------------ invariants
    @staticmethod
    def _check(attn, hidden_states, attention_mask):
        if attention_mask is not None:
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "attention_mask is not supported (SVD blocks never pass one)")
        for name in ("group_norm", "spatial_norm"):
            if getattr(attn, name, None) is not None:
                raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"attn.{name} is set; the libtvi path covers SVD blocks only")
        if getattr(attn, "norm_cross", None):
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "attn.norm_cross is set; not an SVD block")
        if getattr(attn, "residual_connection", False):
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "attn.residual_connection=True is not supported")
        if getattr(attn, "rescale_output_factor", 1.0) != 1.0:
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "attn.rescale_output_factor != 1 is not supported")
        if getattr(attn, "upcast_attention", False) or getattr(attn, "upcast_softmax", False):
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "upcast_attention / upcast_softmax are not supported "
                                                          "(softmax is always fp32 inside the kernels)")
        if hidden_states.dim() != 3:
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "hidden_states must be [batch, tokens, channels]")
        if not hidden_states.is_cuda:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "hidden_states is not on a CUDA device (there is no CPU fallback)")
        drop = attn.to_out[1] if len(attn.to_out) > 1 else None
        if drop is not None and getattr(drop, "p", 0.0) != 0.0 and getattr(attn, "training", False):
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "attention output dropout with p > 0 is not supported")

    # ------------------------------------------------------------------ the call
    def __call__(self, attn, hidden_states: torch.Tensor, encoder_hidden_states: Optional[torch.Tensor] = None,
                 attention_mask: Optional[torch.Tensor] = None, temb: Optional[torch.Tensor] = None,
                 scale: float = 1.0, *args, residual: Optional[torch.Tensor] = None, frames: Optional[int] = None,
                 pre_norm=None, **kwargs) -> torch.Tensor:
        """libtvi extensions used by the block-level forwards of blocks.py:
        `residual` — a tensor of the output's shape added in the out-projection epilogue, i.e. the caller's
        `attn_output + hidden_states`;
        `frames` = T — hidden_states is [B*T, S, C] in the spatial token order and attention runs over the T tokens of
        every (b, s) (encoder_hidden_states: the temporal [B*S, L, Ckv] context): the temporal block without its two
        permute copies (src/models/attention.py:436-438, :474-476).  Inference only.
        `pre_norm` — the nn.LayerNorm the block applies in front of this attention: hidden_states is then the
        UN-normalised input and the LayerNorm is folded into the query-side projection (lnfold.py)."""
        self._check(attn, hidden_states, attention_mask)
        in_dtype = hidden_states.dtype
        if in_dtype != torch.bfloat16:
            if not self.cast_inputs:
                raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED,
                                    f"hidden_states is {in_dtype}; the libtvi path is bf16 (pass cast_inputs=True to convert)")
            hidden_states = hidden_states.to(torch.bfloat16)
            if encoder_hidden_states is not None:
                encoder_hidden_states = encoder_hidden_states.to(torch.bfloat16)
        x = hidden_states.contiguous()
        is_cross = encoder_hidden_states is not None
        heads = int(attn.heads)
        bo = attn.to_out[0].bias
        wo = _linear_weight(attn.to_out[0], "to_out[0]")

        bias = stats = None
        if is_cross:
            e = encoder_hidden_states.contiguous()
            L = e.shape[1]
            Tq = x.shape[1] if frames is None else int(frames)   # bias / statistics rows are the query's frame index
            if self.ramp_bias:
                if L != 2:
                    raise _lib.TviError(_lib.TVI_ERR_INVALID, "ramp_bias expects exactly 2 context tokens "
                                                              "(reference asserts num_kv_frames == 2)")
                bias = self._ramp(Tq, x.device)
            if self.stats_sink is not None:
                stats = torch.empty(Tq, L, dtype=torch.float32, device=x.device)
        else:
            e = None

        if residual is not None:
            if residual.shape != hidden_states.shape:
                raise _lib.TviError(_lib.TVI_ERR_INVALID, "residual must have the shape of hidden_states")
            residual = residual.to(torch.bfloat16).contiguous()
        if frames is not None and (frames < 1 or x.shape[0] % frames != 0):
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "frames must divide the batch dimension of hidden_states")
        y = attention_forward(self, attn, x, e, heads, float(attn.scale), wo, bo, bias, stats, residual, frames, pre_norm)

        if stats is not None:
            self.stats_sink(stats, attn)
        if in_dtype != torch.bfloat16:
            y = y.to(in_dtype)
        return y


# ---------------------------------------------------------------------- install / uninstall
def _is_attention(m) -> bool:
    return all(hasattr(m, a) for a in ("to_q", "to_k", "to_v", "to_out", "heads", "scale")) and hasattr(m, "processor")


def _is_svd_attention(m) -> bool:
    return (getattr(m, "group_norm", None) is None and getattr(m, "spatial_norm", None) is None
            and not getattr(m, "norm_cross", None) and getattr(m, "added_kv_proj_dim", None) is None)


def install(model: torch.nn.Module, processor: Optional[TviAttnProcessor] = None, temporal_ramp_bias: bool = False,
            temporal_stats_sink=None, strict: bool = True) -> int:
    """Installs the libtvi processor on every Attention module of `model`; returns how many.

    Call it AFTER `enable_xformers_memory_efficient_attention()` (reference src/test.py:95, which
    overwrites processors) and AFTER `register_attention_control` (reference
    src/train_decouples.py:545): instance-level `forward` overrides found on `*temporal*` modules
    are replaced by a forward that dispatches to the processor with the same bias / statistics
    semantics (options `temporal_ramp_bias`, `temporal_stats_sink`).
    """
    caches.invalidate_caches()
    shared = processor or TviAttnProcessor()
    temporal = None
    n = 0
    for name, m in model.named_modules():
        if not _is_attention(m):
            continue
        if not _is_svd_attention(m):
            if strict:
                raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"{name}: not an SVD-style Attention module")
            continue
        p = shared
        if "forward" in m.__dict__:  # monkey-patched instance forward (D4): re-own it
            if not hasattr(m, "_tvi_saved_forward"):
                m._tvi_saved_forward = m.__dict__["forward"]
            if "temporal" in name and (temporal_ramp_bias or temporal_stats_sink is not None):
                if temporal is None:
                    temporal = TviAttnProcessor(ramp_bias=temporal_ramp_bias, stats_sink=temporal_stats_sink,
                                                cast_inputs=shared.cast_inputs)
                p = temporal

            def fwd(self, hidden_states, encoder_hidden_states=None, attention_mask=None, temb=None, scale=1.0, **kw):
                return self.processor(self, hidden_states, encoder_hidden_states=encoder_hidden_states,
                                      attention_mask=attention_mask, temb=temb, scale=scale, **kw)

            m.forward = types.MethodType(fwd, m)
        if not hasattr(m, "_tvi_saved_processor"):
            m._tvi_saved_processor = m.processor
        if hasattr(m, "set_processor"):
            m.set_processor(p)
        else:
            m.processor = p
        n += 1
    return n


def uninstall(model: torch.nn.Module) -> int:
    """Restores the processors / forwards that `install` replaced."""
    caches.invalidate_caches()
    n = 0
    for _, m in model.named_modules():
        if hasattr(m, "_tvi_saved_processor"):
            m.processor = m._tvi_saved_processor
            del m._tvi_saved_processor
            n += 1
        if hasattr(m, "_tvi_saved_forward"):
            m.forward = m._tvi_saved_forward
            del m._tvi_saved_forward
    return n
