"""FeedForward / GEGLU / LayerNorm of the transformer blocks on libtvi.

Drop-in for `FeedForward.forward(hidden_states, scale=1.0)` (reference src/models/attention.py:481-535; its
`net[0]` is upstream diffusers' `GEGLU`: `proj = nn.Linear(dim, 2*inner)`, `hidden, gate = proj(x).chunk(2, -1)`,
`hidden * F.gelu(gate)`; `net[2] = nn.Linear(inner, dim_out)`) and for the `nn.LayerNorm` calls in front of it
(:321, :441, :462).  Two tcgen05 GEMM launches per call:

    h = GEGLU(x)     one GEMM over the re-ordered [2*inner, dim] projection weight, bias + value * gelu(gate)
                     fused in the epilogue: the [tokens, 2*inner] intermediate never reaches HBM
    y = h W2^T + b2 (+ residual)   bias and the block's residual add fused in the epilogue

`install_feed_forward(model)` re-binds `forward` on every FeedForward-shaped module (duck-typed like the attention
plugin).  A call that needs gradients goes on the autograd tape through `_FeedForwardFn` / `_LayerNormFn` (backward
kernels: geglu_bwd, layernorm_bwd, GEMM dgrad / wgrad); the no-grad path uses the fused GEGLU / residual epilogues.
"""
from __future__ import annotations

import types
import weakref
from typing import Optional

import torch

from . import _lib, caches, lnfold, ops

__all__ = ["feed_forward", "layer_norm", "linear", "install_feed_forward", "uninstall_feed_forward"]

_cache: weakref.WeakKeyDictionary = weakref.WeakKeyDictionary()  # weak: an id() key could be recycled by a later module


def _is_feed_forward(m) -> bool:
    net = getattr(m, "net", None)
    return (net is not None and len(net) >= 3 and hasattr(net[0], "proj") and hasattr(net[0].proj, "weight")
            and hasattr(net[2], "weight"))


def _geglu_weights(proj):
    """proj.weight [2*inner, dim] (value rows then gate rows, the `.chunk(2, -1)` order) -> rows interleaved in
    blocks of H = tvi_geglu_block(inner); cached per (cache generation, data_ptr, version) so state_dict keys stay
    untouched (caches.py explains the generation: `.data` writes do not bump `_version`)."""
    w, b = proj.weight, proj.bias
    sig = (caches.generation(), w.data_ptr(), w._version, w.dtype, None if b is None else (b.data_ptr(), b._version))
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


def _cast_grad(g, like):
    return None if g is None else (g if g.dtype == like.dtype else g.to(like.dtype)).contiguous().view(like.shape)


# Bias gradient of the first FeedForward projection without a pass of its own: the LayerNorm in front writes its rows into
# a buffer _PAD columns wider, with a 1.0 (and zeros) behind every row; `dW1 = d(pre)^T [x | 1 0 ... 0]` then carries
# db1 = d(pre)^T 1 as column C of the weight-gradient GEMM, whose last column tile was padded to 128 anyway.  Saves the
# [M, 2*inner] column-sum launch (0.25 ms at L0) per trainable FeedForward.
_PAD = 16
_PAD_USED = 8
_pad_rows = {}


def _pad_pattern(device):
    t = _pad_rows.get(device)
    if t is None:
        t = torch.zeros(_PAD, dtype=torch.bfloat16, device=device)
        t[0] = 1.0
        _pad_rows[device] = t
    return t


class _FeedForwardFn(torch.autograd.Function):
    """Tape entry of `Linear(C -> 2*inner) -> GEGLU -> Linear(inner -> C_out)` (reference src/models/attention.py:526-535).
    Saves only x: the backward re-runs the first projection with the GEGLU-backward epilogue (`ops.gemm_geglu_bwd`: h and
    d(pre) straight from the fp32 accumulator; the 2*inner-wide projection is never written, forward or backward) and
    runs the dgrad / wgrad / column-sum kernels of the projections."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, residual=None, x_padded=False):
        if x_padded:  # rows of a [M, C + _PAD] buffer written by _LayerNormResFn(pad_ones=True): keep the row stride
            x2 = x.detach().as_strided((x.numel() // x.shape[-1], x.shape[-1]), (x.stride(-2), 1), x.storage_offset())
        else:
            x2 = x.detach().contiguous().view(x.numel() // max(x.shape[-1], 1), x.shape[-1])
        ctx.x_padded = bool(x_padded)
        with torch.no_grad():
            w1b, w2b = _bf16(w1), _bf16(w2)
            inner = w1b.shape[0] // 2
            H = ops.geglu_block(inner)
            if H == 0:
                raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"GEGLU inner width {inner} is not a multiple of 32")
            wi = torch.stack([w1b[:inner].view(inner // H, H, -1), w1b[inner:].view(inner // H, H, -1)], dim=1).reshape(2 * inner, -1)
            bi = None
            if b1 is not None:
                bb = _bf16(b1)
                bi = torch.stack([bb[:inner].view(inner // H, H), bb[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous()
            wi = wi.contiguous()
            h = ops.gemm_bf16_tn_ex(x2, wi, bi, geglu=True)
            if residual is None:
                y = ops.gemm_bf16_tn(h, w2b, _bf16(b2))
            else:  # `ff_output + hidden_states` in the second GEMM's epilogue; its gradient is dy itself
                y = ops.gemm_bf16_tn_ex(h, w2b, _bf16(b2), residual=residual.detach().contiguous().view(x2.shape[0], -1))
        ctx.save_for_backward(x2, w1, b1, w2, b2)
        ctx.wi, ctx.bi = wi, bi          # the interleaved bf16 copies (1.6 MB at C = 320): the backward re-runs the same GEMM
        ctx.x_shape = x.shape
        return y.view(*x.shape[:-1], y.shape[-1])

    @staticmethod
    def backward(ctx, dy):
        x2, w1, b1, w2, b2 = ctx.saved_tensors
        need_x, need_w1, need_b1, need_w2, need_b2 = ctx.needs_input_grad[:5]
        need_res = len(ctx.needs_input_grad) > 5 and ctx.needs_input_grad[5]
        dy2 = dy.contiguous().view(-1, dy.shape[-1]).to(torch.bfloat16)
        w1b, w2b = _bf16(w1), _bf16(w2)
        dh = ops.gemm_bf16_dgrad(dy2, w2b)                               # [M, inner]
        # the first projection again, its fp32 accumulator turned into h and d(pre) in the epilogue (never written itself)
        h, dpre = ops.gemm_geglu_bwd(x2, ctx.wi, ctx.bi, dh, need_h=need_w2)
        dw2 = ops.gemm_bf16_wgrad(dy2, h) if need_w2 else None
        db2 = ops.colsum_bf16(dy2, dy2.shape[0]).view(-1) if (need_b2 and b2 is not None) else None
        need_b1 = need_b1 and b1 is not None
        if ctx.x_padded and need_w1 and need_b1:
            C = x2.shape[1]
            xa = x2.as_strided((x2.shape[0], C + _PAD_USED), (x2.stride(0), 1), x2.storage_offset())  # [x | 1 0 ... 0]
            dwa = ops.gemm_bf16_wgrad(dpre, xa)
            dw1, db1 = dwa[:, :C], dwa[:, C]
        else:
            dw1 = ops.gemm_bf16_wgrad(dpre, x2) if need_w1 else None
            db1 = ops.colsum_bf16(dpre, dpre.shape[0]).view(-1) if need_b1 else None
        dx = ops.gemm_bf16_dgrad(dpre, w1b).view(ctx.x_shape) if need_x else None
        return dx, _cast_grad(dw1, w1), _cast_grad(db1, b1) if b1 is not None else None, _cast_grad(dw2, w2), \
            _cast_grad(db2, b2) if b2 is not None else None, (dy if need_res else None), None


class _LinearFn(torch.autograd.Function):
    """Tape entry of `F.linear(x, w, b)` on the tcgen05 GEMM: dx = dy W (dgrad), dW = dy^T x (wgrad, fixed summation order),
    db = column sums.  Used for proj_in / proj_out of the module-level forward when it trains."""

    @staticmethod
    def forward(ctx, x, w, b):
        x2 = x.detach().contiguous().view(x.numel() // max(x.shape[-1], 1), x.shape[-1])
        y = ops.gemm_bf16_tn(x2, _bf16(w), _bf16(b))
        ctx.save_for_backward(x2, w, b)
        ctx.xshape = x.shape
        return y.view(*x.shape[:-1], y.shape[-1])

    @staticmethod
    def backward(ctx, dy):
        x2, w, b = ctx.saved_tensors
        need_x, need_w, need_b = ctx.needs_input_grad
        dy2 = dy.contiguous().to(torch.bfloat16).view(x2.shape[0], -1)
        dx = ops.gemm_bf16_dgrad(dy2, _bf16(w)).view(ctx.xshape) if need_x else None
        dw = _cast_grad(ops.gemm_bf16_wgrad(dy2, x2), w) if need_w else None
        db = _cast_grad(ops.colsum_bf16(dy2).view(-1), b) if (need_b and b is not None) else None
        return dx, dw, db


def linear(lin, x: torch.Tensor) -> torch.Tensor:
    """`lin(x)` for an nn.Linear-shaped module on the libtvi GEMM; on the autograd tape when a gradient is needed."""
    if getattr(lin, "lora_layer", None) is not None:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "linear: LoRA layers are not supported by the libtvi path")
    if torch.is_grad_enabled() and (x.requires_grad or lin.weight.requires_grad or (lin.bias is not None and lin.bias.requires_grad)):
        return _LinearFn.apply(x, lin.weight, lin.bias)
    x2 = x.contiguous().view(x.numel() // max(x.shape[-1], 1), x.shape[-1])
    return ops.gemm_bf16_tn(x2, _bf16(lin.weight), _bf16(lin.bias)).view(*x.shape[:-1], lin.weight.shape[0])


class _LayerNormFn(torch.autograd.Function):
    """Tape entry of nn.LayerNorm over the last dimension; the backward recomputes the statistics from x."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        xc = x.detach().contiguous()
        y = ops.layernorm_bf16(xc, _bf16(gamma), _bf16(beta), eps)
        ctx.save_for_backward(xc, gamma, beta)
        ctx.eps = eps
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, gamma, beta = ctx.saved_tensors
        need_x, need_g, need_b, _ = ctx.needs_input_grad
        affine = (need_g and gamma is not None) or (need_b and beta is not None)
        dx, dgamma, dbeta = ops.layernorm_bwd(xc, dy.contiguous().to(torch.bfloat16), _bf16(gamma), ctx.eps, need_dx=need_x,
                                              need_affine=affine)
        return dx, _cast_grad(dgamma, gamma) if (need_g and gamma is not None) else None, \
            _cast_grad(dbeta, beta) if (need_b and beta is not None) else None, None


class _LayerNormResFn(torch.autograd.Function):
    """Entry of a residual branch `x + f(LayerNorm(x))` (every sub-layer of the two blocks): returns (LayerNorm(x), x).
    The second output is what the caller adds back (in the epilogue of f's last GEMM); its gradient arrives here together
    with the gradient of the normalised rows, and the LayerNorm backward kernel adds it to dx in the same pass
    (`tvi_layernorm_bwd_res_bf16`) — autograd would otherwise sum the two with one more elementwise kernel per sub-layer."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps, pad_ones=False):
        xc = x.detach().contiguous()
        if pad_ones:  # rows land in a [M, C + _PAD] buffer with [1 0 ... 0] behind each (see _PAD)
            C = xc.shape[-1]
            buf = torch.empty(xc.numel() // C, C + _PAD, dtype=torch.bfloat16, device=xc.device)
            buf[:, C:] = _pad_pattern(xc.device)
            ops.layernorm_bf16(xc.view(-1, C), _bf16(gamma), _bf16(beta), eps, out=buf[:, :C])
            y = buf[:, :C].view(*xc.shape)
        else:
            y = ops.layernorm_bf16(xc, _bf16(gamma), _bf16(beta), eps)
        ctx.save_for_backward(xc, gamma, beta)
        ctx.eps = eps
        ctx.set_materialize_grads(False)
        return y, x.view_as(x)

    @staticmethod
    def backward(ctx, dn, dres):
        xc, gamma, beta = ctx.saved_tensors
        need_x, need_g, need_b = ctx.needs_input_grad[:3]
        need_g, need_b = need_g and gamma is not None, need_b and beta is not None
        if dres is not None:
            dres = dres.contiguous().to(torch.bfloat16)
        if dn is None:  # nothing flowed through f (a one-key cross-attention): the branch passes dres through
            dg = torch.zeros_like(gamma) if need_g else None
            db = torch.zeros_like(beta) if need_b else None
            return (dres if need_x else None), dg, db, None, None
        dx, dgamma, dbeta = ops.layernorm_bwd(xc, dn.contiguous().to(torch.bfloat16), _bf16(gamma), ctx.eps, need_dx=need_x,
                                              need_affine=need_g or need_b, dres=dres if need_x else None)
        return dx, _cast_grad(dgamma, gamma) if need_g else None, _cast_grad(dbeta, beta) if need_b else None, None, None


def layer_norm_residual(norm, x: torch.Tensor, pad_ones: bool = False):
    """(norm(x), x) on the autograd tape for a residual branch (see _LayerNormResFn); training path only.  `pad_ones`: the
    normalised rows are a strided view into a wider buffer with [1 0 ... 0] behind each row (see _PAD)."""
    if tuple(norm.normalized_shape) != (x.shape[-1],):
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "layer_norm: only normalisation over the last dimension is built")
    if x.dtype != torch.bfloat16 or not x.is_cuda:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "layer_norm: expected a CUDA bfloat16 tensor")
    return _LayerNormResFn.apply(x, norm.weight, norm.bias, float(norm.eps), bool(pad_ones))


def feed_forward(ff, hidden_states: torch.Tensor, scale: float = 1.0, residual: Optional[torch.Tensor] = None,
                 pre_norm=None) -> torch.Tensor:
    """`ff(hidden_states)` of a FeedForward module with GEGLU activation; `residual` (same shape as the output)
    is added in the second GEMM's epilogue (the blocks' `ff_output + hidden_states`).
    `pre_norm` (an nn.LayerNorm): computes `ff(pre_norm(hidden_states))`; on the no-grad path the LayerNorm is folded
    into the first GEMM (lnfold.py) and the normalised rows never reach HBM."""
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
    on_tape = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in ff.parameters()))
    caches.note_call(on_tape)
    if pre_norm is not None and torch.is_grad_enabled() and any(p.requires_grad for p in pre_norm.parameters()):
        on_tape = True
    if on_tape:
        # training: the call goes on the autograd tape.  The residual rides in the second GEMM's epilogue (its gradient is
        # dy itself); when it is the LayerNorm's own input — the blocks' `x + ff(norm(x))` — its gradient is added inside the
        # LayerNorm backward kernel
        proj, out_lin = ff.net[0].proj, ff.net[2]
        padded = False
        if pre_norm is not None and residual is x:
            # the bias gradient of `proj` rides in its weight-gradient GEMM when both are wanted (see _PAD)
            padded = (proj.bias is not None and proj.bias.requires_grad and proj.weight.requires_grad and x.shape[-1] % 16 == 0
                      and x.numel() > 0)
            x, residual = layer_norm_residual(pre_norm, x, pad_ones=padded)
        elif pre_norm is not None:
            x = layer_norm(pre_norm, x)
        if residual is not None and residual.shape[-1] != out_lin.weight.shape[0]:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "feed_forward: residual must have the shape of the output")
        return _FeedForwardFn.apply(x, proj.weight, proj.bias, out_lin.weight, out_lin.bias, residual, padded)
    x = x.contiguous()
    wi, bi = _geglu_weights(ff.net[0].proj)
    x2 = x.view(x.numel() // x.shape[-1], x.shape[-1])
    if pre_norm is not None and lnfold.supported(pre_norm, x):
        wg, colsum, cb = lnfold.folded(ff.net[0].proj, "geglu", wi, bi, pre_norm)
        h = ops.gemm_bf16_tn_ln(x2, wg, cb, colsum, ops.layernorm_stats(x2, float(pre_norm.eps)), geglu=True)
    else:
        if pre_norm is not None:
            x2 = layer_norm(pre_norm, x2)
        h = ops.gemm_bf16_tn_ex(x2, wi, bi, geglu=True)
    out_lin = ff.net[2]
    r2 = None
    if residual is not None:
        r2 = residual.contiguous().view(residual.numel() // residual.shape[-1], residual.shape[-1])
    y = ops.gemm_bf16_tn_ex(h, _bf16(out_lin.weight), _bf16(out_lin.bias), residual=r2)
    return y.view(*x.shape[:-1], y.shape[-1])


def layer_norm(norm, x: torch.Tensor) -> torch.Tensor:
    """`norm(x)` for an nn.LayerNorm over the last dimension (fp32 statistics in the kernel)."""
    if tuple(norm.normalized_shape) != (x.shape[-1],):
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "layer_norm: only normalisation over the last dimension is built")
    if x.dtype != torch.bfloat16 or not x.is_cuda:
        raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, "layer_norm: expected a CUDA bfloat16 tensor")
    if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in norm.parameters())):
        return _LayerNormFn.apply(x, norm.weight, norm.bias, float(norm.eps))
    return ops.layernorm_bf16(x.contiguous(), _bf16(norm.weight), _bf16(norm.bias), norm.eps)


def install_feed_forward(model: torch.nn.Module) -> int:
    """Re-binds `forward` of every FeedForward-shaped module of `model` to the libtvi path; returns how many."""
    caches.invalidate_caches()
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
    caches.invalidate_caches()
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
