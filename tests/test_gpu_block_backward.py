"""GPU: gradients through the block-level swap (training config 5): LayerNorm, FeedForward (GEGLU) and whole-block backward
against fp32 autograd of the same math (tolerance 2e-2 of max |grad|, as for the attention backward)."""
import pytest
import torch
import torch.nn.functional as F

import timelapse_video_interpolation_b200 as tvi
from timelapse_video_interpolation_b200 import blocks, ops
from timelapse_video_interpolation_b200.modules import BasicTransformerBlock, FeedForward, TemporalBasicTransformerBlock

pytestmark = pytest.mark.gpu
GTOL = 2e-2


def _close(name, got, ref):
    ref = ref.float().cpu()
    err = float((got.float().cpu() - ref).abs().max())
    scale = max(float(ref.abs().max()), 1e-6)
    print(f"  {name}: max abs err {err:.3e} / max |ref| {scale:.3e}")
    assert err <= GTOL * scale, name


@pytest.mark.parametrize("rows,C", [(300, 320), (130, 640), (77, 1280), (64, 96), (40, 2048)])
def test_layer_norm_backward(cuda, rows, C):
    g = torch.Generator().manual_seed(rows + C)
    norm = torch.nn.LayerNorm(C)
    with torch.no_grad():
        norm.weight.copy_(1 + 0.2 * torch.randn(C, generator=g))
        norm.bias.copy_(0.1 * torch.randn(C, generator=g))
    norm = norm.to(cuda, torch.bfloat16)
    x = (torch.randn(rows, C, generator=g) * 1.5 + 0.3).to(torch.bfloat16).to(cuda).requires_grad_(True)
    wgt = torch.randn(rows, C, generator=g).to(torch.bfloat16).to(cuda)
    y = tvi.layer_norm(norm, x)
    assert y.requires_grad
    (y.float() * wgt.float()).sum().backward()
    xr = x.detach().float().cpu().requires_grad_(True)
    gr, br = norm.weight.detach().float().cpu().requires_grad_(True), norm.bias.detach().float().cpu().requires_grad_(True)
    (F.layer_norm(xr, (C,), gr, br, norm.eps) * wgt.float().cpu()).sum().backward()
    _close("dx", x.grad, xr.grad)
    _close("dgamma", norm.weight.grad, gr.grad)
    _close("dbeta", norm.bias.grad, br.grad)


@pytest.mark.parametrize("rows,C,with_res,strided", [(40000, 320, True, False), (20001, 640, True, True), (19000, 320, False, True),
                                                    (5000, 1280, True, False), (7, 320, True, False)])
def test_layer_norm_backward_many_rows(cuda, rows, C, with_res, strided):
    """The row ring of the LayerNorm backward (bulk copies `depth` rows ahead of the arithmetic) over several grid strides,
    with the residual gradient folded in and with row strides wider than the row (column slices of wider buffers)."""
    g = torch.Generator(device="cuda").manual_seed(rows + C)
    ld = C + 64 if strided else C
    xb = (torch.randn(rows, ld, generator=g, device=cuda) * 1.5 + 0.3).to(torch.bfloat16)
    dyb = torch.randn(rows, ld, generator=g, device=cuda).to(torch.bfloat16)
    rb = torch.randn(rows, ld, generator=g, device=cuda).to(torch.bfloat16) if with_res else None
    x, dy, r = xb[:, :C], dyb[:, :C], (rb[:, :C] if with_res else None)
    gamma = (1 + 0.2 * torch.randn(C, generator=g, device=cuda)).to(torch.bfloat16)
    dx, dgm, dbt = ops.layernorm_bwd(x, dy, gamma, 1e-5, dres=r)
    xr = x.float().requires_grad_(True)
    gr, br = gamma.float().requires_grad_(True), torch.zeros(C, device=cuda, requires_grad=True)
    (F.layer_norm(xr, (C,), gr, br, 1e-5) * dy.float()).sum().backward()
    ref_dx = xr.grad + (r.float() if with_res else 0.0)
    _close("dx", dx, ref_dx)
    # sums over up to 40 000 rows, rounded to bf16 once at the end
    _close("dgamma", dgm, gr.grad)
    _close("dbeta", dbt, br.grad)
    # two calls give the same bits (fixed summation order, no atomics)
    dx2, dgm2, dbt2 = ops.layernorm_bwd(x, dy, gamma, 1e-5, dres=r)
    assert torch.equal(dx, dx2) and torch.equal(dgm, dgm2) and torch.equal(dbt, dbt2)


def _ff_ref(x, w1, b1, w2, b2):
    hidden, gate = F.linear(x, w1, b1).chunk(2, dim=-1)
    return F.linear(hidden * F.gelu(gate), w2, b2)


@pytest.mark.parametrize("rows,dim", [(300, 320), (1000, 640), (64, 96)])
def test_feed_forward_backward(cuda, rows, dim):
    g = torch.Generator().manual_seed(rows)
    ff = FeedForward(dim)
    with torch.no_grad():
        for p in ff.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.1))
    ff = ff.to(cuda, torch.bfloat16)
    x = torch.randn(rows, dim, generator=g).to(torch.bfloat16).to(cuda).requires_grad_(True)
    wgt = torch.randn(rows, dim, generator=g).to(torch.bfloat16).to(cuda)
    launches0 = ops.LAUNCHES
    y = tvi.feed_forward(ff, x)
    (y.float() * wgt.float()).sum().backward()
    assert ops.LAUNCHES - launches0 >= 9      # 2 forward GEMMs, recompute, dgrad x2, geglu_bwd, wgrad x2, column sums
    names = ["net.0.proj.weight", "net.0.proj.bias", "net.2.weight", "net.2.bias"]
    ps = dict(ff.named_parameters())
    ref = [ps[n].detach().float().cpu().requires_grad_(True) for n in names]
    xr = x.detach().float().cpu().requires_grad_(True)
    (_ff_ref(xr, *ref) * wgt.float().cpu()).sum().backward()
    _close("dx", x.grad, xr.grad)
    for n, r in zip(names, ref):
        _close(n, ps[n].grad, r.grad)
    # frozen weights: only dx is produced
    for p in ff.parameters():
        p.requires_grad_(False)
        p.grad = None
    x2 = x.detach().clone().requires_grad_(True)
    (tvi.feed_forward(ff, x2).float() * wgt.float()).sum().backward()
    assert torch.equal(x2.grad, x.grad) and all(p.grad is None for p in ff.parameters())


@pytest.mark.parametrize("shape,dim,freeze", [((2, 151, 320), 320, ()), ((1000, 640), 640, ()), ((2, 151, 320), 320, ("net.0.proj.bias",)),
                                              ((333, 320), 320, ("net.2.weight", "net.2.bias"))])
def test_residual_feed_forward_branch_backward(cuda, shape, dim, freeze):
    """`x + ff(norm(x))` as the blocks call it (pre_norm=, residual=x): on the tape the first projection's bias gradient
    comes out of its weight-gradient GEMM (a ones column behind the normalised rows, feedforward._PAD) when both are
    trainable, from the column-sum kernel otherwise; a frozen second projection drops the h output of the fused backward."""
    g = torch.Generator().manual_seed(dim + len(freeze))
    ff, norm = FeedForward(dim), torch.nn.LayerNorm(dim)
    with torch.no_grad():
        for p in ff.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.1))
        norm.weight.copy_(1 + 0.2 * torch.randn(dim, generator=g))
        norm.bias.copy_(0.1 * torch.randn(dim, generator=g))
    ff, norm = ff.to(cuda, torch.bfloat16), norm.to(cuda, torch.bfloat16)
    for n, p in ff.named_parameters():
        p.requires_grad_(n not in freeze)
    x = torch.randn(*shape, generator=g).to(torch.bfloat16).to(cuda).requires_grad_(True)
    wgt = torch.randn(*shape, generator=g).to(torch.bfloat16).to(cuda)
    y = tvi.feed_forward(ff, x, residual=x, pre_norm=norm)
    (y.float() * wgt.float()).sum().backward()
    names = ["net.0.proj.weight", "net.0.proj.bias", "net.2.weight", "net.2.bias"]
    ps = dict(ff.named_parameters())
    ref = [ps[n].detach().float().cpu().requires_grad_(True) for n in names]
    xr = x.detach().float().cpu().requires_grad_(True)
    gr, br = norm.weight.detach().float().cpu().requires_grad_(True), norm.bias.detach().float().cpu().requires_grad_(True)
    yr = xr + _ff_ref(F.layer_norm(xr, (dim,), gr, br, norm.eps), *ref)
    assert float((y.detach().float().cpu() - yr.detach()).abs().max()) <= 2e-2 * max(1.0, float(yr.detach().abs().max()))
    (yr * wgt.float().cpu()).sum().backward()
    _close("dx", x.grad, xr.grad)
    _close("dgamma", norm.weight.grad, gr.grad)
    _close("dbeta", norm.bias.grad, br.grad)
    for n, r in zip(names, ref):
        if n in freeze:
            assert ps[n].grad is None
        else:
            _close(n, ps[n].grad, r.grad)
            assert ps[n].grad.is_contiguous() and ps[n].grad.shape == ps[n].shape


def _attn_ref(x, e, wq, wk, wv, wo, bo, heads):
    B, S, C = x.shape
    ctx = x if e is None else e
    q, k, v = F.linear(x, wq), F.linear(ctx, wk), F.linear(ctx, wv)
    d = C // heads
    sp = lambda t: t.view(B, -1, heads, d).transpose(1, 2)
    o = torch.softmax(sp(q) @ sp(k).transpose(-1, -2) * d ** -0.5, dim=-1) @ sp(v)
    return F.linear(o.transpose(1, 2).reshape(B, S, C), wo, bo)


def _block_ref(P, x, e, heads, temporal=False, T=None):
    ln = lambda n, t: F.layer_norm(t, (t.shape[-1],), P[n + ".weight"], P[n + ".bias"], 1e-5)
    at = lambda n, t, c: _attn_ref(t, c, P[n + ".to_q.weight"], P[n + ".to_k.weight"], P[n + ".to_v.weight"], P[n + ".to_out.0.weight"],
                                   P[n + ".to_out.0.bias"], heads)
    ffn = lambda n, t: _ff_ref(t, P[n + ".net.0.proj.weight"], P[n + ".net.0.proj.bias"], P[n + ".net.2.weight"], P[n + ".net.2.bias"])
    if temporal:
        BF, S, C = x.shape
        B = BF // T
        x = x.reshape(B, T, S, C).permute(0, 2, 1, 3).reshape(B * S, T, C)
        x = ffn("ff_in", ln("norm_in", x)) + x
    x = at("attn1", ln("norm1", x), None) + x
    x = at("attn2", ln("norm2", x), e) + x
    x = ffn("ff", ln("norm3", x)) + x
    if temporal:
        x = x.reshape(B, S, T, C).permute(0, 2, 1, 3).reshape(B * T, S, C)
    return x


def _init_block(blk, seed, cuda):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in blk.named_parameters():
            if "norm" in name:
                p.copy_(1 + 0.1 * torch.randn(p.shape, generator=g) if name.endswith("weight") else 0.05 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.05))
    return blk.to(cuda, torch.bfloat16)


CHECK = ["norm1.weight", "norm3.bias", "attn1.to_q.weight", "attn2.to_v.weight", "attn2.to_out.0.bias", "ff.net.0.proj.weight",
         "ff.net.0.proj.bias", "ff.net.2.weight"]


@pytest.mark.parametrize("temporal", [False, True], ids=["spatial", "temporal"])
def test_block_backward_matches_fp32_autograd(cuda, temporal):
    C, heads, T = 320, 5, 14
    if temporal:
        blk = _init_block(TemporalBasicTransformerBlock(C, C, heads, 1024), 8, cuda)
        B, S, L = 1, 24, 2
        x_shape, e_shape = (B * T, S, C), (B * S, L, 1024)
    else:
        blk = _init_block(BasicTransformerBlock(C, heads, 1024), 7, cuda)
        x_shape, e_shape = (2, 144, C), (2, 2, 1024)
    blocks.install_blocks(blk)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(x_shape, generator=g).to(torch.bfloat16).to(cuda).requires_grad_(True)
    e = torch.randn(e_shape, generator=g).to(torch.bfloat16).to(cuda).requires_grad_(True)
    wgt = torch.randn(x_shape, generator=g).to(torch.bfloat16).to(cuda)
    y = blk(x, num_frames=T, encoder_hidden_states=e) if temporal else blk(x, encoder_hidden_states=e)
    assert y.requires_grad and y.shape == x.shape
    (y.float() * wgt.float()).sum().backward()
    P = {n: p.detach().float().cpu().requires_grad_(True) for n, p in blk.named_parameters()}
    xr, er = x.detach().float().cpu().requires_grad_(True), e.detach().float().cpu().requires_grad_(True)
    yr = _block_ref(P, xr, er, heads, temporal, T)
    _close("forward", y.detach(), yr.detach())
    (yr * wgt.float().cpu()).sum().backward()
    _close("dx", x.grad, xr.grad)
    _close("d encoder_hidden_states", e.grad, er.grad)
    ps = dict(blk.named_parameters())
    for n in CHECK + (["norm_in.weight", "ff_in.net.2.bias"] if temporal else []):
        _close(n, ps[n].grad, P[n].grad)


@pytest.mark.parametrize("L,ramp", [(1, False), (2, False), (2, True)])
def test_temporal_block_trains_in_the_spatial_token_order(cuda, L, ramp):
    # N2 in training: the temporal block stays in [B*T, S, C] in BOTH directions (interleaved temporal / few-keys backward cores,
    # interleaved column sums) — same gradients as the reference's permute -> block -> permute (src/models/attention.py:436-438,
    # :474-476).  dx is bit-identical (every kernel on its path works row by row); weight gradients sum the tokens in another order.
    from timelapse_video_interpolation_b200 import TviAttnProcessor
    B, T, S, C = 2, 14, 40, 320
    blk = _init_block(TemporalBasicTransformerBlock(C, C, 5, 1024), 23, cuda)
    if ramp:
        blk.attn2.processor = TviAttnProcessor(ramp_bias=True)
    blocks.install_blocks(blk)
    g = torch.Generator().manual_seed(L)
    x = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(cuda)
    e = torch.randn(B * S, L, 1024, generator=g).to(torch.bfloat16).to(cuda)
    wgt = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(cuda)

    def run(fold):
        for p in blk.parameters():
            p.grad = None
        xi, ei = x.clone().requires_grad_(True), e.clone().requires_grad_(True)
        seen = []
        ops.EVENT_HOOK = lambda name, start: seen.append(name) if start else None
        try:
            y = blk(xi, num_frames=T, encoder_hidden_states=ei, fold_layout=fold)
            (y.float() * wgt.float()).sum().backward()
        finally:
            ops.EVENT_HOOK = None
        return y.detach(), xi.grad, ei.grad, {n: p.grad.clone() for n, p in blk.named_parameters()}, seen

    y0, dx0, de0, g0, _ = run(False)
    y1, dx1, de1, g1, seen = run(True)
    assert "attn_temporal_bwd" in seen and ("attn_fewkeys_bwd" in seen or L == 1)
    assert torch.equal(y0, y1) and torch.equal(dx0, dx1)
    _close("d encoder_hidden_states", de1, de0)
    for n in g0:
        _close(n, g1[n], g0[n])


def test_frozen_spatial_block_under_no_grad_still_uses_the_fast_path(cuda):
    blk = _init_block(BasicTransformerBlock(320, 5, 1024), 9, cuda)
    blocks.install_blocks(blk)
    x = torch.randn(2, 64, 320, device=cuda).to(torch.bfloat16)
    e = torch.randn(2, 1, 1024, device=cuda).to(torch.bfloat16)
    with torch.no_grad():
        y0 = blk(x, encoder_hidden_states=e)
    y1 = blk(x.clone().requires_grad_(True), encoder_hidden_states=e)     # tape path
    assert not y0.requires_grad and y1.requires_grad
    assert float((y0.float() - y1.detach().float()).abs().max()) <= 2e-2 * max(1.0, float(y0.float().abs().max()))


def test_block_under_gradient_checkpointing_gives_identical_gradients(cuda):
    # the reference wraps the blocks in torch.utils.checkpoint(..., use_reentrant=False) when training
    # (src/models/transformer_temporal.py:329-336); the libtvi Functions are deterministic, so the gradients are bit-identical
    from torch.utils.checkpoint import checkpoint
    blk = _init_block(TemporalBasicTransformerBlock(320, 320, 5, 1024), 12, cuda)
    blocks.install_blocks(blk)
    g = torch.Generator().manual_seed(6)
    T, S = 14, 16
    x = torch.randn(T, S, 320, generator=g).to(torch.bfloat16).to(cuda)
    e = torch.randn(S, 1, 1024, generator=g).to(torch.bfloat16).to(cuda)
    wgt = torch.randn(T, S, 320, generator=g).to(torch.bfloat16).to(cuda)

    def run(use_ckpt):
        for p in blk.parameters():
            p.grad = None
        xi = x.clone().requires_grad_(True)
        y = checkpoint(blk, xi, T, e, use_reentrant=False) if use_ckpt else blk(xi, T, e)
        (y.float() * wgt.float()).sum().backward()
        return xi.grad.clone(), {n: p.grad.clone() for n, p in blk.named_parameters()}

    dx0, g0 = run(False)
    dx1, g1 = run(True)
    assert torch.equal(dx0, dx1)
    for n in g0:
        assert torch.equal(g0[n], g1[n]), n
