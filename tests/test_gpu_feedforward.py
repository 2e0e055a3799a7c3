"""GPU: FeedForward (GEGLU) / LayerNorm / fused-epilogue GEMM through the C ABI vs the CPU oracle on the same
bf16-rounded inputs and weights (tolerance of the attention path: 2e-2 max-abs)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import timelapse_video_interpolation_b200 as tvi
from oracle import feedforward_ref
from timelapse_video_interpolation_b200 import _lib, ops
from timelapse_video_interpolation_b200.modules import FeedForward

pytestmark = pytest.mark.gpu
from conftest import assert_block_close

TOL = 2e-2


def _ff(dim, dim_out, seed, device):
    g = torch.Generator().manual_seed(seed)
    ff = FeedForward(dim, dim_out)
    with torch.no_grad():
        for p in ff.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.1))
    return ff.to(device=device, dtype=torch.bfloat16).eval()


@pytest.mark.parametrize("rows,dim,dim_out", [(300, 320, None), (4032, 320, None), (1000, 640, None), (777, 1280, None),
                                              (130, 320, 640), (64, 96, None)])
def test_feed_forward_matches_oracle(cuda, rows, dim, dim_out):
    ff = _ff(dim, dim_out, rows + dim, cuda)
    g = torch.Generator().manual_seed(rows)
    x = torch.randn(2, rows // 2, dim, generator=g).to(torch.bfloat16).to(cuda)
    with torch.no_grad():
        y = tvi.feed_forward(ff, x)
    torch.cuda.synchronize()
    ref = feedforward_ref.feed_forward_ref(ff, x)
    assert y.shape == ref.shape and y.dtype == torch.bfloat16
    assert_block_close(f"feed_forward rows={rows} dim={dim}", y, ref)


def test_feed_forward_with_fused_residual_and_install(cuda):
    ff = _ff(320, None, 5, cuda)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(3, 200, 320, generator=g).to(torch.bfloat16).to(cuda)
    res = torch.randn(3, 200, 320, generator=g).to(torch.bfloat16).to(cuda)
    with torch.no_grad():
        y = tvi.feed_forward(ff, x, residual=res)
        assert tvi.install_feed_forward(ff) == 1
        y2 = ff(x)           # the re-bound forward
        assert tvi.uninstall_feed_forward(ff) == 1
        y3 = ff(x)           # stock eager forward again (torch bf16)
    torch.cuda.synchronize()
    ref = feedforward_ref.feed_forward_ref(ff, x, residual=res)
    assert_block_close("feed_forward + fused residual", y, ref)
    ref2 = feedforward_ref.feed_forward_ref(ff, x)
    assert_block_close("feed_forward via install", y2, ref2)
    assert_block_close("stock torch bf16 forward (comparison)", y3, ref2, factor=2.0)


def test_feed_forward_rejects_bad_inputs(cuda):
    ff = _ff(320, None, 6, cuda)
    x = torch.randn(2, 16, 320, device=cuda).to(torch.bfloat16)
    with torch.no_grad(), pytest.raises(tvi.TviError):
        tvi.feed_forward(ff, x.float())
    with torch.no_grad(), pytest.raises(tvi.TviError):
        tvi.feed_forward(ff, x.cpu())


@pytest.mark.parametrize("rows,C", [(1000, 320), (4032, 640), (333, 1280), (50, 2048), (7, 8)])
def test_layer_norm_matches_oracle(cuda, rows, C):
    g = torch.Generator().manual_seed(rows + C)
    norm = torch.nn.LayerNorm(C)
    with torch.no_grad():
        norm.weight.copy_(1 + 0.1 * torch.randn(C, generator=g))
        norm.bias.copy_(0.1 * torch.randn(C, generator=g))
    norm = norm.to(cuda, torch.bfloat16)
    x = (torch.randn(rows, C, generator=g) * 3 + 5).to(torch.bfloat16).to(cuda)  # large mean: two-pass variance
    y = tvi.layer_norm(norm, x)
    torch.cuda.synchronize()
    ref = feedforward_ref.layer_norm_ref(norm, x)
    err = float((y.detach().float().cpu() - ref.detach()).abs().max())
    print(f"layer_norm rows={rows} C={C}: max abs err {err:.3e}")
    assert err <= 2 ** -7 * max(1.0, float(ref.abs().max()))  # one bf16 rounding of the result


@pytest.mark.parametrize("rows,C", [(1003, 320), (3, 320), (1, 320), (4033, 640), (1, 640), (5, 1280)])
def test_layer_norm_row_tails_and_strided_rows(cuda, rows, C):
    """The row-group kernels of the SVD widths (32 / G rows per warp): row counts that leave a partly filled warp, and rows
    that are column slices of wider buffers (ldx, ldy > C); the padding columns of the output stay untouched."""
    g = torch.Generator(device="cuda").manual_seed(rows * 7 + C)
    xb = (torch.randn(rows, C + 64, generator=g, device=cuda) * 2 + 1).to(torch.bfloat16)
    yb = torch.full((rows, C + 32), 7.0, device=cuda, dtype=torch.bfloat16)
    gamma = (1 + 0.1 * torch.randn(C, generator=g, device=cuda)).to(torch.bfloat16)
    beta = (0.1 * torch.randn(C, generator=g, device=cuda)).to(torch.bfloat16)
    x = xb[:, 32:32 + C]
    ops.layernorm_bf16(x, gamma, beta, 1e-5, out=yb[:, :C])
    ref = torch.nn.functional.layer_norm(x.float(), (C,), gamma.float(), beta.float(), 1e-5)
    assert float((yb[:, :C].float() - ref).abs().max()) <= 2 ** -7 * max(1.0, float(ref.abs().max()))
    assert bool((yb[:, C:] == 7.0).all())
    stats = ops.layernorm_stats(x, 1e-5)
    xf = x.float()
    assert float((stats[:, 0] - xf.mean(-1)).abs().max()) <= 1e-4 * max(1.0, float(xf.abs().max()))
    rstd = (xf.var(-1, unbiased=False) + 1e-5).rsqrt()
    assert float(((stats[:, 1] - rstd) / rstd).abs().max()) <= 1e-4


def test_gemm_residual_epilogue(cuda):
    g = torch.Generator().manual_seed(11)
    M, N, K = 1000, 320, 1280
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(N, K, generator=g) * K ** -0.5).to(torch.bfloat16).to(cuda)
    b = torch.randn(N, generator=g).to(torch.bfloat16).to(cuda)
    r = torch.randn(M, N, generator=g).to(torch.bfloat16).to(cuda)
    y = ops.gemm_bf16_tn_ex(x, w, b, residual=r)
    torch.cuda.synchronize()
    ref = x.float().cpu() @ w.float().cpu().t() + b.float().cpu() + r.float().cpu()
    assert float((y.float().cpu() - ref).abs().max()) <= 2 ** -7 * float(ref.abs().max()) + 1e-2


# ------------------------------------------------------------------------------------------------
# block-level forwards vs the oracle's restatement of the reference blocks
# ------------------------------------------------------------------------------------------------
from oracle import blocks_ref  # noqa: E402
from timelapse_video_interpolation_b200 import blocks  # noqa: E402
from timelapse_video_interpolation_b200.modules import BasicTransformerBlock, TemporalBasicTransformerBlock  # noqa: E402


def _init(module, seed, device):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in module.named_parameters():
            if "norm" in name:
                p.copy_((1.0 if name.endswith("weight") else 0.0) + 0.1 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.05))
    return module.to(device=device, dtype=torch.bfloat16).eval()


@pytest.mark.parametrize("BF,S,C,heads,L", [(4, 576, 320, 5, 2), (2, 1152, 320, 5, 1), (4, 144, 640, 10, 2)])
def test_basic_transformer_block_matches_oracle(cuda, BF, S, C, heads, L):
    blk = _init(BasicTransformerBlock(C, heads, 1024), 3, cuda)
    assert blocks.install_blocks(blk) == 1
    g = torch.Generator().manual_seed(S)
    x = torch.randn(BF, S, C, generator=g).to(torch.bfloat16).to(cuda)
    e = torch.randn(BF, L, 1024, generator=g).to(torch.bfloat16).to(cuda)
    with torch.no_grad():
        y = blk(x, encoder_hidden_states=e)
    torch.cuda.synchronize()
    ref = blocks_ref.basic_transformer_block_ref(blk, x, e)
    assert_block_close(f"basic block S={S} C={C}", y, ref, stores=3)
    assert blocks.uninstall_blocks(blk) == 1


@pytest.mark.parametrize("B,T,S,C,heads,L", [(2, 14, 144, 320, 5, 1), (1, 25, 64, 640, 10, 2)])
def test_temporal_basic_transformer_block_matches_oracle(cuda, B, T, S, C, heads, L):
    blk = _init(TemporalBasicTransformerBlock(C, C, heads, 1024), 4, cuda)
    assert blocks.install_blocks(blk) == 1
    g = torch.Generator().manual_seed(T)
    x = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(cuda)
    e = torch.randn(B * S, L, 1024, generator=g).to(torch.bfloat16).to(cuda)
    with torch.no_grad():
        y = blk(x, num_frames=T, encoder_hidden_states=e)
    torch.cuda.synchronize()
    ref = blocks_ref.temporal_basic_transformer_block_ref(blk, x, T, e)
    assert y.shape == x.shape
    assert_block_close(f"temporal block T={T} C={C}", y, ref, stores=4)


def test_temporal_block_layout_fold_is_bit_exact(cuda):
    # N2: the in-place [B*T, S, C] path (interleaved temporal cores) == the reference's permute -> block -> permute
    B, T, S, C = 2, 14, 96, 320
    blk = _init(TemporalBasicTransformerBlock(C, C, 5, 1024), 6, cuda)
    blocks.install_blocks(blk)
    g = torch.Generator().manual_seed(9)
    x = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(cuda)
    e = torch.randn(B * S, 1, 1024, generator=g).to(torch.bfloat16).to(cuda)
    with torch.no_grad():
        launches0 = ops.LAUNCHES
        y_fold = blk(x, num_frames=T, encoder_hidden_states=e)
        y_perm = blk(x, num_frames=T, encoder_hidden_states=e, fold_layout=False)
    assert ops.LAUNCHES > launches0
    assert torch.equal(y_fold, y_perm)


@pytest.mark.parametrize("L,ramp", [(2, False), (2, True), (1, True), (5, False)])
def test_temporal_block_layout_fold_with_several_keys_bias_and_statistics(cuda, L, ramp):
    # the patched temporal attn2 of src/train_decouples.py:377-495 (two key frames, ramp bias :437-457, logit statistics
    # :460-471) runs in the spatial token order too: `tvi_attn_fewkeys_fwd_interleaved`
    from timelapse_video_interpolation_b200 import TviAttnProcessor
    B, T, S, C = 2, 14, 40, 320
    blk = _init(TemporalBasicTransformerBlock(C, C, 5, 1024), 16, cuda)
    got = []
    if ramp:
        blk.attn2.processor = TviAttnProcessor(ramp_bias=(L == 2), stats_sink=lambda st, a: got.append(st.clone()))
    blocks.install_blocks(blk)
    g = torch.Generator().manual_seed(L)
    x = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(cuda)
    e = torch.randn(B * S, L, 1024, generator=g).to(torch.bfloat16).to(cuda)
    with torch.no_grad():
        y_fold = blk(x, num_frames=T, encoder_hidden_states=e)
        y_perm = blk(x, num_frames=T, encoder_hidden_states=e, fold_layout=False)
    assert torch.equal(y_fold, y_perm)
    if ramp:
        assert len(got) == 2 and got[0].shape == (T, L)
        assert float((got[0] - got[1]).abs().max()) <= 1e-4 * max(1.0, float(got[1].abs().max()))  # atomics: order differs


def test_fewkeys_interleaved_groups(cuda):
    g = torch.Generator().manual_seed(12)
    B, T, S, heads, L = 3, 7, 11, 5, 2
    C = heads * 64
    q = torch.randn(B, T, S, C, generator=g).to(torch.bfloat16).to(cuda)            # spatial token order
    k = torch.randn(B * S * L, C, generator=g).to(torch.bfloat16).to(cuda)
    v = torch.randn(B * S * L, C, generator=g).to(torch.bfloat16).to(cuda)
    o_int = ops.attn_fewkeys(q.view(-1, C), k, v, B * T * S, T, heads, L, 0.125, group_inner=S)
    q_perm = q.permute(0, 2, 1, 3).contiguous()                                   # [B, S, T, C]
    o_ref = ops.attn_fewkeys(q_perm.view(-1, C), k, v, B * T * S, T, heads, L, 0.125)
    assert torch.equal(o_int.view(B, T, S, C), o_ref.view(B, S, T, C).permute(0, 2, 1, 3))
    with pytest.raises(_lib.TviError):
        ops.attn_fewkeys(q.view(-1, C), k, v, B * T * S, T, heads, L, 0.125, group_inner=S + 1)


@pytest.mark.parametrize("M,C,N,geglu,res", [(300, 320, 960, False, False), (1000, 640, 5120, True, False), (257, 1280, 1280, False, True),
                                             (4096, 320, 2560, True, False), (130, 320, 320, False, True)])
def test_layernorm_folded_into_the_gemm(cuda, M, C, N, geglu, res):
    # SURVEY 8.6 N1 "LN prologue": LN(x) W^T + b from the raw rows = rstd * (x Wg^T - mean * colsum) + (W beta + b)
    from timelapse_video_interpolation_b200 import lnfold
    g = torch.Generator().manual_seed(M + N)
    norm = torch.nn.LayerNorm(C)
    lin = torch.nn.Linear(C, N)
    with torch.no_grad():
        norm.weight.copy_(1 + 0.3 * torch.randn(C, generator=g))
        norm.bias.copy_(0.2 * torch.randn(C, generator=g))
        lin.weight.copy_(torch.randn(N, C, generator=g) * C ** -0.5)
        lin.bias.copy_(0.1 * torch.randn(N, generator=g))
    norm, lin = norm.to(cuda, torch.bfloat16), lin.to(cuda, torch.bfloat16)
    # rows with a mean several times their spread: the mean term must cancel, not round away
    x = (torch.randn(M, C, generator=g) * 0.7 + 3.0 * torch.randn(M, 1, generator=g)).to(torch.bfloat16).to(cuda)
    r = torch.randn(M, N // 2 if geglu else N, generator=g).to(torch.bfloat16).to(cuda) if res else None
    w, b = lin.weight.detach(), lin.bias.detach()
    if geglu:  # value | gate rows interleaved in blocks of H, as feedforward._geglu_weights does
        inner, H = N // 2, ops.geglu_block(N // 2)
        w = torch.stack([w[:inner].view(inner // H, H, -1), w[inner:].view(inner // H, H, -1)], dim=1).reshape(N, -1).contiguous()
        b = torch.stack([b[:inner].view(inner // H, H), b[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous()
    stats = ops.layernorm_stats(x, norm.eps)
    xf = x.float().cpu()
    assert float((stats[:, 0].cpu() - xf.mean(1)).abs().max()) < 1e-5 * 10
    assert float((stats[:, 1].cpu() * torch.sqrt(xf.var(1, unbiased=False) + norm.eps) - 1).abs().max()) < 1e-4
    wg, colsum, cb = lnfold.folded(lin, "t", w, b, norm)
    y = ops.gemm_bf16_tn_ln(x, wg, cb, colsum, stats, residual=r, geglu=geglu)
    n_ref = F.layer_norm(xf, (C,), norm.weight.detach().float().cpu(), norm.bias.detach().float().cpu(), norm.eps)
    ref = F.linear(n_ref, lin.weight.detach().float().cpu(), lin.bias.detach().float().cpu())
    if geglu:
        ref = ref[:, :N // 2] * F.gelu(ref[:, N // 2:])
    if res:
        ref = ref + r.float().cpu()
    # against the two-kernel path (bf16 normalised rows in between): the same bound holds for both
    y2 = ops.gemm_bf16_tn_ex(ops.layernorm_bf16(x, norm.weight, norm.bias, norm.eps), w, b, residual=r, geglu=geglu)
    e_fold, e_two = float((y.float().cpu() - ref).abs().max()), float((y2.float().cpu() - ref).abs().max())
    print(f"  folded {e_fold:.3e}  two-kernel {e_two:.3e}  max |ref| {float(ref.abs().max()):.2f}")
    assert_block_close("LN folded GEMM", y, ref, stores=1)
    assert e_fold <= 2.0 * e_two + 1e-3
    assert lnfold.folded(lin, "t", w, b, norm)[0] is wg          # cached
    with torch.no_grad():
        norm.weight.mul_(2.0)                                        # in-place update bumps _version -> rebuilt
    assert lnfold.folded(lin, "t", w, b, norm)[0] is not wg


def test_blocks_with_the_layernorm_fold_enabled(cuda, monkeypatch):
    # TVI_LN_FOLD=1: norm1 -> QKV, norm3 / norm_in -> GEGLU projections go through tvi_gemm_bf16_tn_ln; same bound as the default path
    monkeypatch.setenv("TVI_LN_FOLD", "1")
    B, T, S, C = 1, 14, 96, 320
    g = torch.Generator().manual_seed(31)
    sb = _init(BasicTransformerBlock(C, 5, 1024), 21, cuda)
    tb = _init(TemporalBasicTransformerBlock(C, C, 5, 1024), 22, cuda)
    blocks.install_blocks(sb)
    blocks.install_blocks(tb)
    x = (torch.randn(B * T, S, C, generator=g) + 0.5).to(torch.bfloat16).to(cuda)
    es = torch.randn(B * T, 2, 1024, generator=g).to(torch.bfloat16).to(cuda)
    et = torch.randn(B * S, 1, 1024, generator=g).to(torch.bfloat16).to(cuda)
    seen = []
    ops.EVENT_HOOK = lambda name, start: seen.append(name) if start else None
    try:
        with torch.no_grad():
            ys = sb(x, encoder_hidden_states=es)
            yt = tb(x, num_frames=T, encoder_hidden_states=et)
    finally:
        ops.EVENT_HOOK = None
    assert seen.count("gemm_bf16_tn_ln") == 3 + 3 and seen.count("layernorm_stats") == 6 and "layernorm_bf16" not in seen
    assert_block_close("spatial block, LN folded", ys, blocks_ref.basic_transformer_block_ref(sb, x, es), stores=3)
    assert_block_close("temporal block, LN folded", yt, blocks_ref.temporal_basic_transformer_block_ref(tb, x, T, et), stores=4)
    monkeypatch.setenv("TVI_LN_FOLD", "0")
    with torch.no_grad():
        ys0 = sb(x, encoder_hidden_states=es)
    assert_block_close("fold vs default", ys, ys0.float().cpu(), stores=3)


def test_bcast_rows_add_group_inner(cuda):
    B, T, S, C = 2, 5, 7, 64
    g = torch.Generator().manual_seed(2)
    v = torch.randn(B * S, C, generator=g).to(torch.bfloat16).to(cuda)
    r = torch.randn(B * T * S, C, generator=g).to(torch.bfloat16).to(cuda)
    got = ops.bcast_rows_add(v, r, T, group_inner=S)
    want = (v.float().view(B, 1, S, C) + r.float().view(B, T, S, C)).to(torch.bfloat16).view(-1, C)
    assert torch.equal(got, want)
    got = ops.bcast_rows_add(v, None, T, group_inner=S)
    assert torch.equal(got, v.view(B, 1, S, C).expand(B, T, S, C).reshape(-1, C))
    got = ops.bcast_rows_add(v, None, T)  # contiguous groups
    assert torch.equal(got, v.view(B * S, 1, C).expand(B * S, T, C).reshape(-1, C))


def test_empty_and_ragged_inputs(cuda):
    # zero rows are a no-op (no launch), ragged row counts (not a multiple of any tile) are covered by the TMA tails
    ff = _ff(320, None, 8, cuda)
    norm = torch.nn.LayerNorm(320).to(cuda, torch.bfloat16)
    empty = torch.empty(0, 7, 320, dtype=torch.bfloat16, device=cuda)
    with torch.no_grad():
        assert tvi.feed_forward(ff, empty).shape == (0, 7, 320)
        assert tvi.layer_norm(norm, empty).shape == (0, 7, 320)
        assert ops.blend_rows(empty, empty, torch.empty(0, device=cuda), 7).shape == (0, 7, 320)
        for rows in (1, 127, 129, 4099):
            x = torch.randn(rows, 320, device=cuda).to(torch.bfloat16)
            y = tvi.feed_forward(ff, x, residual=x)
            ref = feedforward_ref.feed_forward_ref(ff, x, residual=x)
            assert_block_close(f"feed_forward rows={rows}", y, ref)


@pytest.mark.gpu
@pytest.mark.parametrize("M,K,inner,bias", [(300, 320, 1280, True), (1000, 640, 2560, True), (257, 320, 96, True), (129, 64, 64, False),
                                            (4096, 1280, 5120, True), (31, 320, 32, True)])
def test_gemm_geglu_bwd_matches_fp32_reference(cuda, M, K, inner, bias):
    """ops.gemm_geglu_bwd (the first FeedForward projection re-run with the GEGLU-backward epilogue) against the fp32
    arithmetic autograd derives from `hidden, gate = proj(x).chunk(2, -1); hidden * F.gelu(gate)` (reference
    src/models/attention.py:526-535 via diffusers GEGLU), and against the two-step path it replaces."""
    import torch.nn.functional as F

    g = torch.Generator().manual_seed(M + inner)
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(2 * inner, K, generator=g) / K ** 0.5).to(torch.bfloat16).to(cuda)
    b = (torch.randn(2 * inner, generator=g) * 0.5).to(torch.bfloat16).to(cuda) if bias else None
    dh = torch.randn(M, inner, generator=g).to(torch.bfloat16).to(cuda)
    H = ops.geglu_block(inner)
    wi = torch.stack([w[:inner].view(inner // H, H, K), w[inner:].view(inner // H, H, K)], dim=1).reshape(2 * inner, K).contiguous()
    bi = torch.stack([b[:inner].view(inner // H, H), b[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous() if bias else None
    h, dpre = ops.gemm_geglu_bwd(x, wi, bi, dh)
    torch.cuda.synchronize()

    pre = x.float() @ w.float().t() + (b.float() if bias else 0.0)
    pre = pre.detach().requires_grad_(True)
    val, gate = pre.chunk(2, dim=-1)
    href = val * F.gelu(gate)
    href.backward(dh.float())
    tol_h = 2e-2 * href.abs().max().item()
    tol_d = 2e-2 * pre.grad.abs().max().item()
    assert (h.float() - href.detach()).abs().max().item() <= tol_h
    assert (dpre.float() - pre.grad).abs().max().item() <= tol_d
    # bf16-level agreement in the mean as well (a wrong column order or a missing bias would pass a loose max test on few rows)
    assert (h.float() - href.detach()).abs().mean().item() <= 4e-3 * href.abs().mean().item() + 1e-6
    assert (dpre.float() - pre.grad).abs().mean().item() <= 4e-3 * pre.grad.abs().mean().item() + 1e-6
    # the forward's fused GEGLU GEMM computes the same h from the same accumulator
    h_fwd = ops.gemm_bf16_tn_ex(x, wi, bi, geglu=True)
    assert torch.equal(h_fwd, h)
    # without the h output (a frozen second projection): the same d(pre), nothing else written
    h3, dpre3 = ops.gemm_geglu_bwd(x, wi, bi, dh, need_h=False)
    assert h3 is None and torch.equal(dpre3, dpre)
    # the path it replaces (bf16-rounded pre-activation, erff): same numbers to bf16 rounding
    h2, dpre2 = ops.geglu_bwd(ops.gemm_bf16_tn(x, w, b), dh)
    assert (dpre.float() - dpre2.float()).abs().max().item() <= tol_d
    assert (h.float() - h2.float()).abs().max().item() <= tol_h
