"""GPU: the three attention cores (through the C ABI) vs the oracle's explicit fp32 softmax(QK^T)V."""
import math

import pytest
import torch

from oracle.attention_ref import batch_to_head_dim, get_attention_scores, head_to_batch_dim
from timelapse_video_interpolation_b200 import _lib, ops

pytestmark = pytest.mark.gpu
TOL = 2e-2  # bf16 attention bar from BASELINE.json (max-abs); typical error is ~4e-3, printed below


def _core_ref(q, k, v, heads, scale, bias=None):
    """q [B,Sq,h*d], k/v [B,Sk,h*d] fp32 CPU -> (out [B,Sq,h*d], logits [B*h,Sq,Sk])."""
    p, logits = get_attention_scores(head_to_batch_dim(q, heads), head_to_batch_dim(k, heads), scale, bias)
    return batch_to_head_dim(torch.bmm(p, head_to_batch_dim(v, heads)), heads), logits


def _rand(shape, g, cuda, mul=1.0):
    t = (torch.randn(*shape, generator=g) * mul).to(torch.bfloat16)
    return t.to(cuda), t.float()


@pytest.mark.parametrize("nseq,heads,Tq,Tk", [
    (64, 5, 14, 14), (37, 5, 25, 25), (300, 1, 1, 1), (19, 2, 16, 16), (11, 3, 17, 17),
    (9, 2, 40, 40), (5, 2, 64, 64), (7, 2, 14, 3), (2000, 10, 14, 14),
])
def test_temporal_core(cuda, nseq, heads, Tq, Tk):
    g = torch.Generator().manual_seed(nseq * 100 + Tq)
    C = heads * 64
    q, qf = _rand((nseq, Tq, C), g, cuda)
    k, kf = _rand((nseq, Tk, C), g, cuda)
    v, vf = _rand((nseq, Tk, C), g, cuda)
    scale = 0.125
    lse = torch.empty(nseq * Tq, heads, dtype=torch.float32, device=cuda)
    o = ops.attn_temporal(q.view(-1, C), k.view(-1, C), v.view(-1, C), nseq, heads, Tq, Tk, scale, lse=lse)
    torch.cuda.synchronize()
    ref, logits = _core_ref(qf, kf, vf, heads, scale)
    err = (o.float().cpu().view(nseq, Tq, C) - ref).abs().max().item()
    print(f"temporal core nseq={nseq} T={Tq}x{Tk}: max abs err {err:.3e}")
    assert err < TOL
    ref_lse = torch.logsumexp(logits, dim=-1).view(nseq, heads, Tq).permute(0, 2, 1).reshape(nseq * Tq, heads)
    assert (lse.cpu() - ref_lse).abs().max().item() < 1e-3


@pytest.mark.parametrize("nseq,heads,Tq,Tk", [(64, 10, 14, 14), (37, 3, 25, 25), (300, 1, 1, 1), (9, 2, 40, 40), (7, 2, 14, 3),
                                              (2000, 10, 14, 14)])
def test_temporal_core_head_dim_128(cuda, nseq, heads, Tq, Tk):
    # the class default num_attention_heads=(5, 10, 10, 20) gives d = 1280 / 10 = 128 at level 2
    # (src/models/unet_spatio_temporal_condition.py:94)
    g = torch.Generator().manual_seed(nseq + Tq)
    C = heads * 128
    q, qf = _rand((nseq, Tq, C), g, cuda)
    k, kf = _rand((nseq, Tk, C), g, cuda)
    v, vf = _rand((nseq, Tk, C), g, cuda)
    scale = 128 ** -0.5
    lse = torch.empty(nseq * Tq, heads, dtype=torch.float32, device=cuda)
    o = ops.attn_temporal(q.view(-1, C), k.view(-1, C), v.view(-1, C), nseq, heads, Tq, Tk, scale, lse=lse)
    torch.cuda.synchronize()
    ref, logits = _core_ref(qf, kf, vf, heads, scale)
    assert (o.float().cpu().view(nseq, Tq, C) - ref).abs().max().item() < TOL
    ref_lse = torch.logsumexp(logits, dim=-1).view(nseq, heads, Tq).permute(0, 2, 1).reshape(nseq * Tq, heads)
    assert (lse.cpu() - ref_lse).abs().max().item() < 1e-3
    # interleaved sequences (N2) at d = 128: bit-identical to the contiguous kernel on permuted copies
    if Tq == Tk and nseq % 4 == 0:
        B, S = 4, nseq // 4
        qkv, _ = _rand((B, Tq, S, 3 * C), g, cuda)
        flat = qkv.view(-1, 3 * C)
        o_i = ops.attn_temporal(flat[:, :C], flat[:, C:2 * C], flat[:, 2 * C:], B * S, heads, Tq, Tq, scale, seq_inner=S)
        perm = qkv.permute(0, 2, 1, 3).contiguous().view(-1, 3 * C)
        o_p = ops.attn_temporal(perm[:, :C], perm[:, C:2 * C], perm[:, 2 * C:], B * S, heads, Tq, Tq, scale)
        assert torch.equal(o_i.view(B, Tq, S, C), o_p.view(B, S, Tq, C).permute(0, 2, 1, 3))


@pytest.mark.parametrize("B,heads,S,Skv", [(2, 2, 80, 80), (1, 10, 576, 576), (1, 2, 129, 63), (2, 3, 1000, 1000), (1, 1, 2304, 2304)])
def test_spatial_core_head_dim_128(cuda, B, heads, S, Skv):
    g = torch.Generator().manual_seed(S + Skv + 1)
    C = heads * 128
    q, qf = _rand((B, S, C), g, cuda)
    k, kf = _rand((B, Skv, C), g, cuda)
    v, vf = _rand((B, Skv, C), g, cuda)
    scale = 128 ** -0.5
    lse = torch.empty(B * S, heads, dtype=torch.float32, device=cuda)
    o = ops.attn_spatial(q.view(-1, C), k.view(-1, C), v.view(-1, C), B, heads, S, Skv, scale, lse=lse)   # AUTO -> mma.sync
    torch.cuda.synchronize()
    ref, logits = _core_ref(qf, kf, vf, heads, scale)
    err = (o.float().cpu().view(B, S, C) - ref).abs().max().item()
    print(f"spatial core d=128 B={B} h={heads} S={S}x{Skv}: max abs err {err:.3e}")
    assert err < TOL
    ref_lse = torch.logsumexp(logits, dim=-1).view(B, heads, S).permute(0, 2, 1).reshape(B * S, heads)
    assert (lse.cpu() - ref_lse).abs().max().item() < 1e-3
    with pytest.raises(_lib.TviError):      # the tcgen05 kernel is built for 64-wide heads: asking for it explicitly is an error
        ops.attn_spatial(q.view(-1, C), k.view(-1, C), v.view(-1, C), B, heads, S, Skv, scale, algo=_lib.SPATIAL_TCGEN05)


def test_temporal_core_reads_packed_qkv_in_place(cuda):
    g = torch.Generator().manual_seed(3)
    nseq, heads, T = 50, 5, 14
    C = heads * 64
    qkv, qkvf = _rand((nseq * T, 3 * C), g, cuda)
    o = ops.attn_temporal(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], nseq, heads, T, T, 0.125)
    torch.cuda.synchronize()
    f = qkvf.view(nseq, T, 3 * C)
    ref, _ = _core_ref(f[..., :C], f[..., C:2 * C], f[..., 2 * C:], heads, 0.125)
    assert (o.float().cpu().view(nseq, T, C) - ref).abs().max().item() < TOL


@pytest.mark.parametrize("B,S,heads,T", [(2, 37, 5, 14), (1, 64, 2, 25), (3, 5, 1, 40), (2, 1, 2, 14)])
def test_temporal_core_interleaved_sequences(cuda, B, S, heads, T):
    # N2: sequences read straight out of the [B, T, S] token order == the contiguous kernel on permuted copies (bit-exact)
    g = torch.Generator().manual_seed(S * 10 + T)
    C = heads * 64
    qkv, _ = _rand((B, T, S, 3 * C), g, cuda)
    flat = qkv.view(-1, 3 * C)
    lse = torch.full((B * T * S, heads), float("nan"), dtype=torch.float32, device=cuda)
    o = ops.attn_temporal(flat[:, :C], flat[:, C:2 * C], flat[:, 2 * C:], B * S, heads, T, T, 0.125, lse=lse, seq_inner=S)
    perm = qkv.permute(0, 2, 1, 3).contiguous().view(-1, 3 * C)          # [B*S*T, 3C]
    lse_p = torch.empty(B * S * T, heads, dtype=torch.float32, device=cuda)
    o_p = ops.attn_temporal(perm[:, :C], perm[:, C:2 * C], perm[:, 2 * C:], B * S, heads, T, T, 0.125, lse=lse_p)
    torch.cuda.synchronize()
    assert torch.equal(o.view(B, T, S, C), o_p.view(B, S, T, C).permute(0, 2, 1, 3))
    assert torch.equal(lse.view(B, T, S, heads), lse_p.view(B, S, T, heads).permute(0, 2, 1, 3))
    with pytest.raises(_lib.TviError):
        ops.attn_temporal(flat[:, :C], flat[:, C:2 * C], flat[:, 2 * C:], B * S, heads, T, T, 0.125, seq_inner=B * S + 1)


@pytest.mark.parametrize("B,heads,S,Skv", [
    (2, 2, 80, 80), (1, 5, 144, 144), (3, 1, 576, 576), (1, 2, 100, 100), (2, 2, 65, 130),
    (1, 1, 2304, 2304), (1, 2, 129, 63), (2, 5, 1000, 1000),
])
@pytest.mark.parametrize("algo", [_lib.SPATIAL_MMA_SYNC, _lib.SPATIAL_TCGEN05], ids=["mma_sync", "tcgen05"])
def test_spatial_core(cuda, B, heads, S, Skv, algo):
    g = torch.Generator().manual_seed(S + Skv)
    C = heads * 64
    q, qf = _rand((B, S, C), g, cuda)
    k, kf = _rand((B, Skv, C), g, cuda)
    v, vf = _rand((B, Skv, C), g, cuda)
    lse = torch.empty(B * S, heads, dtype=torch.float32, device=cuda)
    o = ops.attn_spatial(q.view(-1, C), k.view(-1, C), v.view(-1, C), B, heads, S, Skv, 0.125, lse=lse, algo=algo)
    torch.cuda.synchronize()
    ref, logits = _core_ref(qf, kf, vf, heads, 0.125)
    err = (o.float().cpu().view(B, S, C) - ref).abs().max().item()
    print(f"spatial core algo={algo} B={B} h={heads} S={S}x{Skv}: max abs err {err:.3e}")
    assert err < TOL
    ref_lse = torch.logsumexp(logits, dim=-1).view(B, heads, S).permute(0, 2, 1).reshape(B * S, heads)
    assert (lse.cpu() - ref_lse).abs().max().item() < 1e-3


@pytest.mark.parametrize("algo", [_lib.SPATIAL_MMA_SYNC, _lib.SPATIAL_TCGEN05], ids=["mma_sync", "tcgen05"])
def test_spatial_core_large_logits_stay_finite(cuda, algo):
    # online-softmax rescaling: a late key block dominates an early one by a huge margin (the tcgen05
    # kernel's lazy reference maximum must rescale its TMEM accumulator here)
    g = torch.Generator().manual_seed(9)
    B, heads, S = 1, 1, 640
    q, qf = _rand((B, S, 64), g, cuda, 6.0)
    k, kf = _rand((B, S, 64), g, cuda, 6.0)
    v, vf = _rand((B, S, 64), g, cuda)
    o = ops.attn_spatial(q.view(-1, 64), k.view(-1, 64), v.view(-1, 64), B, heads, S, S, 0.125, algo=algo)
    torch.cuda.synchronize()
    ref, _ = _core_ref(qf, kf, vf, heads, 0.125)
    assert torch.isfinite(o).all() and (o.float().cpu().view(B, S, 64) - ref).abs().max().item() < 3e-2


def test_spatial_core_tcgen05_is_deterministic_under_load(cuda):
    # regression: the epilogue once read O before the last two PV products had landed (mbarrier phase
    # aliasing), which showed up as rare, run-to-run different tiles.  Many CTAs, repeated launches.
    g = torch.Generator().manual_seed(7)
    B, heads, S = 2, 5, 4096
    C = heads * 64
    q, qf = _rand((B, S, C), g, cuda)
    k, kf = _rand((B, S, C), g, cuda)
    v, vf = _rand((B, S, C), g, cuda)
    ref = torch.nn.functional.scaled_dot_product_attention(
        *(t.to(cuda).view(B, S, heads, 64).transpose(1, 2) for t in (qf, kf, vf)), scale=0.125).transpose(1, 2).reshape(B, S, C)
    first = None
    for _ in range(8):
        o = ops.attn_spatial(q.view(-1, C), k.view(-1, C), v.view(-1, C), B, heads, S, S, 0.125, algo=_lib.SPATIAL_TCGEN05)
        torch.cuda.synchronize()
        assert (o.float().view(B, S, C) - ref).abs().max().item() < 5e-3
        if first is None:
            first = o.clone()
        assert torch.equal(o, first)


def test_spatial_core_reads_packed_qkv_in_place_tcgen05(cuda):
    # the processor hands the kernel three column slices of one [tokens, 3C] buffer (row stride 3C)
    g = torch.Generator().manual_seed(21)
    B, heads, S = 2, 5, 1152
    C = heads * 64
    qkv, qkvf = _rand((B * S, 3 * C), g, cuda)
    o = ops.attn_spatial(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], B, heads, S, S, 0.125, algo=_lib.SPATIAL_TCGEN05)
    o2 = ops.attn_spatial(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], B, heads, S, S, 0.125, algo=_lib.SPATIAL_MMA_SYNC)
    torch.cuda.synchronize()
    f = qkvf.view(B, S, 3 * C)
    ref, _ = _core_ref(f[..., :C], f[..., C:2 * C], f[..., 2 * C:], heads, 0.125)
    assert (o.float().cpu().view(B, S, C) - ref).abs().max().item() < TOL
    assert (o.float() - o2.float()).abs().max().item() < TOL


@pytest.mark.parametrize("groups,S,heads,L,d", [
    (3, 14, 2, 1, 64), (3, 14, 5, 2, 64), (2, 40, 5, 2, 64), (4, 9, 2, 3, 64), (2, 33, 1, 8, 64),
    (3, 14, 2, 2, 128), (2, 1000, 10, 2, 64), (5, 14, 20, 2, 64),
])
def test_fewkeys_core(cuda, groups, S, heads, L, d):
    g = torch.Generator().manual_seed(groups * 10 + L)
    C = heads * d
    q, qf = _rand((groups, S, C), g, cuda)
    k, kf = _rand((groups, L, C), g, cuda)
    v, vf = _rand((groups, L, C), g, cuda)
    scale = d ** -0.5
    o = ops.attn_fewkeys(q.view(-1, C), k.view(-1, C), v.view(-1, C), groups * S, S, heads, L, scale)
    torch.cuda.synchronize()
    p, _ = get_attention_scores(head_to_batch_dim(qf, heads), head_to_batch_dim(kf, heads), scale)
    ref = batch_to_head_dim(torch.bmm(p, head_to_batch_dim(vf, heads)), heads)
    assert (o.float().cpu().view(groups, S, C) - ref).abs().max().item() < TOL
    if L == 1:  # exact broadcast, and Q may be omitted
        o2 = ops.attn_fewkeys(None, k.view(-1, C), v.view(-1, C), groups * S, S, heads, 1, scale)
        torch.cuda.synchronize()
        assert torch.equal(o2.view(groups, S, C), v.expand(groups, S, C))


def test_fewkeys_ramp_bias_and_statistics(cuda):
    # src/train_decouples.py:437-471: (T,2) linspace bias and per-frame mean logits
    g = torch.Generator().manual_seed(4)
    groups, T, heads, L = 37, 14, 5, 2
    C = heads * 64
    q, qf = _rand((groups, T, C), g, cuda)
    k, kf = _rand((groups, L, C), g, cuda)
    v, vf = _rand((groups, L, C), g, cuda)
    w = torch.linspace(0, 1, T)
    bias = torch.stack([1 - w, w], dim=1).contiguous()
    stats = torch.empty(T, L, dtype=torch.float32, device=cuda)
    o = ops.attn_fewkeys(q.view(-1, C), k.view(-1, C), v.view(-1, C), groups * T, T, heads, L, 0.125, Tq=T,
                         bias=bias.to(cuda), stats=stats)
    torch.cuda.synchronize()
    p, logits = get_attention_scores(head_to_batch_dim(qf, heads), head_to_batch_dim(kf, heads), 0.125, bias[None])
    ref = batch_to_head_dim(torch.bmm(p, head_to_batch_dim(vf, heads)), heads)
    assert (o.float().cpu().view(groups, T, C) - ref).abs().max().item() < TOL
    assert (stats.cpu() - logits.mean(dim=0)).abs().max().item() < 1e-4
