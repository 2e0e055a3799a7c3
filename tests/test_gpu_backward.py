"""GPU: backward kernels (training config 5) through the C ABI vs fp32 CPU references of the same
bf16 inputs: projection dgrad / wgrad / bias column sums, attention-core gradients, and whole
processor calls against torch autograd of the oracle."""
import pytest
import torch

from timelapse_video_interpolation_b200 import ops

pytestmark = pytest.mark.gpu


def _close(got, ref, K, what=""):
    err = (got.float().cpu() - ref).abs()
    tol = ref.abs() * 2 ** -8 + 1e-3 * (K ** 0.5) * 2 ** -6 + 1e-6
    assert bool((err <= tol).all()), f"{what}: max err {err.max():.4g}, worst ratio {(err / tol).max():.3g}"


@pytest.mark.parametrize("M,N,K", [
    (128, 64, 64),
    (1000, 320, 320),
    (4032, 960, 320),      # d(qkv) -> dx at L0 width
    (4032, 3840, 1280),    # L2 width
    (77, 72, 40),          # tails everywhere
    (56, 640, 1024),       # context rows: dx of the fused K/V projection
    (33000, 320, 320),
])
def test_dgrad_matches_fp32_reference(cuda, M, N, K):
    g = torch.Generator().manual_seed(M + N + K)
    dy = torch.randn(M, N, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(N, K, generator=g) * N ** -0.5).to(torch.bfloat16).to(cuda)
    dx = ops.gemm_bf16_dgrad(dy, w)
    torch.cuda.synchronize()
    assert dx.shape == (M, K)
    _close(dx, dy.float().cpu() @ w.float().cpu(), N, "dgrad")


@pytest.mark.parametrize("M,N,K", [
    (64, 128, 64),
    (1000, 320, 320),
    (4032, 960, 320),
    (4032, 3840, 1280),
    (77, 72, 40),
    (56, 640, 1024),
    (258048, 960, 320),    # full L0 token count: split reduction over all SMs
])
def test_wgrad_matches_fp32_reference(cuda, M, N, K):
    g = torch.Generator().manual_seed(M + N + K + 1)
    dy = (torch.randn(M, N, generator=g) * M ** -0.5).to(torch.bfloat16).to(cuda)
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda)
    dw = ops.gemm_bf16_wgrad(dy, x)
    dw2 = ops.gemm_bf16_wgrad(dy, x)
    torch.cuda.synchronize()
    assert dw.shape == (N, K)
    assert torch.equal(dw, dw2)  # fixed-order split reduction
    ref = (dy.float().t() @ x.float()).cpu()  # fp32 on the GPU for the big case (torch is only the checker here)
    _close(dw, ref, M, "wgrad")


def test_dgrad_wgrad_on_strided_views(cuda):
    g = torch.Generator().manual_seed(3)
    M, C = 900, 320
    dqkv = torch.randn(M, 3 * C, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(3 * C, C, generator=g) * C ** -0.5).to(torch.bfloat16).to(cuda)
    x = torch.randn(M, C, generator=g).to(torch.bfloat16).to(cuda)
    dk = dqkv[:, C:2 * C]
    dx = ops.gemm_bf16_dgrad(dk, w[C:2 * C])
    dwbuf = torch.full((3 * C, C), 5.0, dtype=torch.bfloat16, device=cuda)
    ops.gemm_bf16_wgrad(dk, x, out=dwbuf[C:2 * C])
    torch.cuda.synchronize()
    _close(dx, dk.float().cpu() @ w[C:2 * C].float().cpu(), C, "dgrad view")
    _close(dwbuf[C:2 * C], dk.float().cpu().t() @ x.float().cpu(), M, "wgrad view")
    assert bool((dwbuf[:C] == 5.0).all()) and bool((dwbuf[2 * C:] == 5.0).all())


@pytest.mark.parametrize("nrows,rpg,C", [(14 * 300, 14, 320), (9216 * 3, 9216, 320), (5000, 5000, 1280), (64, 1, 72),
                                         (258048, 258048, 320)])
def test_colsum_matches_fp32_reference(cuda, nrows, rpg, C):
    g = torch.Generator().manual_seed(nrows + C)
    x = torch.randn(nrows, C, generator=g).to(torch.bfloat16).to(cuda)
    out = ops.colsum_bf16(x, rpg)
    torch.cuda.synchronize()
    ref = x.float().view(nrows // rpg, rpg, C).sum(1).cpu()
    _close(out, ref, rpg, "colsum")


# ------------------------------------------------------------------------------------------------
# attention cores: gradients vs fp32 autograd of the explicit softmax attention
# (reference math: src/models/attention_processor.py:598-615 + :750) on the same bf16-rounded operands
# ------------------------------------------------------------------------------------------------
def _ref_attention_grads(q, k, v, dout, nseq, heads, Tq, Tk, scale, bias=None, groups_rows=None):
    """q [nseq*Tq, C], k/v [nseq*Tk, C] (fp32, requires_grad) -> (o, dq, dk, dv) by autograd."""
    C = q.shape[-1]
    d = C // heads
    q = q.detach().float().requires_grad_(True)
    k = k.detach().float().requires_grad_(True)
    v = v.detach().float().requires_grad_(True)
    qh = q.view(nseq, Tq, heads, d).transpose(1, 2)
    kh = k.view(nseq, Tk, heads, d).transpose(1, 2)
    vh = v.view(nseq, Tk, heads, d).transpose(1, 2)
    s = qh @ kh.transpose(-1, -2) * scale
    if bias is not None:
        s = s + bias
    o = (s.softmax(-1) @ vh).transpose(1, 2).reshape(nseq * Tq, C)
    o.backward(dout.float())
    return o.detach(), q.grad, k.grad, v.grad


def _grad_close(got, ref, what, tol=2e-2):
    ref = ref.float()
    err = float((got.float() - ref).abs().max())
    scale = float(ref.abs().max()) + 1e-12
    print(f"{what}: max abs err {err:.3e}, max |ref| {scale:.3e}, rel {err / scale:.3e}")
    assert err <= tol * scale, f"{what}: {err:.4g} > {tol} * {scale:.4g}"


def _packed_qkv(rows_q, rows_kv, C, g, dev, std=1.0):
    """q, k, v as column slices of packed buffers (strided rows), like the fused projection output."""
    if rows_q == rows_kv:
        buf = (torch.randn(rows_q, 3 * C, generator=g) * std).to(torch.bfloat16).to(dev)
        return buf[:, :C], buf[:, C:2 * C], buf[:, 2 * C:]
    qb = (torch.randn(rows_q, C, generator=g) * std).to(torch.bfloat16).to(dev)
    kvb = (torch.randn(rows_kv, 2 * C, generator=g) * std).to(torch.bfloat16).to(dev)
    return qb, kvb[:, :C], kvb[:, C:]


@pytest.mark.parametrize("nseq,heads,Tq,Tk", [(300, 5, 14, 14), (97, 5, 25, 25), (40, 2, 64, 64), (33, 3, 7, 7),
                                              (21, 2, 16, 16), (19, 1, 14, 9), (11, 2, 33, 50), (4608, 5, 14, 14)])
def test_temporal_bwd_matches_autograd(cuda, nseq, heads, Tq, Tk):
    g = torch.Generator().manual_seed(nseq + Tq)
    C = heads * 64
    q, k, v = _packed_qkv(nseq * Tq, nseq * Tk, C, g, cuda)
    dout = torch.randn(nseq * Tq, C, generator=g).to(torch.bfloat16).to(cuda)
    scale = 0.125
    dqkv = torch.zeros(nseq * max(Tq, Tk), 3 * C, dtype=torch.bfloat16, device=cuda)
    if Tq == Tk:  # gradients written in place into a packed [rows, 3C] buffer
        dq, dk, dv = ops.attn_temporal_bwd(q, k, v, dout, nseq, heads, Tq, Tk, scale, dq=dqkv[:, :C], dk=dqkv[:, C:2 * C],
                                           dv=dqkv[:, 2 * C:])
    else:
        dq, dk, dv = ops.attn_temporal_bwd(q, k, v, dout, nseq, heads, Tq, Tk, scale)
    torch.cuda.synchronize()
    _, rq, rk, rv = _ref_attention_grads(q, k, v, dout, nseq, heads, Tq, Tk, scale)
    _grad_close(dq, rq, "temporal dq")
    _grad_close(dk, rk, "temporal dk")
    _grad_close(dv, rv, "temporal dv")


@pytest.mark.parametrize("nseq,heads,Tq,Tk", [(300, 10, 14, 14), (97, 3, 25, 25), (12, 2, 64, 64), (19, 1, 14, 9), (11, 2, 33, 50)])
def test_temporal_bwd_head_dim_128(cuda, nseq, heads, Tq, Tk):
    # d = 128: the class default num_attention_heads=(5, 10, 10, 20) at C = 1280 (src/models/unet_spatio_temporal_condition.py:94)
    g = torch.Generator().manual_seed(nseq + Tq + 1)
    C = heads * 128
    q, k, v = _packed_qkv(nseq * Tq, nseq * Tk, C, g, cuda)
    dout = torch.randn(nseq * Tq, C, generator=g).to(torch.bfloat16).to(cuda)
    scale = 128 ** -0.5
    dq, dk, dv = ops.attn_temporal_bwd(q, k, v, dout, nseq, heads, Tq, Tk, scale)
    torch.cuda.synchronize()
    _, rq, rk, rv = _ref_attention_grads(q, k, v, dout, nseq, heads, Tq, Tk, scale)
    _grad_close(dq, rq, "temporal dq d=128")
    _grad_close(dk, rk, "temporal dk d=128")
    _grad_close(dv, rv, "temporal dv d=128")


@pytest.mark.parametrize("batch,heads,S,Skv", [(2, 2, 144, 144), (1, 10, 576, 576), (2, 1, 300, 200), (1, 3, 130, 65), (1, 2, 1111, 1111)])
def test_spatial_bwd_head_dim_128(cuda, batch, heads, S, Skv):
    from timelapse_video_interpolation_b200 import _lib
    g = torch.Generator().manual_seed(S + Skv + 7)
    C = heads * 128
    q, k, v = _packed_qkv(batch * S, batch * Skv, C, g, cuda)
    dout = torch.randn(batch * S, C, generator=g).to(torch.bfloat16).to(cuda)
    scale = 128 ** -0.5
    lse = torch.empty(batch * S, heads, dtype=torch.float32, device=cuda)
    o = ops.attn_spatial(q, k, v, batch, heads, S, Skv, scale, lse=lse)
    dq, dk, dv = ops.attn_spatial_bwd(q, k, v, o, dout, lse, batch, heads, S, Skv, scale)      # AUTO -> the mma.sync passes
    torch.cuda.synchronize()
    ro, rq, rk, rv = _ref_attention_grads(q, k, v, dout, batch, heads, S, Skv, scale)
    _grad_close(o, ro, "spatial o d=128")
    _grad_close(dq, rq, "spatial dq d=128")
    _grad_close(dk, rk, "spatial dk d=128")
    _grad_close(dv, rv, "spatial dv d=128")
    with pytest.raises(_lib.TviError):   # the tcgen05 passes are built for 64-wide heads
        ops.attn_spatial_bwd(q, k, v, o, dout, lse, batch, heads, S, Skv, scale, algo=_lib.SPATIAL_TCGEN05)


@pytest.mark.parametrize("algo", ["mma_sync", "tcgen05"])
@pytest.mark.parametrize("batch,heads,S,Skv", [(2, 5, 144, 144), (2, 5, 576, 576), (1, 2, 1000, 1000), (2, 1, 300, 200),
                                               (1, 3, 130, 65), (1, 5, 2304, 2304), (3, 2, 1111, 1111), (1, 1, 128, 256)])
def test_spatial_bwd_matches_autograd(cuda, batch, heads, S, Skv, algo):
    """Both implementations at every shape (AUTO picks tcgen05 from 1024 tokens): the tcgen05 / TMEM / TMA passes must also
    be right on ragged tiles (row / column counts that are not multiples of 128) and when S != Skv."""
    from timelapse_video_interpolation_b200 import _lib
    algo_id = {"mma_sync": _lib.SPATIAL_MMA_SYNC, "tcgen05": _lib.SPATIAL_TCGEN05}[algo]
    g = torch.Generator().manual_seed(S + Skv)
    C = heads * 64
    q, k, v = _packed_qkv(batch * S, batch * Skv, C, g, cuda)
    dout = torch.randn(batch * S, C, generator=g).to(torch.bfloat16).to(cuda)
    scale = 0.125
    lse = torch.empty(batch * S, heads, dtype=torch.float32, device=cuda)
    o = ops.attn_spatial(q, k, v, batch, heads, S, Skv, scale, lse=lse)
    dq, dk, dv = ops.attn_spatial_bwd(q, k, v, o, dout, lse, batch, heads, S, Skv, scale, algo=algo_id)
    dq2, dk2, dv2 = ops.attn_spatial_bwd(q, k, v, o, dout, lse, batch, heads, S, Skv, scale, algo=algo_id)
    torch.cuda.synchronize()
    assert torch.equal(dq, dq2) and torch.equal(dk, dk2) and torch.equal(dv, dv2)  # no atomics: repeatable
    ro, rq, rk, rv = _ref_attention_grads(q, k, v, dout, batch, heads, S, Skv, scale)
    _grad_close(o, ro, "spatial o")
    _grad_close(dq, rq, f"spatial dq ({algo})")
    _grad_close(dk, rk, f"spatial dk ({algo})")
    _grad_close(dv, rv, f"spatial dv ({algo})")


def test_spatial_bwd_tcgen05_in_place_packed_gradients(cuda):
    """dq | dk | dv written as the three column slices of one packed [tokens, 3C] buffer (how autograd._AttentionFn
    calls it), at a shape AUTO sends to the tcgen05 passes."""
    batch, heads, S = 2, 5, 1152
    C = heads * 64
    g = torch.Generator().manual_seed(77)
    q, k, v = _packed_qkv(batch * S, batch * S, C, g, cuda)
    dout = torch.randn(batch * S, C, generator=g).to(torch.bfloat16).to(cuda)
    lse = torch.empty(batch * S, heads, dtype=torch.float32, device=cuda)
    o = ops.attn_spatial(q, k, v, batch, heads, S, S, 0.125, lse=lse)
    dqkv = torch.zeros(batch * S, 3 * C, dtype=torch.bfloat16, device=cuda)
    dq, dk, dv = ops.attn_spatial_bwd(q, k, v, o, dout, lse, batch, heads, S, S, 0.125, dq=dqkv[:, :C], dk=dqkv[:, C:2 * C],
                                      dv=dqkv[:, 2 * C:])
    torch.cuda.synchronize()
    _, rq, rk, rv = _ref_attention_grads(q, k, v, dout, batch, heads, S, S, 0.125)
    _grad_close(dq, rq, "packed dq")
    _grad_close(dk, rk, "packed dk")
    _grad_close(dv, rv, "packed dv")


@pytest.mark.parametrize("ngroups,rpg,heads,L,use_bias", [(50, 14, 5, 2, False), (50, 14, 5, 2, True), (3, 9216, 5, 2, False),
                                                          (4, 1000, 10, 1, False), (6, 144, 20, 4, False), (5, 25, 5, 8, False),
                                                          (2, 2304, 10, 2, False)])
def test_fewkeys_bwd_matches_autograd(cuda, ngroups, rpg, heads, L, use_bias):
    g = torch.Generator().manual_seed(ngroups + rpg + L)
    C = heads * 64
    nrows = ngroups * rpg
    q = torch.randn(nrows, C, generator=g).to(torch.bfloat16).to(cuda)
    kv = torch.randn(ngroups * L, 2 * C, generator=g).to(torch.bfloat16).to(cuda)
    k, v = kv[:, :C], kv[:, C:]
    dout = torch.randn(nrows, C, generator=g).to(torch.bfloat16).to(cuda)
    scale = 0.125
    bias = None
    if use_bias:
        w = torch.linspace(0, 1, rpg, device=cuda)
        bias = torch.stack([1 - w, w], 1).contiguous()
    dq, dk, dv = ops.attn_fewkeys_bwd(q, k, v, dout, nrows, rpg, heads, L, scale, Tq=rpg if use_bias else 1, bias=bias)
    torch.cuda.synchronize()
    _, rq, rk, rv = _ref_attention_grads(q, k, v, dout, ngroups, heads, rpg, L, scale,
                                         bias=None if bias is None else bias.view(1, 1, rpg, L))
    _grad_close(dq, rq, "fewkeys dq")
    _grad_close(dk, rk, "fewkeys dk")
    _grad_close(dv, rv, "fewkeys dv")


# ------------------------------------------------------------------------------------------------
# whole processor calls on the autograd tape vs torch autograd of the oracle (gradients
# within 2e-2 of max|grad| against the fp32 oracle fed the same bf16-rounded inputs and weights)
# ------------------------------------------------------------------------------------------------
from oracle import attention_ref  # noqa: E402
from util import make_attn  # noqa: E402
from timelapse_video_interpolation_b200 import TviAttnProcessor  # noqa: E402


def _processor_grads(attn, x, e, dy):
    for p in attn.parameters():
        p.grad = None
    xg = x.detach().clone().requires_grad_(True)
    eg = e.detach().clone().requires_grad_(True) if e is not None else None
    y = attn(xg, encoder_hidden_states=eg)
    y.backward(dy)
    torch.cuda.synchronize()
    return y.detach(), {"x": xg.grad, "e": eg.grad if eg is not None else None, "to_q": attn.to_q.weight.grad,
                        "to_k": attn.to_k.weight.grad, "to_v": attn.to_v.weight.grad,
                        "to_out_w": attn.to_out[0].weight.grad, "to_out_b": attn.to_out[0].bias.grad}


def _compare_grads(got, ref, what, zero_ok=()):
    for key, r in ref.items():
        gk = got[key]
        if r is None:
            assert gk is None, f"{what}: unexpected gradient for {key}"
            continue
        assert gk is not None, f"{what}: missing gradient for {key}"
        assert gk.shape == r.shape, (what, key, gk.shape, r.shape)
        if float(r.abs().max()) == 0.0 or key in zero_ok:
            assert float(gk.float().abs().max()) <= 1e-6 + 2e-2 * float(r.abs().max()), f"{what}: {key} should vanish"
            continue
        _grad_close(gk.cpu(), r, f"{what} d{key}")


@pytest.mark.parametrize("name,Bp,S,C,heads,cross,L,ramp", [
    ("temporal self T=14", 288, 14, 320, 5, None, None, False),
    ("temporal self T=25", 64, 25, 640, 10, None, None, False),
    ("spatial self S=144", 4, 144, 1280, 20, None, None, False),
    ("spatial self S=576", 2, 576, 320, 5, None, None, False),
    ("spatial self S=1152 (tcgen05 forward)", 2, 1152, 320, 5, None, None, False),
    ("spatial cross L=2", 4, 576, 320, 5, 1024, 2, False),
    ("spatial cross L=1", 4, 576, 320, 5, 1024, 1, False),
    ("temporal cross L=2 + ramp bias", 144, 14, 320, 5, 1024, 2, True),
    ("temporal cross L=1", 144, 14, 640, 10, 1024, 1, False),
])
def test_processor_gradients_match_oracle_autograd(cuda, name, Bp, S, C, heads, cross, L, ramp):
    attn = make_attn(C, cross, heads, seed=7, device=cuda, dtype=torch.bfloat16,
                     processor=TviAttnProcessor(ramp_bias=ramp))
    attn.train()
    g = torch.Generator().manual_seed(S + C)
    x = torch.randn(Bp, S, C, generator=g).to(torch.bfloat16).to(cuda)
    e = torch.randn(Bp, L, cross, generator=g).to(torch.bfloat16).to(cuda) if cross else None
    dy = (torch.randn(Bp, S, C, generator=g) / S ** 0.5).to(torch.bfloat16).to(cuda)
    y, got = _processor_grads(attn, x, e, dy)
    ry, ref = attention_ref.attn_processor_grads_ref(attn, x, e, dy, use_attentionmask=ramp)
    _grad_close(y.cpu(), ry, f"{name} y")
    # one key: the softmax is constant, so x / to_q / to_k receive exactly zero
    _compare_grads(got, ref, name, zero_ok=("x", "to_q", "to_k") if L == 1 else ())


def test_processor_backward_with_frozen_spatial_weights_and_checkpoint(cuda):
    # train_decouples freezes everything but temporal_transformer_block*: spatial modules only pass
    # gradients through to x; blocks run under torch.utils.checkpoint(use_reentrant=False)
    # (src/models/transformer_temporal.py:322-327)
    from torch.utils.checkpoint import checkpoint
    attn = make_attn(320, None, 5, seed=3, device=cuda, dtype=torch.bfloat16, processor=TviAttnProcessor())
    for p in attn.parameters():
        p.requires_grad_(False)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 576, 320, generator=g).to(torch.bfloat16).to(cuda).requires_grad_(True)
    dy = (torch.randn(2, 576, 320, generator=g) / 24).to(torch.bfloat16).to(cuda)
    y = checkpoint(lambda t: attn(t), x, use_reentrant=False)
    y.backward(dy)
    torch.cuda.synchronize()
    assert all(p.grad is None for p in attn.parameters())
    _, ref = attention_ref.attn_processor_grads_ref(attn, x, None, dy)
    _grad_close(x.grad.cpu(), ref["x"], "frozen-weights dx under checkpoint")


def test_processor_backward_fp32_parameters_cast_inputs(cuda):
    # the reference's shipped training configs keep fp32 parameters: gradients come back in fp32
    attn = make_attn(320, None, 5, seed=4, device=cuda, dtype=torch.float32, processor=TviAttnProcessor(cast_inputs=True))
    g = torch.Generator().manual_seed(6)
    x = torch.randn(36, 14, 320, generator=g).to(cuda).requires_grad_(True)
    dy = torch.randn(36, 14, 320, generator=g).to(cuda) / 4
    y = attn(x)
    assert y.dtype == torch.float32
    y.backward(dy)
    torch.cuda.synchronize()
    assert attn.to_q.weight.grad.dtype == torch.float32 and x.grad.dtype == torch.float32
    ref_attn = make_attn(320, None, 5, seed=4)
    _, ref = attention_ref.attn_processor_grads_ref(ref_attn, x, None, dy)
    _grad_close(x.grad.cpu(), ref["x"], "fp32 dx", tol=3e-2)
    _grad_close(attn.to_v.weight.grad.cpu(), ref["to_v"], "fp32 dWv", tol=3e-2)


def test_backward_as_first_cuda_work_of_the_autograd_thread():
    # regression: the autograd engine's worker thread reaches cuTensorMapEncodeTiled (a driver call) with no
    # context bound when a libtvi GEMM is its very first CUDA work -> CUDA_ERROR_INVALID_CONTEXT
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    code = ("import torch, sys; sys.path.insert(0, %r)\n"
            "from timelapse_video_interpolation_b200.modules import Attention\n"
            "a = Attention(320, None, heads=5).to('cuda', torch.bfloat16)\n"
            "x = torch.randn(4, 14, 320, device='cuda').to(torch.bfloat16).requires_grad_(True)\n"
            "a(x).float().sum().backward(); torch.cuda.synchronize()\n"
            "assert bool(torch.isfinite(x.grad.float()).all()); print('ok')\n") % str(root)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]
