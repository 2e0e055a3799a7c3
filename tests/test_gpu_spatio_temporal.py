"""GPU parity of the model-level forward (TransformerSpatioTemporalModel) and its layout kernels against the oracle."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import spatio_temporal_ref
from timelapse_video_interpolation_b200 import _lib, ops, spatio_temporal
from timelapse_video_interpolation_b200.modules import TransformerSpatioTemporalModel

pytestmark = pytest.mark.gpu
from conftest import assert_block_close

TOL = 2e-2


def _bf(t, cuda):
    return t.to(torch.bfloat16).to(cuda)


@pytest.mark.parametrize("N,C,H,W,groups", [(3, 320, 8, 16, 32), (2, 64, 4, 6, 32), (1, 640, 24, 40, 32), (2, 96, 10, 12, 4)])
def test_groupnorm_tokens(cuda, N, C, H, W, groups):
    g = torch.Generator().manual_seed(C + H)
    x = _bf(torch.randn(N, C, H, W, generator=g) * 2 + 0.7, cuda)
    gamma, beta = _bf(1 + 0.3 * torch.randn(C, generator=g), cuda), _bf(0.2 * torch.randn(C, generator=g), cuda)
    y = ops.groupnorm_tokens(x, gamma, beta, groups, 1e-6)
    ref = F.group_norm(x.float().cpu(), groups, gamma.float().cpu(), beta.float().cpu(), 1e-6).permute(0, 2, 3, 1).reshape(N, H * W, C)
    assert y.shape == (N, H * W, C)
    err = float((y.float().cpu() - ref).abs().max())
    assert err <= 2 ** -7 * max(1.0, float(ref.abs().max())), err  # one bf16 rounding of the result
    y2 = ops.groupnorm_tokens(x, None, None, groups, 1e-6)
    ref2 = F.group_norm(x.float().cpu(), groups, None, None, 1e-6).permute(0, 2, 3, 1).reshape(N, H * W, C)
    assert float((y2.float().cpu() - ref2).abs().max()) <= 2 ** -7 * max(1.0, float(ref2.abs().max()))


def test_groupnorm_tokens_large_mean_does_not_cancel(cuda):
    x = _bf(torch.randn(1, 64, 16, 16) * 0.05 + 30.0, cuda)
    y = ops.groupnorm_tokens(x, None, None, 32, 1e-6)
    ref = F.group_norm(x.float().cpu(), 32, None, None, 1e-6).permute(0, 2, 3, 1).reshape(1, 256, 64)
    assert float((y.float().cpu() - ref).abs().max()) < 5e-2


@pytest.mark.parametrize("N,C,H,W", [(3, 320, 8, 16), (2, 64, 4, 6), (1, 72, 10, 12)])
def test_tokens_to_nchw_add(cuda, N, C, H, W):
    g = torch.Generator().manual_seed(C)
    y = _bf(torch.randn(N, H * W, C, generator=g), cuda)
    r = _bf(torch.randn(N, C, H, W, generator=g), cuda)
    got = ops.tokens_to_nchw_add(y, r, H, W)
    want = (y.float().view(N, H, W, C).permute(0, 3, 1, 2) + r.float()).to(torch.bfloat16)
    assert torch.equal(got, want)
    assert torch.equal(ops.tokens_to_nchw_add(y, None, H, W), y.view(N, H, W, C).permute(0, 3, 1, 2).contiguous())


def test_blend_rows(cuda):
    g = torch.Generator().manual_seed(4)
    BF, S, C = 6, 50, 320
    xs, xt = _bf(torch.randn(BF, S, C, generator=g), cuda), _bf(torch.randn(BF, S, C, generator=g), cuda)
    xt[2] = float("nan")                                  # an image-only frame never reads the temporal branch
    alpha = torch.tensor([0.3, 0.5, 1.0, 0.0, 0.7310, 0.25], device=cuda)
    got = ops.blend_rows(xs, xt, alpha, S)
    a = alpha.view(BF, 1, 1)
    want = torch.addcmul(a * xs.float(), (1 - a), xt.float())
    ok = torch.ones(BF, dtype=torch.bool)
    ok[2] = False
    assert float((got.float() - want)[ok].abs().max()) <= 2 ** -7 * 4
    assert torch.equal(got[2], xs[2]) and torch.equal(got[3], xt[3])
    with pytest.raises(_lib.TviError):
        ops.blend_rows(xs, xt, alpha[:5].contiguous(), S)


@pytest.mark.parametrize("N,C,H,W,groups", [(3, 320, 8, 16, 32), (2, 64, 4, 6, 32), (1, 640, 24, 40, 32), (2, 96, 10, 12, 4)])
def test_groupnorm_tokens_backward(cuda, N, C, H, W, groups):
    g = torch.Generator().manual_seed(C + W)
    x = _bf(torch.randn(N, C, H, W, generator=g) * 1.5 + 0.4, cuda)
    gamma, beta = _bf(1 + 0.3 * torch.randn(C, generator=g), cuda), _bf(0.2 * torch.randn(C, generator=g), cuda)
    dn = _bf(torch.randn(N, H * W, C, generator=g), cuda)
    dx, dg, db = ops.groupnorm_tokens_bwd(x, dn, gamma, groups, 1e-6)
    xr = x.float().cpu().requires_grad_(True)
    gr, br = gamma.float().cpu().requires_grad_(True), beta.float().cpu().requires_grad_(True)
    y = F.group_norm(xr, groups, gr, br, 1e-6).permute(0, 2, 3, 1).reshape(N, H * W, C)
    (y * dn.float().cpu()).sum().backward()
    for name, got, ref in (("dx", dx, xr.grad), ("dgamma", dg, gr.grad), ("dbeta", db, br.grad)):
        err, scale = float((got.float().cpu() - ref).abs().max()), float(ref.abs().max())
        print(f"  {name}: max abs err {err:.3e} / max |ref| {scale:.3e}")
        assert err <= 2e-2 * scale, name
    # only what is asked for is produced; identical values
    dx2, dg2, db2 = ops.groupnorm_tokens_bwd(x, dn, gamma, groups, 1e-6, need_affine=False)
    assert dg2 is None and db2 is None and torch.equal(dx2, dx)
    dx3, dg3, _ = ops.groupnorm_tokens_bwd(x, dn, gamma, groups, 1e-6, need_dx=False)
    assert dx3 is None and torch.equal(dg3, dg)
    assert torch.equal(ops.nchw_to_tokens(x), x.permute(0, 2, 3, 1).reshape(N, H * W, C))


def test_blend_rows_backward(cuda):
    g = torch.Generator().manual_seed(8)
    BF, S, C = 5, 70, 320
    xs, xt, dy = (_bf(torch.randn(BF * S, C, generator=g), cuda) for _ in range(3))
    alpha = torch.tensor([0.3, 1.0, 0.0, 0.7310, 0.5], device=cuda)
    dxs, dxt, da = ops.blend_rows_bwd(dy, xs, xt, alpha, S, True)
    a = alpha.view(BF, 1, 1).cpu()
    d3 = dy.float().cpu().view(BF, S, C)
    assert float((dxs.float().cpu().view(BF, S, C) - a * d3).abs().max()) <= 2 ** -7 * 4
    assert float((dxt.float().cpu().view(BF, S, C) - (1 - a) * d3).abs().max()) <= 2 ** -7 * 4
    ref = (d3 * (xs.float().cpu() - xt.float().cpu()).view(BF, S, C)).sum(dim=(1, 2))
    assert float((da.cpu() - ref).abs().max()) <= 1e-3 * float(ref.abs().max())
    dxs2, dxt2, da2 = ops.blend_rows_bwd(dy, None, None, alpha, S, False)
    assert da2 is None and torch.equal(dxs2, dxs) and torch.equal(dxt2, dxt)
    assert torch.equal(dxs[S:2 * S], dy[S:2 * S]) and not dxt[S:2 * S].any()      # image-only frame: all of dy goes to the spatial branch


def _model(C, heads, cross, seed, cuda, head_dim=64):
    m = TransformerSpatioTemporalModel(heads, head_dim, in_channels=C, num_layers=1, cross_attention_dim=cross)
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("mix_factor"):
                p.fill_(0.4)
            elif "norm" in name:
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g) if name.endswith("weight") else 0.05 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.05))
    return m.to(cuda, torch.bfloat16).eval()


@pytest.mark.parametrize("B,T,H,W,C,heads", [(2, 14, 8, 16, 320, 5), (1, 25, 8, 8, 640, 10)])
def test_transformer_spatio_temporal_model_matches_oracle(cuda, B, T, H, W, C, heads):
    m = _model(C, heads, 1024, 11, cuda)
    assert spatio_temporal.install_spatio_temporal(m) == 1
    g = torch.Generator().manual_seed(T)
    x = _bf(torch.randn(B * T, C, H, W, generator=g), cuda)
    e = _bf(torch.randn(B * T, 1, 1024, generator=g), cuda)
    ind = torch.zeros(B, T, device=cuda)
    ind[0, 2] = 1.0
    launches0 = ops.LAUNCHES
    with torch.no_grad():
        out = m(x, encoder_hidden_states=e, image_only_indicator=ind)
        y = out.sample
        y_tuple = m(x, encoder_hidden_states=(e, e), image_only_indicator=ind, return_dict=False)[0]
    torch.cuda.synchronize()
    assert ops.LAUNCHES - launches0 >= 2 * 28
    assert torch.equal(y, y_tuple) and y.shape == x.shape and y.dtype == torch.bfloat16
    ref = spatio_temporal_ref.transformer_spatio_temporal_ref(m, x, e, ind, alpha_dtype=torch.bfloat16)
    assert_block_close(f"spatio-temporal model C={C} T={T}", y, ref, stores=9)  # proj_in, 3 + 4 block stores, blend
    assert spatio_temporal.uninstall_spatio_temporal(m) == 1
    with pytest.raises(RuntimeError):
        m(x, encoder_hidden_states=e, image_only_indicator=ind)


def test_model_with_128_wide_heads_matches_oracle(cuda):
    # level 2 of a UNet built from the class defaults: C = 1280, 10 heads of 128 (src/models/unet_spatio_temporal_condition.py:94)
    B, T, H, W, C, heads = 1, 14, 6, 8, 1280, 10
    m = _model(C, heads, 1024, 13, cuda, head_dim=128)
    assert m.transformer_blocks[0].attn1.scale == 128 ** -0.5
    spatio_temporal.install_spatio_temporal(m)
    g = torch.Generator().manual_seed(2)
    x = _bf(torch.randn(B * T, C, H, W, generator=g), cuda)
    e = _bf(torch.randn(B * T, 1, 1024, generator=g), cuda)
    ind = torch.zeros(B, T, device=cuda)
    with torch.no_grad():
        y = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
    ref = spatio_temporal_ref.transformer_spatio_temporal_ref(m, x, e, ind, alpha_dtype=torch.bfloat16)
    assert_block_close("spatio-temporal model d=128", y, ref, stores=9)


def test_model_with_128_wide_heads_trains(cuda):
    # forward + backward through a d = 128 module (mma.sync spatial / temporal kernels in both directions): finite gradients that
    # match fp32 autograd of the same math
    B, T, H, W, C, heads = 1, 4, 12, 12, 256, 2          # S = 144: the mma.sync spatial kernels; T = 4: the temporal ones
    m = _model(C, heads, 1024, 19, cuda, head_dim=128).train()
    spatio_temporal.install_spatio_temporal(m)
    g = torch.Generator().manual_seed(4)
    x = _bf(torch.randn(B * T, C, H, W, generator=g), cuda).requires_grad_(True)
    e = _bf(torch.randn(B * T, 1, 1024, generator=g), cuda)
    wgt = _bf(torch.randn(B * T, C, H, W, generator=g), cuda)
    ind = torch.zeros(B, T, device=cuda)
    y = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
    (y.float() * wgt.float()).sum().backward()
    P = {n: p.detach().float().cpu().requires_grad_(True) for n, p in m.named_parameters()}
    xr = x.detach().float().cpu().requires_grad_(True)
    yr = _model_ref(P, xr, e.float().cpu(), ind.cpu(), heads, T)
    (yr * wgt.float().cpu()).sum().backward()
    ps = dict(m.named_parameters())
    for name, got, ref in [("dx", x.grad, xr.grad)] + [(n, ps[n].grad, P[n].grad) for n in (
            "transformer_blocks.0.attn1.to_q.weight", "temporal_transformer_blocks.0.attn1.to_v.weight", "proj_in.weight")]:
        err, scale = float((got.float().cpu() - ref).abs().max()), max(float(ref.abs().max()), 1e-6)
        assert err <= 2e-2 * scale, (name, err, scale)


def test_model_forward_matches_reference_golden(cuda, golden_dir):
    # the reference's own module output (fp32, CPU) vs the libtvi forward in bf16 on the same weights
    z = np.load(golden_dir / "spatio_temporal.npz")
    m = TransformerSpatioTemporalModel(1, 64, in_channels=64, num_layers=1, cross_attention_dim=48)
    m.load_state_dict({k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd.")}, strict=True)
    m = m.to(cuda, torch.bfloat16).eval()
    spatio_temporal.install_spatio_temporal(m)
    with torch.no_grad():
        y = m(_bf(torch.from_numpy(z["x"]), cuda), encoder_hidden_states=_bf(torch.from_numpy(z["e"]), cuda),
              image_only_indicator=torch.from_numpy(z["ind"]).to(cuda), return_dict=False)[0]
    ref = torch.from_numpy(z["y"])
    # bf16 inputs AND bf16 weights against the reference's fp32 weights: twice the bound
    assert_block_close("vs reference golden", y, ref, factor=2.0, stores=9)


def test_model_forward_refuses_cpu_tensors(cuda):
    m = _model(320, 5, 1024, 3, cuda)
    spatio_temporal.install_spatio_temporal(m)
    x = torch.randn(2, 320, 8, 8, device=cuda).to(torch.bfloat16)
    e = torch.randn(2, 1, 1024, device=cuda).to(torch.bfloat16)
    ind = torch.zeros(1, 2, device=cuda)
    with torch.no_grad(), pytest.raises(_lib.TviError):
        m(x.detach().cpu(), encoder_hidden_states=e, image_only_indicator=ind)


def _model_ref(P, x, e, ind, heads, T):
    """fp32 autograd restatement of the forward (:253-360) over a dict of leaf parameters (gradients flow to them)."""
    from test_gpu_block_backward import _block_ref
    BF, C, H, W = x.shape
    B = BF // T
    sub = lambda pre: {k[len(pre):]: v for k, v in P.items() if k.startswith(pre)}
    tc = e.reshape(B, T, -1, e.shape[-1])[:, 0]
    tc = tc[None, :].broadcast_to(H * W, B, tc.shape[-2], tc.shape[-1]).reshape(H * W * B, tc.shape[-2], tc.shape[-1])
    h = F.group_norm(x, 32, P["norm.weight"], P["norm.bias"], 1e-6).permute(0, 2, 3, 1).reshape(BF, H * W, C)
    h = F.linear(h, P["proj_in.weight"], P["proj_in.bias"])
    t_emb = spatio_temporal_ref.timesteps_ref(torch.arange(T).repeat(B, 1).reshape(-1), C)
    emb = F.linear(F.silu(F.linear(t_emb, P["time_pos_embed.linear_1.weight"], P["time_pos_embed.linear_1.bias"])),
                   P["time_pos_embed.linear_2.weight"], P["time_pos_embed.linear_2.bias"])[:, None, :]
    alpha = torch.where(ind.bool().reshape(-1), torch.ones(1), torch.sigmoid(P["time_mixer.mix_factor"])).reshape(-1, 1, 1)
    h = _block_ref(sub("transformer_blocks.0."), h, e, heads)
    hm = _block_ref(sub("temporal_transformer_blocks.0."), h + emb, tc, heads, True, T)
    h = alpha * h + (1.0 - alpha) * hm
    h = F.linear(h, P["proj_out.weight"], P["proj_out.bias"])
    return h.reshape(BF, H, W, C).permute(0, 3, 1, 2) + x


MODEL_CHECK = ["norm.weight", "norm.bias", "proj_in.weight", "proj_in.bias", "proj_out.weight", "proj_out.bias",
               "time_pos_embed.linear_1.weight", "time_pos_embed.linear_2.bias", "time_mixer.mix_factor",
               "transformer_blocks.0.attn1.to_q.weight", "transformer_blocks.0.ff.net.2.weight",
               "temporal_transformer_blocks.0.ff_in.net.0.proj.weight", "temporal_transformer_blocks.0.attn2.to_v.weight",
               "temporal_transformer_blocks.0.norm3.weight"]


@pytest.mark.parametrize("ckpt", [False, True], ids=["plain", "checkpointed"])
def test_model_forward_trains(cuda, ckpt):
    # src/train.py back-propagates through this module; with gradient_checkpointing the spatial block runs under
    # torch.utils.checkpoint(..., use_reentrant=False) (src/models/transformer_temporal.py:321-327)
    B, T, H, W, C, heads = 1, 7, 6, 8, 320, 5
    m = _model(C, heads, 1024, 17, cuda).train()
    m.gradient_checkpointing = ckpt
    spatio_temporal.install_spatio_temporal(m)
    g = torch.Generator().manual_seed(3)
    x = _bf(torch.randn(B * T, C, H, W, generator=g), cuda).requires_grad_(True)
    e = _bf(torch.randn(B * T, 1, 1024, generator=g), cuda)
    wgt = _bf(torch.randn(B * T, C, H, W, generator=g), cuda)
    ind = torch.zeros(B, T, device=cuda)
    ind[0, 1] = 1.0
    launches0 = ops.LAUNCHES
    seen = []
    ops.EVENT_HOOK = lambda name, start: seen.append(name) if start else None
    try:
        y = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
        assert y.requires_grad and y.shape == x.shape
        (y.float() * wgt.float()).sum().backward()
    finally:
        ops.EVENT_HOOK = None
    torch.cuda.synchronize()
    for name in ("groupnorm_tokens", "groupnorm_tokens_bwd", "tokens_to_nchw_add", "nchw_to_tokens", "blend_rows", "blend_rows_bwd",
                 "bcast_rows_add", "colsum_bf16"):
        assert name in seen, name                   # the glue runs on libtvi kernels in both directions
    assert ops.LAUNCHES - launches0 >= 100          # the GEMMs, LayerNorms, attention cores and their backward kernels
    P = {n: p.detach().float().cpu().requires_grad_(True) for n, p in m.named_parameters()}
    xr = x.detach().float().cpu().requires_grad_(True)
    yr = _model_ref(P, xr, e.float().cpu(), ind.cpu(), heads, T)
    (yr * wgt.float().cpu()).sum().backward()

    def close(name, got, ref):
        err, scale = float((got.float().cpu() - ref).abs().max()), max(float(ref.abs().max()), 1e-6)
        print(f"  {name}: max abs err {err:.3e} / max |ref| {scale:.3e}")
        assert err <= 2e-2 * scale, name

    close("forward", y.detach(), yr.detach())
    close("dx", x.grad, xr.grad)
    ps = dict(m.named_parameters())
    for n in MODEL_CHECK:
        assert ps[n].grad is not None, n
        close(n, ps[n].grad, P[n].grad)
    # the no-grad fast path (fused glue kernels) and the tape path agree
    with torch.no_grad():
        y0 = m(x.detach(), encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
    assert_block_close("fused vs tape", y0, y.detach().float().cpu(), stores=9)


def test_graphed_forward_replays_bit_exact(cuda):
    from timelapse_video_interpolation_b200.graphs import GraphedForward
    m = _model(320, 5, 1024, 21, cuda)
    spatio_temporal.install_spatio_temporal(m)
    g = torch.Generator().manual_seed(1)
    B, T, H, W = 1, 14, 8, 8
    mk = lambda: (_bf(torch.randn(B * T, 320, H, W, generator=g), cuda), _bf(torch.randn(B * T, 1, 1024, generator=g), cuda))
    x0, e0 = mk()
    ind = torch.zeros(B, T, device=cuda)
    fwd = GraphedForward(m, x0, encoder_hidden_states=e0, image_only_indicator=ind, return_dict=False)
    for _ in range(2):  # new inputs: copied into the static buffers, one graph launch
        x, e = mk()
        launches0 = ops.LAUNCHES
        y = fwd(x, encoder_hidden_states=e, image_only_indicator=ind)[0].clone()
        assert ops.LAUNCHES == launches0  # no Python-side launches on replay
        with torch.no_grad():
            want = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
        assert torch.equal(y, want)
    with pytest.raises(_lib.TviError):
        fwd(x[:7], encoder_hidden_states=e, image_only_indicator=ind)


def test_alpha_cache_follows_the_indicator(cuda):
    m = _model(320, 5, 1024, 5, cuda)
    spatio_temporal.install_spatio_temporal(m)
    g = torch.Generator().manual_seed(2)
    x = _bf(torch.randn(4, 320, 8, 8, generator=g), cuda)
    e = _bf(torch.randn(4, 1, 1024, generator=g), cuda)
    ind = torch.zeros(1, 4, device=cuda)
    with torch.no_grad():
        y0 = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
        y0b = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]   # cached alpha
        ind[0, 1] = 1.0                                                                          # in-place edit bumps _version
        y1 = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
    assert torch.equal(y0, y0b)
    assert torch.equal(y0[[0, 2, 3]], y1[[0, 2, 3]]) and not torch.equal(y0[1], y1[1])
    ref = spatio_temporal_ref.transformer_spatio_temporal_ref(m, x, e, ind, alpha_dtype=torch.bfloat16)
    assert_block_close("graphed forward", y1, ref, stores=9)
