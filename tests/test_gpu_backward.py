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
