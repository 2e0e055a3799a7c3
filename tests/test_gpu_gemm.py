"""GPU: tcgen05 projection GEMM (through the C ABI) vs an fp32 CPU contraction of the same bf16 inputs."""
import pytest
import torch

from timelapse_video_interpolation_b200 import ops

pytestmark = pytest.mark.gpu


def _ref(x, w, b):
    y = x.float().cpu() @ w.float().cpu().t()
    if b is not None:
        y = y + b.float().cpu()
    return y


def _check(y, ref, K):
    # one bf16 rounding of an fp32-accumulated result: |err| <= 2^-9 |y| (+ fp32 summation noise)
    err = (y.float().cpu() - ref).abs()
    tol = ref.abs() * 2 ** -8 + 1e-3 * (K ** 0.5) * 2 ** -8 + 1e-6
    assert bool((err <= tol).all()), f"max err {err.max():.4g}, worst ratio {(err / tol).max():.3g}"


@pytest.mark.parametrize("M,N,K,bias", [
    (128, 64, 64, False),        # one tile, one k-block
    (128, 160, 320, True),       # the L0 out-projection tile shape
    (1000, 320, 320, True),      # M tail
    (4032, 960, 320, False),     # fused QKV at L0 width, mid-level token count
    (4032, 3840, 1280, False),   # fused QKV at L2 width
    (77, 72, 40, True),          # N and K tails: N % 32 != 0, K % 64 != 0
    (56, 640, 1024, False),      # cross-attention K/V rows (few rows, K = 1024)
    (2, 2560, 1024, False),      # 2-row context
    (33000, 320, 320, True),     # > 148 tiles per N column: persistent loop + accumulator double buffering
    (20000, 1280, 1280, True),
])
def test_gemm_matches_fp32_reference(cuda, M, N, K, bias):
    g = torch.Generator().manual_seed(M + N + K)
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(N, K, generator=g) * K ** -0.5).to(torch.bfloat16).to(cuda)
    b = torch.randn(N, generator=g).to(torch.bfloat16).to(cuda) if bias else None
    y = ops.gemm_bf16_tn(x, w, b)
    torch.cuda.synchronize()
    assert y.shape == (M, N) and y.dtype == torch.bfloat16
    _check(y, _ref(x, w, b), K)


def test_gemm_strided_views_and_untouched_padding(cuda):
    # X is a column slice of a wider buffer, Y a column slice of a packed QKV-like buffer
    g = torch.Generator().manual_seed(0)
    M, N, K = 700, 320, 320
    xbuf = torch.randn(M, 3 * K, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(N, K, generator=g) * K ** -0.5).to(torch.bfloat16).to(cuda)
    ybuf = torch.full((M, 3 * N), 7.0, dtype=torch.bfloat16, device=cuda)
    x = xbuf[:, K:2 * K]
    ops.gemm_bf16_tn(x, w, None, out=ybuf[:, N:2 * N])
    torch.cuda.synchronize()
    _check(ybuf[:, N:2 * N], _ref(x, w, None), K)
    assert bool((ybuf[:, :N] == 7.0).all()) and bool((ybuf[:, 2 * N:] == 7.0).all())


def test_gemm_is_deterministic_and_linear(cuda):
    # full-size property check (L0: 258 048 tokens x 320 -> 960): repeatability and exact scaling by 2
    g = torch.Generator().manual_seed(1)
    M, N, K = 258048, 960, 320
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(N, K, generator=g) * K ** -0.5).to(torch.bfloat16).to(cuda)
    y1 = ops.gemm_bf16_tn(x, w)
    y2 = ops.gemm_bf16_tn(x, w)
    y3 = ops.gemm_bf16_tn(x * 2, w)
    torch.cuda.synchronize()
    assert torch.equal(y1, y2)
    assert torch.equal(y3, y1 * 2)  # power-of-two scaling commutes with every rounding
    rows = torch.randint(0, M, (512,), generator=g)
    _check(y1[rows.to(cuda)], _ref(x[rows.to(cuda)], w, None), K)


def test_gemm_rejects_bad_arguments(cuda):
    from timelapse_video_interpolation_b200 import TviError
    x = torch.zeros(16, 64, dtype=torch.bfloat16, device=cuda)
    with pytest.raises(TviError):
        ops.gemm_bf16_tn(x, torch.zeros(12, 64, dtype=torch.bfloat16, device=cuda))  # N % 8
    with pytest.raises(TviError):
        ops.gemm_bf16_tn(x.float(), torch.zeros(16, 64, dtype=torch.bfloat16, device=cuda))
    y = ops.gemm_bf16_tn(x[:0], torch.zeros(16, 64, dtype=torch.bfloat16, device=cuda))
    assert y.shape == (0, 16)


@pytest.mark.parametrize("M,inner,K,bias", [(4304, 416, 64, True), (300, 32, 320, False), (129, 96, 72, True), (5000, 160, 640, True)])
def test_geglu_epilogue_narrow_blocks(cuda, M, inner, K, bias):
    # regression: GEGLU tiles with fewer output chunks than epilogue warps per quadrant (value/gate block of 32 -> BN = 64)
    # never released their accumulator stage and the kernel hung (found by tools/gemm_soak.py)
    g = torch.Generator().manual_seed(inner)
    H = ops.geglu_block(inner)
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(2 * inner, K, generator=g) * K ** -0.5).to(torch.bfloat16).to(cuda)
    b = torch.randn(2 * inner, generator=g).to(torch.bfloat16).to(cuda) if bias else None
    wi = torch.stack([w[:inner].view(inner // H, H, -1), w[inner:].view(inner // H, H, -1)], dim=1).reshape(2 * inner, -1).contiguous()
    bi = torch.stack([b[:inner].view(inner // H, H), b[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous() if bias else None
    y = ops.gemm_bf16_tn_ex(x, wi, bi, geglu=True)
    torch.cuda.synchronize()
    pre = x.float() @ w.float().T + (b.float() if bias else 0)
    ref = pre[:, :inner] * torch.nn.functional.gelu(pre[:, inner:])
    assert float((y.float() - ref).abs().max()) <= 2e-2 * max(1.0, float(ref.abs().max()))
