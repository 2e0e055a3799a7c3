"""GPU: the CTA-pair (tcgen05 cta_group::2) variant of the forward GEMM, selected with TVI_GEMM_PAIR=1, against the
single-CTA kernel and an fp32 matmul: two CTAs of a cluster share one 256-row tile, each loads half of the B tile."""
import pytest
import torch
import torch.nn.functional as F

from timelapse_video_interpolation_b200 import ops

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K,bias,res,geglu", [
    (256, 64, 64, False, False, False), (512, 320, 320, True, False, False), (1000, 320, 320, True, True, False),
    (777, 960, 320, False, False, False), (4096, 2560, 320, True, False, True), (2048, 320, 1280, True, True, False),
    (300, 128, 192, True, False, False), (40000, 960, 320, False, False, False),
])
def test_pair_gemm_matches_single_cta_kernel(cuda, monkeypatch, M, N, K, bias, res, geglu):
    g = torch.Generator().manual_seed(M + N + K)
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda)
    w = (torch.randn(N, K, generator=g) * K ** -0.5).to(torch.bfloat16).to(cuda)
    b = torch.randn(N, generator=g).to(torch.bfloat16).to(cuda) if bias else None
    r = torch.randn(M, N // 2 if geglu else N, generator=g).to(torch.bfloat16).to(cuda) if res else None
    ref = F.linear(x.float(), w.float(), b.float() if bias else None)
    if geglu:
        inner, H = N // 2, ops.geglu_block(N // 2)
        ref = ref[:, :inner] * F.gelu(ref[:, inner:])
        w = torch.stack([w[:inner].view(inner // H, H, -1), w[inner:].view(inner // H, H, -1)], dim=1).reshape(N, -1).contiguous()
        b = torch.stack([b[:inner].view(inner // H, H), b[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous()
    if res:
        ref = ref + r.float()

    def run():
        return ops.gemm_bf16_tn_ex(x, w, b, residual=r, geglu=geglu) if (geglu or res) else ops.gemm_bf16_tn(x, w, b)

    monkeypatch.setenv("TVI_GEMM_PAIR", "0")
    y_single = run()
    monkeypatch.setenv("TVI_GEMM_PAIR", "1")
    y_pair = run()
    torch.cuda.synchronize()
    # same products, same fp32 accumulation order along K: the two kernels agree bit for bit
    assert torch.equal(y_pair, y_single)
    assert float((y_pair.float() - ref).abs().max()) <= 2.0 ** -7 * max(1.0, float(ref.abs().max()))
