"""A/B of the LayerNorm fold at the L0 shape: module forward with TVI_LN_FOLD=1 / 0, and the raw kernels."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from timelapse_video_interpolation_b200 import lnfold, ops, spatio_temporal
from timelapse_video_interpolation_b200.modules import TransformerSpatioTemporalModel

dev = torch.device("cuda:0")


def timed(fn, n=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


M, C = 258048, 320
x = (torch.randn(M, C, device=dev) + 0.5).bfloat16()
norm = torch.nn.LayerNorm(C).to(dev, torch.bfloat16)
print(f"layernorm_bf16  {timed(lambda: ops.layernorm_bf16(x, norm.weight, norm.bias, 1e-5)):.4f} ms")
print(f"layernorm_stats {timed(lambda: ops.layernorm_stats(x, 1e-5)):.4f} ms")
stats = ops.layernorm_stats(x, 1e-5)
for N, geglu in [(960, False), (2560, True)]:
    lin = torch.nn.Linear(C, N).to(dev, torch.bfloat16)
    w, b = lin.weight.detach(), lin.bias.detach()
    wg, cs, cb = lnfold.folded(lin, "t", w, b, norm)
    print(f"N={N} geglu={geglu}: gemm_ex {timed(lambda: ops.gemm_bf16_tn_ex(x, w, b, geglu=geglu)):.4f} ms   "
          f"gemm_ln {timed(lambda: ops.gemm_bf16_tn_ln(x, wg, cb, cs, stats, geglu=geglu)):.4f} ms")

m = TransformerSpatioTemporalModel(5, 64, in_channels=320, num_layers=1, cross_attention_dim=1024).to(dev, torch.bfloat16).eval()
spatio_temporal.install_spatio_temporal(m)
B, T, H, W = 2, 14, 72, 128
xx = torch.randn(B * T, 320, H, W, device=dev).bfloat16()
e = torch.randn(B * T, 1, 1024, device=dev).bfloat16()
ind = torch.zeros(B, T, device=dev)
for flag in ("1", "0", "1"):
    os.environ["TVI_LN_FOLD"] = flag
    with torch.no_grad():
        l0 = ops.LAUNCHES
        m(xx, encoder_hidden_states=e, image_only_indicator=ind)
        n = ops.LAUNCHES - l0
        ms = timed(lambda: m(xx, encoder_hidden_states=e, image_only_indicator=ind))
    print(f"L0 module forward TVI_LN_FOLD={flag}: {ms:.3f} ms, {n} launches")
