"""Per-shape timing of the training-only kernels at the L0 token count (M = 258 048): wgrad / dgrad GEMMs, GEGLU backward,
LayerNorm backward, bias-gradient column sums.  Prints ms, TFLOP/s (GEMMs) and GB/s of algorithmic traffic."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from timelapse_video_interpolation_b200 import ops

dev = torch.device("cuda:0")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 258048


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


def rnd(*shape):
    return (torch.randn(*shape, device=dev) * 0.5).bfloat16()


for N, K in [(2560, 320), (320, 1280), (960, 320), (320, 320), (640, 1024)]:
    dy, x, w = rnd(M if N != 640 else 9216, N), rnd(M if N != 640 else 9216, K), rnd(N, K)
    rows = dy.shape[0]
    fl = 2.0 * rows * N * K
    ms = timed(lambda: ops.gemm_bf16_wgrad(dy, x))
    print(f"wgrad dW[{N},{K}] M={rows}: {ms:.4f} ms  {fl / ms / 1e9:.0f} TFLOP/s  {2.0 * rows * (N + K) / ms / 1e6:.0f} GB/s")
    ms = timed(lambda: ops.gemm_bf16_dgrad(dy, w))
    print(f"dgrad dX[M,{K}] from dY[M,{N}]: {ms:.4f} ms  {fl / ms / 1e9:.0f} TFLOP/s  {2.0 * rows * (N + K) / ms / 1e6:.0f} GB/s")
    ms = timed(lambda: ops.colsum_bf16(dy))
    print(f"colsum [{rows},{N}]: {ms:.4f} ms  {2.0 * rows * N / ms / 1e6:.0f} GB/s")
    del dy, x, w

pre, dh = rnd(M, 2560), rnd(M, 1280)
ms = timed(lambda: ops.geglu_bwd(pre, dh, need_h=True))
print(f"geglu_bwd (+h): {ms:.4f} ms  {2.0 * M * (2560 + 1280 + 1280 + 2560) / ms / 1e6:.0f} GB/s")
ms = timed(lambda: ops.geglu_bwd(pre, dh, need_h=False))
print(f"geglu_bwd     : {ms:.4f} ms  {2.0 * M * (2560 + 1280 + 2560) / ms / 1e6:.0f} GB/s")
x, wi, bi = rnd(M, 320), rnd(2560, 320) * 0.1, rnd(2560)
ms0 = timed(lambda: ops.gemm_bf16_tn(x, wi, bi))
ms = timed(lambda: ops.gemm_geglu_bwd(x, wi, bi, dh))
print(f"gemm_geglu_bwd (recompute + h + dpre in one kernel): {ms:.4f} ms  {2.0 * M * (320 + 1280 + 1280 + 2560) / ms / 1e6:.0f} GB/s"
      f"   [replaces plain GEMM {ms0:.4f} ms + geglu_bwd (+h)]")
del pre, dh, x, wi, bi
x, dy, g = rnd(M, 320), rnd(M, 320), rnd(320)
ms = timed(lambda: ops.layernorm_bwd(x, dy, g, 1e-5))
print(f"layernorm_bwd (dx, dgamma, dbeta): {ms:.4f} ms  {2.0 * M * 320 * 3 / ms / 1e6:.0f} GB/s")
dres = rnd(M, 320)
ms = timed(lambda: ops.layernorm_bwd(x, dy, g, 1e-5, dres=dres))
print(f"layernorm_bwd (+ dres: dx, dgamma, dbeta): {ms:.4f} ms  {2.0 * M * 320 * 4 / ms / 1e6:.0f} GB/s")
ms = timed(lambda: ops.layernorm_bwd(x, dy, g, 1e-5, need_affine=False))
print(f"layernorm_bwd (dx only): {ms:.4f} ms  {2.0 * M * 320 * 3 / ms / 1e6:.0f} GB/s")
