"""Times the fused GEGLU-backward GEMM (ops.gemm_geglu_bwd) at the L0 FeedForward shape (M = 258 048, K = 320,
inner = 1 280); the one kernel a `ncu -k regex:gemm_bf16_kernel` capture of this script sees."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from timelapse_video_interpolation_b200 import ops

dev = torch.device("cuda:0")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 258048
K, inner = 320, 1280
x = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
wi = (torch.randn(2 * inner, K, device=dev) * 0.05).bfloat16()
bi = (torch.randn(2 * inner, device=dev) * 0.5).bfloat16()
dh = (torch.randn(M, inner, device=dev) * 0.5).bfloat16()
for _ in range(3):
    ops.gemm_geglu_bwd(x, wi, bi, dh)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.gemm_geglu_bwd(x, wi, bi, dh)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"gemm_geglu_bwd M={M} K={K} inner={inner}: {ms:.4f} ms  {2.0 * M * (K + 4 * inner) / ms / 1e6:.0f} GB/s algorithmic")
