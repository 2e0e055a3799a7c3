"""QKV-projection launch of the tcgen05 GEMM at the L0 shape for ncu / timing (M=258048, N=960, K=320)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from timelapse_video_interpolation_b200 import ops
dev = torch.device("cuda:0")
M, N, K = 258048, int(sys.argv[1]) if len(sys.argv) > 1 else 960, 320
x = torch.randn(M, K, device=dev).bfloat16()
w = (torch.randn(N, K, device=dev) * K ** -0.5).bfloat16()
b = torch.randn(N, device=dev).bfloat16() if len(sys.argv) > 2 else None
y = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
for _ in range(3):
    ops.gemm_bf16_tn(x, w, b, out=y)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    ops.gemm_bf16_tn(x, w, b, out=y)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
byt = 2.0 * (M * K + M * N + N * K)
print(f"N={N} bias={b is not None}: {ms:.4f} ms, {2.0 * M * N * K / ms / 1e9:.0f} TFLOP/s, {byt / ms / 1e6:.0f} GB/s")
