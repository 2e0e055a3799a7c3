"""One timing line for the tcgen05 spatial kernel at the L0 shape (env TVI_TC_* select the variant)."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from timelapse_video_interpolation_b200 import _lib, ops
dev = torch.device("cuda:0")
B, heads, S = 28, 5, 9216
C = heads * 64
g = torch.Generator(device=dev).manual_seed(0)
qkv = (torch.randn(B * S, 3 * C, device=dev, generator=g) * float(os.environ.get("TVI_TC_STD", "1"))).bfloat16()
out = torch.empty(B * S, C, dtype=torch.bfloat16, device=dev)
fl = 4.0 * B * heads * S * S * 64
run = lambda: ops.attn_spatial(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], B, heads, S, S, 0.125, out=out, algo=_lib.SPATIAL_TCGEN05)
for _ in range(3):
    run()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"POLY={os.environ.get('TVI_TC_POLY')} STAGGER={os.environ.get('TVI_TC_STAGGER')}: {ms:.3f} ms {fl / ms / 1e9:.1f} TFLOP/s checksum {out.float().abs().mean().item():.6f}", flush=True)
