"""Small launch of the spatial attention kernel for ncu (L0 shape, 4 frames)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from timelapse_video_interpolation_b200 import _lib, ops
dev = torch.device("cuda:0")
B, heads, S = 4, 5, 9216
C = heads * 64
qkv = torch.randn(B * S, 3 * C, device=dev).bfloat16()
out = torch.empty(B * S, C, dtype=torch.bfloat16, device=dev)
algo = int(sys.argv[1]) if len(sys.argv) > 1 else _lib.SPATIAL_TCGEN05
for _ in range(3):
    ops.attn_spatial(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], B, heads, S, S, 0.125, out=out, algo=algo)
torch.cuda.synchronize()
print("ok")
