"""Small launch set of the flow kernels for ncu (8 pairs of 1080p)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from timelapse_video_interpolation_b200 import ops
dev = torch.device("cuda:0")
B, H, W = 8, 1080, 1920
g = torch.Generator().manual_seed(0)
img = torch.rand(B, 3, H, W, generator=g).to(dev)
coarse = torch.randn(B, 2, H // 30, W // 30, generator=g).to(dev) * 5
f = torch.nn.functional.interpolate(coarse, size=(H, W), mode="bicubic", align_corners=False).contiguous()
fb = -f + torch.randn(B, 2, H, W, generator=g).to(dev) * 0.2
for _ in range(3):
    ops.backward_warp_f32(img, f, fb, 0.01, 0.5)
torch.cuda.synchronize()
print("ok")
