import sys
sys.path.insert(0, '/root/repo')
import torch
from timelapse_video_interpolation_b200 import ops
dev = torch.device('cuda:0')
B, S, heads, L = 28, 9216, 5, 2
C = heads * 64
g = torch.Generator(device=dev).manual_seed(0)
q = torch.randn(B * S, C, device=dev, generator=g).bfloat16()
do = torch.randn(B * S, C, device=dev, generator=g).bfloat16()
k = torch.randn(B * L, C, device=dev, generator=g).bfloat16()
v = torch.randn(B * L, C, device=dev, generator=g).bfloat16()
for _ in range(5):
    ops.attn_fewkeys_bwd(q, k, v, do, B * S, S, heads, L, 0.125)
    ops.attn_fewkeys(q, k, v, B * S, S, heads, L, 0.125)
torch.cuda.synchronize()
