"""Quick GPU check of the tcgen05 spatial kernel against torch SDPA (fp32) + timing vs mma.sync and torch."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from timelapse_video_interpolation_b200 import _lib, ops

dev = torch.device("cuda:0")
def ref(q, k, v, heads, scale):
    B, S, C = q.shape
    f = lambda t: t.float().view(t.shape[0], t.shape[1], heads, 64).transpose(1, 2)
    o = torch.nn.functional.scaled_dot_product_attention(f(q), f(k), f(v), scale=scale)
    return o.transpose(1, 2).reshape(B, S, C)

for (B, heads, S, Skv) in [(1, 1, 256, 128), (1, 1, 256, 256), (1, 2, 512, 640), (2, 5, 1000, 1100), (1, 5, 2304, 2304)]:
    g = torch.Generator().manual_seed(S)
    C = heads * 64
    q = torch.randn(B, S, C, generator=g).bfloat16().to(dev)
    k = torch.randn(B, Skv, C, generator=g).bfloat16().to(dev)
    v = torch.randn(B, Skv, C, generator=g).bfloat16().to(dev)
    o = ops.attn_spatial(q.view(-1, C), k.view(-1, C), v.view(-1, C), B, heads, S, Skv, 0.125, algo=_lib.SPATIAL_TCGEN05)
    torch.cuda.synchronize()
    r = ref(q, k, v, heads, 0.125)
    err = (o.float().view(B, S, C) - r).abs().max().item()
    print(f"tc B={B} h={heads} S={S} Skv={Skv} max abs err {err:.3e}", flush=True)

# timing at L0 shape
B, heads, S = 28, 5, 9216
C = heads * 64
qkv = torch.randn(B * S, 3 * C, device=dev).bfloat16()
fl = 4.0 * B * heads * S * S * 64
for name, algo in [("mma_sync", _lib.SPATIAL_MMA_SYNC), ("tcgen05", _lib.SPATIAL_TCGEN05)]:
    out = torch.empty(B * S, C, dtype=torch.bfloat16, device=dev)
    for _ in range(2):
        ops.attn_spatial(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], B, heads, S, S, 0.125, out=out, algo=algo)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ops.attn_spatial(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], B, heads, S, S, 0.125, out=out, algo=algo)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"{name}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s", flush=True)
    if name == "mma_sync":
        base = out.clone()
    else:
        print("tc vs mma_sync max abs diff", (out.float() - base.float()).abs().max().item())
q4 = qkv.view(B, S, 3, heads, 64).permute(2, 0, 3, 1, 4).contiguous()
for _ in range(2):
    torch.nn.functional.scaled_dot_product_attention(q4[0], q4[1], q4[2])
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    torch.nn.functional.scaled_dot_product_attention(q4[0], q4[1], q4[2])
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"torch SDPA bf16: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s")
