"""Shared helpers for the parity tests."""
import numpy as np
import torch

from timelapse_video_interpolation_b200.modules import Attention


class _NoProcessor:
    def __call__(self, *a, **k):
        raise AssertionError("oracle-side module: processor must not be called")


def attn_from_golden(z, name: str, device="cpu", dtype=torch.float32, processor=None) -> Attention:
    """Rebuilds an Attention container carrying the golden case's weights."""
    wq, wk = z[f"{name}.wq"], z[f"{name}.wk"]
    heads = int(z[f"{name}.heads"]) if f"{name}.heads" in z else 2
    attn = Attention(query_dim=wq.shape[1], cross_attention_dim=wk.shape[1], heads=heads, dim_head=wq.shape[0] // heads,
                     processor=processor if processor is not None else _NoProcessor())
    with torch.no_grad():
        attn.to_q.weight.copy_(torch.from_numpy(wq))
        attn.to_k.weight.copy_(torch.from_numpy(wk))
        attn.to_v.weight.copy_(torch.from_numpy(z[f"{name}.wv"]))
        attn.to_out[0].weight.copy_(torch.from_numpy(z[f"{name}.wo"]))
        attn.to_out[0].bias.copy_(torch.from_numpy(z[f"{name}.bo"]))
    return attn.to(device=device, dtype=dtype).eval()


def make_attn(query_dim, cross_dim, heads, seed, device="cpu", dtype=torch.float32, processor=None) -> Attention:
    g = torch.Generator().manual_seed(seed)
    attn = Attention(query_dim=query_dim, cross_attention_dim=cross_dim, heads=heads, dim_head=64,
                     processor=processor if processor is not None else _NoProcessor())
    with torch.no_grad():
        for p in attn.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5))
    return attn.to(device=device, dtype=dtype).eval()


def bf16_round(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


def max_abs(a, b) -> float:
    return float((torch.as_tensor(a, dtype=torch.float64) - torch.as_tensor(b, dtype=torch.float64)).abs().max())
