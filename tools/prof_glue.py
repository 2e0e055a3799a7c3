"""Times the layout / glue kernels of TransformerSpatioTemporalModel.forward at the L0 shape (28 frames, C = 320, 72 x 128):
GroupNorm -> tokens, tokens -> NCHW + residual, AlphaBlender rows, one-key broadcast + residual, frame-embedding add.
Prints ms and GB/s of algorithmic traffic."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from timelapse_video_interpolation_b200 import ops

dev = torch.device("cuda:0")
N, C, H, W = 28, 320, 72, 128
S = H * W
M = N * S


def timed(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def line(name, ms, byt):
    print(f"{name:44s} {ms:.4f} ms  {byt / ms / 1e6:6.0f} GB/s", flush=True)


x = torch.randn(N, C, H, W, device=dev).bfloat16()
tok = torch.randn(N, S, C, device=dev).bfloat16()
tok2 = torch.randn(N, S, C, device=dev).bfloat16()
g, b = torch.randn(C, device=dev).bfloat16(), torch.randn(C, device=dev).bfloat16()
out_tok = torch.empty(N, S, C, device=dev, dtype=torch.bfloat16)
out_img = torch.empty(N, C, H, W, device=dev, dtype=torch.bfloat16)
line("groupnorm_tokens (NCHW -> tokens)", timed(lambda: ops.groupnorm_tokens(x, g, b, 32, 1e-6, out=out_tok)), 2.0 * 2 * M * C + 2.0 * M * C)
line("tokens_to_nchw_add (+ residual)", timed(lambda: ops.tokens_to_nchw_add(tok, x, H, W, out=out_img)), 3 * 2.0 * M * C)
alpha = torch.rand(N, device=dev)
line("blend_rows", timed(lambda: ops.blend_rows(tok, tok2, alpha, S, out=out_tok)), 3 * 2.0 * M * C)
v = torch.randn(N, C, device=dev).bfloat16()
line("bcast_rows_add (one key, + residual)", timed(lambda: ops.bcast_rows_add(v, tok.view(M, C), S, out=out_tok.view(M, C))), 2 * 2.0 * M * C)
vt = torch.randn(2 * S, C, device=dev).bfloat16()
line("bcast_rows_add (interleaved groups)", timed(lambda: ops.bcast_rows_add(vt, tok.view(M, C), 14, out=out_tok.view(M, C), group_inner=S)), 2 * 2.0 * M * C)
