#!/usr/bin/env python
"""Times tvi_attn_spatial_bwd at a UNet level's shape (default L0: 28 frames x 5 heads x 9216 tokens) for both
implementations and prints ms, TFLOP/s of the 7 GEMM units actually executed and of the 5 units of a single-pass
backward.  python tools/time_spatial_bwd.py [batch heads S]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402

from timelapse_video_interpolation_b200 import _lib, ops  # noqa: E402

batch, heads, S = (int(a) for a in sys.argv[1:4]) if len(sys.argv) >= 4 else (28, 5, 9216)
dev = torch.device("cuda", 0)
C = heads * 64
g = torch.Generator().manual_seed(0)
qkv = torch.randn(batch * S, 3 * C, generator=g).to(torch.bfloat16).to(dev)
q, k, v = qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:]
dout = torch.randn(batch * S, C, generator=g).to(torch.bfloat16).to(dev)
lse = torch.empty(batch * S, heads, dtype=torch.float32, device=dev)
o = ops.attn_spatial(q, k, v, batch, heads, S, S, 0.125, lse=lse)
dqkv = torch.empty_like(qkv)
unit = 2.0 * batch * heads * S * S * 64
res = {}
import os  # noqa: E402
cases = [("tcgen05", _lib.SPATIAL_TCGEN05, 10), ("mma_sync", _lib.SPATIAL_MMA_SYNC, 3)]
if os.environ.get("TVI_BWD_ONLY_TC"):  # profiling runs
    cases = [("tcgen05", _lib.SPATIAL_TCGEN05, 3)]
for name, algo, reps in cases:
    def run():
        ops.attn_spatial_bwd(q, k, v, o, dout, lse, batch, heads, S, S, 0.125, dq=dqkv[:, :C], dk=dqkv[:, C:2 * C], dv=dqkv[:, 2 * C:],
                             algo=algo)
    for _ in range(2):
        run()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        run()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    res[name] = dqkv.clone()
    print(f"{name}: {ms:.3f} ms  ({7 * unit / ms / 1e9:.1f} TFLOP/s over the 7 units executed, {5 * unit / ms / 1e9:.1f} over 5)", flush=True)
if "mma_sync" not in res:
    sys.exit(0)
d = (res["tcgen05"].float() - res["mma_sync"].float()).abs().max().item()
print(f"max |tcgen05 - mma_sync| over dq|dk|dv: {d:.4e} (max |grad| {res['mma_sync'].float().abs().max().item():.4e})")
