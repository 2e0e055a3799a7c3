"""Per-kernel timing of the N1 (FeedForward / LayerNorm / out-projection) launches at the L0 block shape:
M = 258048 tokens, C = 320.  One line per kernel: ms, TFLOP/s, algorithmic GB/s."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from timelapse_video_interpolation_b200 import ops, _lib
import os
if os.environ.get("PROF_LIB"):  # a tuning build (tools/micro/build_tune.sh)
    _lib.LIB_PATH = Path(os.environ["PROF_LIB"]).resolve()
dev = torch.device("cuda:0")
import os
M, C = int(os.environ.get("PROF_M", 258048)), int(os.environ.get("PROF_C", 320))
g = torch.Generator(device=dev).manual_seed(0)
rn = lambda *s: torch.randn(*s, device=dev, generator=g)
x = rn(M, C).bfloat16()
res = rn(M, C).bfloat16()
h = rn(M, 4 * C).bfloat16()
w1 = (rn(8 * C, C) * C ** -0.5).bfloat16()
b1 = rn(8 * C).bfloat16()
w2 = (rn(C, 4 * C) * (4 * C) ** -0.5).bfloat16()
b2 = rn(C).bfloat16()
wo = (rn(C, C) * C ** -0.5).bfloat16()
gam, bet = rn(C).bfloat16(), rn(C).bfloat16()
y = torch.empty(M, C, dtype=torch.bfloat16, device=dev)
hh = torch.empty(M, 4 * C, dtype=torch.bfloat16, device=dev)
big = torch.empty(M, 8 * C, dtype=torch.bfloat16, device=dev)


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


def line(name, ms, flops, byt):
    print(f"{name:34s} {ms:.4f} ms  {flops / ms / 1e9:7.0f} TFLOP/s  {byt / ms / 1e6:6.0f} GB/s", flush=True)


print(f"M={M} C={C}")
only = sys.argv[1].split(",") if len(sys.argv) > 1 else ["ln", "ffin", "ffin_plain", "ffout", "ffout_res", "qkv", "oproj", "oproj_res"]
if "ln" in only:
    line(f"layernorm C={C}", timed(lambda: ops.layernorm_bf16(x, gam, bet, 1e-5, out=y)), 0, 4.0 * M * C)
if "ffin" in only:
    line("ff_in GEGLU (N=2560->1280,K=320)", timed(lambda: ops.gemm_bf16_tn_ex(x, w1, b1, geglu=True, out=hh)),
         2.0 * M * 8 * C * C, 2.0 * (M * C + M * 4 * C))
if "ffin_plain" in only:
    line("plain GEMM (N=2560,K=320)", timed(lambda: ops.gemm_bf16_tn(x, w1, b1, out=big)), 2.0 * M * 8 * C * C, 2.0 * (M * C + M * 8 * C))
if "ffout" in only:
    line("ff_out (N=320,K=1280)", timed(lambda: ops.gemm_bf16_tn(h, w2, b2, out=y)), 2.0 * M * 4 * C * C, 2.0 * (M * 4 * C + M * C))
if "ffout_res" in only:
    line("ff_out + residual", timed(lambda: ops.gemm_bf16_tn_ex(h, w2, b2, residual=res, out=y)), 2.0 * M * 4 * C * C, 2.0 * (M * 4 * C + 2 * M * C))
if "qkv" in only:
    w3 = (rn(3 * C, C) * C ** -0.5).bfloat16()
    y3 = torch.empty(M, 3 * C, dtype=torch.bfloat16, device=dev)
    line("qkv (N=3C,K=C)", timed(lambda: ops.gemm_bf16_tn(x, w3, None, out=y3)), 2.0 * M * 3 * C * C, 2.0 * (M * C + 3 * M * C))
if "oproj" in only:
    line("out-proj (N=320,K=320)", timed(lambda: ops.gemm_bf16_tn(x, wo, b2, out=y)), 2.0 * M * C * C, 4.0 * M * C)
if "oproj_res" in only:
    line("out-proj + residual", timed(lambda: ops.gemm_bf16_tn_ex(x, wo, b2, residual=res, out=y)), 2.0 * M * C * C, 6.0 * M * C)
