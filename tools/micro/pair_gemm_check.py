"""CTA-pair (cta_group::2) forward GEMM: correctness against a torch fp32 matmul and timing at the L0 shapes.
Run once with TVI_GEMM_PAIR=1 and once with TVI_GEMM_PAIR=0 (the mode is read once per process)."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import torch
import torch.nn.functional as F

from timelapse_video_interpolation_b200 import ops

dev = torch.device("cuda:0")
mode = os.environ.get("TVI_GEMM_PAIR", "unset")
g = torch.Generator(device=dev).manual_seed(0)
rn = lambda *s: torch.randn(*s, device=dev, generator=g)


def check(M, N, K, bias=False, res=False, geglu=False):
    x = rn(M, K).bfloat16()
    w = (rn(N, K) * K ** -0.5).bfloat16()
    b = rn(N).bfloat16() if bias else None
    n_out = N // 2 if geglu else N
    r = rn(M, n_out).bfloat16() if res else None
    wi, bi = w, b
    if geglu:
        inner, H = N // 2, ops.geglu_block(N // 2)
        wi = torch.stack([w[:inner].view(inner // H, H, -1), w[inner:].view(inner // H, H, -1)], dim=1).reshape(N, -1).contiguous()
        bi = torch.stack([b[:inner].view(inner // H, H), b[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous() if bias else None
    y = ops.gemm_bf16_tn_ex(x, wi, bi, residual=r, geglu=geglu) if (geglu or res) else ops.gemm_bf16_tn(x, w, b)
    torch.cuda.synchronize()
    ref = F.linear(x.float(), w.float(), b.float() if bias else None)
    if geglu:
        ref = ref[:, :N // 2] * F.gelu(ref[:, N // 2:])
    if res:
        ref = ref + r.float()
    err = float((y.float() - ref).abs().max())
    tol = 2.0 ** -7 * max(1.0, float(ref.abs().max()))
    print(f"pair={mode} M={M} N={N} K={K} bias={bias} res={res} geglu={geglu}: max err {err:.3e} (tol {tol:.3e}) {'OK' if err <= tol else 'FAIL'}", flush=True)
    return err <= tol


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


ok = True
if "check" in sys.argv[1:] or len(sys.argv) == 1:
    for args in [(256, 64, 64), (512, 320, 320), (1000, 320, 320, True), (777, 960, 320), (4096, 2560, 320, True, False, True),
                 (2048, 320, 1280, True, True), (300, 128, 192, True), (258048, 960, 320), (65536, 2560, 320, True, False, True)]:
        ok &= check(*args)
if "time" in sys.argv[1:] or len(sys.argv) == 1:
    M, C = 258048, 320
    x, h = rn(M, C).bfloat16(), rn(M, 4 * C).bfloat16()
    res = rn(M, C).bfloat16()
    w1, b1 = (rn(8 * C, C) * C ** -0.5).bfloat16(), rn(8 * C).bfloat16()
    w2, b2 = (rn(C, 4 * C) * (4 * C) ** -0.5).bfloat16(), rn(C).bfloat16()
    w3, wo = (rn(3 * C, C) * C ** -0.5).bfloat16(), (rn(C, C) * C ** -0.5).bfloat16()
    for name, fn, fl in [("ff_in GEGLU N=2560 K=320", lambda: ops.gemm_bf16_tn_ex(x, w1, b1, geglu=True), 2.0 * M * 8 * C * C),
                         ("plain N=2560 K=320", lambda: ops.gemm_bf16_tn(x, w1, b1), 2.0 * M * 8 * C * C),
                         ("qkv N=960 K=320", lambda: ops.gemm_bf16_tn(x, w3), 2.0 * M * 3 * C * C),
                         ("out-proj + res N=320 K=320", lambda: ops.gemm_bf16_tn_ex(x, wo, b2, residual=res), 2.0 * M * C * C),
                         ("ff_out + res N=320 K=1280", lambda: ops.gemm_bf16_tn_ex(h, w2, b2, residual=res), 2.0 * M * 4 * C * C)]:
        ms = timed(fn)
        print(f"pair={mode} {name:30s} {ms:.4f} ms  {fl / ms / 1e9:6.0f} TFLOP/s", flush=True)
sys.exit(0 if ok else 1)
