"""Randomised shape sweep of the tcgen05 GEMM entry points against fp32 torch (bias / residual / GEGLU / dgrad / wgrad /
the GEGLU-backward epilogue)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from timelapse_video_interpolation_b200 import ops
dev = torch.device("cuda:0")
g = torch.Generator(device="cpu").manual_seed(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n_cases = int(sys.argv[2]) if len(sys.argv) > 2 else 80
worst = 0.0
for case in range(n_cases):
    M = int(torch.randint(1, 5000, (1,), generator=g)) if case % 4 else int(torch.randint(1, 40, (1,), generator=g))
    N = 8 * int(torch.randint(1, 200, (1,), generator=g))
    K = 8 * int(torch.randint(1, 180, (1,), generator=g))
    mode = case % 6
    if "-v" in sys.argv:
        print(f"case {case} mode {mode} M={M} N={N} K={K} bias={case % 2}", flush=True)
    x = torch.randn(M, K, generator=g).bfloat16().to(dev)
    w = (torch.randn(N, K, generator=g) * K ** -0.5).bfloat16().to(dev)
    b = torch.randn(N, generator=g).bfloat16().to(dev) if case % 2 else None
    xf, wf = x.float(), w.float()
    if mode == 0:
        y = ops.gemm_bf16_tn(x, w, b)
        ref = xf @ wf.T + (b.float() if b is not None else 0)
        what = "tn"
    elif mode == 1:
        r = torch.randn(M, N, generator=g).bfloat16().to(dev)
        y = ops.gemm_bf16_tn_ex(x, w, b, residual=r)
        ref = xf @ wf.T + (b.float() if b is not None else 0) + r.float()
        what = "tn+res"
    elif mode == 2:
        inner = 32 * max(1, N // 64)
        H = ops.geglu_block(inner)
        w2 = (torch.randn(2 * inner, K, generator=g) * K ** -0.5).bfloat16().to(dev)
        b2 = torch.randn(2 * inner, generator=g).bfloat16().to(dev) if case % 2 else None
        wi = torch.stack([w2[:inner].view(inner // H, H, -1), w2[inner:].view(inner // H, H, -1)], dim=1).reshape(2 * inner, -1).contiguous()
        bi = torch.stack([b2[:inner].view(inner // H, H), b2[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous() if b2 is not None else None
        y = ops.gemm_bf16_tn_ex(x, wi, bi, geglu=True)
        pre = xf @ w2.float().T + (b2.float() if b2 is not None else 0)
        ref = pre[:, :inner] * torch.nn.functional.gelu(pre[:, inner:])
        what = f"geglu inner={inner}"
    elif mode == 3:
        dy = torch.randn(M, N, generator=g).bfloat16().to(dev)
        y = ops.gemm_bf16_dgrad(dy, w)
        ref = dy.float() @ wf
        what = "dgrad"
    elif mode == 4:
        dy = torch.randn(M, N, generator=g).bfloat16().to(dev)
        y = ops.gemm_bf16_wgrad(dy, x)
        ref = dy.float().T @ xf
        what = "wgrad"
    else:
        inner = 32 * max(1, N // 64)
        H = ops.geglu_block(inner)
        w2 = (torch.randn(2 * inner, K, generator=g) * K ** -0.5).bfloat16().to(dev)
        b2 = torch.randn(2 * inner, generator=g).bfloat16().to(dev) if case % 2 else None
        wi = torch.stack([w2[:inner].view(inner // H, H, -1), w2[inner:].view(inner // H, H, -1)], dim=1).reshape(2 * inner, -1).contiguous()
        bi = torch.stack([b2[:inner].view(inner // H, H), b2[inner:].view(inner // H, H)], dim=1).reshape(-1).contiguous() if b2 is not None else None
        dh = torch.randn(M, inner, generator=g).bfloat16().to(dev)
        h, dpre = ops.gemm_geglu_bwd(x, wi, bi, dh)
        pre = (xf @ w2.float().T + (b2.float() if b2 is not None else 0)).requires_grad_(True)
        href = pre[:, :inner] * torch.nn.functional.gelu(pre[:, inner:])
        href.backward(dh.float())
        torch.cuda.synchronize()
        eh = float((h.float() - href.detach()).abs().max())
        th = 2e-2 * max(1.0, float(href.abs().max()))
        if not eh <= th:
            print(f"FAIL case {case} geglu_bwd h M={M} inner={inner} K={K}: err {eh:.3e} tol {th:.3e}")
            sys.exit(1)
        y, ref = dpre, pre.grad
        what = f"geglu_bwd inner={inner}"
    torch.cuda.synchronize()
    err = float((y.float() - ref).abs().max())
    tol = 2e-2 * max(1.0, float(ref.abs().max()))
    worst = max(worst, err / tol)
    if not err <= tol:
        print(f"FAIL case {case} {what} M={M} N={N} K={K} bias={b is not None}: err {err:.3e} tol {tol:.3e}")
        sys.exit(1)
print(f"gemm soak: {n_cases} random cases OK, worst err/tol {worst:.3f}")
