"""Randomised shape sweep of the row-wise / layout kernels (LayerNorm fwd+bwd, GroupNorm->tokens, tokens->NCHW, blend,
broadcast-add, column sums) against fp32 torch.  Run under `timeout`."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import torch.nn.functional as F
from timelapse_video_interpolation_b200 import ops
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n_cases = int(sys.argv[2]) if len(sys.argv) > 2 else 60
ri = lambda lo, hi: int(torch.randint(lo, hi + 1, (1,), generator=g))
bf = lambda *s: torch.randn(*s, generator=g).bfloat16().to(dev)


def check(name, got, want, rel=2 ** -7):
    want = want.reshape(got.shape)
    err = float((got.float().cpu() - want.float().cpu()).abs().max()) if got.numel() else 0.0
    tol = rel * max(1.0, float(want.abs().max()) if want.numel() else 1.0)
    if not err <= tol:
        print(f"FAIL {name}: err {err:.3e} tol {tol:.3e}")
        sys.exit(1)


for case in range(n_cases):
    kind = case % 6
    if kind == 0:
        M, C = ri(1, 3000), 8 * ri(1, 256)
        C = [320, 640, 1280, C][ri(0, 3)]
        x, dy, gm, bt = bf(M, C) * 2 + 0.5, bf(M, C), bf(C), bf(C)
        name = f"layernorm M={M} C={C}"
        print(f"case {case}: {name}", flush=True)
        y = ops.layernorm_bf16(x, gm, bt, 1e-5)
        xr, gr, br = x.float().requires_grad_(True), gm.float().requires_grad_(True), bt.float().requires_grad_(True)
        ref = F.layer_norm(xr, (C,), gr, br, 1e-5)
        (ref * dy.float()).sum().backward()
        check(name, y, ref.detach(), 2e-2)
        dx, dg, db = ops.layernorm_bwd(x, dy, gm, 1e-5)
        for nm, a, b in (("dx", dx, xr.grad), ("dgamma", dg, gr.grad), ("dbeta", db, br.grad)):
            check(name + " " + nm, a, b, 2e-2 * max(1.0, float(b.abs().max())) / max(1.0, float(b.abs().max())))
    elif kind == 1:
        groups = [1, 2, 4, 8, 32][ri(0, 4)]
        N, C, H, W = ri(1, 5), 8 * groups * ri(1, 6), ri(1, 20), 8 * ri(1, 12)
        name = f"groupnorm N={N} C={C} H={H} W={W} groups={groups}"
        print(f"case {case}: {name}", flush=True)
        x, gm, bt = bf(N, C, H, W) + 0.3, bf(C), bf(C)
        y = ops.groupnorm_tokens(x, gm, bt, groups, 1e-6)
        ref = F.group_norm(x.float(), groups, gm.float(), bt.float(), 1e-6).permute(0, 2, 3, 1).reshape(N, H * W, C)
        check(name, y, ref, 2e-2)
    elif kind == 2:
        N, C, H, W = ri(1, 5), 8 * ri(1, 60), ri(1, 20), 8 * ri(1, 12)
        name = f"tokens_to_nchw N={N} C={C} H={H} W={W}"
        print(f"case {case}: {name}", flush=True)
        y, r = bf(N, H * W, C), bf(N, C, H, W)
        check(name, ops.tokens_to_nchw_add(y, r, H, W), y.float().view(N, H, W, C).permute(0, 3, 1, 2) + r.float())
    elif kind == 3:
        BF, S, C = ri(1, 30), ri(1, 200), 8 * ri(1, 100)
        name = f"blend BF={BF} S={S} C={C}"
        print(f"case {case}: {name}", flush=True)
        xs, xt, a = bf(BF, S, C), bf(BF, S, C), torch.rand(BF, generator=g).to(dev)
        check(name, ops.blend_rows(xs, xt, a, S), a.view(BF, 1, 1) * xs.float() + (1 - a.view(BF, 1, 1)) * xt.float())
    elif kind == 4:
        B, T, S, C = ri(1, 4), ri(1, 30), ri(1, 60), 8 * ri(1, 100)
        name = f"bcast B={B} T={T} S={S} C={C}"
        print(f"case {case}: {name}", flush=True)
        v, r = bf(B * S, C), bf(B * T * S, C)
        check(name, ops.bcast_rows_add(v, r, T, group_inner=S), v.float().view(B, 1, S, C) + r.float().view(B, T, S, C))
        r2 = bf(B * S * T, C)
        check(name + " contiguous", ops.bcast_rows_add(v, r2, T), v.float().view(B * S, 1, C) + r2.float().view(B * S, T, C))
    else:
        groups, rpg, C = ri(1, 6), ri(1, 3000), 8 * ri(1, 200)
        name = f"colsum groups={groups} rows={rpg} C={C}"
        print(f"case {case}: {name}", flush=True)
        x = bf(groups * rpg, C)
        want = x.float().view(groups, rpg, C).sum(1)
        check(name, ops.colsum_bf16(x, rpg), want, 2e-2)
    torch.cuda.synchronize()
print(f"misc soak: {n_cases} random cases OK")
