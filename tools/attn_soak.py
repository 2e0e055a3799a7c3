"""Randomised shape sweep of the attention cores (spatial tcgen05 / mma.sync, temporal incl. interleaved, few-keys) and their
backward kernels against fp32 torch.  Run under `timeout`: a hang is a finding."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from timelapse_video_interpolation_b200 import _lib, ops
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n_cases = int(sys.argv[2]) if len(sys.argv) > 2 else 40
verbose = "-v" in sys.argv
ri = lambda lo, hi: int(torch.randint(lo, hi + 1, (1,), generator=g))


def ref_core(q, k, v, heads, scale):
    B, S, C = q.shape
    sp = lambda t: t.view(B, -1, heads, 64).transpose(1, 2)
    o = torch.softmax(sp(q) @ sp(k).transpose(-1, -2) * scale, dim=-1) @ sp(v)
    return o.transpose(1, 2).reshape(B, S, C)


worst = 0.0
for case in range(n_cases):
    kind = case % 4
    heads = ri(1, 5)
    C = heads * 64
    if kind == 0:      # spatial, auto dispatch (tcgen05 for long sequences)
        B, S = ri(1, 3), ri(1, 2600)
        Skv = S if case % 8 else ri(1, 2600)
        name = f"spatial B={B} h={heads} S={S} Skv={Skv}"
    elif kind == 1:    # spatial forced tcgen05 where it accepts the shape, else mma.sync
        B, S, Skv = ri(1, 2), ri(1000, 2400), ri(1000, 2400)
        name = f"spatial-tc B={B} h={heads} S={S} Skv={Skv}"
    elif kind == 2:    # temporal, optionally interleaved
        B, S, Skv = ri(1, 400), ri(1, 64), 0
        Skv = S
        name = f"temporal nseq={B} h={heads} T={S}"
    else:              # few keys
        B, S, Skv = ri(1, 40), ri(1, 300), ri(1, 8)
        name = f"fewkeys groups={B} h={heads} rows={S} L={Skv}"
    if verbose:
        print(f"case {case}: {name}", flush=True)
    q = torch.randn(B, S, C, generator=g).bfloat16().to(dev).requires_grad_(False)
    k = torch.randn(B, Skv, C, generator=g).bfloat16().to(dev)
    v = torch.randn(B, Skv, C, generator=g).bfloat16().to(dev)
    do = torch.randn(B, S, C, generator=g).bfloat16().to(dev)
    scale = 0.125
    qf, kf, vf = (t.float().requires_grad_(True) for t in (q, k, v))
    ref = ref_core(qf, kf, vf, heads, scale)
    (ref * do.float()).sum().backward()
    q2, k2, v2, do2 = (t.view(-1, C) for t in (q, k, v, do))
    if kind in (0, 1):
        lse = torch.empty(B * S, heads, dtype=torch.float32, device=dev)
        o = ops.attn_spatial(q2, k2, v2, B, heads, S, Skv, scale, lse=lse, algo=_lib.SPATIAL_AUTO)
        dq, dk, dv = ops.attn_spatial_bwd(q2, k2, v2, o, do2, lse, B, heads, S, Skv, scale)
    elif kind == 2:
        o = ops.attn_temporal(q2, k2, v2, B, heads, S, S, scale)
        dq, dk, dv = ops.attn_temporal_bwd(q2, k2, v2, do2, B, heads, S, S, scale)
        inner = ri(1, 6)
        if B % inner == 0:   # the same sequences interleaved: [B/inner, T, inner] token order
            perm = lambda t: t.view(B // inner, inner, S, C).transpose(1, 2).contiguous().view(-1, C)
            oi = ops.attn_temporal(perm(q), perm(k), perm(v), B, heads, S, S, scale, seq_inner=inner)
            assert torch.equal(oi.view(B // inner, S, inner, C).transpose(1, 2).reshape(-1, C), o), name + f" inner={inner}"
    else:
        o = ops.attn_fewkeys(q2, k2, v2, B * S, S, heads, Skv, scale)
        dq, dk, dv = ops.attn_fewkeys_bwd(q2, k2, v2, do2, B * S, S, heads, Skv, scale)
    torch.cuda.synchronize()
    for nm, got, want, tol in (("o", o, ref.detach(), 2e-2), ("dq", dq, qf.grad, None), ("dk", dk, kf.grad, None), ("dv", dv, vf.grad, None)):
        if got is None:
            continue
        want = want.reshape(got.shape)
        err = float((got.float() - want).abs().max())
        t = tol if tol is not None else 2e-2 * max(1e-6, float(want.abs().max()))
        worst = max(worst, err / t)
        if not err <= t:
            print(f"FAIL case {case} {name} {nm}: err {err:.3e} tol {t:.3e}")
            sys.exit(1)
print(f"attn soak: {n_cases} random cases OK, worst err/tol {worst:.3f}")
