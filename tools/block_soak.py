"""Random configurations of the block-level and module-level forwards against the oracle (run under `timeout`)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from oracle import blocks_ref, spatio_temporal_ref
from timelapse_video_interpolation_b200 import blocks, spatio_temporal
from timelapse_video_interpolation_b200.modules import BasicTransformerBlock, TemporalBasicTransformerBlock, TransformerSpatioTemporalModel
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n = int(sys.argv[2]) if len(sys.argv) > 2 else 24
ri = lambda lo, hi: int(torch.randint(lo, hi + 1, (1,), generator=g))


def init(m):
    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("mix_factor"):
                p.fill_(float(torch.randn(1, generator=g)))
            elif "norm" in name:
                p.copy_(1 + 0.1 * torch.randn(p.shape, generator=g) if name.endswith("weight") else 0.05 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * (p.shape[-1] ** -0.5 if p.dim() > 1 else 0.05))
    return m.to(dev, torch.bfloat16).eval()


bf = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16).to(dev)
worst = 0.0
for case in range(n):
    heads = [1, 2, 5, 10, 20][ri(0, 4)] if case % 3 else ri(1, 6)
    C = 64 * heads
    cross = 32 * ri(1, 32)
    kind = case % 3
    with torch.no_grad():
        if kind == 0:
            BF, S, L = ri(1, 6), ri(1, 700), ri(1, 3)
            print(f"case {case}: spatial block C={C} BF={BF} S={S} L={L} cross={cross}", flush=True)
            m = init(BasicTransformerBlock(C, heads, cross))
            blocks.install_blocks(m)
            x, e = bf(BF, S, C), bf(BF, L, cross)
            y = m(x, encoder_hidden_states=e)
            ref = blocks_ref.basic_transformer_block_ref(m, x, e)
        elif kind == 1:
            B, T, S, L = ri(1, 3), ri(1, 32), ri(1, 120), ri(1, 2)
            print(f"case {case}: temporal block C={C} B={B} T={T} S={S} L={L} cross={cross}", flush=True)
            m = init(TemporalBasicTransformerBlock(C, C, heads, cross))
            blocks.install_blocks(m)
            x, e = bf(B * T, S, C), bf(B * S, L, cross)
            y = m(x, num_frames=T, encoder_hidden_states=e)
            ref = blocks_ref.temporal_basic_transformer_block_ref(m, x, T, e)
        else:
            B, T, H, W = ri(1, 2), ri(1, 16), ri(1, 12), 8 * ri(1, 3)
            print(f"case {case}: module C={C} B={B} T={T} H={H} W={W} cross={cross}", flush=True)
            m = init(TransformerSpatioTemporalModel(heads, 64, in_channels=C, num_layers=ri(1, 2), cross_attention_dim=cross))
            spatio_temporal.install_spatio_temporal(m)
            x, e = bf(B * T, C, H, W), bf(B * T, 1, cross)
            ind = (torch.rand(B, T, generator=g) < 0.2).float().to(dev)
            y = m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]
            ref = spatio_temporal_ref.transformer_spatio_temporal_ref(m, x, e, ind, alpha_dtype=torch.bfloat16)
    torch.cuda.synchronize()
    err = float((y.float().cpu() - ref).abs().max())
    tol = 2e-2 * max(1.0, float(ref.abs().max()))
    worst = max(worst, err / tol)
    if not (y.shape == x.shape and err <= tol):
        print(f"FAIL: err {err:.3e} tol {tol:.3e}")
        sys.exit(1)
print(f"block soak: {n} random cases OK, worst err/tol {worst:.3f}")
