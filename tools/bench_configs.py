#!/usr/bin/env python
"""Secondary measurements for the parity-test configs of BASELINE.json: bench.py keeps the
headline line (configs[1]); this script times the other configurations on one B200 and prints one JSON
line per config (copied under profiles/ per round).

    python tools/bench_configs.py [--only slerp,flow,train,loop25,kernels] [--steps K]

  slerp   config 1: latent slerp, clips in {1, 64, 4096}, T in {14, 25}          -> GB/s vs HBM peak
  flow    config 4: 1080p flow cosine + backward warp / consistency, B in {1,8,32} -> GB/s vs HBM peak
  train   config 5: forward + backward of one L0 block's four attention calls, 2 clips x T=14 (no CFG),
          temporal modules trainable, spatial modules frozen (train_decouples)     -> ms, clip-frames/s
  loop25  config 3: 25-frame, 25-step loop over the 23 block instances of UNet + ControlNet branch,
          attention path only, 1 clip (CFG 2)                                      -> clip-frames/s
  kernels per-kernel-family device time of one L0 forward and forward+backward step (CUDA events)
"""
from __future__ import annotations

import argparse
import json
import sys
from collections import defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

import bench  # noqa: E402
import timelapse_video_interpolation_b200 as tvi  # noqa: E402
from timelapse_video_interpolation_b200 import ops  # noqa: E402
from timelapse_video_interpolation_b200.modules import Attention  # noqa: E402
from timelapse_video_interpolation_b200.processor import TviAttnProcessor  # noqa: E402

DEV = torch.device("cuda", 0)
PK = bench.peaks()


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def timed_graph(fn, launches=10, replays=5):
    """ms per call of `fn` with the CPU out of the picture: `launches` calls captured in one CUDA graph, replayed; for kernels
    of ~0.1 ms, where the Python / ctypes cost of a call is of the same order as the kernel."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()                                   # allocator warm-up on the capture stream
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(launches):
            fn()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(replays):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (launches * replays)


def emit(d):
    print(json.dumps(d), flush=True)


def flush_l2():
    # write a buffer larger than the 126 MB L2 between timed launches of small problems
    if not hasattr(flush_l2, "buf"):
        flush_l2.buf = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    flush_l2.buf.fill_(1)


def cfg_slerp(steps):
    n = 4 * 72 * 128
    for T in (14, 25):
        for clips in (1, 64, 4096):
            g = torch.Generator().manual_seed(0)
            v0 = torch.randn(clips, 4, 72, 128, generator=g).to(DEV)
            v1 = torch.randn(clips, 4, 72, 128, generator=g).to(DEV)
            out = torch.empty(clips, T, 4, 72, 128, device=DEV)
            ms = timed(lambda: ops.slerp_f32(v0, v1, T, 0.9995, out=out), steps)
            nbytes = (2 + T) * n * 4 * clips
            gbs = nbytes / (ms * 1e-3) / 1e9
            emit({"config": "slerp", "clips": clips, "T": T, "ms": ms, "clip_frames_per_s": clips * T / (ms * 1e-3),
                  "algorithmic_bytes": nbytes, "gbs": gbs, "hbm_frac": gbs / PK["hbm"],
                  "note": "working set < L2 for clips <= 64 (launch-bound)" if clips <= 64 else "5 GB working set"})


def cfg_flow(steps):
    H, W = 1080, 1920
    for B in (1, 8, 32):
        g = torch.Generator().manual_seed(B)
        img = torch.rand(B, 3, H, W, generator=g).to(DEV)
        # smooth flow fields (N(0, 5 px) blurred at ~15 px): a coarse random field,
        # bicubically upsampled; the backward flow is the negated forward flow plus N(0, 0.2) noise
        coarse = torch.randn(B, 2, H // 30, W // 30, generator=g).to(DEV) * 5
        f = torch.nn.functional.interpolate(coarse, size=(H, W), mode="bicubic", align_corners=False).contiguous()
        fb = (-f + torch.randn(B, 2, H, W, generator=g).to(DEV) * 0.2)
        fa_hwc = f.permute(0, 2, 3, 1).contiguous()
        fb_hwc = fb.permute(0, 2, 3, 1).contiguous()
        warped = torch.empty_like(img)
        mask = torch.empty(B, H, W, dtype=torch.uint8, device=DEV)
        ms_w = timed(lambda: ops.backward_warp_f32(img, f, fb, 0.01, 0.5, warped=warped, mask=mask), steps)
        ms_c = timed(lambda: ops.flow_cosine_f32(fa_hwc, fb_hwc, True), steps)
        bw = ((3 + 2 + 2) * 4 + 3 * 4 + 1) * H * W * B
        bc = 2 * 2 * 4 * H * W * B
        emit({"config": "flow_warp_check_1080p", "pairs": B, "ms": ms_w, "pairs_per_s": B / (ms_w * 1e-3), "algorithmic_bytes": bw,
              "gbs": bw / (ms_w * 1e-3) / 1e9, "hbm_frac": bw / (ms_w * 1e-3) / 1e9 / PK["hbm"]})
        emit({"config": "flow_cosine_1080p", "pairs": B, "ms": ms_c, "pairs_per_s": B / (ms_c * 1e-3), "algorithmic_bytes": bc,
              "gbs": bc / (ms_c * 1e-3) / 1e9, "hbm_frac": bc / (ms_c * 1e-3) / 1e9 / PK["hbm"]})


class KernelTimes:
    """CUDA-event bracket around every libtvi launch (ops.EVENT_HOOK), aggregated per entry point."""

    def __init__(self):
        self.pairs = []
        self.open = None

    def hook(self, name, is_start):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        if is_start:
            self.open = (name, ev)
        else:
            self.pairs.append((self.open[0], self.open[1], ev))

    def summary(self, steps):
        torch.cuda.synchronize()
        agg = defaultdict(lambda: [0, 0.0])
        for name, a, b in self.pairs:
            agg[name][0] += 1
            agg[name][1] += a.elapsed_time(b)
        return {k: {"launches_per_step": v[0] / steps, "ms_per_step": v[1] / steps} for k, v in sorted(agg.items())}


def make_block(level, clips, cfg, T, trainable_temporal=False):
    C, heads, h, w = bench.LEVELS[level]
    S = h * w
    g = torch.Generator().manual_seed(2)
    proc = TviAttnProcessor()

    def mk(cross):
        a = Attention(C, cross, heads=heads, dim_head=64, processor=proc)
        with torch.no_grad():
            for p in a.parameters():
                bound = p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5
                p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * bound)
        return a.to(device=DEV, dtype=torch.bfloat16)

    mods = [mk(None), mk(1024), mk(None), mk(1024)]
    for i, m in enumerate(mods):
        for p in m.parameters():
            p.requires_grad_(trainable_temporal and i >= 2)
    BF, nseq = clips * cfg * T, clips * cfg * S
    x = torch.randn(BF, S, C, generator=g).to(torch.bfloat16).to(DEV)
    ctx_s = torch.randn(BF, 2, 1024, generator=g).to(torch.bfloat16).to(DEV)
    ctx_t = torch.randn(nseq, 1, 1024, generator=g).to(torch.bfloat16).to(DEV)

    def step(xin):
        y1 = mods[0](xin)
        y2 = mods[1](y1, encoder_hidden_states=ctx_s)
        y3 = mods[2](y2.view(nseq, T, C))
        return mods[3](y3, encoder_hidden_states=ctx_t)

    return mods, x, step, (BF, S, C, nseq)


def cfg_train(steps):
    # accelerate config of the reference: per_gpu_batch_size 2, T = 14, no CFG -> BF = 28 frames at 576x1024
    clips, T = 2, 14
    mods, x, step, (BF, S, C, nseq) = make_block("L0", clips, 1, T, trainable_temporal=True)
    xg = x.clone().requires_grad_(True)  # gradient flows on to the blocks below
    dy = (torch.randn(nseq, T, C, device=DEV) / 8).to(torch.bfloat16)

    def fwd():
        with torch.no_grad():
            step(x)

    def fwd_bwd():
        for m in mods:
            for p in m.parameters():
                p.grad = None
        xg.grad = None
        step(xg).backward(dy)

    ms_f = timed(fwd, steps)
    ms_fb = timed(fwd_bwd, steps)
    kt = KernelTimes()
    ops.EVENT_HOOK = kt.hook
    for _ in range(3):
        fwd_bwd()
    ops.EVENT_HOOK = None
    ngrad = sum(p.numel() for m in mods for p in m.parameters() if p.requires_grad)
    emit({"config": "train_fwd_bwd_L0_block_attention", "clips_per_gpu": clips, "T": T, "resolution": "576x1024",
          "trainable": "temporal attn1/attn2 (spatial frozen)", "trainable_params": ngrad, "ms_fwd": ms_f, "ms_fwd_bwd": ms_fb,
          "clip_frames_per_s_fwd_bwd": clips * T / (ms_fb * 1e-3), "kernels": kt.summary(3)})


def cfg_train_block(steps):
    """Config 5 through the block-level swap: spatial block frozen, temporal block trainable (src/train_decouples.py:707-720),
    2 clips x T = 14 at 576x1024, no CFG; forward + backward of LayerNorm / attention / FeedForward on libtvi."""
    from timelapse_video_interpolation_b200 import blocks
    from timelapse_video_interpolation_b200.modules import BasicTransformerBlock, TemporalBasicTransformerBlock
    C, heads, h, w = bench.LEVELS["L0"]
    S, T, B = h * w, 14, 2
    g = torch.Generator().manual_seed(3)

    def init(m, trainable):
        with torch.no_grad():
            for name, p in m.named_parameters():
                if "norm" in name:
                    p.fill_(1.0 if name.endswith("weight") else 0.0)
                else:
                    p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5))
                p.requires_grad_(trainable)
        return m.to(DEV, torch.bfloat16)

    sb, tb = init(BasicTransformerBlock(C, heads, 1024), False), init(TemporalBasicTransformerBlock(C, C, heads, 1024), True)
    blocks.install_blocks(sb)
    blocks.install_blocks(tb)
    x = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(DEV)
    xg = x.clone().requires_grad_(True)
    ctx_s = torch.randn(B * T, 2, 1024, generator=g).to(torch.bfloat16).to(DEV)
    ctx_t = torch.randn(B * S, 1, 1024, generator=g).to(torch.bfloat16).to(DEV)
    dy = (torch.randn(B * T, S, C, generator=g) / 8).to(torch.bfloat16).to(DEV)

    def fwd_bwd():
        for p in tb.parameters():
            p.grad = None
        xg.grad = None
        y = tb(sb(xg, encoder_hidden_states=ctx_s), num_frames=T, encoder_hidden_states=ctx_t)
        y.backward(dy)

    def fwd():
        with torch.no_grad():
            tb(sb(x, encoder_hidden_states=ctx_s), num_frames=T, encoder_hidden_states=ctx_t)

    ms_f = timed(fwd, steps)
    ms_fb = timed(fwd_bwd, max(3, steps // 2))
    kt = KernelTimes()
    ops.EVENT_HOOK = kt.hook
    for _ in range(2):
        fwd_bwd()
    ops.EVENT_HOOK = None
    emit({"config": "train_fwd_bwd_L0_block_pair", "clips_per_gpu": B, "T": T, "resolution": "576x1024",
          "trainable": "temporal block (spatial block frozen, gradient flows through it to the input)",
          "trainable_params": n_trainable, "ms_fwd": ms_f, "ms_fwd_bwd": ms_fb,
          "clip_frames_per_s_fwd_bwd": B * T / (ms_fb * 1e-3), "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30,
          "kernels": kt.summary(2)})


def cfg_loop25(steps):
    # 16 UNet + 7 ControlNet TransformerSpatioTemporalModel instances per denoising step:
    # UNet: L0 x5, L1 x5, L2 x5, mid x1; ControlNet branch (down + mid): L0 x2, L1 x2, L2 x2, mid x1
    T, n_steps = 25, 25
    inst = {"L0": 7, "L1": 7, "L2": 7, "mid": 2}
    blocks = {lv: make_block(lv, 1, 2, T) for lv in inst}
    per_level = {}
    with torch.no_grad():
        for lv, (mods, x, step, _) in blocks.items():
            per_level[lv] = timed(lambda: step(x), max(3, steps // 2))

        def unet_step():
            for lv, n in inst.items():
                _, x, step, _ = blocks[lv]
                for _ in range(n):
                    step(x)

        ms_step = timed(unet_step, 3, warmup=1)
    total_ms = ms_step * n_steps
    emit({"config": "loop25_attention_path", "frames": T, "denoise_steps": n_steps, "instances_per_step": sum(inst.values()),
          "clips": 1, "cfg": 2, "ms_per_level_block": per_level, "ms_per_denoise_step": ms_step, "ms_per_clip": total_ms,
          "clip_frames_per_s": T / (total_ms * 1e-3),
          "note": "attention path only (the four processor calls of every block instance); resnets/convs/VAE are out of scope"})


def cfg_kernels(steps):
    mods, x, step, (BF, S, C, nseq) = make_block("L0", 1, 2, 14)
    kt = KernelTimes()
    with torch.no_grad():
        for _ in range(3):
            step(x)
        ops.EVENT_HOOK = kt.hook
        for _ in range(steps):
            step(x)
        ops.EVENT_HOOK = None
    emit({"config": "kernels_L0_forward", "per_entry_point": kt.summary(steps)})


def cfg_block(steps):
    """N1: one full L0 spatial + temporal transformer block pair (LayerNorms, attention, FeedForward, residuals) at the
    bench shape (1 clip, CFG 2, T = 14, 576x1024), through blocks.install_blocks."""
    from timelapse_video_interpolation_b200 import blocks
    from timelapse_video_interpolation_b200.modules import BasicTransformerBlock, TemporalBasicTransformerBlock
    C, heads, h, w = bench.LEVELS["L0"]
    S, T, B = h * w, 14, 2
    g = torch.Generator().manual_seed(3)

    def init(m):
        with torch.no_grad():
            for name, p in m.named_parameters():
                if "norm" in name:
                    p.fill_(1.0 if name.endswith("weight") else 0.0)
                else:
                    p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5))
        return m.to(DEV, torch.bfloat16).eval()

    sb, tb = init(BasicTransformerBlock(C, heads, 1024)), init(TemporalBasicTransformerBlock(C, C, heads, 1024))
    blocks.install_blocks(sb)
    blocks.install_blocks(tb)
    x = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(DEV)
    ctx_s = torch.randn(B * T, 2, 1024, generator=g).to(torch.bfloat16).to(DEV)
    ctx_t = torch.randn(B * S, 1, 1024, generator=g).to(torch.bfloat16).to(DEV)

    def step():
        y = sb(x, encoder_hidden_states=ctx_s)
        return tb(y, num_frames=T, encoder_hidden_states=ctx_t)

    with torch.no_grad():
        ms = timed(step, steps)
        kt = KernelTimes()
        ops.EVENT_HOOK = kt.hook
        for _ in range(3):
            step()
        ops.EVENT_HOOK = None
    N = B * T * S
    ff_flops = 3 * (2.0 * N * C * 8 * C + 2.0 * N * 4 * C * C)  # ff (spatial), ff_in + ff (temporal)
    fl = bench.algorithmic_flops("L0", 1)["total"] + ff_flops
    emit({"config": "block_L0_spatial_plus_temporal", "clips": 1, "cfg": 2, "T": T, "ms": ms, "clip_frames_per_s": T / (ms * 1e-3),
          "tflop": fl / 1e12, "tflops": fl / 1e12 / (ms * 1e-3), "ff_tflop": ff_flops / 1e12, "kernels": kt.summary(3),
          "note": "temporal block runs in the spatial token order (N2 fold): no permute copies"})


def cfg_model(steps):
    """The whole L0 TransformerSpatioTemporalModel.forward (GroupNorm -> proj_in -> spatial block -> +frame embedding ->
    temporal block -> AlphaBlender -> proj_out -> NCHW + residual) at the bench shape, NCHW in, NCHW out."""
    from timelapse_video_interpolation_b200 import spatio_temporal
    from timelapse_video_interpolation_b200.modules import TransformerSpatioTemporalModel
    C, heads, h, w = bench.LEVELS["L0"]
    T, B = 14, 2
    g = torch.Generator().manual_seed(5)
    m = TransformerSpatioTemporalModel(heads, 64, in_channels=C, num_layers=1, cross_attention_dim=1024)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("mix_factor"):
                p.fill_(0.0)
            elif "norm" in name:
                p.fill_(1.0 if name.endswith("weight") else 0.0)
            else:
                p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5))
    m = m.to(DEV, torch.bfloat16).eval()
    spatio_temporal.install_spatio_temporal(m)
    x = torch.randn(B * T, C, h, w, generator=g).to(torch.bfloat16).to(DEV)
    e = torch.randn(B * T, 1, 1024, generator=g).to(torch.bfloat16).to(DEV)
    ind = torch.zeros(B, T, device=DEV)

    def step():
        return m(x, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)[0]

    with torch.no_grad():
        ms = timed(step, steps)
        kt = KernelTimes()
        ops.EVENT_HOOK = kt.hook
        for _ in range(3):
            step()
        ops.EVENT_HOOK = None
    N = B * T * h * w
    fl = bench.algorithmic_flops("L0", 1)["total"] + 3 * (2.0 * N * C * 8 * C + 2.0 * N * 4 * C * C) + 2 * 2.0 * N * C * C
    emit({"config": "model_L0_TransformerSpatioTemporalModel", "clips": 1, "cfg": 2, "T": T, "ms": ms,
          "clip_frames_per_s": T / (ms * 1e-3), "tflop": fl / 1e12, "tflops": fl / 1e12 / (ms * 1e-3), "kernels": kt.summary(3)})


def _st_model(level, T, B, seed=5):
    from timelapse_video_interpolation_b200 import spatio_temporal
    from timelapse_video_interpolation_b200.modules import TransformerSpatioTemporalModel
    C, heads, h, w = bench.LEVELS[level]
    g = torch.Generator().manual_seed(seed)
    m = TransformerSpatioTemporalModel(heads, 64, in_channels=C, num_layers=1, cross_attention_dim=1024)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("mix_factor"):
                p.fill_(0.0)
            elif "norm" in name:
                p.fill_(1.0 if name.endswith("weight") else 0.0)
            else:
                p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5))
    m = m.to(DEV, torch.bfloat16).eval()
    spatio_temporal.install_spatio_temporal(m)
    x = torch.randn(B * T, C, h, w, generator=g).to(torch.bfloat16).to(DEV)
    e = torch.randn(B * T, 1, 1024, generator=g).to(torch.bfloat16).to(DEV)
    ind = torch.zeros(B, T, device=DEV)
    return m, x, e, ind


def cfg_levels(steps):
    """Config 3 shape (T = 25, CFG 2): the whole TransformerSpatioTemporalModel.forward at every UNet level, eager vs one
    CUDA graph per module (graphs.GraphedForward); 23 instances per denoising step (16 UNet + 7 ControlNet)."""
    from timelapse_video_interpolation_b200.graphs import GraphedForward
    T, B = 25, 2
    inst = {"L0": 7, "L1": 7, "L2": 7, "mid": 2}
    eager, graphed = {}, {}
    with torch.no_grad():
        for lv in inst:
            m, x, e, ind = _st_model(lv, T, B)
            kw = dict(encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)
            eager[lv] = timed(lambda: m(x, **kw), steps)
            gf = GraphedForward(m, x, **kw)
            y_ref = m(x, **kw)[0]
            assert torch.equal(gf(x, **kw)[0], y_ref)
            graphed[lv] = timed(lambda: gf(x, **kw), steps)
            del m, gf
            torch.cuda.empty_cache()
    step_eager = sum(eager[lv] * n for lv, n in inst.items())
    step_graph = sum(graphed[lv] * n for lv, n in inst.items())
    emit({"config": "levels_T25_TransformerSpatioTemporalModel", "frames": T, "cfg": 2, "instances_per_step": inst,
          "ms_eager": eager, "ms_graphed": graphed, "ms_per_denoise_step_eager": step_eager, "ms_per_denoise_step_graphed": step_graph,
          "clip_frames_per_s_25_steps_graphed": T / (25 * step_graph * 1e-3),
          "note": "transformer modules only (resnets / convs / VAE are out of scope); graphed = one cudaGraphLaunch per module"})


def cfg_farneback(steps):
    """N3: Farneback flow of 1080p BGR frame pairs (device-resident uint8), and OpenCV's CPU implementation beside it."""
    import numpy as np
    H, W = 1080, 1920
    rng = np.random.default_rng(0)
    base = torch.from_numpy(rng.random((H // 8 + 2, W // 8 + 2, 3)).astype(np.float32)).permute(2, 0, 1)[None]
    img = torch.nn.functional.interpolate(base, size=(H + 8, W + 8), mode="bicubic")[0].permute(1, 2, 0)
    img = (img.clamp(0, 1) * 255).to(torch.uint8)
    a, b = img[4:-4, 4:-4].contiguous(), img[2:-6, 7:-1].contiguous()
    da, db = a.to(DEV), b.to(DEV)
    out = torch.empty(H, W, 2, dtype=torch.float32, device=DEV)
    ms = timed(lambda: ops.farneback(da, db, out=out), steps)
    rec = {"config": "farneback_1080p_pair", "ms_per_pair": ms, "pairs_per_s": 1e3 / ms, "median_flow": [float(out[..., 0].median()), float(out[..., 1].median())]}
    clip = torch.stack([torch.roll(da, (i, 2 * i), dims=(0, 1)) for i in range(14)])          # a 14-frame clip, 13 pairs
    cflows = torch.empty(13, H, W, 2, dtype=torch.float32, device=DEV)
    ms_clip = timed(lambda: ops.farneback_clip(clip, out=cflows), max(3, steps // 2))
    rec["clip14_ms_per_pair"] = ms_clip / 13
    rec["clip14_pairs_per_s"] = 13e3 / ms_clip
    try:
        import cv2
        import time
        an, bn = a.numpy(), b.numpy()
        t0 = time.perf_counter()
        for _ in range(3):
            ga, gb = cv2.cvtColor(an, cv2.COLOR_BGR2GRAY), cv2.cvtColor(bn, cv2.COLOR_BGR2GRAY)
            ref = cv2.calcOpticalFlowFarneback(ga, gb, None, 0.5, 3, 15, 3, 5, 1.2, 0)
        rec["opencv_cpu_ms_per_pair"] = (time.perf_counter() - t0) / 3 * 1e3
        rec["opencv_threads"] = cv2.getNumThreads()
        rec["max_abs_diff_vs_opencv"] = float(np.abs(out.cpu().numpy() - ref).max())
    except ImportError:
        rec["opencv_cpu_ms_per_pair"] = None
    emit(rec)


# ---------------------------------------------------------------------------------------------------
# bench.py's extra keys (`secondary`, `unet_step`, `train_step`): short runs after its timed regions, rank 0 / N = 1
# ---------------------------------------------------------------------------------------------------
def _frac(nbytes, ms, hbm):
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"ms": ms, "algorithmic_bytes": int(nbytes), "gbs": gbs, "frac": gbs / hbm, "bound": "hbm"}


def _secondary(pk, steps=5):
    hbm = pk["hbm"]
    sec = {"peak_hbm_gbs": hbm, "peak_source": pk["src"]}
    # config 1: slerp, 4096 clips, T = 14
    n, T, clips = 4 * 72 * 128, 14, 4096
    g = torch.Generator().manual_seed(0)
    v0 = torch.randn(clips, 4, 72, 128, generator=g).to(DEV)
    v1 = torch.randn(clips, 4, 72, 128, generator=g).to(DEV)
    out = torch.empty(clips, T, 4, 72, 128, device=DEV)
    ms = timed(lambda: ops.slerp_f32(v0, v1, T, 0.9995, out=out), steps)
    sec["slerp_4096clips_T14"] = dict(_frac((2 + T) * n * 4 * clips, ms, hbm), clip_frames_per_s=clips * T / (ms * 1e-3))
    del v0, v1, out
    torch.cuda.empty_cache()
    # config 4: flow cosine + backward warp / consistency, 32 pairs of 1080p
    H, W, B = 1080, 1920, 32
    g = torch.Generator().manual_seed(B)
    img = torch.rand(B, 3, H, W, generator=g).to(DEV)
    coarse = torch.randn(B, 2, H // 30, W // 30, generator=g).to(DEV) * 5
    f = torch.nn.functional.interpolate(coarse, size=(H, W), mode="bicubic", align_corners=False).contiguous()
    fb = (-f + torch.randn(B, 2, H, W, generator=g).to(DEV) * 0.2)
    fa_hwc, fb_hwc = f.permute(0, 2, 3, 1).contiguous(), fb.permute(0, 2, 3, 1).contiguous()
    warped, mask = torch.empty_like(img), torch.empty(B, H, W, dtype=torch.uint8, device=DEV)
    ms = timed(lambda: ops.backward_warp_f32(img, f, fb, 0.01, 0.5, warped=warped, mask=mask), steps)
    sec["flow_warp_check_32x1080p"] = dict(_frac(((3 + 2 + 2) * 4 + 3 * 4 + 1) * H * W * B, ms, hbm), pairs_per_s=B / (ms * 1e-3))
    ms = timed(lambda: ops.flow_cosine_f32(fa_hwc, fb_hwc, True), steps)
    sec["flow_cosine_32x1080p"] = dict(_frac(2 * 2 * 4 * H * W * B, ms, hbm), pairs_per_s=B / (ms * 1e-3))
    del img, coarse, f, fb, fa_hwc, fb_hwc, warped, mask
    torch.cuda.empty_cache()
    # N3: Farneback, one 1080p pair and a 14-frame clip (13 pairs)
    import numpy as np
    rng = np.random.default_rng(0)
    base = torch.from_numpy(rng.random((H // 8 + 2, W // 8 + 2, 3)).astype(np.float32)).permute(2, 0, 1)[None]
    im = torch.nn.functional.interpolate(base, size=(H + 8, W + 8), mode="bicubic")[0].permute(1, 2, 0)
    im = (im.clamp(0, 1) * 255).to(torch.uint8)
    da, db = im[4:-4, 4:-4].contiguous().to(DEV), im[2:-6, 7:-1].contiguous().to(DEV)
    fo = torch.empty(H, W, 2, dtype=torch.float32, device=DEV)
    ms = timed(lambda: ops.farneback(da, db, out=fo), steps)
    clip = torch.stack([torch.roll(da, (i, 2 * i), dims=(0, 1)) for i in range(14)])
    cflows = torch.empty(13, H, W, 2, dtype=torch.float32, device=DEV)
    ms_clip = timed(lambda: ops.farneback_clip(clip, out=cflows), 3)
    sec["farneback_1080p"] = {"ms_per_pair": ms, "pairs_per_s": 1e3 / ms, "clip14_ms_per_pair": ms_clip / 13,
                              "algorithmic_bytes": int(FARNEBACK_BYTES_1080P), "gbs": FARNEBACK_BYTES_1080P / (ms * 1e-3) / 1e9,
                              "frac": FARNEBACK_BYTES_1080P / (ms * 1e-3) / 1e9 / hbm, "bound": "hbm"}
    del da, db, fo, clip, cflows
    torch.cuda.empty_cache()
    # the streaming attention cores at the L0 bench shape (N = 258 048 tokens, C = 320, 5 heads); ~0.1 ms kernels: timed as
    # CUDA-graph replays so that the number is the kernels', not the Python call's (sec["small_kernel_timing"])
    sec["small_kernel_timing"] = "cuda-graph replay, 10 launches x 5"
    C, heads, S, Tn, Bc = 320, 5, 9216, 14, 2
    N, nseq = Bc * Tn * S, Bc * S
    g = torch.Generator().manual_seed(4)
    qkv = torch.randn(N, 3 * C, generator=g).to(torch.bfloat16).to(DEV)
    q, k, v = qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:]
    o = torch.empty(N, C, dtype=torch.bfloat16, device=DEV)
    do = torch.randn(N, C, generator=g).to(torch.bfloat16).to(DEV)
    scale = 0.125
    ms = timed_graph(lambda: ops.attn_temporal(q, k, v, nseq, heads, Tn, Tn, scale, out=o))
    sec["temporal_core_fwd_L0"] = _frac(4 * N * C * 2, ms, hbm)
    dqkv = torch.empty_like(qkv)
    ms = timed_graph(lambda: ops.attn_temporal_bwd(q, k, v, do, nseq, heads, Tn, Tn, scale, dq=dqkv[:, :C], dk=dqkv[:, C:2 * C],
                                                   dv=dqkv[:, 2 * C:]))
    sec["temporal_core_bwd_L0"] = _frac(7 * N * C * 2, ms, hbm)
    del dqkv
    # few keys, L = 2: spatial cross-attention (one context group per frame)
    L = 2
    kv = torch.randn(Bc * Tn * L, 2 * C, generator=g).to(torch.bfloat16).to(DEV)
    qc = qkv[:, :C].contiguous()
    ms = timed_graph(lambda: ops.attn_fewkeys(qc, kv[:, :C], kv[:, C:], N, S, heads, L, scale, out=o))
    sec["fewkeys_L2_fwd_L0"] = _frac(2 * N * C * 2, ms, hbm)
    dq = torch.empty(N, C, dtype=torch.bfloat16, device=DEV)
    dkv = torch.empty(Bc * Tn * L, 2 * C, dtype=torch.bfloat16, device=DEV)
    ms = timed_graph(lambda: ops.attn_fewkeys_bwd(qc, kv[:, :C], kv[:, C:], do, N, S, heads, L, scale, dq=dq, dk=dkv[:, :C], dv=dkv[:, C:]))
    sec["fewkeys_L2_bwd_L0"] = _frac(3 * N * C * 2, ms, hbm)
    del qkv, o, do, kv, qc, dq, dkv
    torch.cuda.empty_cache()
    return sec


# Farneback at 1080p, (pyr_scale 0.5, 3 levels, winsize 15, 3 iterations, poly_n 5): algorithmic HBM bytes per pair of the
# unfused pass structure: levels 1, 1/4, 1/16 of 2.07 Mpx ~ 2.72 Mpx-equivalents, per pixel and level:
# blur + resize 2 frames (16 B), polynomial expansion 2 x (4 read + 20 written), per iteration matrix update
# (5 R1 gathered + 20 + 8 read, 20 written), box sums (20 read, 20 written), solve (20 read, 8 written) -> ~ 740 B.
FARNEBACK_BYTES_1080P = 2.72e6 * 740


def _unet_step(steps=5):
    """Whole TransformerSpatioTemporalModel.forward per UNet level, CFG 2, 1 clip: config 2 topology (T = 14, 16 instances per
    UNet step: 5/5/5/1) and config 3 topology (T = 25, 23 instances incl. the ControlNet branch: 7/7/7/2)."""
    out = {}
    for name, T, inst in (("T14_unet16", 14, {"L0": 5, "L1": 5, "L2": 5, "mid": 1}),
                          ("T25_unet16_controlnet7", 25, {"L0": 7, "L1": 7, "L2": 7, "mid": 2})):
        per, tfl = {}, {}
        with torch.no_grad():
            for lv in inst:
                m, x, e, ind = _st_model(lv, T, 2)
                kw = dict(encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)
                per[lv] = timed(lambda: m(x, **kw), steps)
                C, heads, h, w = bench.LEVELS[lv]
                S, BF = h * w, 2 * T
                N = BF * S
                fl = (4.0 * BF * heads * S * S * 64 + 4.0 * 2 * S * heads * T * T * 64      # attention cores
                      + 2.0 * N * C * C * (4 + 2 + 4 + 1) + 2 * 2.0 * N * C * C            # q/k/v/o projections, proj_in/out
                      + 3 * (2.0 * N * C * 8 * C + 2.0 * N * 4 * C * C))                    # three GEGLU feed-forwards
                tfl[lv] = fl / 1e12 / (per[lv] * 1e-3)
                del m, x, e, ind
                torch.cuda.empty_cache()
        ms_step = sum(per[lv] * n for lv, n in inst.items())
        out[name] = {"frames": T, "cfg_batch": 2, "instances": inst, "ms_per_module": per, "tflops_per_module": tfl,
                     "ms_per_unet_step": ms_step, "clip_frames_per_s_per_step": T / (ms_step * 1e-3),
                     "note": "transformer modules only; resnets / convs / VAE / scheduler are out of scope (SURVEY 8)"}
    return out


def _train_step(steps=3):
    """Config 5 at one GPU: forward + backward through the L0 block pair, spatial block frozen (gradient flows through it),
    temporal block trainable (src/train_decouples.py:707-720), per_gpu_batch_size 2, T = 14, no CFG, 576x1024."""
    from timelapse_video_interpolation_b200 import blocks
    from timelapse_video_interpolation_b200.modules import BasicTransformerBlock, TemporalBasicTransformerBlock
    C, heads, h, w = bench.LEVELS["L0"]
    S, T, B = h * w, 14, 2
    g = torch.Generator().manual_seed(3)

    def init(m, trainable):
        with torch.no_grad():
            for name, p in m.named_parameters():
                if "norm" in name:
                    p.fill_(1.0 if name.endswith("weight") else 0.0)
                else:
                    p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5))
                p.requires_grad_(trainable)
        return m.to(DEV, torch.bfloat16)

    sb, tb = init(BasicTransformerBlock(C, heads, 1024), False), init(TemporalBasicTransformerBlock(C, C, heads, 1024), True)
    blocks.install_blocks(sb)
    blocks.install_blocks(tb)
    x = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(DEV)
    xg = x.clone().requires_grad_(True)
    ctx_s = torch.randn(B * T, 2, 1024, generator=g).to(torch.bfloat16).to(DEV)
    ctx_t = torch.randn(B * S, 1, 1024, generator=g).to(torch.bfloat16).to(DEV)
    dy = (torch.randn(B * T, S, C, generator=g) / 8).to(torch.bfloat16).to(DEV)

    def fwd_bwd():
        for p in tb.parameters():
            p.grad = None
        xg.grad = None
        tb(sb(xg, encoder_hidden_states=ctx_s), num_frames=T, encoder_hidden_states=ctx_t).backward(dy)

    def fwd():
        with torch.no_grad():
            tb(sb(x, encoder_hidden_states=ctx_s), num_frames=T, encoder_hidden_states=ctx_t)

    ms_f = timed(fwd, steps)
    ms_fb = timed(fwd_bwd, steps)
    kt = KernelTimes()
    ops.EVENT_HOOK = kt.hook
    fwd_bwd()
    ops.EVENT_HOOK = None
    ks = kt.summary(1)
    n_trainable = sum(p.numel() for p in tb.parameters())
    del sb, tb, x, xg, dy
    torch.cuda.empty_cache()
    module = _train_step_module(B, T, C, heads, h, w, g, steps)
    return {"config": "train_fwd_bwd_L0_block_pair_576x1024", "clips_per_gpu": B, "frames": T, "cfg_batch": 1, "module": module,
            "trainable": "temporal block (spatial block frozen, gradient flows through it to the input)",
            "trainable_params": n_trainable, "ms_fwd": ms_f, "ms_fwd_bwd": ms_fb,
            "clip_frames_per_s_fwd_bwd": B * T / (ms_fb * 1e-3),
            "ms_spatial_attention_bwd": ks.get("attn_spatial_bwd", {}).get("ms_per_step"),
            "ms_by_entry_point": {k: round(v["ms_per_step"], 4) for k, v in ks.items()}}


def _train_step_module(B, T, C, heads, h, w, g, steps):
    """The same step through the MODULE-level swap (`install_spatio_temporal`): TransformerSpatioTemporalModel.forward at L0
    (GroupNorm, proj_in, the block pair, AlphaBlender, proj_out, residual) trained the way src/train.py does — temporal
    block + time_mixer trainable, everything else frozen, plain and with the spatial block under
    torch.utils.checkpoint(use_reentrant=False) (src/models/transformer_temporal.py:321-327)."""
    from timelapse_video_interpolation_b200 import spatio_temporal
    from timelapse_video_interpolation_b200.modules import TransformerSpatioTemporalModel
    m = TransformerSpatioTemporalModel(heads, C // heads, in_channels=C, num_layers=1, cross_attention_dim=1024)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("mix_factor"):
                p.fill_(0.5)
            elif "norm" in name:
                p.fill_(1.0 if name.endswith("weight") else 0.0)
            else:
                p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5))
            p.requires_grad_("temporal_transformer_blocks" in name or "time_mixer" in name)
    m = m.to(DEV, torch.bfloat16).train()
    spatio_temporal.install_spatio_temporal(m)
    x = torch.randn(B * T, C, h, w, generator=g).to(torch.bfloat16).to(DEV).requires_grad_(True)
    ctx = torch.randn(B * T, 1, 1024, generator=g).to(torch.bfloat16).to(DEV)
    ind = torch.zeros(B, T, device=DEV)
    dy = (torch.randn(B * T, C, h, w, generator=g) / 8).to(torch.bfloat16).to(DEV)

    def fwd_bwd():
        for p in m.parameters():
            p.grad = None
        x.grad = None
        m(x, encoder_hidden_states=ctx, image_only_indicator=ind, return_dict=False)[0].backward(dy)

    out = {"config": "train_fwd_bwd_L0_module_576x1024", "trainable_params": sum(p.numel() for p in m.parameters() if p.requires_grad)}
    for key, ck in (("ms_fwd_bwd", False), ("ms_fwd_bwd_checkpointed", True)):
        m.gradient_checkpointing = ck
        out[key] = timed(fwd_bwd, steps)
    out["clip_frames_per_s_fwd_bwd"] = B * T / (out["ms_fwd_bwd"] * 1e-3)
    return out


def driver_extras(dev, pk):
    """The extra keys of bench.py's JSON line (VERDICT r1 task 4): every config the driver cannot otherwise see."""
    global DEV
    DEV = dev
    out = {"secondary": _secondary(pk)}
    torch.cuda.empty_cache()
    out["unet_step"] = _unet_step()
    torch.cuda.empty_cache()
    out["train_step"] = _train_step()
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="slerp,flow,train,train_block,loop25,kernels,block,model,levels,farneback")
    ap.add_argument("--steps", type=int, default=10)
    args = ap.parse_args()
    torch.cuda.set_device(0)
    emit({"config": "env", "device": torch.cuda.get_device_name(0), "libtvi": tvi.version(), "peaks": PK})
    table = {"slerp": cfg_slerp, "flow": cfg_flow, "train": cfg_train, "loop25": cfg_loop25, "kernels": cfg_kernels, "block": cfg_block, "model": cfg_model, "levels": cfg_levels, "farneback": cfg_farneback, "train_block": cfg_train_block}
    for name in args.only.split(","):
        table[name.strip()](args.steps)


if __name__ == "__main__":
    main()
