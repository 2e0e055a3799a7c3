#!/usr/bin/env python
"""Config 5 (BASELINE.json): forward + backward of the L0 transformer block pair under DistributedDataParallel, one process
per GPU, NCCL gradient all-reduce (DDP's own - no kernel on this path is directly followed by a collective).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/train_ddp_bench.py [--res 576x1024|512x320] [--param-dtype bf16|fp32] [--steps K]

Mirrors src/train_decouples.py: the spatial block is frozen (its gradient still flows to the block input, :707-720), the
temporal block is trainable, per_gpu_batch_size 2, T = 14, no CFG (accelerate_config/acc_0_1_2_3.yaml, train_decouples.sh).
`--res 512x320 --param-dtype fp32` is the reference's own operating point (latent 64x40, fp32 master parameters; the
kernels compute in bf16).  Prints ONE JSON line from rank 0: step ms under DDP, step ms without the gradient exchange
(`no_sync`), the stand-alone all-reduce of the same gradient bytes, the overlap these imply (when the exchange is long enough to
measure against step-to-step noise), and whether every rank ends
the step with bit-identical gradients.
"""
import argparse
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch.nn.parallel import DistributedDataParallel as DDP  # noqa: E402

from timelapse_video_interpolation_b200 import blocks  # noqa: E402
from timelapse_video_interpolation_b200.modules import BasicTransformerBlock, TemporalBasicTransformerBlock  # noqa: E402
from timelapse_video_interpolation_b200.processor import TviAttnProcessor  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--res", default="576x1024")
ap.add_argument("--param-dtype", default="bf16", choices=["bf16", "fp32"])
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--clips", type=int, default=2)
ap.add_argument("--graph", action="store_true", help="also capture forward + backward of the (un-wrapped) module into ONE CUDA "
                "graph and time its replay (+ one all-reduce of the flattened gradients when world > 1)")
ap.add_argument("--module", action="store_true", help="train through the module-level swap (TransformerSpatioTemporalModel.forward, "
                "NCHW in / out: GroupNorm, proj_in, block pair, AlphaBlender, proj_out) instead of the bare block pair")
args = ap.parse_args()

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)

Hpx, Wpx = (int(v) for v in args.res.split("x"))
h, w = Hpx // 8, Wpx // 8
C, heads, T, B = 320, 5, 14, args.clips
S = h * w
pdt = torch.bfloat16 if args.param_dtype == "bf16" else torch.float32


class Pair(torch.nn.Module):
    def __init__(self):
        super().__init__()
        g = torch.Generator().manual_seed(3)  # identical initial weights on every rank
        self.spatial = BasicTransformerBlock(C, heads, 1024)
        self.temporal = TemporalBasicTransformerBlock(C, C, heads, 1024)
        with torch.no_grad():
            for m, trainable in ((self.spatial, False), (self.temporal, True)):
                for name, p in m.named_parameters():
                    if "norm" in name:
                        p.fill_(1.0 if name.endswith("weight") else 0.0)
                    else:
                        p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5))
                    p.requires_grad_(trainable)

    def forward(self, x, ctx_s, ctx_t):
        return self.temporal(self.spatial(x, encoder_hidden_states=ctx_s), num_frames=T, encoder_hidden_states=ctx_t)


class Module(torch.nn.Module):
    """src/train.py's unit: the whole TransformerSpatioTemporalModel, temporal block + time_mixer trainable."""

    def __init__(self):
        super().__init__()
        from timelapse_video_interpolation_b200.modules import TransformerSpatioTemporalModel
        g = torch.Generator().manual_seed(3)
        self.m = TransformerSpatioTemporalModel(heads, C // heads, in_channels=C, num_layers=1, cross_attention_dim=1024)
        with torch.no_grad():
            for name, p in self.m.named_parameters():
                if name.endswith("mix_factor"):
                    p.fill_(0.5)
                elif "norm" in name:
                    p.fill_(1.0 if name.endswith("weight") else 0.0)
                else:
                    p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * (p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5))
                p.requires_grad_("temporal_transformer_blocks" in name or "time_mixer" in name)

    def forward(self, x, ctx_s, ind):
        return self.m(x, encoder_hidden_states=ctx_s, image_only_indicator=ind, return_dict=False)[0]


proc = TviAttnProcessor(cast_inputs=True)
g = torch.Generator().manual_seed(100 + rank)  # different clips on every rank
if args.module:
    from timelapse_video_interpolation_b200 import spatio_temporal
    pair = Module().to(dev, pdt).train()
    spatio_temporal.install_spatio_temporal(pair, processor=proc)
    x = torch.randn(B * T, C, h, w, generator=g).to(torch.bfloat16).to(dev).requires_grad_(True)
    ctx_s = torch.randn(B * T, 1, 1024, generator=g).to(torch.bfloat16).to(dev)
    ctx_t = torch.zeros(B, T, device=dev)                      # image_only_indicator
    dy = (torch.randn(B * T, C, h, w, generator=g) / 8).to(torch.bfloat16).to(dev)
else:
    pair = Pair().to(dev, pdt)
    blocks.install_blocks(pair, processor=proc)
    x = torch.randn(B * T, S, C, generator=g).to(torch.bfloat16).to(dev).requires_grad_(True)
    ctx_s = torch.randn(B * T, 2, 1024, generator=g).to(torch.bfloat16).to(dev)
    ctx_t = torch.randn(B * S, 1, 1024, generator=g).to(torch.bfloat16).to(dev)
    dy = (torch.randn(B * T, S, C, generator=g) / 8).to(torch.bfloat16).to(dev)


def timed(fn, n):
    for _ in range(2):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / n], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)  # time of the slowest rank
    return float(t)


# Every libtvi entry is capture-safe (caller's stream, no host sync; torch's allocations come from the graph's pool), so the
# whole forward + backward records into one CUDA graph: at the reference's 512x320 operating point the eager step is bound by
# the Python / launch path (~190 launches of a few microseconds each), not by the GPU.  Measured before DDP wraps the module
# (its reducer hooks expect DDP.forward); data parallelism is then ONE all-reduce of the flattened gradients after the replay.
graph_ms = graph_same = None
if args.graph:
    gparams = [p for p in pair.parameters() if p.requires_grad]

    def eager():
        pair.zero_grad(set_to_none=True)
        x.grad = None
        pair(x, ctx_s, ctx_t).backward(dy)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            eager()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    want = [p.grad.clone() for p in gparams] + [x.grad.clone()]
    pair.zero_grad(set_to_none=True)
    x.grad = None
    cg = torch.cuda.CUDAGraph()
    with torch.cuda.graph(cg):
        pair(x, ctx_s, ctx_t).backward(dy)   # .grad tensors are created inside the capture: static across replays

    def gstep():
        cg.replay()
        if world > 1:
            flatg = torch.cat([p.grad.flatten() for p in gparams])
            dist.all_reduce(flatg)

    graph_ms = timed(gstep, args.steps)
    cg.replay()
    torch.cuda.synchronize()
    graph_same = all(torch.equal(a, b) for a, b in zip(want, [p.grad for p in gparams] + [x.grad]))
    del cg
    pair.zero_grad(set_to_none=True)
    x.grad = None
model = DDP(pair, device_ids=[local], gradient_as_bucket_view=True)


def step(sync=True):
    model.zero_grad(set_to_none=True)
    x.grad = None
    if sync:
        model(x, ctx_s, ctx_t).backward(dy)
    else:
        with model.no_sync():
            model(x, ctx_s, ctx_t).backward(dy)


ms_ddp = timed(lambda: step(True), args.steps)
ms_local = timed(lambda: step(False), args.steps)
params = [p for p in pair.parameters() if p.requires_grad]
nbytes = sum(p.numel() * p.element_size() for p in params)
flat = torch.zeros(nbytes // params[0].element_size(), dtype=params[0].dtype, device=dev)
ms_allreduce = timed(lambda: dist.all_reduce(flat), 20)
# identical gradients everywhere after a synchronised step
step(True)
torch.cuda.synchronize()
grads = torch.cat([p.grad.float().flatten() for p in params])
finite = bool(torch.isfinite(grads).all())
ref = grads.clone()
dist.broadcast(ref, 0)
same = torch.tensor([1 if torch.equal(ref, grads) else 0], device=dev)
dist.all_reduce(same, op=dist.ReduceOp.MIN)
# what DDP adds to a step: bucket bookkeeping + whatever part of the all-reduce is not hidden under the backward
ddp_extra = ms_ddp - ms_local
if rank == 0:
    print(json.dumps({
        "config": "train_ddp_L0_module" if args.module else "train_ddp_L0_block_pair", "n_gpus": world, "resolution": args.res, "tokens_per_frame": S, "frames": T,
        "clips_per_gpu": B, "param_dtype": args.param_dtype, "trainable_params": sum(p.numel() for p in params),
        "gradient_bytes": nbytes, "ms_step_ddp": ms_ddp, "ms_step_no_sync": ms_local, "ms_allreduce_alone": ms_allreduce,
        "ms_ddp_minus_no_sync": ddp_extra,
        "ms_step_cuda_graph": graph_ms, "cuda_graph_gradients_identical_to_eager": graph_same,
        # meaningful only when the exchange itself is long against the run-to-run noise of a step (a few 0.1 ms)
        "overlap_fraction": (max(0.0, min(1.0, 1.0 - max(0.0, ddp_extra) / ms_allreduce)) if world > 1 and ms_allreduce >= 1.0 else None),
        "clip_frames_per_s": world * B * T / (ms_ddp * 1e-3), "gradients_identical_on_all_ranks": bool(int(same)), "finite": finite,
        "timing": "CUDA events per rank, max over ranks"}), flush=True)
dist.barrier()
dist.destroy_process_group()
