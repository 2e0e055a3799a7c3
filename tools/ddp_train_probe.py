"""Config 5 probe: forward + backward of one L1 block's four attention calls under DistributedDataParallel
(NCCL gradient all-reduce), one process per GPU: checks that every rank ends with identical, finite gradients
and prints step time.  torchrun --nproc-per-node N tools/ddp_train_probe.py"""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel as DDP
from timelapse_video_interpolation_b200.modules import Attention
from timelapse_video_interpolation_b200.processor import TviAttnProcessor

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
C, heads, S, T, clips = 640, 10, 36 * 64, 14, 2


class Block(torch.nn.Module):
    def __init__(self):
        super().__init__()
        proc = TviAttnProcessor()
        torch.manual_seed(0)  # same initial weights on every rank
        self.attn1_s, self.attn2_s = Attention(C, None, heads, processor=proc), Attention(C, 1024, heads, processor=proc)
        self.attn1_t, self.attn2_t = Attention(C, None, heads, processor=proc), Attention(C, 1024, heads, processor=proc)
        for m in (self.attn1_s, self.attn2_s):  # train_decouples: only temporal_transformer_block* is trainable
            for p in m.parameters():
                p.requires_grad_(False)

    def forward(self, x, cs, ct):
        y = x + self.attn1_s(x)
        y = y + self.attn2_s(y, encoder_hidden_states=cs)
        v = y.view(clips * S, T, C)
        v = v + self.attn1_t(v)
        return v + self.attn2_t(v, encoder_hidden_states=ct)


model = DDP(Block().to(dev, torch.bfloat16), device_ids=[local])
g = torch.Generator().manual_seed(100 + rank)  # different clips per rank
x = torch.randn(clips * T, S, C, generator=g).to(torch.bfloat16).to(dev).requires_grad_(True)
cs = torch.randn(clips * T, 2, 1024, generator=g).to(torch.bfloat16).to(dev)
ct = torch.randn(clips * S, 1, 1024, generator=g).to(torch.bfloat16).to(dev)
times = []
for it in range(5):
    model.zero_grad(set_to_none=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loss = model(x, cs, ct).float().pow(2).mean()
    loss.backward()
    e1.record()
    torch.cuda.synchronize()
    times.append(e0.elapsed_time(e1))
flat = torch.cat([p.grad.float().flatten() for p in model.parameters() if p.grad is not None])
assert bool(torch.isfinite(flat).all())
ref = flat.clone()
dist.broadcast(ref, 0)
same = bool(torch.equal(ref, flat))
t = torch.tensor([min(times[1:])], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"ddp probe: world {world}, {flat.numel()} gradient elements, identical on all ranks: {same}, "
          f"fwd+bwd+allreduce {t.item():.2f} ms/step, {world * clips * T / (t.item() * 1e-3):.0f} clip-frames/s", flush=True)
assert same
dist.destroy_process_group()
