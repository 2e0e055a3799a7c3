"""HBM bandwidth for write-only and 1:3 read:write traffic (the QKV projection's mix) vs a 1:1 copy."""
import torch
dev = torch.device("cuda:0")
def t(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
n = 256 << 20  # bf16 elements: 512 MB
a = torch.empty(n, dtype=torch.bfloat16, device=dev).normal_()
b = torch.empty(n, dtype=torch.bfloat16, device=dev)
c = torch.empty(3, n, dtype=torch.bfloat16, device=dev)
ms = t(lambda: b.copy_(a)); print(f"copy 1:1        {2 * n * 2 / ms / 1e6:8.0f} GB/s")
ms = t(lambda: b.fill_(1.0)); print(f"write only      {n * 2 / ms / 1e6:8.0f} GB/s")
ms = t(lambda: c.copy_(a.expand(3, n))); print(f"read 1 : write 3 {4 * n * 2 / ms / 1e6:8.0f} GB/s")
ms = t(lambda: a.sum()); print(f"read only       {n * 2 / ms / 1e6:8.0f} GB/s")
