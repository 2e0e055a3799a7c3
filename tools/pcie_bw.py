import torch
dev = torch.device("cuda:0")
n = 200 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device=dev)
h2 = torch.empty(n, dtype=torch.uint8).pin_memory(); d2 = torch.empty(n, dtype=torch.uint8, device=dev)
def t(fn, it=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
ms = t(lambda: d.copy_(h, non_blocking=True)); print("H2D", ms, n / ms / 1e6, "GB/s")
ms = t(lambda: h.copy_(d, non_blocking=True)); print("D2H", ms, n / ms / 1e6, "GB/s")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def both():
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur); s2.wait_stream(cur)
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
    cur.wait_stream(s1); cur.wait_stream(s2)
ms = t(both); print("both", ms, 2 * n / ms / 1e6, "GB/s total")
