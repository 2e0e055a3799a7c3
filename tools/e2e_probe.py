"""Probe: where does the pipelined host<->device loop of bench.py lose its overlap?"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import bench
from timelapse_video_interpolation_b200.processor import TviAttnProcessor
dev = torch.device("cuda:0")
wl = bench.Workload("L0", 1, dev, torch.bfloat16, TviAttnProcessor)
x_pin, cs_pin, ct_pin = wl.x_host.pin_memory(), wl.ctx_s_host.pin_memory(), wl.ctx_t_host.pin_memory()
x, cs, ct = x_pin.to(dev), cs_pin.to(dev), ct_pin.to(dev)
out_pins = [torch.empty(wl.nseq, 14, wl.C, dtype=torch.bfloat16).pin_memory() for _ in range(2)]
dbuf = [(torch.empty_like(x), torch.empty_like(cs), torch.empty_like(ct)) for _ in range(2)]
cur = torch.cuda.current_stream()
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
ev_in = [torch.cuda.Event() for _ in range(2)]
ev_free = [torch.cuda.Event() for _ in range(2)]

def run(n, h2d, d2h, compute=True):
    for k in range(n):
        b = k & 1
        if h2d:
            with torch.cuda.stream(s_in):
                if k >= 2:
                    s_in.wait_event(ev_free[b])
                dbuf[b][0].copy_(x_pin, non_blocking=True)
                dbuf[b][1].copy_(cs_pin, non_blocking=True)
                dbuf[b][2].copy_(ct_pin, non_blocking=True)
                ev_in[b].record(s_in)
            cur.wait_event(ev_in[b])
        y = wl.step(*dbuf[b]) if compute else dbuf[b][0].view(wl.nseq, 14, wl.C)
        ev_free[b].record(cur)
        if d2h:
            s_out.wait_stream(cur)
            with torch.cuda.stream(s_out):
                out_pins[b].copy_(y, non_blocking=True)
            y.record_stream(s_out)
    cur.wait_stream(s_out)

def t(name, **kw):
    with torch.no_grad():
        run(3, **kw)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(10, **kw)
        e1.record()
        torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 10:.3f} ms/step", flush=True)

for b in range(2):
    for d, s in zip(dbuf[b], (x_pin, cs_pin, ct_pin)):
        d.copy_(s)
t("compute only        ", h2d=False, d2h=False)
t("h2d only            ", h2d=True, d2h=False, compute=False)
t("d2h only            ", h2d=False, d2h=True, compute=False)
t("h2d + d2h, no kernels", h2d=True, d2h=True, compute=False)
t("h2d + compute       ", h2d=True, d2h=False)
t("compute + d2h       ", h2d=False, d2h=True)
t("h2d + compute + d2h ", h2d=True, d2h=True)
