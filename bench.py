#!/usr/bin/env python
"""bench.py — clip-frames/sec of the SVD-XT attention hot path on B200 (BASELINE.json metric).

Workload (BASELINE.json configs[1]): the four attention-processor calls of ONE
L0 `TransformerSpatioTemporalModel` instance of SVD-XT at 576x1024 (latent 72x128 -> S = 9216 tokens,
C = 320, 5 heads of 64), T = 14 frames, CFG batch 2 (= 1 clip), bf16, through the AttnProcessor plugin:

    y1 = attn1_spatial(x[28, 9216, 320])                 fused-QKV GEMM -> flash core -> out GEMM
    y2 = attn2_spatial(y1, ctx_spatial[28, 2, 1024])     KV GEMM, Q GEMM -> 2-key core -> out GEMM
    y3 = attn1_temporal(y2 viewed as [18432, 14, 320])   fused-QKV GEMM -> short-sequence core -> out GEMM
    y4 = attn2_temporal(y3, ctx_temporal[18432, 1, 1024]) KV GEMM -> out GEMM on 18432 rows -> broadcast

A "step" is one such pass for `--clips` clips per GPU (default 1); clip-frames = clips x T (CFG duplicates
are not counted).  Multi-GPU: one process per GPU, clips sharded, no collective on the data path
(weak scaling); time = max over ranks.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl tvi|reference] [--level L0|L1|L2|mid]
"""
from __future__ import annotations

import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

LEVELS = {  # name: (C, heads, h, w)
    "L0": (320, 5, 72, 128), "L1": (640, 10, 36, 64), "L2": (1280, 20, 18, 32), "mid": (1280, 20, 9, 16),
}
T_FRAMES = 14
CFG = 2
L_SPATIAL, L_TEMPORAL, CTX_DIM = 2, 1, 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="tvi", choices=["tvi", "reference"])
    ap.add_argument("--level", default="L0", choices=list(LEVELS))
    ap.add_argument("--clips", type=int, default=1, help="clips per GPU per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the comparator and the secondary / unet_step / train_step keys")
    return ap.parse_args()


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sust": d["bf16_tflops_sustained"], "src": "measured"}
    return {"hbm": 6650.0, "tf_burst": 1590.0, "tf_sust": 1400.0, "src": "fallback"}


def measured_traffic(level, clips):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of this very
    command (profiles/dominant_kernel_traffic.json, written by tools/ncu_summary.py), or None."""
    p = ROOT / "profiles" / "dominant_kernel_traffic.json"
    if not p.exists():
        return None, None
    d = json.loads(p.read_text())
    if d.get("workload") != workload_name(level, clips):
        return None, None
    return d.get("dram_bytes_per_launch"), d.get("source")


def workload_name(level, clips):
    C, heads, h, w = LEVELS[level]
    return f"svd_xt_{level}_block_attention_576x1024_T{T_FRAMES}_cfg{CFG}_clips{clips}"


def algorithmic_flops(level, clips):
    """FLOPs of one step (GEMMs 2MNK, cores 4*B*h*Sq*Sk*d), listed per kernel family."""
    C, heads, h, w = LEVELS[level]
    S = h * w
    BF = clips * CFG * T_FRAMES
    N = BF * S
    nseq = clips * CFG * S
    f = {}
    f["spatial_core"] = 4.0 * BF * heads * S * S * 64
    f["gemm_spatial_self"] = 2.0 * N * C * (3 * C) + 2.0 * N * C * C
    f["gemm_spatial_cross"] = 2.0 * N * C * C * 2 + 2.0 * BF * L_SPATIAL * CTX_DIM * 2 * C
    f["gemm_temporal_self"] = 2.0 * N * C * (3 * C) + 2.0 * N * C * C
    f["temporal_core"] = 4.0 * nseq * heads * T_FRAMES * T_FRAMES * 64
    f["gemm_temporal_cross"] = 2.0 * nseq * L_TEMPORAL * CTX_DIM * 2 * C + 2.0 * nseq * C * C
    f["total"] = sum(f.values())
    return f


def config_dict(level, clips, world):
    """`config` of the JSON line — the same dict in both arms (the reference arm adds `sample_fraction`)."""
    C, heads, h, w = LEVELS[level]
    fl = algorithmic_flops(level, clips)
    return {"workload": workload_name(level, clips), "level": level, "clips_per_gpu": clips,
            "frames": T_FRAMES, "cfg_batch": CFG, "tokens_per_frame": h * w, "channels": C, "heads": heads,
            "ctx_tokens": [L_SPATIAL, L_TEMPORAL], "parallelism": f"clip-sharded x{world}, no collective",
            "l2": "inputs (165 MB activations per call at L0) exceed the 126 MB L2; no explicit flush",
            "step_tflop": fl["total"] / 1e12}


def bind_rank_to_cores(local_rank, world):
    """One NUMA node, every GPU attached to CPUs 0-31 (nvidia-smi topo): give each rank its own slice of the cores
    BEFORE it allocates pinned memory, so that first-touch pages and the copy-engine doorbells of different ranks do
    not share cores.  Returns the core list (or None when the platform has no affinity API)."""
    if world <= 1 or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        avail = sorted(os.sched_getaffinity(0))
        per = max(1, len(avail) // world)
        mine = avail[local_rank * per:(local_rank + 1) * per] or avail
        os.sched_setaffinity(0, mine)
        return mine
    except OSError:
        return None


class ClockSampler:
    """Samples SM clocks / throttle reasons / power during the timed regions (B200_PROFILING.md's clocks line).

    NVML is queried in-process (the counters nvidia-smi prints), initialised BEFORE the warm-up: spawning nvidia-smi
    next to the timed region stalled kernel launches for tens of ms while its own NVML start-up held the driver
    (one run in three lost 30 % of `value` to it).  nvidia-smi is the fallback when pynvml is missing.
    """

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int, period_s: float = 0.02):
        self.index, self.period = index, period_s
        self.proc = self.nvml = self.handle = self.thread = None
        self.samples = []   # (sm_mhz, power_w, reason bits)
        self.lines = []
        self.active = threading.Event()
        self.done = threading.Event()
        self.smax = None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self._read()  # first query pays NVML's lazy set-up outside the timed region
            self.samples.clear()
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
        except Exception:
            self.nvml = None

    def _read(self):
        import pynvml
        t0 = time.perf_counter()
        sm = float(pynvml.nvmlDeviceGetClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        t1 = time.perf_counter()
        try:
            bits = int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        except Exception:
            bits = 0
        try:
            power = pynvml.nvmlDeviceGetPowerUsage(self.handle) / 1000.0
        except Exception:
            power = None
        t2 = time.perf_counter()
        self.query_ms_max = max(getattr(self, "query_ms_max", 0.0), (t2 - t0) * 1e3)
        self.clock_query_ms_max = max(getattr(self, "clock_query_ms_max", 0.0), (t1 - t0) * 1e3)
        self.samples.append((sm, power, bits))

    def _poll(self):
        while not self.done.is_set():
            if self.active.is_set():
                try:
                    self._read()
                except Exception:
                    pass
            time.sleep(self.period)

    def start(self):
        """Called right before a timed region (may be called again after `pause`)."""
        if os.environ.get("TVI_BENCH_NO_SAMPLER"):
            return
        if self.nvml is not None:
            self.active.set()
            return
        if self.proc is not None:
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def pause(self):
        self.active.clear()

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml is not None:
            self.active.clear()
            self.done.set()
            self.thread.join(timeout=2)
            sm = sorted(s[0] for s in self.samples)
            power = [s[1] for s in self.samples if s[1] is not None]
            bits = 0
            for s in self.samples:
                bits |= s[2]
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.smax,
                    "reasons": sorted(n for b, n in self.REASONS.items() if bits & b), "samples": len(sm),
                    "power_w_max": max(power) if power else None, "source": "nvml",
                    "query_ms_max": round(getattr(self, "query_ms_max", 0.0), 3),
                    "clock_query_ms_max": round(getattr(self, "clock_query_ms_max", 0.0), 3)}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons, power = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                smax = float(parts[1])
                power.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(power) if power else None, "source": "nvidia-smi"}


class Workload:
    """The four Attention modules of one block instance + synthetic operands (seeded)."""

    def __init__(self, level: str, clips: int, device, dtype, processor_factory):
        from timelapse_video_interpolation_b200.modules import Attention
        C, heads, h, w = LEVELS[level]
        self.C, self.heads, self.S = C, heads, h * w
        self.BF = clips * CFG * T_FRAMES
        self.nseq = clips * CFG * self.S
        self.clips = clips
        g = torch.Generator().manual_seed(2)
        proc = processor_factory()

        def mk(cross):
            a = Attention(C, cross, heads=heads, dim_head=64, processor=proc)
            with torch.no_grad():
                for p in a.parameters():  # nn.Linear default init U(+-1/sqrt(in)), seeded
                    bound = p.shape[-1] ** -0.5 if p.dim() > 1 else C ** -0.5
                    p.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * bound)
            return a.to(device=device, dtype=dtype).eval()

        self.attn1_s, self.attn2_s, self.attn1_t, self.attn2_t = mk(None), mk(CTX_DIM), mk(None), mk(CTX_DIM)
        g0, g1 = torch.Generator().manual_seed(0), torch.Generator().manual_seed(1)
        self.x_host = torch.randn(self.BF, self.S, C, generator=g0).to(dtype)
        self.ctx_s_host = torch.randn(self.BF, L_SPATIAL, CTX_DIM, generator=g1).to(dtype)
        # time_context is the first frame's tokens broadcast over the h*w pixels of a clip
        # (reference src/models/transformer_temporal.py:279-297): materialise the broadcast like .reshape does
        ctx_t = torch.randn(clips * CFG, L_TEMPORAL, CTX_DIM, generator=g1).to(dtype)
        self.ctx_t_small_host = ctx_t.contiguous()   # [B, Ltm, 1024]: what the pipeline actually hands the module
        self.ctx_t_host = ctx_t[:, None].expand(clips * CFG, self.S, L_TEMPORAL, CTX_DIM).reshape(self.nseq, L_TEMPORAL, CTX_DIM).contiguous()

    def expand_time_context(self, ctx_small):
        """[B, Ltm, 1024] -> [B*S, Ltm, 1024] on the device, the reference's broadcast_to + reshape
        (src/models/transformer_temporal.py:279-297) in this workload's (b, s) sequence order."""
        B = ctx_small.shape[0]
        return ctx_small[:, None].expand(B, self.S, L_TEMPORAL, CTX_DIM).reshape(self.nseq, L_TEMPORAL, CTX_DIM)

    def step(self, x, ctx_s, ctx_t):
        y1 = self.attn1_s(x)
        y2 = self.attn2_s(y1, encoder_hidden_states=ctx_s)
        v = y2.view(self.nseq, T_FRAMES, self.C)
        y3 = self.attn1_t(v)
        return self.attn2_t(y3, encoder_hidden_states=ctx_t)


# ---------------------------------------------------------------------------------------------------
# CPU arm: the oracle port (oracle/attention_ref.py, fp32 eager math of the reference processor) on the
# host cores.  The reference itself cannot be imported on the GPU box (diffusers is absent).
# ---------------------------------------------------------------------------------------------------
def cpu_step_sample(wl_cpu, x, ctx_s, ctx_t):
    from oracle import attention_ref as ar
    y1 = ar.attn_processor_ref(wl_cpu.attn1_s, x)
    y2 = ar.attn_processor_ref(wl_cpu.attn2_s, y1, ctx_s)
    y3 = ar.attn_processor_ref(wl_cpu.attn1_t, wl_cpu.xt_sample)
    return y2, ar.attn_processor_ref(wl_cpu.attn2_t, y3, ctx_t)


def run_cpu(level, steps, warmup):
    """Bounded sample = the work of ONE clip-frame: one CFG pair of frames through the spatial calls
    (2 x S tokens) and 2*S/T temporal sequences through the temporal calls."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)

    class _NoProc:
        def __call__(self, *a, **k):
            raise AssertionError("CPU arm must not call a processor")

    wl = Workload(level, 1, "cpu", torch.float32, _NoProc)
    nseq_sample = max(1, round(CFG * wl.S / T_FRAMES))
    x = wl.x_host[:CFG].float()
    ctx_s = wl.ctx_s_host[:CFG].float()
    ctx_t = wl.ctx_t_host[:nseq_sample].float()
    wl.xt_sample = torch.randn(nseq_sample, T_FRAMES, wl.C, generator=torch.Generator().manual_seed(5))
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            cpu_step_sample(wl, x, ctx_s, ctx_t)
            if i >= warmup:
                times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    sample = (f"1 clip-frame of the workload per step: {CFG} frames x {wl.S} tokens through both spatial calls + "
              f"{nseq_sample} temporal sequences of {T_FRAMES} through both temporal calls, fp32 eager oracle, torch {cores} threads")
    return 1.0 / sec, sec * 1e3, cores, sample


def comparators(level, clips, dev):
    """COMPARATOR ONLY, never on the product path and outside every timed region: torch's own
    scaled_dot_product_attention (the call the plugin replaces, reference src/models/attention_processor.py:1243) on the
    shape of the dominant kernel, same box, same run."""
    C, heads, h, w = LEVELS[level]
    S, BF = h * w, clips * CFG * T_FRAMES
    g = torch.Generator().manual_seed(7)
    q, k, v = (torch.randn(BF, heads, S, 64, generator=g).to(torch.bfloat16).to(dev) for _ in range(3))
    out = {}
    try:
        with torch.no_grad():
            for _ in range(3):
                torch.nn.functional.scaled_dot_product_attention(q, k, v)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 10
            a.record()
            for _ in range(n):
                torch.nn.functional.scaled_dot_product_attention(q, k, v)
            b.record()
            torch.cuda.synchronize()
        ms = a.elapsed_time(b) / n
        out["torch_sdpa_ms"] = ms
        out["torch_sdpa_tflops"] = 4.0 * BF * heads * S * S * 64 / (ms * 1e-3) / 1e12
        out["torch"] = torch.__version__
    except Exception as exc:
        out["torch_sdpa_error"] = f"{type(exc).__name__}: {exc}"
    return out


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return 0
        # a sample step is ~0.7 s of CPU work: the driver's K / W are honoured up to a cap that keeps the run in minutes
        steps, warmup = max(1, min(args.steps, 40)), max(1, min(args.warmup, 10))
        value, ms, cores, sample = run_cpu(args.level, steps, warmup)
        cfg = config_dict(args.level, args.clips, args.gpus)
        line = {"impl": "reference", "sample_fraction": 1.0 / (args.clips * T_FRAMES), "metric": "clip-frames/sec", "value": value, "unit": "clip-frames/s", "n_gpus": args.gpus,
                "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
                "cpu_baseline": {"value": value, "unit": "clip-frames/s", "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": value, "unit": "clip-frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl tvi needs a CUDA device (there is no CPU fallback)")
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    import timelapse_video_interpolation_b200 as tvi
    from timelapse_video_interpolation_b200 import ops
    from timelapse_video_interpolation_b200.processor import TviAttnProcessor

    cores_bound = bind_rank_to_cores(local_rank, world)  # before any pinned allocation
    wl = Workload(args.level, args.clips, dev, torch.bfloat16, TviAttnProcessor)
    # e2e inputs in pinned host memory: hidden states, the spatial context and the [B, Ltm, 1024] temporal context the
    # pipeline really holds (its broadcast over the h*w pixels is done on the device, as the reference does)
    x_pin, cs_pin, ct_pin = wl.x_host.pin_memory(), wl.ctx_s_host.pin_memory(), wl.ctx_t_small_host.pin_memory()
    x, ctx_s, ctx_t = x_pin.to(dev), cs_pin.to(dev), wl.ctx_t_host.to(dev)
    out_pin = torch.empty(wl.nseq, T_FRAMES, wl.C, dtype=torch.bfloat16).pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    dom = "attn_spatial"
    dom_events = []

    hook_t = []  # (host time, kernel, is_start) of every libtvi launch in timed region 1: locates host-side stalls

    def hook(name, is_start):
        hook_t.append((time.perf_counter(), name, is_start))
        if name == dom:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()  # torch's current stream == the stream libtvi launches on
            dom_events.append(ev)

    sampler = ClockSampler(local_rank)  # NVML set-up happens here, before the warm-up
    with torch.no_grad():
        for _ in range(max(3, args.warmup)):
            # hold the previous result across the next step exactly as the timed loop does, so that the caching
            # allocator already owns every block the loop needs: a first-use cudaMalloc inside the timed region
            # cost 0.6 ms on a good day and 24-32 ms on a bad one (one run in six)
            y = wl.step(x, ctx_s, ctx_t)
        barrier()
        # ---------------- timed region 1: device-resident inputs (`value`) ----------------
        ops.EVENT_HOOK = hook
        launches0 = ops.LAUNCHES
        barrier()
        # a generation-2 collection of the interpreter's heap takes tens of ms and, landing inside a 40 ms timed
        # region, stalls the launches of several steps: collect now, keep the collector off while timing
        gc.collect()
        gc.disable()
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        host_t = [time.perf_counter()]
        e0.record()
        for _ in range(args.steps):
            y = wl.step(x, ctx_s, ctx_t)
            host_t.append(time.perf_counter())
        e1.record()
        barrier()
        sampler.pause()
        gc.enable()
        host_gap_ms = max(b - a for a, b in zip(host_t, host_t[1:])) * 1e3  # longest host-side launch loop of one step
        gaps = [(b[0] - a[0], a, b) for a, b in zip(hook_t, hook_t[1:])]
        g = max(gaps, key=lambda v: v[0]) if gaps else (0.0, (0, "", True), (0, "", True))
        host_stall = {"ms": round(g[0] * 1e3, 3), "from": f"{g[1][1]}:{'enter' if g[1][2] else 'exit'}",
                      "to": f"{g[2][1]}:{'enter' if g[2][2] else 'exit'}",
                      "step": sum(1 for t in host_t[1:] if t < g[2][0])}
        ops.EVENT_HOOK = None
        launches = ops.LAUNCHES - launches0
        ms_total = e0.elapsed_time(e1)
        dom_ms = [dom_events[i].elapsed_time(dom_events[i + 1]) for i in range(0, len(dom_events), 2)]
        # ---------------- timed region 2: host buffers, H2D + D2H inside (`e2e`) ----------------
        # Every step copies ITS inputs from pinned host memory and reads ITS result back to pinned host memory;
        # the copies run on their own streams against double-buffered device inputs, so the H2D of step k+1 and
        # the D2H of step k-1 overlap the kernels of step k (PCIe is full duplex).  All of it is inside f0..f1.
        cur = torch.cuda.current_stream()
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        dbuf = [(torch.empty_like(x), torch.empty_like(ctx_s), torch.empty(ct_pin.shape, dtype=ct_pin.dtype, device=dev))
                for _ in range(2)]
        out_pins = [out_pin, torch.empty_like(out_pin).pin_memory()]
        ev_in = [torch.cuda.Event() for _ in range(2)]
        ev_free = [torch.cuda.Event() for _ in range(2)]
        ev_out = [torch.cuda.Event() for _ in range(2)]
        ys = [None, None]  # results whose D2H copy may still be in flight (kept alive; no record_stream, so the
        #                    allocator's reuse pattern is the same every step and settles during the warm-up)

        step_done = []  # compute-stream events after every e2e step: (last - first) / (K - 1) is the steady-state step

        def e2e_steps(n):
            step_done.clear()
            for k in range(n):
                b = k & 1
                with torch.cuda.stream(s_in):
                    if k >= 2:
                        s_in.wait_event(ev_free[b])  # the kernels of step k-2 are done with input set b
                    dbuf[b][0].copy_(x_pin, non_blocking=True)
                    dbuf[b][1].copy_(cs_pin, non_blocking=True)
                    dbuf[b][2].copy_(ct_pin, non_blocking=True)
                    ev_in[b].record(s_in)
                cur.wait_event(ev_in[b])
                if k >= 2:
                    cur.wait_event(ev_out[b])  # result k-2 has left the device: its block may be reused from here on
                ys[b] = None
                ys[b] = y = wl.step(dbuf[b][0], dbuf[b][1], wl.expand_time_context(dbuf[b][2]))
                ev_free[b].record(cur)
                step_done.append(torch.cuda.Event(enable_timing=True))
                step_done[-1].record(cur)
                s_out.wait_stream(cur)
                with torch.cuda.stream(s_out):
                    out_pins[b].copy_(y, non_blocking=True)
                    ev_out[b].record(s_out)
            cur.wait_stream(s_out)  # the last result has reached host memory

        e2e_steps(4)
        barrier()
        gc.collect()
        gc.disable()
        sampler.start()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        e2e_steps(args.steps)
        f1.record()
        barrier()
        ms_e2e = f0.elapsed_time(f1)
        ms_e2e_steady = step_done[0].elapsed_time(step_done[-1]) / max(1, len(step_done) - 1)
        gc.enable()
        clocks = sampler.stop()  # samples cover both timed regions

        # ---------------- outside the timed regions: the host path alone, comparator, secondary configs ----------------
        # copies only: the same per-step H2D and D2H bytes on the same two streams, all ranks at once -> the per-N
        # ceiling of the shared host memory / PCIe path that bounds `e2e` (reported as e2e.host_ceiling_ms)
        def copy_only(n):
            for k in range(n):
                b = k & 1
                with torch.cuda.stream(s_in):
                    dbuf[b][0].copy_(x_pin, non_blocking=True)
                    dbuf[b][1].copy_(cs_pin, non_blocking=True)
                    dbuf[b][2].copy_(ct_pin, non_blocking=True)
                with torch.cuda.stream(s_out):
                    out_pins[b].copy_(y, non_blocking=True)
            cur.wait_stream(s_in)
            cur.wait_stream(s_out)

        copy_only(2)
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        copy_only(args.steps)
        c1.record()
        barrier()
        ms_copy = c0.elapsed_time(c1)

    from timelapse_video_interpolation_b200 import sharding
    ms_total, ms_e2e = sharding.max_over_ranks(ms_total, dev), sharding.max_over_ranks(ms_e2e, dev)
    ms_copy = sharding.max_over_ranks(ms_copy, dev)
    # every rank owns `--clips` whole clips (weak scaling); the job's units are the sum over ranks
    frames_per_step = int(sharding.sum_over_ranks(args.clips * T_FRAMES, dev))
    value = frames_per_step * args.steps / (ms_total * 1e-3)
    e2e_value = frames_per_step * args.steps / (ms_e2e * 1e-3)

    if rank == 0:
        pk = peaks()
        fl = algorithmic_flops(args.level, args.clips)
        dom_avg_ms = sum(dom_ms) / len(dom_ms)
        traffic, traffic_src = measured_traffic(args.level, args.clips)
        achieved = fl["spatial_core"] / (dom_avg_ms * 1e-3) / 1e12
        h2d = x_pin.numel() * 2 + cs_pin.numel() * 2 + ct_pin.numel() * 2
        line = {
            "metric": "clip-frames/sec", "value": value, "unit": "clip-frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": config_dict(args.level, args.clips, world),
            "e2e": {"value": e2e_value, "unit": "clip-frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": out_pin.numel() * 2,
                    "ms_per_step": ms_e2e / args.steps,
                    # the copies of a step alone, all ranks at once (outside the timed regions): the floor the shared host
                    # memory / PCIe path sets for this N; e2e within ~15 % of it is copy-bound, not a software stall
                    "host_ceiling_ms": ms_copy / args.steps,
                    "kernel_ms_per_step": ms_total / args.steps,
                    # the timed region is bracketed by synchronisations, so it also holds the pipeline's fill (the first
                    # step's H2D, nothing to overlap it with) and drain (the last step's D2H): ~2 copies / K per step.
                    # steady_ms_per_step is the spacing of consecutive steps' completions inside the same region (rank 0).
                    "steady_ms_per_step": ms_e2e_steady,
                    "rank_cores": cores_bound,
                    "note": "temporal context travels as [B, Ltm, 1024] and is broadcast on the device "
                            "(reference src/models/transformer_temporal.py:279-297)"},
            "gpu_launches": launches,
            "host_launch_ms_max_step": round(host_gap_ms, 3), "host_stall_max": host_stall,
            "clocks": clocks,
            "roofline": {"kernel": "attn_spatial_tc_kernel", "bound": "tensor", "achieved": achieved, "peak": pk["tf_sust"],
                         "unit": "TFLOP/s", "frac": achieved / pk["tf_sust"], "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": pk["src"] + " (sustained)",
                         "ms_per_launch": dom_avg_ms, "share_of_step": dom_avg_ms / (ms_total / args.steps),
                         "algorithmic_flops_per_launch": fl["spatial_core"]},
            "step_tflops": fl["total"] / 1e12 / (ms_total / args.steps * 1e-3),
        }
        if world == 1 and not args.no_extras:
            line["comparators"] = comparators(args.level, args.clips, dev)
            try:
                from tools import bench_configs
                line.update(bench_configs.driver_extras(dev, pk))
            except Exception as exc:  # the headline line must not depend on the secondary measurements
                line["secondary_error"] = f"{type(exc).__name__}: {exc}"
        if world == 1 and not args.no_cpu_baseline:
            v, ms, cores, sample = run_cpu(args.level, 2, 1)
            line["cpu_baseline"] = {"value": v, "unit": "clip-frames/s", "cores": cores, "kind": "port", "sample": sample,
                                    "ms_per_sample": ms}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
