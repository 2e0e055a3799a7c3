"""CUDA-graph capture of a libtvi forward ("CUDA-graph capture of the whole UNet step").

A TransformerSpatioTemporalModel.forward is 32 kernel launches; at the UNet's inner levels (C = 1280, 576 or 144 tokens
per frame) every kernel runs for a few microseconds and the step is bound by the host (ctypes marshalling, tensor-map
encoding, torch allocations), not by the GPU.  libtvi entry points never allocate, never synchronise and launch on the
caller's stream, so a whole forward records into one CUDA graph: `GraphedForward` captures it once over static
input buffers and replays it with one `cudaGraphLaunch` per call.

    fwd = GraphedForward(module, hidden_states, encoder_hidden_states=e, image_only_indicator=ind, return_dict=False)
    y = fwd(hidden_states, encoder_hidden_states=e, image_only_indicator=ind)     # copies into the static buffers, replays

Rules (checked): tensor arguments must keep shape / dtype / device; non-tensor arguments are frozen at capture.  The
returned tensors are the graph's static outputs — valid until the next call (clone to keep).  Weights are read through
the pointers captured at record time: re-capture after an optimizer step or a `.to()`.
"""
from __future__ import annotations

from typing import Any

import torch

from . import _lib

__all__ = ["GraphedForward"]


def _map(obj: Any, fn):
    if isinstance(obj, torch.Tensor):
        return fn(obj)
    if isinstance(obj, (tuple, list)):
        return type(obj)(_map(o, fn) for o in obj)
    if isinstance(obj, dict):
        return {k: _map(v, fn) for k, v in obj.items()}
    return obj


def _flat(obj: Any, out: list):
    if isinstance(obj, torch.Tensor):
        out.append(obj)
    elif isinstance(obj, (tuple, list)):
        for o in obj:
            _flat(o, out)
    elif isinstance(obj, dict):
        for v in obj.values():
            _flat(v, out)
    return out


class GraphedForward:
    def __init__(self, module, *args, warmup: int = 2, **kwargs):
        tensors = _flat((args, kwargs), [])
        if not tensors or not all(t.is_cuda for t in tensors):
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "GraphedForward: every tensor argument must live on a CUDA device")
        self._module = module
        self._static_args = _map(args, lambda t: t.detach().clone())
        self._static_kwargs = _map(kwargs, lambda t: t.detach().clone())
        self._static_flat = _flat((self._static_args, self._static_kwargs), [])
        dev = tensors[0].device
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.no_grad(), torch.cuda.stream(side):
            for _ in range(max(1, warmup)):  # first-call work (cudaFuncSetAttribute, weight caches) stays out of the graph
                module(*self._static_args, **self._static_kwargs)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self._static_out = module(*self._static_args, **self._static_kwargs)

    def __call__(self, *args, **kwargs):
        new = _flat((args, kwargs), [])
        if len(new) != len(self._static_flat):
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "GraphedForward: argument structure differs from the captured call")
        for dst, src in zip(self._static_flat, new):
            if dst.shape != src.shape or dst.dtype != src.dtype or dst.device != src.device:
                raise _lib.TviError(_lib.TVI_ERR_INVALID, f"GraphedForward: tensor argument changed ({tuple(src.shape)} {src.dtype} "
                                                          f"vs captured {tuple(dst.shape)} {dst.dtype})")
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self._static_out
