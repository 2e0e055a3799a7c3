"""GPU: the no-allocation entry point `tvi_attn_fewkeys_fwd_static` (include/tvi.h) against `tvi_attn_fewkeys_fwd`.

Collected last on purpose.  Both entry points are called straight through the C ABI
(ctypes); the tensors are created before the calls, so the calls themselves see no allocator."""
import pytest
import torch

from timelapse_video_interpolation_b200 import _lib

pytestmark = pytest.mark.gpu
TOL = 2e-2        # the bf16 attention bar of tests/test_gpu_attn_cores.py (max-abs); the two paths are in fact bit-identical
STATS_TOL = 1e-4  # fp32 atomics into the per-CTA partial sums arrive in a different order from run to run


def _ptr(t):
    return None if t is None else t.data_ptr()


def _case(cuda, groups, T, heads, L, seed, with_bias):
    g = torch.Generator().manual_seed(seed)
    C = heads * 64
    mk = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16).to(cuda)
    q, k, v = mk(groups * T, C), mk(groups * L, C), mk(groups * L, C)
    bias = None
    if with_bias:
        w = torch.linspace(0, 1, T)
        bias = torch.stack([1 - w] + [w] * (L - 1), dim=1).contiguous().to(cuda)      # [T, L] fp32
    return q, k, v, bias, C


def _run_existing(lib, q, k, v, bias, groups, T, heads, L, C, stream):
    o = torch.empty(groups * T, C, dtype=torch.bfloat16, device=q.device)
    stats = torch.empty(T, L, dtype=torch.float32, device=q.device)
    ws = torch.empty(T * L, dtype=torch.float64, device=q.device)                      # the caller-provided scratch
    _lib.check(lib.tvi_attn_fewkeys_fwd(_ptr(q), _ptr(k), _ptr(v), _ptr(o), groups * T, T, heads, L, 64, C, C, C, C, 0.125, T,
                                        _ptr(bias), _ptr(stats), _ptr(ws), stream))
    return o, stats


def _run_static(lib, q, k, v, bias, o, stats, groups, T, heads, L, C, stream):
    _lib.check(lib.tvi_attn_fewkeys_fwd_static(_ptr(q), _ptr(k), _ptr(v), _ptr(o), groups * T, T, heads, L, 64, C, C, C, C, 0.125, T,
                                               _ptr(bias), _ptr(stats), stream))


def test_static_scratch_entry_point_matches_the_workspace_one_twice_in_a_row(cuda):
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    # two calls in a row with DIFFERENT shapes and data: the second must not see the first one's accumulators
    cases = [(37, 14, 5, 2, 1, True), (53, 25, 5, 3, 2, False)]
    for groups, T, heads, L, seed, with_bias in cases:
        q, k, v, bias, C = _case(cuda, groups, T, heads, L, seed, with_bias)
        o_ref, stats_ref = _run_existing(lib, q, k, v, bias, groups, T, heads, L, C, stream)
        o = torch.full_like(o_ref, float("nan"))
        stats = torch.full_like(stats_ref, float("nan"))
        torch.cuda.synchronize()
        # nothing is allocated between these two readings but what the call itself would allocate (its kernels were
        # already loaded by the workspace entry point above: same device functions)
        allocated0, free0 = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        _run_static(lib, q, k, v, bias, o, stats, groups, T, heads, L, C, stream)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == allocated0 and torch.cuda.mem_get_info()[0] >= free0
        assert torch.isfinite(o.float()).all() and torch.isfinite(stats).all()
        assert float((o.float() - o_ref.float()).abs().max()) <= TOL
        assert float((stats - stats_ref).abs().max()) <= STATS_TOL * max(1.0, float(stats_ref.abs().max()))
    # same inputs again, straight after: the array keeps the last call's contents, the call clears what it uses
    groups, T, heads, L, seed, with_bias = cases[0]
    q, k, v, bias, C = _case(cuda, groups, T, heads, L, seed, with_bias)
    o_ref, stats_ref = _run_existing(lib, q, k, v, bias, groups, T, heads, L, C, stream)
    outs = []
    for _ in range(2):
        o, stats = torch.empty_like(o_ref), torch.empty_like(stats_ref)
        _run_static(lib, q, k, v, bias, o, stats, groups, T, heads, L, C, stream)
        outs.append((o, stats))
    torch.cuda.synchronize()
    for o, stats in outs:
        assert torch.equal(o, o_ref)
        assert float((stats - stats_ref).abs().max()) <= STATS_TOL * max(1.0, float(stats_ref.abs().max()))


def test_static_scratch_entry_point_without_statistics_and_its_limit(cuda):
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    groups, T, heads, L = 11, 14, 5, 2
    q, k, v, bias, C = _case(cuda, groups, T, heads, L, 5, False)
    o_ref = torch.empty(groups * T, C, dtype=torch.bfloat16, device=cuda)
    _lib.check(lib.tvi_attn_fewkeys_fwd(_ptr(q), _ptr(k), _ptr(v), _ptr(o_ref), groups * T, T, heads, L, 64, C, C, C, C, 0.125, 1,
                                        None, None, None, stream))
    o = torch.empty_like(o_ref)
    _lib.check(lib.tvi_attn_fewkeys_fwd_static(_ptr(q), _ptr(k), _ptr(v), _ptr(o), groups * T, T, heads, L, 64, C, C, C, C, 0.125, 1,
                                               None, None, stream))
    torch.cuda.synchronize()
    assert torch.equal(o, o_ref)
    # Tq * L beyond the static array: a clear error, nothing launched
    Tq = _lib.FEWKEYS_STATIC_MAX_STATS // L + 1
    stats = torch.empty(Tq, L, dtype=torch.float32, device=cuda)
    with pytest.raises(_lib.TviError, match="exceed the static scratch"):
        _lib.check(lib.tvi_attn_fewkeys_fwd_static(_ptr(q), _ptr(k), _ptr(v), _ptr(o), groups * T, T, heads, L, 64, C, C, C, C, 0.125,
                                                   Tq, None, _ptr(stats), stream))
