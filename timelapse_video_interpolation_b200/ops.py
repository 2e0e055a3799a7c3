"""Thin tensor-level wrappers over the C ABI: pointer/stride marshalling only, no arithmetic.

Every function takes CUDA tensors, launches on `torch.cuda.current_stream()`, never synchronises
the host and never falls back to torch math.  `LAUNCHES` counts libtvi kernel launches (bench.py
reports it as `gpu_launches`).
"""
from __future__ import annotations

import torch

from . import _lib

LAUNCHES = 0  # kernels launched through libtvi since import (per process)
# Optional profiling hook: callable(kernel_name, is_start) invoked right before / after a launch is
# enqueued (bench.py uses it to bracket one kernel with CUDA events on the launching stream).
EVENT_HOOK = None


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts):
    """Every tensor of a call must live on the CURRENT CUDA device: the launch goes to torch's current stream of the
    current device, so a tensor of another GPU would be read through a stream / context that does not own it."""
    cur = None
    for t in ts:
        if t is None:
            continue
        if not t.is_cuda:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "libtvi operates on CUDA tensors only (there is no CPU fallback)")
        if cur is None:
            cur = torch.cuda.current_device()
        if t.device.index != cur:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, f"tensor on cuda:{t.device.index} while the current device is cuda:{cur}: "
                                                      "call under torch.cuda.device(tensor.device) (one process per GPU sets it once)")


def _rows2d(t: torch.Tensor, what: str):
    """Returns (tensor, ld) for a tensor whose last dim is contiguous and whose leading dims
    collapse to one row stride (e.g. a column slice of a packed QKV buffer)."""
    if t.dtype != torch.bfloat16:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, f"{what}: expected bfloat16, got {t.dtype}")
    if t.dim() < 2 or t.stride(-1) != 1:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, f"{what}: last dimension must be contiguous")
    ld = t.stride(-2)
    exp = ld
    for i in range(t.dim() - 2, 0, -1):
        exp *= t.shape[i]
        if t.shape[i - 1] != 1 and t.stride(i - 1) != exp:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, f"{what}: leading dimensions do not collapse to one row stride")
    return t, ld


def gemm_bf16_tn(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None,
                 out: torch.Tensor | None = None) -> torch.Tensor:
    """y[..., N] = x[..., K] @ w[N, K]^T (+ bias) — tcgen05 GEMM (replaces nn.Linear on the path)."""
    global LAUNCHES
    _need_cuda(x, w, bias)
    x, ldx = _rows2d(x, "x")
    w, ldw = _rows2d(w, "w")
    K = x.shape[-1]
    N = w.shape[0]
    if w.shape[1] != K:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, f"gemm: x has K={K}, w has K={w.shape[1]}")
    M = x.numel() // K if K else 0
    if out is None:
        out = torch.empty(*x.shape[:-1], N, dtype=torch.bfloat16, device=x.device)
    out, ldy = _rows2d(out, "out")
    if bias is not None and (bias.dtype != torch.bfloat16 or not bias.is_contiguous() or bias.numel() != N):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm: bias must be a contiguous bf16 vector of length N")
    if EVENT_HOOK is not None:
        EVENT_HOOK("gemm_bf16_tn", True)
    _lib.check(_lib.load().tvi_gemm_bf16_tn(x.data_ptr(), w.data_ptr(), bias.data_ptr() if bias is not None else None,
                                            out.data_ptr(), M, N, K, ldx, ldw, ldy, _stream()))
    if EVENT_HOOK is not None:
        EVENT_HOOK("gemm_bf16_tn", False)
    LAUNCHES += 1
    return out


def _hooked(name: str, n_launches: int, fn, *args):
    global LAUNCHES
    if EVENT_HOOK is not None:
        EVENT_HOOK(name, True)
    _lib.check(fn(*args))
    if EVENT_HOOK is not None:
        EVENT_HOOK(name, False)
    LAUNCHES += n_launches


def _workspace(nbytes: int, device) -> torch.Tensor | None:
    return torch.empty(max(1, (nbytes + 3) // 4), dtype=torch.float32, device=device) if nbytes else None


def gemm_bf16_tn_ex(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None, residual: torch.Tensor | None = None,
                    geglu: bool = False, out: torch.Tensor | None = None) -> torch.Tensor:
    """y = x @ w^T (+ bias) (+ residual); with `geglu` w / bias hold value and gate rows interleaved in blocks of
    `geglu_block(inner)` and y[..., inner] = value * gelu(gate) (+ residual) — one tcgen05 GEMM, fused epilogue."""
    _need_cuda(x, w, bias, residual)
    x, ldx = _rows2d(x, "x")
    w, ldw = _rows2d(w, "w")
    K, N = x.shape[-1], w.shape[0]
    if w.shape[1] != K:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, f"gemm: x has K={K}, w has K={w.shape[1]}")
    n_out = N // 2 if geglu else N
    M = x.numel() // K if K else 0
    if out is None:
        out = torch.empty(*x.shape[:-1], n_out, dtype=torch.bfloat16, device=x.device)
    out, ldy = _rows2d(out, "out")
    ldr = 0
    if residual is not None:
        residual, ldr = _rows2d(residual, "residual")
        if residual.shape[-1] != n_out or residual.numel() != M * n_out:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm: residual must have the shape of the output")
    if bias is not None and (bias.dtype != torch.bfloat16 or not bias.is_contiguous() or bias.numel() != N):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm: bias must be a contiguous bf16 vector of length N")
    _hooked("gemm_bf16_tn_ex", 1, _lib.load().tvi_gemm_bf16_tn_ex, x.data_ptr(), w.data_ptr(),
            bias.data_ptr() if bias is not None else None, residual.data_ptr() if residual is not None else None,
            out.data_ptr(), M, N, K, ldx, ldw, ldr, ldy, _lib.EPILOGUE_GEGLU if geglu else _lib.EPILOGUE_NONE, _stream())
    return out


def layernorm_stats_supported(C: int) -> bool:
    return bool(_lib.load().tvi_layernorm_stats_supported(int(C)))


def layernorm_stats(x: torch.Tensor, eps: float) -> torch.Tensor:
    """(mean, rstd) of every row of x[..., C] as fp32 [M, 2]: the read-only half of a LayerNorm whose affine map runs in
    the epilogue of gemm_bf16_tn_ln."""
    _need_cuda(x)
    x, ldx = _rows2d(x, "x")
    C = x.shape[-1]
    M = x.numel() // C if C else 0
    stats = torch.empty(M, 2, dtype=torch.float32, device=x.device)
    _hooked("layernorm_stats", 1, _lib.load().tvi_layernorm_stats_f32, x.data_ptr(), stats.data_ptr(), M, C, ldx, float(eps), _stream())
    return stats


def gemm_bf16_tn_ln(x: torch.Tensor, wg: torch.Tensor, bias: torch.Tensor, colsum: torch.Tensor, stats: torch.Tensor,
                    residual: torch.Tensor | None = None, geglu: bool = False, out: torch.Tensor | None = None) -> torch.Tensor:
    """y = LayerNorm(x) @ w^T + b computed from the RAW rows x: wg = gamma * w (bf16), colsum = wg.float().sum(1),
    bias = w @ beta + b, stats = layernorm_stats(x); optional GEGLU / residual as gemm_bf16_tn_ex."""
    _need_cuda(x, wg, bias, colsum, stats, residual)
    x, ldx = _rows2d(x, "x")
    wg, ldw = _rows2d(wg, "wg")
    K, N = x.shape[-1], wg.shape[0]
    if wg.shape[1] != K:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, f"gemm: x has K={K}, wg has K={wg.shape[1]}")
    n_out = N // 2 if geglu else N
    M = x.numel() // K if K else 0
    if out is None:
        out = torch.empty(*x.shape[:-1], n_out, dtype=torch.bfloat16, device=x.device)
    out, ldy = _rows2d(out, "out")
    ldr = 0
    if residual is not None:
        residual, ldr = _rows2d(residual, "residual")
        if residual.shape[-1] != n_out or residual.numel() != M * n_out:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm: residual must have the shape of the output")
    if bias.dtype != torch.bfloat16 or not bias.is_contiguous() or bias.numel() != N:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm: bias must be a contiguous bf16 vector of length N")
    if colsum.dtype != torch.float32 or not colsum.is_contiguous() or colsum.numel() != N:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm: colsum must be a contiguous fp32 vector of length N")
    if stats.dtype != torch.float32 or not stats.is_contiguous() or stats.numel() != 2 * M:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm: stats must be a contiguous fp32 [M, 2] tensor")
    _hooked("gemm_bf16_tn_ln", 1, _lib.load().tvi_gemm_bf16_tn_ln, x.data_ptr(), wg.data_ptr(), bias.data_ptr(), colsum.data_ptr(),
            stats.data_ptr(), residual.data_ptr() if residual is not None else None, out.data_ptr(), M, N, K, ldx, ldw, ldr, ldy,
            _lib.EPILOGUE_GEGLU if geglu else _lib.EPILOGUE_NONE, _stream())
    return out


def geglu_block(inner: int) -> int:
    return int(_lib.load().tvi_geglu_block(int(inner)))


def layernorm_bf16(x: torch.Tensor, gamma: torch.Tensor | None, beta: torch.Tensor | None, eps: float,
                   out: torch.Tensor | None = None) -> torch.Tensor:
    """LayerNorm over the last dimension (fp32 statistics), bf16 in / out."""
    _need_cuda(x, gamma, beta)
    x, ldx = _rows2d(x, "x")
    C = x.shape[-1]
    M = x.numel() // C if C else 0
    for t in (gamma, beta):
        if t is not None and (t.dtype != torch.bfloat16 or not t.is_contiguous() or t.numel() != C):
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "layernorm: gamma / beta must be contiguous bf16 vectors of length C")
    if out is None:
        out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    out, ldy = _rows2d(out, "out")
    _hooked("layernorm_bf16", 1, _lib.load().tvi_layernorm_bf16, x.data_ptr(), gamma.data_ptr() if gamma is not None else None,
            beta.data_ptr() if beta is not None else None, out.data_ptr(), M, C, ldx, ldy, float(eps), _stream())
    return out


def bcast_rows_add(v: torch.Tensor, residual: torch.Tensor | None, rows_per_group: int, out: torch.Tensor | None = None,
                   group_inner: int = 1, nrows: int | None = None) -> torch.Tensor:
    """out[r] = v[group(r)] (+ residual[r]) — one-key cross-attention output (+ the block's residual);
    group(r) = (r // (group_inner * rows_per_group)) * group_inner + r % group_inner."""
    _need_cuda(v, residual)
    v, ldv = _rows2d(v, "v")
    C = v.shape[-1]
    ldr = 0
    if residual is not None:
        residual, ldr = _rows2d(residual, "residual")
        nrows = residual.numel() // C if C else 0
    elif nrows is None:
        nrows = v.shape[0] * int(rows_per_group)
    if out is None:
        out = torch.empty(nrows, C, dtype=torch.bfloat16, device=v.device)
    out, ldo = _rows2d(out, "out")
    _hooked("bcast_rows_add", 1, _lib.load().tvi_bcast_rows_add_bf16, v.data_ptr(),
            residual.data_ptr() if residual is not None else None, out.data_ptr(), nrows, int(rows_per_group), int(group_inner), C,
            ldv, ldr, ldo, _stream())
    return out


def groupnorm_tokens(x: torch.Tensor, gamma, beta, groups: int, eps: float, out: torch.Tensor | None = None) -> torch.Tensor:
    """GroupNorm of an NCHW bf16 feature map written token-major: [N, C, H, W] -> [N, H*W, C]."""
    _need_cuda(x, gamma, beta)
    if x.dim() != 4 or x.dtype != torch.bfloat16 or not x.is_contiguous():
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "groupnorm_tokens: x must be a contiguous bf16 [N, C, H, W] tensor")
    N, C, H, W = x.shape
    for t in (gamma, beta):
        if t is not None and (t.dtype != torch.bfloat16 or not t.is_contiguous() or t.numel() != C):
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "groupnorm_tokens: gamma / beta must be contiguous bf16 vectors of length C")
    if out is None:
        out = torch.empty(N, H * W, C, dtype=torch.bfloat16, device=x.device)
    o2, ldy = _rows2d(out, "out")
    ws = _workspace(_lib.load().tvi_groupnorm_tokens_workspace_bytes(N, C, H * W, int(groups)), x.device)
    _hooked("groupnorm_tokens", 2, _lib.load().tvi_groupnorm_tokens_bf16, x.data_ptr(),
            gamma.data_ptr() if gamma is not None else None, beta.data_ptr() if beta is not None else None, o2.data_ptr(),
            ws.data_ptr() if ws is not None else None, N, C, H * W, int(groups), ldy, float(eps), _stream())
    return out


def nchw_to_tokens(x: torch.Tensor) -> torch.Tensor:
    """[N, C, H, W] -> token-major [N, H*W, C] (the backward of tokens_to_nchw_add with respect to the tokens)."""
    _need_cuda(x)
    if x.dim() != 4 or x.dtype != torch.bfloat16 or not x.is_contiguous():
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "nchw_to_tokens: x must be a contiguous bf16 [N, C, H, W] tensor")
    N, C, H, W = x.shape
    out = torch.empty(N, H * W, C, dtype=torch.bfloat16, device=x.device)
    _hooked("nchw_to_tokens", 1, _lib.load().tvi_nchw_to_tokens_bf16, x.data_ptr(), out.data_ptr(), N, C, H * W, C, _stream())
    return out


def groupnorm_tokens_bwd(x: torch.Tensor, dn: torch.Tensor, gamma, groups: int, eps: float, need_dx: bool = True,
                         need_affine: bool = True):
    """Backward of groupnorm_tokens: x [N, C, H, W] (the forward input), dn [N, H*W, C] -> (dx [N, C, H, W] or None,
    dgamma [C] or None, dbeta [C] or None), all bf16."""
    _need_cuda(x, dn, gamma)
    if x.dim() != 4 or x.dtype != torch.bfloat16 or not x.is_contiguous():
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "groupnorm_tokens_bwd: x must be a contiguous bf16 [N, C, H, W] tensor")
    N, C, H, W = x.shape
    if dn.shape != (N, H * W, C):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "groupnorm_tokens_bwd: dn must be [N, H*W, C]")
    d2, lddn = _rows2d(dn, "dn")
    if gamma is not None and (gamma.dtype != torch.bfloat16 or not gamma.is_contiguous() or gamma.numel() != C):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "groupnorm_tokens_bwd: gamma must be a contiguous bf16 vector of length C")
    dx = torch.empty_like(x) if need_dx else None
    dgamma = torch.empty(C, dtype=torch.bfloat16, device=x.device) if need_affine else None
    dbeta = torch.empty(C, dtype=torch.bfloat16, device=x.device) if need_affine else None
    lib = _lib.load()
    ws = _workspace(lib.tvi_groupnorm_tokens_bwd_workspace_bytes(N, C, H * W, int(groups)), x.device)
    _hooked("groupnorm_tokens_bwd", 4 + int(need_dx) + int(need_affine), lib.tvi_groupnorm_tokens_bwd_bf16, x.data_ptr(), d2.data_ptr(),
            gamma.data_ptr() if gamma is not None else None, dx.data_ptr() if dx is not None else None,
            dgamma.data_ptr() if dgamma is not None else None, dbeta.data_ptr() if dbeta is not None else None,
            ws.data_ptr(), N, C, H * W, int(groups), lddn, float(eps), _stream())
    return dx, dgamma, dbeta


def blend_rows_bwd(dy: torch.Tensor, xs, xt, alpha: torch.Tensor, rows_per_group: int, need_dalpha: bool):
    """Backward of blend_rows: (dx_spatial, dx_temporal, dalpha [frames] fp32 or None); dy / xs / xt are [frames*rows, C]
    contiguous bf16."""
    _need_cuda(dy, xs, xt, alpha)
    C = dy.shape[-1]
    nrows = dy.numel() // C if C else 0
    frames = nrows // max(1, int(rows_per_group))
    for t, what in ((dy, "dy"), (xs, "x_spatial"), (xt, "x_temporal")):
        if t is not None and (t.dtype != torch.bfloat16 or not t.is_contiguous() or t.numel() != nrows * C):
            raise _lib.TviError(_lib.TVI_ERR_INVALID, f"blend_rows_bwd: {what} must be a contiguous bf16 tensor of dy's shape")
    if alpha.dtype != torch.float32 or not alpha.is_contiguous() or alpha.numel() != frames or frames * rows_per_group != nrows:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "blend_rows_bwd: alpha must be a contiguous fp32 vector with one entry per frame")
    dxs, dxt = torch.empty_like(dy), torch.empty_like(dy)
    lib = _lib.load()
    part = None
    if need_dalpha:
        part = torch.empty(frames, lib.tvi_blend_rows_bwd_blocks(int(rows_per_group), C), dtype=torch.float32, device=dy.device)
    _hooked("blend_rows_bwd", 1, lib.tvi_blend_rows_bwd_bf16, dy.data_ptr(), xs.data_ptr() if need_dalpha else None,
            xt.data_ptr() if need_dalpha else None, alpha.data_ptr(), dxs.data_ptr(), dxt.data_ptr(),
            part.data_ptr() if part is not None else None, frames, int(rows_per_group), C, C, _stream())
    return dxs, dxt, (part.sum(dim=1) if part is not None else None)


def tokens_to_nchw_add(y: torch.Tensor, residual: torch.Tensor | None, H: int, W: int, out: torch.Tensor | None = None) -> torch.Tensor:
    """[N, H*W, C] token-major -> NCHW [N, C, H, W] (+ residual of that shape)."""
    _need_cuda(y, residual)
    if y.dim() != 3 or y.shape[1] != H * W:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "tokens_to_nchw_add: y must be [N, H*W, C]")
    N, HW, C = y.shape
    y2, ldy = _rows2d(y, "y")
    if residual is not None and (residual.shape != (N, C, H, W) or residual.dtype != torch.bfloat16 or not residual.is_contiguous()):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "tokens_to_nchw_add: residual must be a contiguous bf16 [N, C, H, W] tensor")
    if out is None:
        out = torch.empty(N, C, H, W, dtype=torch.bfloat16, device=y.device)
    _hooked("tokens_to_nchw_add", 1, _lib.load().tvi_tokens_to_nchw_add_bf16, y2.data_ptr(),
            residual.data_ptr() if residual is not None else None, out.data_ptr(), N, C, HW, ldy, _stream())
    return out


def blend_rows(x_spatial: torch.Tensor, x_temporal: torch.Tensor, alpha: torch.Tensor, rows_per_group: int,
               out: torch.Tensor | None = None) -> torch.Tensor:
    """out[r] = alpha[r // rows_per_group] * x_spatial[r] + (1 - alpha[...]) * x_temporal[r]; alpha fp32."""
    _need_cuda(x_spatial, x_temporal, alpha)
    xs, lds = _rows2d(x_spatial, "x_spatial")
    xt, ldt = _rows2d(x_temporal, "x_temporal")
    C = xs.shape[-1]
    nrows = xs.numel() // C if C else 0
    if xt.shape != xs.shape:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "blend_rows: x_spatial and x_temporal differ in shape")
    if alpha.dtype != torch.float32 or not alpha.is_contiguous() or alpha.numel() * int(rows_per_group) != nrows:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "blend_rows: alpha must be a contiguous fp32 vector with one entry per row group")
    if out is None:
        out = torch.empty(x_spatial.shape, dtype=torch.bfloat16, device=xs.device)
    o2, ldo = _rows2d(out, "out")
    _hooked("blend_rows", 1, _lib.load().tvi_blend_rows_bf16, xs.data_ptr(), xt.data_ptr(), alpha.data_ptr(), o2.data_ptr(), nrows,
            int(rows_per_group), C, lds, ldt, ldo, _stream())
    return out


def farneback(prev: torch.Tensor, nxt: torch.Tensor, pyr_scale: float = 0.5, levels: int = 3, winsize: int = 15, iterations: int = 3,
              poly_n: int = 5, poly_sigma: float = 1.2, out: torch.Tensor | None = None) -> torch.Tensor:
    """Farneback dense flow of two uint8 CUDA images [H, W] (gray) or [H, W, 3] (BGR) -> [H, W, 2] fp32."""
    _need_cuda(prev, nxt)
    if prev.dtype != torch.uint8 or nxt.dtype != torch.uint8 or prev.shape != nxt.shape or prev.dim() not in (2, 3):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "farneback: expected two uint8 images of the same shape, [H, W] or [H, W, 3]")
    ch = 1 if prev.dim() == 2 else prev.shape[2]
    H, W = prev.shape[:2]
    prev, nxt = prev.contiguous(), nxt.contiguous()
    if out is None:
        out = torch.empty(H, W, 2, dtype=torch.float32, device=prev.device)
    ws = _workspace(_lib.load().tvi_farneback_workspace_bytes(H, W, float(pyr_scale), int(levels)), prev.device)
    _hooked("farneback", 26 * (levels + 1), _lib.load().tvi_farneback_f32, prev.data_ptr(), nxt.data_ptr(), ch, out.data_ptr(),
            ws.data_ptr() if ws is not None else None, H, W, float(pyr_scale), int(levels), int(winsize), int(iterations), int(poly_n),
            float(poly_sigma), _stream())
    return out


def layernorm_bwd(x: torch.Tensor, dy: torch.Tensor, gamma: torch.Tensor | None, eps: float, need_dx: bool = True,
                  need_affine: bool = True, dres: torch.Tensor | None = None):
    """Backward of layernorm_bf16: returns (dx or None, dgamma or None, dbeta or None), all bf16.  `dres` (shape of x): the
    gradient of a residual branch `x + f(LayerNorm(x))`, added to dx in the same pass."""
    _need_cuda(x, dy, gamma, dres)
    x, ldx = _rows2d(x, "x")
    dy, lddy = _rows2d(dy, "dy")
    C = x.shape[-1]
    M = x.numel() // C if C else 0
    lddres = 0
    if dres is not None:
        if not need_dx or dres.numel() != x.numel():
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "layernorm_bwd: dres needs need_dx and the shape of x")
        dres, lddres = _rows2d(dres.reshape(x.shape), "dres")
    dx = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device) if need_dx else None
    dgamma = torch.empty(C, dtype=torch.bfloat16, device=x.device) if need_affine else None
    dbeta = torch.empty(C, dtype=torch.bfloat16, device=x.device) if need_affine else None
    ws = _workspace(_lib.load().tvi_layernorm_bwd_workspace_bytes(M, C), x.device)
    lddx = _rows2d(dx, "dx")[1] if dx is not None else 0
    _hooked("layernorm_bwd", 2 if need_affine else 1, _lib.load().tvi_layernorm_bwd_res_bf16, x.data_ptr(), dy.data_ptr(),
            gamma.data_ptr() if gamma is not None else None, dres.data_ptr() if dres is not None else None,
            dx.data_ptr() if dx is not None else None, dgamma.data_ptr() if dgamma is not None else None,
            dbeta.data_ptr() if dbeta is not None else None, ws.data_ptr() if ws is not None else None, M, C, ldx, lddy, lddres,
            lddx, float(eps), _stream())
    return dx, dgamma, dbeta


def geglu_bwd(pre: torch.Tensor, dh: torch.Tensor, need_h: bool = True):
    """pre [M, 2*inner] = [value | gate], dh [M, inner] -> (h or None [M, inner], dpre [M, 2*inner])."""
    _need_cuda(pre, dh)
    pre, ldpre = _rows2d(pre, "pre")
    dh, lddh = _rows2d(dh, "dh")
    inner = dh.shape[-1]
    if pre.shape[-1] != 2 * inner:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "geglu_bwd: pre must be twice as wide as dh")
    M = dh.numel() // inner if inner else 0
    h = torch.empty(M, inner, dtype=torch.bfloat16, device=pre.device) if need_h else None
    dpre = torch.empty(M, 2 * inner, dtype=torch.bfloat16, device=pre.device)
    _hooked("geglu_bwd", 1, _lib.load().tvi_geglu_bwd_bf16, pre.data_ptr(), dh.data_ptr(), h.data_ptr() if h is not None else None,
            dpre.data_ptr(), M, inner, ldpre, lddh, inner, 2 * inner, _stream())
    return h, dpre


def gemm_geglu_bwd(x: torch.Tensor, wi: torch.Tensor, bias_i: torch.Tensor | None, dh: torch.Tensor, need_h: bool = True):
    """Backward of the fused GEGLU projection in one GEMM: x [M, K], wi [2*inner, K] / bias_i with value / gate rows
    interleaved in blocks of `geglu_block(inner)` (as `gemm_bf16_tn_ex(..., geglu=True)` takes them), dh [M, inner] ->
    (h [M, inner] or None, dpre [M, 2*inner] = [d value | d gate])."""
    _need_cuda(x, wi, dh)
    x, ldx = _rows2d(x, "x")
    wi, ldw = _rows2d(wi, "wi")
    dh, lddh = _rows2d(dh, "dh")
    inner, K = dh.shape[-1], x.shape[-1]
    if wi.shape[0] != 2 * inner or wi.shape[-1] != K:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm_geglu_bwd: wi must be [2 * inner, K]")
    if bias_i is not None:
        _need_cuda(bias_i)
        bias_i = bias_i.contiguous()
    M = x.numel() // K if K else 0
    if dh.numel() != M * inner:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "gemm_geglu_bwd: dh must have one row per row of x")
    h = torch.empty(M, inner, dtype=torch.bfloat16, device=x.device) if need_h else None
    dpre = torch.empty(M, 2 * inner, dtype=torch.bfloat16, device=x.device)
    _hooked("gemm_bf16_geglu_bwd", 1, _lib.load().tvi_gemm_bf16_geglu_bwd, x.data_ptr(), wi.data_ptr(),
            bias_i.data_ptr() if bias_i is not None else None, dh.data_ptr(), h.data_ptr() if h is not None else None,
            dpre.data_ptr(), M, inner, K, ldx, ldw, lddh, inner, 2 * inner, _stream())
    return h, dpre


def farneback_clip(frames: torch.Tensor, pyr_scale: float = 0.5, levels: int = 3, winsize: int = 15, iterations: int = 3,
                   poly_n: int = 5, poly_sigma: float = 1.2, out: torch.Tensor | None = None) -> torch.Tensor:
    """Flows of every consecutive frame pair of a uint8 CUDA clip [N, H, W] / [N, H, W, 3] -> [N - 1, H, W, 2] fp32; every
    frame's pyramid is built once."""
    _need_cuda(frames)
    if frames.dtype != torch.uint8 or frames.dim() not in (3, 4) or frames.shape[0] < 2:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "farneback_clip: expected a uint8 clip [N >= 2, H, W] or [N, H, W, 3]")
    ch = 1 if frames.dim() == 3 else frames.shape[3]
    N, H, W = frames.shape[:3]
    frames = frames.contiguous()
    if out is None:
        out = torch.empty(N - 1, H, W, 2, dtype=torch.float32, device=frames.device)
    ws = _workspace(_lib.load().tvi_farneback_workspace_bytes(H, W, float(pyr_scale), int(levels)), frames.device)
    _hooked("farneback_clip", 20 * (levels + 1) * (N - 1), _lib.load().tvi_farneback_clip_f32, frames.data_ptr(), None, N, ch,
            out.data_ptr(), ws.data_ptr() if ws is not None else None, H, W, float(pyr_scale), int(levels), int(winsize),
            int(iterations), int(poly_n), float(poly_sigma), _stream())
    return out


def gemm_bf16_dgrad(dy: torch.Tensor, w: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """dx[..., K] = dy[..., N] @ w[N, K] — input gradient of nn.Linear; w is consumed in place."""
    _need_cuda(dy, w)
    dy, lddy = _rows2d(dy, "dy")
    w, ldw = _rows2d(w, "w")
    N, K = w.shape
    if dy.shape[-1] != N:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, f"dgrad: dy has N={dy.shape[-1]}, w has N={N}")
    M = dy.numel() // N if N else 0
    if out is None:
        out = torch.empty(*dy.shape[:-1], K, dtype=torch.bfloat16, device=dy.device)
    out, lddx = _rows2d(out, "out")
    _hooked("gemm_bf16_dgrad", 1, _lib.load().tvi_gemm_bf16_dgrad, dy.data_ptr(), w.data_ptr(), out.data_ptr(), M, N, K,
            lddy, ldw, lddx, _stream())
    return out


def gemm_bf16_wgrad(dy: torch.Tensor, x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """dw[N, K] = dy[M, N]^T @ x[M, K] — weight gradient of nn.Linear (split token reduction)."""
    _need_cuda(dy, x)
    dy, lddy = _rows2d(dy, "dy")
    x, ldx = _rows2d(x, "x")
    N, K = dy.shape[-1], x.shape[-1]
    M = dy.numel() // N if N else 0
    if x.numel() != M * K:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "wgrad: dy and x have different row counts")
    if out is None:
        out = torch.empty(N, K, dtype=torch.bfloat16, device=dy.device)
    out, lddw = _rows2d(out, "out")
    lib = _lib.load()
    ws = _workspace(lib.tvi_gemm_bf16_wgrad_workspace_bytes(M, N, K), dy.device)
    _hooked("gemm_bf16_wgrad", 2, lib.tvi_gemm_bf16_wgrad, dy.data_ptr(), x.data_ptr(), out.data_ptr(),
            ws.data_ptr() if ws is not None else None, M, N, K, lddy, ldx, lddw, _stream())
    return out


def colsum_bf16(x: torch.Tensor, rows_per_group: int | None = None, out: torch.Tensor | None = None,
                group_inner: int = 1) -> torch.Tensor:
    """out[g, :] = sum of the rows of group g of x[nrows, C] (fp32 accumulation, bf16 result).  group_inner > 1: interleaved
    groups (logical row i of group (b, s) is physical row (b * rows_per_group + i) * group_inner + s)."""
    _need_cuda(x)
    x, ldx = _rows2d(x, "x")
    C = x.shape[-1]
    nrows = x.numel() // C if C else 0
    rpg = nrows if rows_per_group is None else int(rows_per_group)
    ngroups = nrows // rpg if rpg else 0
    if out is None:
        out = torch.empty(ngroups, C, dtype=torch.bfloat16, device=x.device)
    out, ldo = _rows2d(out, "out")
    lib = _lib.load()
    nb = lib.tvi_colsum_bf16_workspace_bytes(nrows, rpg, C)
    ws = _workspace(nb, x.device)
    _hooked("colsum_bf16", 2 if nb else 1, lib.tvi_colsum_bf16_interleaved, x.data_ptr(), out.data_ptr(),
            ws.data_ptr() if ws is not None else None, nrows, rpg, int(group_inner), C, ldx, ldo, _stream())
    return out


def attn_temporal(q, k, v, nseq: int, heads: int, Tq: int, Tk: int, scale: float, out=None, lse=None, seq_inner: int = 1):
    """Short-sequence attention core; q/k/v are [nseq*T, heads*64] row views (may be strided).  seq_inner > 1: the
    sequences are interleaved (token t of sequence s is row (s // inner) * inner * T + t * inner + s % inner)."""
    global LAUNCHES
    _need_cuda(q, k, v)
    q, ldq = _rows2d(q, "q")
    k, ldk = _rows2d(k, "k")
    v, ldv = _rows2d(v, "v")
    d = q.shape[-1] // heads
    if out is None:
        out = torch.empty(nseq * Tq, heads * d, dtype=torch.bfloat16, device=q.device)
    out, ldo = _rows2d(out, "out")
    if EVENT_HOOK is not None:
        EVENT_HOOK("attn_temporal", True)
    _lib.check(_lib.load().tvi_attn_temporal_fwd_interleaved(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(),
                                                             lse.data_ptr() if lse is not None else None, nseq, heads, Tq, Tk,
                                                             d, int(seq_inner), ldq, ldk, ldv, ldo, float(scale), _stream()))
    if EVENT_HOOK is not None:
        EVENT_HOOK("attn_temporal", False)
    LAUNCHES += 1
    return out


def attn_fewkeys(q, k, v, nrows: int, rows_per_group: int, heads: int, L: int, scale: float, out=None,
                 Tq: int = 1, bias=None, stats=None, group_inner: int = 1):
    """Few-key cross-attention core; k/v are [ngroups*L, heads*d]; q may be None when L == 1.  group_inner > 1: the
    groups are interleaved (logical row i of group (b, s) is physical row (b * rows_per_group + i) * group_inner + s)."""
    global LAUNCHES
    _need_cuda(q, k, v, bias, stats)
    k, ldk = _rows2d(k, "k")
    v, ldv = _rows2d(v, "v")
    C_ = k.shape[-1]
    d = C_ // heads
    ldq = 0
    if q is not None:
        q, ldq = _rows2d(q, "q")
    if out is None:
        out = torch.empty(nrows, C_, dtype=torch.bfloat16, device=k.device)
    out, ldo = _rows2d(out, "out")
    ws = None
    if stats is not None:
        if stats.dtype != torch.float32 or not stats.is_contiguous() or stats.numel() != Tq * L:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "stats must be a contiguous fp32 [Tq, L] tensor")
        ws = torch.empty(Tq * L, dtype=torch.float64, device=k.device)
    if bias is not None and (bias.dtype != torch.float32 or not bias.is_contiguous() or bias.numel() != Tq * L):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "bias must be a contiguous fp32 [Tq, L] tensor")
    if EVENT_HOOK is not None:
        EVENT_HOOK("attn_fewkeys", True)
    _lib.check(_lib.load().tvi_attn_fewkeys_fwd_interleaved(
        q.data_ptr() if q is not None else None, k.data_ptr(), v.data_ptr(), out.data_ptr(), nrows, rows_per_group, heads, L, d,
        int(group_inner), ldq, ldk, ldv, ldo, float(scale), Tq, bias.data_ptr() if bias is not None else None,
        stats.data_ptr() if stats is not None else None, ws.data_ptr() if ws is not None else None, _stream()))
    if EVENT_HOOK is not None:
        EVENT_HOOK("attn_fewkeys", False)
    LAUNCHES += 1 if stats is None else 2
    return out


def attn_spatial(q, k, v, batch: int, heads: int, S: int, Skv: int, scale: float, out=None, lse=None,
                 algo: int = _lib.SPATIAL_AUTO):
    """Flash-style spatial attention core; q is [batch*S, heads*64], k/v [batch*Skv, heads*64].
    `algo` picks the kernel (_lib.SPATIAL_AUTO / SPATIAL_MMA_SYNC / SPATIAL_TCGEN05)."""
    global LAUNCHES
    _need_cuda(q, k, v)
    q, ldq = _rows2d(q, "q")
    k, ldk = _rows2d(k, "k")
    v, ldv = _rows2d(v, "v")
    d = q.shape[-1] // heads
    if out is None:
        out = torch.empty(batch * S, heads * d, dtype=torch.bfloat16, device=q.device)
    out, ldo = _rows2d(out, "out")
    if EVENT_HOOK is not None:
        EVENT_HOOK("attn_spatial", True)
    _lib.check(_lib.load().tvi_attn_spatial_fwd_algo(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(),
                                                     lse.data_ptr() if lse is not None else None, batch, heads, S, Skv,
                                                     d, ldq, ldk, ldv, ldo, float(scale), int(algo), _stream()))
    if EVENT_HOOK is not None:
        EVENT_HOOK("attn_spatial", False)
    LAUNCHES += 1
    return out


def _grad_like(t, rows, width):
    return torch.empty(rows, width, dtype=torch.bfloat16, device=t.device)


def attn_spatial_bwd(q, k, v, o, dout, lse, batch: int, heads: int, S: int, Skv: int, scale: float, dq=None, dk=None, dv=None,
                     algo: int = _lib.SPATIAL_AUTO):
    """Gradients of the flash spatial core; q/k/v/o/dout are [rows, heads*64] row views, lse the forward's
    [batch*S, heads] fp32 log-sum-exp.  Returns (dq, dk, dv).  `algo` as for the forward (AUTO: tcgen05 passes for
    S, Skv >= 1024, mma.sync passes below)."""
    _need_cuda(q, k, v, o, dout, lse)
    q, ldq = _rows2d(q, "q")
    k, ldk = _rows2d(k, "k")
    v, ldv = _rows2d(v, "v")
    o, ldo = _rows2d(o, "o")
    dout, lddo = _rows2d(dout, "dout")
    C_ = q.shape[-1]
    d = C_ // heads
    dq = _grad_like(q, batch * S, C_) if dq is None else dq
    dk = _grad_like(q, batch * Skv, C_) if dk is None else dk
    dv = _grad_like(q, batch * Skv, C_) if dv is None else dv
    dq, lddq = _rows2d(dq, "dq")
    dk, lddk = _rows2d(dk, "dk")
    dv, lddv = _rows2d(dv, "dv")
    if lse.dtype != torch.float32 or not lse.is_contiguous() or lse.numel() != batch * S * heads:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "lse must be a contiguous fp32 [batch*S, heads] tensor")
    lib = _lib.load()
    ws = _workspace(lib.tvi_attn_spatial_bwd_workspace_bytes(batch, heads, S), q.device)
    _hooked("attn_spatial_bwd", 3, lib.tvi_attn_spatial_bwd_algo, q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(),
            dout.data_ptr(), lse.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), ws.data_ptr(), batch, heads, S, Skv,
            d, ldq, ldk, ldv, ldo, lddo, lddq, lddk, lddv, float(scale), int(algo), _stream())
    return dq, dk, dv


def attn_temporal_bwd(q, k, v, dout, nseq: int, heads: int, Tq: int, Tk: int, scale: float, dq=None, dk=None, dv=None,
                      seq_inner: int = 1):
    """Gradients of the short-sequence core (recomputes the T x T softmax; needs neither o nor lse); seq_inner as the forward."""
    _need_cuda(q, k, v, dout)
    q, ldq = _rows2d(q, "q")
    k, ldk = _rows2d(k, "k")
    v, ldv = _rows2d(v, "v")
    dout, lddo = _rows2d(dout, "dout")
    C_ = q.shape[-1]
    d = C_ // heads
    dq = _grad_like(q, nseq * Tq, C_) if dq is None else dq
    dk = _grad_like(q, nseq * Tk, C_) if dk is None else dk
    dv = _grad_like(q, nseq * Tk, C_) if dv is None else dv
    dq, lddq = _rows2d(dq, "dq")
    dk, lddk = _rows2d(dk, "dk")
    dv, lddv = _rows2d(dv, "dv")
    _hooked("attn_temporal_bwd", 1, _lib.load().tvi_attn_temporal_bwd_interleaved, q.data_ptr(), k.data_ptr(), v.data_ptr(),
            dout.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), nseq, heads, Tq, Tk, d, int(seq_inner), ldq, ldk, ldv,
            lddo, lddq, lddk, lddv, float(scale), _stream())
    return dq, dk, dv


def attn_fewkeys_bwd(q, k, v, dout, nrows: int, rows_per_group: int, heads: int, L: int, scale: float, Tq: int = 1, bias=None,
                     dq=None, dk=None, dv=None, group_inner: int = 1):
    """Gradients of the few-keys core; returns (dq or None, dk, dv) with dk/dv [ngroups*L, heads*d]; group_inner as the forward."""
    _need_cuda(q, k, v, dout, bias)
    k, ldk = _rows2d(k, "k")
    v, ldv = _rows2d(v, "v")
    dout, lddo = _rows2d(dout, "dout")
    C_ = k.shape[-1]
    d = C_ // heads
    ngroups = nrows // rows_per_group
    ldq = lddq = 0
    if q is not None:
        q, ldq = _rows2d(q, "q")
        dq = _grad_like(k, nrows, C_) if dq is None else dq
        dq, lddq = _rows2d(dq, "dq")
    dk = _grad_like(k, ngroups * L, C_) if dk is None else dk
    dv = _grad_like(k, ngroups * L, C_) if dv is None else dv
    dk, lddk = _rows2d(dk, "dk")
    dv, lddv = _rows2d(dv, "dv")
    lib = _lib.load()
    nb = lib.tvi_attn_fewkeys_bwd_workspace_bytes(nrows, rows_per_group, heads, L, d)
    ws = _workspace(nb, k.device)
    _hooked("attn_fewkeys_bwd", 2 if nb else 1, lib.tvi_attn_fewkeys_bwd_interleaved, q.data_ptr() if q is not None else None,
            k.data_ptr(), v.data_ptr(), dout.data_ptr(), dq.data_ptr() if q is not None else None, dk.data_ptr(),
            dv.data_ptr(), ws.data_ptr() if ws is not None else None, nrows, rows_per_group, heads, L, d, int(group_inner), ldq,
            ldk, ldv, lddo, lddq, lddk, lddv, float(scale), Tq, bias.data_ptr() if bias is not None else None, _stream())
    return (dq if q is not None else None), dk, dv


def slerp_f32(v0: torch.Tensor, v1: torch.Tensor, T: int, dot_threshold: float, out=None) -> torch.Tensor:
    global LAUNCHES
    _need_cuda(v0, v1)
    if v0.dtype != torch.float32 or v1.dtype != torch.float32 or v0.shape != v1.shape:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "slerp: v0 and v1 must be fp32 tensors of the same shape")
    v0 = v0.contiguous()
    v1 = v1.contiguous()
    nclips = v0.shape[0] if v0.dim() > 0 else 1
    n = v0.numel() // nclips if nclips else 0
    if out is None:
        out = torch.empty((nclips, T) + tuple(v0.shape[1:]), dtype=torch.float32, device=v0.device)
    if EVENT_HOOK is not None:
        EVENT_HOOK("slerp", True)
    _lib.check(_lib.load().tvi_slerp_f32(v0.data_ptr(), v1.data_ptr(), out.data_ptr(), nclips, n, T,
                                         float(dot_threshold), _stream()))
    if EVENT_HOOK is not None:
        EVENT_HOOK("slerp", False)
    LAUNCHES += 1
    return out


def flow_cosine_f32(fa: torch.Tensor, fb: torch.Tensor, hwc: bool, out=None) -> torch.Tensor:
    """fa, fb: [npairs, H, W, 2] (hwc) or [npairs, 2, H, W] fp32 -> [npairs] fp32."""
    global LAUNCHES
    _need_cuda(fa, fb)
    if fa.dtype != torch.float32 or fb.dtype != torch.float32 or fa.shape != fb.shape or fa.dim() != 4:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "flow_cosine: expected two fp32 4-D tensors of the same shape")
    fa = fa.contiguous()
    fb = fb.contiguous()
    npairs = fa.shape[0]
    H, W = (fa.shape[1], fa.shape[2]) if hwc else (fa.shape[2], fa.shape[3])
    if (fa.shape[3] if hwc else fa.shape[1]) != 2:
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "flow_cosine: a flow field has exactly 2 components")
    lib = _lib.load()
    ws = torch.empty(max(1, lib.tvi_flow_cosine_workspace_bytes(npairs, H, W) // 8), dtype=torch.float64, device=fa.device)
    if out is None:
        out = torch.empty(npairs, dtype=torch.float32, device=fa.device)
    if EVENT_HOOK is not None:
        EVENT_HOOK("flow_cosine", True)
    _lib.check(lib.tvi_flow_cosine_f32(fa.data_ptr(), fb.data_ptr(), out.data_ptr(), ws.data_ptr(), npairs, H, W,
                                       1 if hwc else 0, _stream()))
    if EVENT_HOOK is not None:
        EVENT_HOOK("flow_cosine", False)
    LAUNCHES += 2
    return out


def backward_warp_f32(img2, f_fwd, f_bwd, alpha: float, beta: float, warped=None, mask=None):
    """img2 [B,C,H,W], f_fwd/f_bwd [B,2,H,W] fp32 -> (warped [B,C,H,W] fp32, mask [B,H,W] u8)."""
    global LAUNCHES
    _need_cuda(img2, f_fwd, f_bwd)
    for t in (img2, f_fwd, f_bwd):
        if t.dtype != torch.float32 or t.dim() != 4:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "backward_warp: expected fp32 [B,C,H,W] tensors")
    B, Cc, H, W = img2.shape
    if tuple(f_fwd.shape) != (B, 2, H, W) or tuple(f_bwd.shape) != (B, 2, H, W):
        raise _lib.TviError(_lib.TVI_ERR_INVALID, "backward_warp: flows must be [B,2,H,W] matching img2")
    img2, f_fwd, f_bwd = img2.contiguous(), f_fwd.contiguous(), f_bwd.contiguous()
    if warped is None:
        warped = torch.empty_like(img2)
    if mask is None:
        mask = torch.empty(B, H, W, dtype=torch.uint8, device=img2.device)
    if EVENT_HOOK is not None:
        EVENT_HOOK("backward_warp", True)
    _lib.check(_lib.load().tvi_backward_warp_f32(img2.data_ptr(), f_fwd.data_ptr(), f_bwd.data_ptr(), warped.data_ptr(),
                                                 mask.data_ptr(), B, Cc, H, W, float(alpha), float(beta), _stream()))
    if EVENT_HOOK is not None:
        EVENT_HOOK("backward_warp", False)
    LAUNCHES += 1
    return warped, mask
