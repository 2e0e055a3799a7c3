"""ctypes binding of libtvi.so (the C ABI declared in include/tvi.h).

This is the only place Python touches native code.  There is NO fallback: if the shared library
is missing or a call fails, a `TviError` is raised (north star: "no CPU fallback").
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("TVI_LIB", _PKG / "libtvi.so"))

TVI_OK = 0
TVI_ERR_INVALID = -1
TVI_ERR_CUDA = -2
TVI_ERR_UNSUPPORTED = -3


class TviError(RuntimeError):
    """A libtvi call returned a non-zero status; `.code` is the TVI_ERR_* value."""

    def __init__(self, code: int, message: str):
        super().__init__(f"libtvi error {code}: {message}")
        self.code = code


_i64, _i32, _f32, _f64, _ptr, _sz = C.c_int64, C.c_int, C.c_float, C.c_double, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); mirrors include/tvi.h one to one (checked by tests/test_abi.py)
SIGNATURES = {
    "tvi_version": (C.c_char_p, []),
    "tvi_last_error": (C.c_char_p, []),
    "tvi_device_info": (_i32, [C.POINTER(_i32)] * 3),
    "tvi_slerp_f32": (_i32, [_ptr, _ptr, _ptr, _i64, _i64, _i32, _f32, _ptr]),
    "tvi_flow_cosine_workspace_bytes": (_sz, [_i64, _i32, _i32]),
    "tvi_flow_cosine_f32": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i32, _ptr]),
    "tvi_backward_warp_f32": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i32, _f32, _f32, _ptr]),
    "tvi_gemm_bf16_tn": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i64, _i64, _i64, _ptr]),
    "tvi_geglu_block": (_i32, [_i32]),
    "tvi_layernorm_stats_supported": (_i32, [_i32]),
    "tvi_layernorm_stats_f32": (_i32, [_ptr, _ptr, _i64, _i32, _i64, _f32, _ptr]),
    "tvi_gemm_bf16_tn_ln": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i64, _i64, _i64, _i64, _i32, _ptr]),
    "tvi_gemm_bf16_tn_ex": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i64, _i64, _i64, _i64, _i32, _ptr]),
    "tvi_layernorm_bf16": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i32, _i64, _i64, _f32, _ptr]),
    "tvi_bcast_rows_add_bf16": (_i32, [_ptr, _ptr, _ptr, _i64, _i64, _i64, _i32, _i64, _i64, _i64, _ptr]),
    "tvi_groupnorm_tokens_workspace_bytes": (_sz, [_i64, _i32, _i32, _i32]),
    "tvi_groupnorm_tokens_bf16": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i32, _i64, _f32, _ptr]),
    "tvi_tokens_to_nchw_add_bf16": (_i32, [_ptr, _ptr, _ptr, _i64, _i32, _i32, _i64, _ptr]),
    "tvi_nchw_to_tokens_bf16": (_i32, [_ptr, _ptr, _i64, _i32, _i32, _i64, _ptr]),
    "tvi_groupnorm_tokens_bwd_workspace_bytes": (_sz, [_i64, _i32, _i32, _i32]),
    "tvi_groupnorm_tokens_bwd_bf16": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i32, _i64, _f32, _ptr]),
    "tvi_blend_rows_bwd_blocks": (_i32, [_i64, _i32]),
    "tvi_blend_rows_bwd_bf16": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i64, _i32, _i64, _ptr]),
    "tvi_blend_rows_bf16": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i64, _i32, _i64, _i64, _i64, _ptr]),
    "tvi_cfg_euler_step_f32": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i32, _i64, _i64, _f32, _f32, _i32, _ptr]),
    "tvi_cfg_euler_step_round_f32": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i32, _i64, _i64, _f32, _f32, _i32, _i32, _ptr]),
    "tvi_model_input_bf16": (_i32, [_ptr, _ptr, _ptr, _i64, _i32, _i32, _i64, _f32, _ptr]),
    "tvi_farneback_workspace_bytes": (_sz, [_i32, _i32, _f64, _i32]),
    "tvi_farneback_clip_f32": (_i32, [_ptr, _ptr, _i64, _i32, _ptr, _ptr, _i32, _i32, _f64, _i32, _i32, _i32, _i32, _f64, _ptr]),
    "tvi_farneback_f32": (_i32, [_ptr, _ptr, _i32, _ptr, _ptr, _i32, _i32, _f64, _i32, _i32, _i32, _i32, _f64, _ptr]),
    "tvi_layernorm_bwd_workspace_bytes": (_sz, [_i64, _i32]),
    "tvi_layernorm_bwd_res_bf16": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i64, _i64, _i64, _i64, _f32, _ptr]),
    "tvi_layernorm_bwd_bf16": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i64, _i64, _i64, _f32, _ptr]),
    "tvi_geglu_bwd_bf16": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i32, _i64, _i64, _i64, _i64, _ptr]),
    "tvi_gemm_bf16_geglu_bwd": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i64, _i64, _i64, _i64, _i64, _ptr]),
    "tvi_gemm_bf16_dgrad": (_i32, [_ptr, _ptr, _ptr, _i64, _i32, _i32, _i64, _i64, _i64, _ptr]),
    "tvi_gemm_bf16_wgrad_workspace_bytes": (_sz, [_i64, _i32, _i32]),
    "tvi_gemm_bf16_wgrad": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i64, _i64, _i64, _ptr]),
    "tvi_colsum_bf16_workspace_bytes": (_sz, [_i64, _i64, _i32]),
    "tvi_colsum_bf16": (_i32, [_ptr, _ptr, _ptr, _i64, _i64, _i32, _i64, _i64, _ptr]),
    "tvi_colsum_bf16_interleaved": (_i32, [_ptr, _ptr, _ptr, _i64, _i64, _i64, _i32, _i64, _i64, _ptr]),
    "tvi_attn_temporal_fwd": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i32, _i32,
                                     _i64, _i64, _i64, _i64, _f32, _ptr]),
    "tvi_attn_temporal_fwd_interleaved": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _i64, _i32, _i32, _i32, _i32, _i64,
                                                 _i64, _i64, _i64, _i64, _f32, _ptr]),
    "tvi_attn_fewkeys_fwd": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i64, _i32, _i32, _i32,
                                    _i64, _i64, _i64, _i64, _f32, _i32, _ptr, _ptr, _ptr, _ptr]),
    "tvi_attn_fewkeys_fwd_static": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i64, _i32, _i32, _i32,
                                           _i64, _i64, _i64, _i64, _f32, _i32, _ptr, _ptr, _ptr]),
    "tvi_attn_fewkeys_fwd_interleaved": (_i32, [_ptr, _ptr, _ptr, _ptr, _i64, _i64, _i32, _i32, _i32, _i64,
                                                _i64, _i64, _i64, _i64, _f32, _i32, _ptr, _ptr, _ptr, _ptr]),
    "tvi_attn_spatial_fwd": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _i32, _i32, _i32,
                                    _i64, _i64, _i64, _i64, _f32, _ptr]),
    "tvi_attn_spatial_fwd_algo": (_i32, [_ptr, _ptr, _ptr, _ptr, _ptr, _i32, _i32, _i32, _i32, _i32,
                                         _i64, _i64, _i64, _i64, _f32, _i32, _ptr]),
    "tvi_attn_spatial_bwd_workspace_bytes": (_sz, [_i32, _i32, _i32]),
    "tvi_attn_spatial_bwd": (_i32, [_ptr] * 10 + [_i32] * 5 + [_i64] * 8 + [_f32, _ptr]),
    "tvi_attn_spatial_bwd_algo": (_i32, [_ptr] * 10 + [_i32] * 5 + [_i64] * 8 + [_f32, _i32, _ptr]),
    "tvi_attn_temporal_bwd": (_i32, [_ptr] * 7 + [_i64, _i32, _i32, _i32, _i32] + [_i64] * 7 + [_f32, _ptr]),
    "tvi_attn_temporal_bwd_interleaved": (_i32, [_ptr] * 7 + [_i64, _i32, _i32, _i32, _i32] + [_i64] * 8 + [_f32, _ptr]),
    "tvi_attn_fewkeys_bwd_workspace_bytes": (_sz, [_i64, _i64, _i32, _i32, _i32]),
    "tvi_attn_fewkeys_bwd": (_i32, [_ptr] * 8 + [_i64, _i64, _i32, _i32, _i32] + [_i64] * 7 + [_f32, _i32, _ptr, _ptr]),
    "tvi_attn_fewkeys_bwd_interleaved": (_i32, [_ptr] * 8 + [_i64, _i64, _i32, _i32, _i32] + [_i64] * 8 + [_f32, _i32, _ptr, _ptr]),
}
SPATIAL_AUTO, SPATIAL_MMA_SYNC, SPATIAL_TCGEN05 = 0, 1, 2
EPILOGUE_NONE, EPILOGUE_GEGLU = 0, 1
FEWKEYS_STATIC_MAX_STATS = 4096  # TVI_FEWKEYS_STATIC_MAX_STATS of include/tvi.h
PRED_EPSILON, PRED_V = 0, 1
LATENT_ROUND_NONE, LATENT_ROUND_BF16 = 0, 1

_lib = None


def load() -> C.CDLL:
    """dlopen libtvi.so once and attach prototypes.  Works on a machine without a GPU."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise TviError(TVI_ERR_UNSUPPORTED,
                       f"{LIB_PATH} is missing - build it with `python -m timelapse_video_interpolation_b200.build` "
                       "(there is no CPU fallback)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != TVI_OK:
        raise TviError(rc, load().tvi_last_error().decode("utf-8", "replace"))


def version() -> str:
    return load().tvi_version().decode()


def device_info() -> tuple[int, int, int]:
    a, b, c = _i32(), _i32(), _i32()
    check(load().tvi_device_info(C.byref(a), C.byref(b), C.byref(c)))
    return a.value, b.value, c.value
