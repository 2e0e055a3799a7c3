"""CPU: libtvi.so loads without a GPU and exports every symbol include/tvi.h declares; the ctypes
prototypes in _lib.SIGNATURES cover exactly the header's functions."""
import ctypes
import re
from pathlib import Path

import pytest

from timelapse_video_interpolation_b200 import _lib

ROOT = Path(__file__).resolve().parents[1]


def header_functions():
    text = (ROOT / "include" / "tvi.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tvi_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_functions():
    names = header_functions()
    assert "tvi_gemm_bf16_tn" in names and "tvi_attn_spatial_fwd" in names and len(names) >= 11


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    for name in header_functions():
        assert hasattr(lib, name), f"{name} declared in include/tvi.h but not exported by libtvi.so"


def test_ctypes_signatures_match_header():
    assert sorted(_lib.SIGNATURES) == header_functions()
    # argument counts agree with the C prototypes
    text = re.sub(r"/\*.*?\*/", "", (ROOT / "include" / "tvi.h").read_text(), flags=re.S)
    for name, (_, args) in _lib.SIGNATURES.items():
        m = re.search(rf"\b{name}\s*\(([^)]*)\)", text)
        params = [p for p in m.group(1).split(",") if p.strip() and p.strip() != "void"]
        assert len(params) == len(args), name


def test_version_and_error_string_without_gpu():
    assert "sm_100a" in _lib.version()
    lib = _lib.load()
    # argument validation happens before any CUDA call, so it is testable on CPU
    rc = lib.tvi_slerp_f32(None, None, None, 1, 16, 14, 0.9995, None)
    assert rc == _lib.TVI_ERR_INVALID and b"null pointer" in lib.tvi_last_error()
    rc = lib.tvi_attn_temporal_fwd(1, 1, 1, 1, None, 1, 1, 14, 14, 96, 64, 64, 64, 64, 0.125, None)
    assert rc == _lib.TVI_ERR_INVALID or rc == _lib.TVI_ERR_UNSUPPORTED
    rc = lib.tvi_gemm_bf16_tn(16, 16, None, 16, 4, 12, 64, 64, 64, 16, None)
    assert rc == _lib.TVI_ERR_INVALID and b"multiples of 8" in lib.tvi_last_error()
    with pytest.raises(_lib.TviError):
        _lib.check(rc)
    # empty problems are a no-op success
    assert lib.tvi_slerp_f32(None, None, None, 0, 16, 14, 0.9995, None) == 0
    assert lib.tvi_gemm_bf16_tn(None, None, None, None, 0, 8, 8, 8, 8, 8, None) == 0
