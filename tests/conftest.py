import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("this test is marked gpu but no CUDA device is visible")
    return torch.device("cuda:0")


def block_output_tolerance(ref_max: float, factor: float = 1.0, stores: int = 1) -> float:
    """North star (BASELINE.json): bf16 attention-block outputs within 2e-2 MAX-ABS of the fp32 oracle.  That absolute
    bound is what the block / module tests assert for the COMPUTATION.  On top of it comes only what no bf16 path can
    avoid and the fp32 oracle does not pay: the storage grid of the residual stream, half a bf16 ulp at the largest output
    (2^-9 |y| rounded down to a power of two: 0.004 below 2, 0.0156 at [4, 8), 0.031 at [8, 16)) for each time the stream
    is stored in bf16 on the way (`stores`: 1 for a single sub-layer, 3 for a spatial block = attn1 / attn2 / ff each
    followed by its residual add, 4 for a temporal block, ...).  Residual-carrying block outputs reach |y| ~ 6-8 on N(0, 1)
    inputs, where one grid term alone is most of the 2e-2."""
    import math
    tol = 2e-2 * factor
    if ref_max > 0.0:
        tol += stores * 2.0 ** (math.floor(math.log2(ref_max)) - 8)  # half a bf16 ulp at the largest output, per store
    return tol


def assert_block_close(name: str, y, ref, factor: float = 1.0, stores: int = 1) -> float:
    """Asserts the absolute bound above and prints the scale-relative figure next to it."""
    err = float((y.float().cpu() - ref.float()).abs().max())
    ref_max = float(ref.abs().max())
    tol = block_output_tolerance(ref_max, factor, stores)
    print(f"{name}: max abs err {err:.3e} (bound {tol:.3e}), max |ref| {ref_max:.3f}, "
          f"relative to max(1, |ref|) {err / max(1.0, ref_max):.3e}")
    assert err <= tol, f"{name}: {err:.3e} > {tol:.3e}"
    return err
