"""Random image sizes / parameters for the Farneback kernels against the numpy restatement (run under `timeout`)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
from oracle import farneback_ref as fr
from timelapse_video_interpolation_b200 import ops
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
worst = 0.0
for case in range(n):
    h, w = int(rng.integers(1, 260)), int(rng.integers(1, 300))
    levels = int(rng.integers(0, 5))
    winsize = int(rng.choice([5, 9, 15, 21]))
    iterations = int(rng.integers(1, 4))
    poly_n, poly_sigma = ((5, 1.2), (7, 1.5), (3, 0.9))[int(rng.integers(0, 3))]
    pyr = float(rng.choice([0.5, 0.7]))
    base = np.kron(rng.random((h // 8 + 2, w // 8 + 2)), np.ones((8, 8)))[:h, :w]
    a = np.clip(base * 220 + 20 * rng.random((h, w)), 0, 255).astype(np.uint8)
    shift = (int(rng.integers(-2, 3)), int(rng.integers(-3, 4)))
    if shift == (0, 0):
        # identical frames are a degenerate input: the flow is ~1e-8 at the border and FarnebackUpdateMatrices takes
        # floor(x + dx) there, so a last-bit difference flips the in-bounds branch (OpenCV itself is compiler-dependent here)
        shift = (1, 0)
    b = np.roll(a, shift, axis=(0, 1))
    print(f"case {case}: {h}x{w} levels={levels} pyr={pyr} win={winsize} it={iterations} poly={poly_n}", flush=True)
    got = ops.farneback(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), pyr, levels, winsize, iterations, poly_n, poly_sigma).cpu().numpy()
    want = fr.calc_optical_flow_farneback(a, b, pyr, levels, winsize, iterations, poly_n, poly_sigma)
    if min(h, w) < 12:
        # a side shorter than the 11-tap polynomial window: the problem is degenerate (the restatement and OpenCV itself differ
        # by 0.3 px on a 4-row image); only require a finite result of the right shape
        assert got.shape == want.shape and np.isfinite(got).all()
        continue
    d = np.abs(got - want).max(-1)
    err = float(d.max())
    tol = 5e-3 * max(1.0, float(np.abs(want).max()))
    # FarnebackUpdateMatrices branches on floor(x + dx) at the image border: a last-bit difference in dx flips single pixels
    # there, so the bulk (99.5th percentile) is held to the tolerance and isolated pixels to 10x
    p995 = float(np.quantile(d, 0.995))
    worst = max(worst, p995 / tol)
    if not (got.shape == want.shape and p995 <= tol and err <= 10 * tol):
        iy, ix = np.unravel_index(d.argmax(), d.shape)
        print(f"FAIL: max err {err:.3e} at ({iy}, {ix}) p99.5 {p995:.3e} tol {tol:.3e}")
        sys.exit(1)
print(f"farneback soak: {n} random cases OK, worst err/tol {worst:.3f}")
