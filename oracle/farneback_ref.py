"""CPU restatement of cv2.calcOpticalFlowFarneback as called by the reference (test infrastructure).

  call site     dataset_construction/step4_flow_check.py:13-22
                (pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5, poly_sigma=1.2, flags=0; BGR -> gray first)
  algorithm     OpenCV 4.x modules/video/src/optflowgf.cpp (third-party, NOT in the reference tree; version unpinned by the
                reference, 4.13.0 in this image): FarnebackPrepareGaussian / FarnebackPolyExp / FarnebackUpdateMatrices /
                FarnebackUpdateFlow_Blur and the pyramid loop of FarnebackOpticalFlowImpl::calc, restated in numpy.
Pinned by tests/golden/farneback.npz = outputs of cv2.calcOpticalFlowFarneback itself (tests/golden/make_golden.py:
make_farneback); tolerance stated in tests/test_oracle_golden.py (OpenCV accumulates in float SIMD, this file in double).
"""
from __future__ import annotations

import numpy as np

BORDER = np.array([0.14, 0.14, 0.4472, 0.4472, 0.4472], dtype=np.float32)


def cv_round(v: float) -> int:
    return int(np.rint(v))  # round-half-to-even, as cvRound


def gray_from_bgr(img: np.ndarray) -> np.ndarray:
    """cv2.cvtColor(BGR2GRAY) for uint8 (OpenCV 4.x): 15-bit fixed point (B*3735 + G*19235 + R*9798 + 16384) >> 15."""
    b, g, r = (img[..., i].astype(np.int64) for i in range(3))
    return ((b * 3735 + g * 19235 + r * 9798 + 16384) >> 15).astype(np.uint8)


def gaussian_kernel(ksize: int, sigma: float) -> np.ndarray:
    """cv::getGaussianKernel(ksize, sigma, CV_32F)."""
    small = {1: [1.0], 3: [0.25, 0.5, 0.25], 5: [0.0625, 0.25, 0.375, 0.25, 0.0625],
             7: [0.03125, 0.109375, 0.21875, 0.28125, 0.21875, 0.109375, 0.03125]}
    if sigma <= 0 and ksize in small:
        return np.array(small[ksize], dtype=np.float32)
    s = sigma if sigma > 0 else ((ksize - 1) * 0.5 - 1) * 0.3 + 0.8
    x = np.arange(ksize, dtype=np.float64) - (ksize - 1) * 0.5
    k = np.exp(-0.5 / (s * s) * x * x)
    return (k / k.sum()).astype(np.float32)


def _reflect101(idx: np.ndarray, n: int) -> np.ndarray:
    if n == 1:
        return np.zeros_like(idx)
    p = 2 * (n - 1)
    idx = np.mod(idx, p)
    return np.where(idx >= n, p - idx, idx)


def gaussian_blur(img: np.ndarray, ksize: int, sigma: float) -> np.ndarray:
    """cv2.GaussianBlur(float32 image, (ksize, ksize), sigma) with BORDER_REFLECT_101 (rows, then columns)."""
    k = gaussian_kernel(ksize, sigma).astype(np.float64)
    h, w = img.shape
    r = ksize // 2
    src = img.astype(np.float64)
    xi = _reflect101(np.arange(-r, w + r), w)
    tmp = sum(k[t] * src[:, xi[t:t + w]] for t in range(ksize)).astype(np.float32).astype(np.float64)
    yi = _reflect101(np.arange(-r, h + r), h)
    return sum(k[t] * tmp[yi[t:t + h], :] for t in range(ksize)).astype(np.float32)


def resize_linear(img: np.ndarray, width: int, height: int) -> np.ndarray:
    """cv2.resize(..., INTER_LINEAR) for float32 (half-pixel centres, edge clamp); img is [H, W] or [H, W, C]."""
    h, w = img.shape[:2]

    def coords(n_dst, n_src):
        scale = n_src / n_dst
        f = (np.arange(n_dst, dtype=np.float64) + 0.5) * scale - 0.5
        i0 = np.floor(f).astype(np.int64)
        frac = (f - i0).astype(np.float32)
        lo = i0 < 0
        frac[lo] = 0
        i0[lo] = 0
        hi = i0 >= n_src - 1
        frac[hi] = 0
        i0[hi] = n_src - 1
        return i0, np.minimum(i0 + 1, n_src - 1), frac

    x0, x1, fx = coords(width, w)
    y0, y1, fy = coords(height, h)
    src = img.astype(np.float32)
    if src.ndim == 3:
        fx = fx[:, None]
    top = src[y0][:, x0] * (1 - fx) + src[y0][:, x1] * fx
    bot = src[y1][:, x0] * (1 - fx) + src[y1][:, x1] * fx
    fyb = fy[:, None, None] if src.ndim == 3 else fy[:, None]
    return (top * (1 - fyb) + bot * fyb).astype(np.float32)


def prepare_gaussian(n: int, sigma: float):
    if sigma < np.finfo(np.float32).eps:
        sigma = n * 0.3
    x = np.arange(-n, n + 1, dtype=np.float64)
    g = np.exp(-x * x / (2 * sigma * sigma)).astype(np.float32).astype(np.float64)
    g = (g * (1.0 / g.sum())).astype(np.float32)
    xg = (x * g).astype(np.float32)
    xxg = (x * x * g).astype(np.float32)
    G = np.zeros((6, 6))
    gd = g.astype(np.float64)
    gy, gx = np.meshgrid(gd, gd, indexing="ij")
    yy, xx = np.meshgrid(x, x, indexing="ij")
    G[0, 0] = (gy * gx).sum()
    G[1, 1] = (gy * gx * xx * xx).sum()
    G[3, 3] = (gy * gx * xx ** 4).sum()
    G[5, 5] = (gy * gx * xx * xx * yy * yy).sum()
    G[2, 2] = G[0, 3] = G[0, 4] = G[3, 0] = G[4, 0] = G[1, 1]
    G[4, 4] = G[3, 3]
    G[3, 4] = G[4, 3] = G[5, 5]
    inv = np.linalg.inv(G)
    return g, xg, xxg, inv[1, 1], inv[0, 3], inv[3, 3], inv[5, 5]


def poly_exp(src: np.ndarray, n: int, sigma: float) -> np.ndarray:
    """FarnebackPolyExp: [H, W] float32 -> [H, W, 5] = (y, x, yy, xx, xy) coefficients of the local quadratic."""
    g, xg, xxg, ig11, ig03, ig33, ig55 = prepare_gaussian(n, sigma)
    h, w = src.shape
    s = src.astype(np.float32)
    rows = np.arange(h)
    r0 = s * g[n]
    r1 = np.zeros_like(s)
    r2 = np.zeros_like(s)
    for k in range(1, n + 1):  # vertical pass, float32 as in OpenCV
        a, b = s[np.maximum(rows - k, 0)], s[np.minimum(rows + k, h - 1)]
        p = a + b
        r0 = r0 + g[n + k] * p
        r1 = r1 + xg[n + k] * (b - a)
        r2 = r2 + xxg[n + k] * p
    cols = np.arange(w)

    def at(arr, d):  # replicate border on the row buffer
        return arr[:, np.clip(cols + d, 0, w - 1)].astype(np.float64)

    b1 = at(r0, 0) * g[n]
    b3 = at(r1, 0) * g[n]
    b5 = at(r2, 0) * g[n]
    b2 = np.zeros_like(b1)
    b4 = np.zeros_like(b1)
    b6 = np.zeros_like(b1)
    for k in range(1, n + 1):  # horizontal pass, double accumulators
        tg = at(r0, k) + at(r0, -k)
        b1 += tg * g[n + k]
        b4 += tg * xxg[n + k]
        b2 += (at(r0, k) - at(r0, -k)) * xg[n + k]
        b3 += (at(r1, k) + at(r1, -k)) * g[n + k]
        b6 += (at(r1, k) - at(r1, -k)) * xg[n + k]
        b5 += (at(r2, k) + at(r2, -k)) * g[n + k]
    out = np.empty((h, w, 5), dtype=np.float32)
    out[..., 1] = b2 * ig11
    out[..., 0] = b3 * ig11
    out[..., 3] = b1 * ig03 + b4 * ig33
    out[..., 2] = b1 * ig03 + b5 * ig33
    out[..., 4] = b6 * ig55
    return out


def update_matrices(R0: np.ndarray, R1: np.ndarray, flow: np.ndarray) -> np.ndarray:
    """FarnebackUpdateMatrices over the whole image: [H, W, 5] float32 (G11, G12, G22, h1, h2)."""
    h, w = flow.shape[:2]
    f32 = np.float32
    ys, xs = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    dx, dy = flow[..., 0].astype(f32), flow[..., 1].astype(f32)
    fx = xs.astype(f32) + dx
    fy = ys.astype(f32) + dy
    x1 = np.floor(fx).astype(np.int64)
    y1 = np.floor(fy).astype(np.int64)
    fx = (fx - x1.astype(f32)).astype(f32)
    fy = (fy - y1.astype(f32)).astype(f32)
    inside = (x1 >= 0) & (x1 < w - 1) & (y1 >= 0) & (y1 < h - 1)
    xc, yc = np.clip(x1, 0, max(w - 2, 0)), np.clip(y1, 0, max(h - 2, 0))
    one = f32(1)
    a00, a01, a10, a11 = (one - fx) * (one - fy), fx * (one - fy), (one - fx) * fy, fx * fy
    if w > 1 and h > 1:
        samp = (a00[..., None] * R1[yc, xc] + a01[..., None] * R1[yc, xc + 1] + a10[..., None] * R1[yc + 1, xc]
                + a11[..., None] * R1[yc + 1, xc + 1]).astype(f32)
    else:
        samp = np.zeros_like(R0)
    r2 = np.where(inside, samp[..., 0], 0).astype(f32)
    r3 = np.where(inside, samp[..., 1], 0).astype(f32)
    r4 = np.where(inside, (R0[..., 2] + samp[..., 2]) * f32(0.5), R0[..., 2]).astype(f32)
    r5 = np.where(inside, (R0[..., 3] + samp[..., 3]) * f32(0.5), R0[..., 3]).astype(f32)
    r6 = np.where(inside, (R0[..., 4] + samp[..., 4]) * f32(0.25), R0[..., 4] * f32(0.5)).astype(f32)
    r2 = (R0[..., 0] - r2) * f32(0.5)
    r3 = (R0[..., 1] - r3) * f32(0.5)
    r2 = r2 + r4 * dy + r6 * dx
    r3 = r3 + r6 * dy + r5 * dx
    nb = len(BORDER)

    def edge(idx, n):
        sc = np.ones(n, dtype=f32)
        for i in range(n):
            if i < nb:
                sc[i] *= BORDER[i]
            if i >= n - nb:
                sc[i] *= BORDER[n - i - 1]
        return sc[idx]

    # OpenCV tests `(unsigned)(x - BORDER) >= (unsigned)(width - BORDER*2) || (unsigned)(y - BORDER) >= (unsigned)(height - BORDER*2)`:
    # with a dimension below 10 the right-hand side wraps around and the attenuation is skipped for most pixels
    u32 = lambda v: np.mod(v, 1 << 32)
    on_border = (u32(xs - nb) >= u32(w - 2 * nb)) | (u32(ys - nb) >= u32(h - 2 * nb))
    scale = np.where(on_border, edge(xs, w) * edge(ys, h), f32(1)).astype(f32)
    r2, r3, r4, r5, r6 = (v * scale for v in (r2, r3, r4, r5, r6))
    M = np.empty((h, w, 5), dtype=f32)
    M[..., 0] = r4 * r4 + r6 * r6
    M[..., 1] = (r4 + r5) * r6
    M[..., 2] = r5 * r5 + r6 * r6
    M[..., 3] = r4 * r2 + r6 * r3
    M[..., 4] = r6 * r2 + r5 * r3
    return M


def blur_solve(M: np.ndarray, block_size: int) -> np.ndarray:
    """FarnebackUpdateFlow_Blur: (2m+1)^2 box sums of M with replicated borders (double), then the 2x2 solve."""
    h, w = M.shape[:2]
    m = block_size // 2
    Md = M.astype(np.float64)
    yi = np.clip(np.arange(-m, h + m), 0, h - 1)
    v = sum(Md[yi[t:t + h]] for t in range(2 * m + 1))
    xi = np.clip(np.arange(-m, w + m), 0, w - 1)
    s = sum(v[:, xi[t:t + w]] for t in range(2 * m + 1)) * (1.0 / (block_size * block_size))
    g11, g12, g22, h1, h2 = (s[..., i] for i in range(5))
    idet = 1.0 / (g11 * g22 - g12 * g12 + 1e-3)
    flow = np.empty((h, w, 2), dtype=np.float32)
    flow[..., 0] = (g11 * h2 - g12 * h1) * idet
    flow[..., 1] = (g22 * h1 - g12 * h2) * idet
    return flow


def calc_optical_flow_farneback(prev: np.ndarray, nxt: np.ndarray, pyr_scale=0.5, levels=3, winsize=15, iterations=3, poly_n=5,
                                poly_sigma=1.2) -> np.ndarray:
    """uint8 / float32 single-channel images [H, W] -> flow [H, W, 2] float32 (flags = 0)."""
    min_size = 32
    h0, w0 = prev.shape
    k, scale = 0, 1.0
    while k < levels:
        scale *= pyr_scale
        if w0 * scale < min_size or h0 * scale < min_size:
            break
        k += 1
    levels = k
    imgs = [prev.astype(np.float32), nxt.astype(np.float32)]
    prev_flow = None
    for k in range(levels, -1, -1):
        scale = pyr_scale ** k
        sigma = (1.0 / scale - 1) * 0.5
        smooth = max(cv_round(sigma * 5) | 1, 3)
        width, height = cv_round(w0 * scale), cv_round(h0 * scale)
        if prev_flow is None:
            flow = np.zeros((height, width, 2), dtype=np.float32)
        else:
            flow = resize_linear(prev_flow, width, height) * np.float32(1.0 / pyr_scale)
        R = []
        for img in imgs:
            I = resize_linear(gaussian_blur(img, smooth, sigma), width, height)
            R.append(poly_exp(I, poly_n, poly_sigma))
        M = update_matrices(R[0], R[1], flow)
        for i in range(iterations):
            flow = blur_solve(M, winsize)
            if i < iterations - 1:
                M = update_matrices(R[0], R[1], flow)
        prev_flow = flow
    return prev_flow
