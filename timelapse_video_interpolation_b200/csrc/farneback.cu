// Dense optical flow after Farnebäck, as cv2.calcOpticalFlowFarneback computes it for the reference's call
// (dataset_construction/step4_flow_check.py:13-22: pyr_scale 0.5, levels 3, winsize 15, iterations 3, poly_n 5,
// poly_sigma 1.2, flags 0) — row N3 of SURVEY.md 8.6.  The algorithm is OpenCV's (modules/video/src/optflowgf.cpp,
// third-party, not in the reference tree); oracle/farneback_ref.py restates it and is pinned against cv2 itself.
// Same arithmetic types as OpenCV where it matters: fp32 images / polynomial coefficients / matrices, fp64 accumulators
// in the horizontal polynomial pass and in the box sums, fp64 2x2 solve.
//
// Per pyramid level (coarse to fine), everything stays in HBM, one thread per pixel, no host synchronisation:
//   gauss_h / gauss_v   GaussianBlur(full-resolution image, sigma = (1/scale - 1)/2, BORDER_REFLECT_101)
//   resize1             bilinear resize to the level size (half-pixel centres)
//   polyexp_v / _h      separable projection on the quadratic basis -> 5 coefficients per pixel (y, x, yy, xx, xy)
//   resize_flow         previous level's flow, bilinear, x 1/pyr_scale
//   update_matrices     warp R1 by the flow (bilinear), average with R0, border attenuation -> G11 G12 G22 h1 h2
//   box_v / box_h_solve (winsize x winsize) box sums with replicated borders, then flow = G^-1 h per pixel
// All kernels are bandwidth-trivial at 1080p (2 M pixels); the point of the GPU version is the 5x cost ratio against the
// consumer of its output (0.96 s per pair for OpenCV on 8 cores vs 0.20 s for the cosine check).
#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kMaxTaps = 63;
struct Taps {
  float k[kMaxTaps];
  int n;
};
struct PolyTaps {
  float g[16], xg[16], xxg[16];  // index 0..n (symmetric / antisymmetric halves)
  int n;
  double ig11, ig03, ig33, ig55;
};

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  const int p = 2 * (n - 1);
  i %= p;
  if (i < 0) i += p;
  return i >= n ? p - i : i;
}

__global__ void bgr_to_gray_kernel(const uint8_t* __restrict__ bgr, float* __restrict__ gray, int64_t npix) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const int b = bgr[3 * i], g = bgr[3 * i + 1], r = bgr[3 * i + 2];
  gray[i] = (float)((b * 3735 + g * 19235 + r * 9798 + 16384) >> 15);  // cv::cvtColor(BGR2GRAY), 15-bit fixed point
}

__global__ void u8_to_f32_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = (float)src[i];
}

template <bool kVertical>
__global__ void gauss_kernel(const float* __restrict__ src, float* __restrict__ dst, int h, int w, const Taps taps) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  const int r = taps.n / 2;
  const int c = kVertical ? y : x, n = kVertical ? h : w;
  const int64_t stride = kVertical ? w : 1;
  float acc = 0.f;
  if (c >= r && c + r < n) {  // interior: no border arithmetic (the reflect path costs two integer divisions per tap)
    const float* p = src + (int64_t)y * w + x - (int64_t)r * stride;
    for (int t = 0; t < taps.n; ++t) acc = fmaf(taps.k[t], p[(int64_t)t * stride], acc);
  } else {
    for (int t = 0; t < taps.n; ++t) {
      const float v = kVertical ? src[(int64_t)reflect101(y + t - r, h) * w + x] : src[(int64_t)y * w + reflect101(x + t - r, w)];
      acc = fmaf(taps.k[t], v, acc);
    }
  }
  dst[(int64_t)y * w + x] = acc;
}

__device__ __forceinline__ void lin_coord(int d, int n_dst, int n_src, int& i0, int& i1, float& f) {
  const double scale = (double)n_src / n_dst;
  const double c = (d + 0.5) * scale - 0.5;
  int a = (int)floor(c);
  float fr = (float)(c - a);
  if (a < 0) {
    a = 0;
    fr = 0.f;
  }
  if (a >= n_src - 1) {
    a = n_src - 1;
    fr = 0.f;
  }
  i0 = a;
  i1 = min(a + 1, n_src - 1);
  f = fr;
}

// cv::resize(INTER_LINEAR) of a CH-channel float image, times `mul`
template <int CH>
__global__ void resize_kernel(const float* __restrict__ src, float* __restrict__ dst, int hs, int ws, int hd, int wd, float mul) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= wd) return;
  int x0, x1, y0, y1;
  float fx, fy;
  lin_coord(x, wd, ws, x0, x1, fx);
  lin_coord(y, hd, hs, y0, y1, fy);
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    const float a = src[((int64_t)y0 * ws + x0) * CH + c], b = src[((int64_t)y0 * ws + x1) * CH + c];
    const float d = src[((int64_t)y1 * ws + x0) * CH + c], e = src[((int64_t)y1 * ws + x1) * CH + c];
    const float top = a * (1.f - fx) + b * fx, bot = d * (1.f - fx) + e * fx;
    dst[((int64_t)y * wd + x) * CH + c] = (top * (1.f - fy) + bot * fy) * mul;
  }
}

// FarnebackPolyExp, vertical part: row[x] = (sum g, sum xg, sum xxg) over the column, replicated rows, fp32
__global__ void polyexp_v_kernel(const float* __restrict__ src, float* __restrict__ row3, int h, int w, const PolyTaps t) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  const float s0 = src[(int64_t)y * w + x];
  float r0 = s0 * t.g[0], r1 = 0.f, r2 = 0.f;
  for (int k = 1; k <= t.n; ++k) {
    const float a = src[(int64_t)max(y - k, 0) * w + x], b = src[(int64_t)min(y + k, h - 1) * w + x];
    const float p = a + b;
    r0 = r0 + t.g[k] * p;
    r1 = r1 + t.xg[k] * (b - a);
    r2 = r2 + t.xxg[k] * p;
  }
  float* o = row3 + ((int64_t)y * w + x) * 3;
  o[0] = r0;
  o[1] = r1;
  o[2] = r2;
}

// horizontal part (fp64 accumulators, replicated columns) -> (y, x, yy, xx, xy)
__global__ void polyexp_h_kernel(const float* __restrict__ row3, float* __restrict__ R, int h, int w, const PolyTaps t) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  const float* row = row3 + (int64_t)y * w * 3;
  const double g0 = t.g[0];
  double b1 = row[x * 3] * g0, b2 = 0, b3 = row[x * 3 + 1] * g0, b4 = 0, b5 = row[x * 3 + 2] * g0, b6 = 0;
  for (int k = 1; k <= t.n; ++k) {
    const float* p = row + min(x + k, w - 1) * 3;
    const float* m = row + max(x - k, 0) * 3;
    const double tg = (double)p[0] + (double)m[0];
    const double gk = t.g[k], xgk = t.xg[k], xxgk = t.xxg[k];
    b1 += tg * gk;
    b4 += tg * xxgk;
    b2 += ((double)p[0] - (double)m[0]) * xgk;
    b3 += ((double)p[1] + (double)m[1]) * gk;
    b6 += ((double)p[1] - (double)m[1]) * xgk;
    b5 += ((double)p[2] + (double)m[2]) * gk;
  }
  float* o = R + ((int64_t)y * w + x) * 5;
  o[1] = (float)(b2 * t.ig11);
  o[0] = (float)(b3 * t.ig11);
  o[3] = (float)(b1 * t.ig03 + b4 * t.ig33);
  o[2] = (float)(b1 * t.ig03 + b5 * t.ig33);
  o[4] = (float)(b6 * t.ig55);
}

__device__ __forceinline__ float border_scale(int i, int n) {
  const float tab[5] = {0.14f, 0.14f, 0.4472f, 0.4472f, 0.4472f};
  float s = 1.f;
  if (i < 5) s *= tab[i];
  if (i >= n - 5) s *= tab[n - i - 1];
  return s;
}

// FarnebackUpdateMatrices
__global__ void update_matrices_kernel(const float* __restrict__ R0, const float* __restrict__ R1, const float* __restrict__ flow,
                                       float* __restrict__ M, int h, int w) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  const int64_t pix = (int64_t)y * w + x;
  const float dx = flow[pix * 2], dy = flow[pix * 2 + 1];
  float fx = (float)x + dx, fy = (float)y + dy;
  const int x1 = (int)floorf(fx), y1 = (int)floorf(fy);
  fx -= (float)x1;
  fy -= (float)y1;
  const float* r0 = R0 + pix * 5;
  float r2, r3, r4, r5, r6;
  if ((unsigned)x1 < (unsigned)(w - 1) && (unsigned)y1 < (unsigned)(h - 1)) {
    const float a00 = (1.f - fx) * (1.f - fy), a01 = fx * (1.f - fy), a10 = (1.f - fx) * fy, a11 = fx * fy;
    const float* p = R1 + ((int64_t)y1 * w + x1) * 5;
    const float* q = p + (int64_t)w * 5;
    r2 = a00 * p[0] + a01 * p[5] + a10 * q[0] + a11 * q[5];
    r3 = a00 * p[1] + a01 * p[6] + a10 * q[1] + a11 * q[6];
    r4 = a00 * p[2] + a01 * p[7] + a10 * q[2] + a11 * q[7];
    r5 = a00 * p[3] + a01 * p[8] + a10 * q[3] + a11 * q[8];
    r6 = a00 * p[4] + a01 * p[9] + a10 * q[4] + a11 * q[9];
    r4 = (r0[2] + r4) * 0.5f;
    r5 = (r0[3] + r5) * 0.5f;
    r6 = (r0[4] + r6) * 0.25f;
  } else {
    r2 = r3 = 0.f;
    r4 = r0[2];
    r5 = r0[3];
    r6 = r0[4] * 0.5f;
  }
  r2 = (r0[0] - r2) * 0.5f;
  r3 = (r0[1] - r3) * 0.5f;
  r2 += r4 * dy + r6 * dx;
  r3 += r6 * dy + r5 * dx;
  if ((unsigned)(x - 5) >= (unsigned)(w - 10) || (unsigned)(y - 5) >= (unsigned)(h - 10)) {
    const float s = border_scale(x, w) * border_scale(y, h);
    r2 *= s;
    r3 *= s;
    r4 *= s;
    r5 *= s;
    r6 *= s;
  }
  float* o = M + pix * 5;
  o[0] = r4 * r4 + r6 * r6;
  o[1] = (r4 + r5) * r6;
  o[2] = r5 * r5 + r6 * r6;
  o[3] = r4 * r2 + r6 * r3;
  o[4] = r6 * r2 + r5 * r3;
}

// vertical box sums (rows y-m .. y+m, replicated), fp64.  A thread owns one (column, channel) over a strip of kBoxRows
// rows and slides the window: two loads per output instead of 2m+1 (OpenCV slides one fp64 sum down the whole column).
constexpr int kBoxRows = 32;
__global__ void box_v_kernel(const float* __restrict__ M, double* __restrict__ V, int h, int w, int m) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // over w * 5
  if (i >= w * 5) return;
  const int y0 = blockIdx.y * kBoxRows;
  const int64_t pitch = (int64_t)w * 5;
  double s = 0.0;
  for (int j = -m; j <= m; ++j) s += (double)M[(int64_t)min(max(y0 + j, 0), h - 1) * pitch + i];
  const int y1 = min(y0 + kBoxRows, h);
  for (int y = y0; y < y1; ++y) {
    V[(int64_t)y * pitch + i] = s;
    s += (double)M[(int64_t)min(y + 1 + m, h - 1) * pitch + i] - (double)M[(int64_t)max(y - m, 0) * pitch + i];
  }
}

// horizontal box sums + the 2x2 solve of FarnebackUpdateFlow_Blur.  The row segment of the block (kBoxCols + 2m pixels x 5
// fp64 sums, replicated at the row ends) is staged in shared memory with coalesced loads; each thread then sums its
// 2m+1 neighbours from there (75 strided L1 hits per pixel before).
constexpr int kBoxCols = 128;
constexpr int kBoxMaxM = 15;
__global__ void __launch_bounds__(kBoxCols)
box_h_solve_kernel(const double* __restrict__ V, float* __restrict__ flow, int h, int w, int m, double scale) {
  __shared__ double seg[(kBoxCols + 2 * kBoxMaxM) * 5];
  const int x0 = blockIdx.x * kBoxCols, y = blockIdx.y;
  const double* row = V + (int64_t)y * w * 5;
  const int npix = kBoxCols + 2 * m;
  for (int i = threadIdx.x; i < npix * 5; i += kBoxCols) {
    const int px = i / 5, c = i - px * 5;
    seg[i] = row[min(max(x0 - m + px, 0), w - 1) * 5 + c];
  }
  __syncthreads();
  const int x = x0 + threadIdx.x;
  if (x >= w) return;
  double g11 = 0, g12 = 0, g22 = 0, h1 = 0, h2 = 0;
  const double* p = seg + threadIdx.x * 5;
  for (int j = 0; j <= 2 * m; ++j, p += 5) {
    g11 += p[0];
    g12 += p[1];
    g22 += p[2];
    h1 += p[3];
    h2 += p[4];
  }
  g11 *= scale;
  g12 *= scale;
  g22 *= scale;
  h1 *= scale;
  h2 *= scale;
  const double idet = 1.0 / (g11 * g22 - g12 * g12 + 1e-3);
  float* o = flow + ((int64_t)y * w + x) * 2;
  o[0] = (float)((g11 * h2 - g12 * h1) * idet);
  o[1] = (float)((g22 * h1 - g12 * h2) * idet);
}

Taps gaussian_taps(int ksize, double sigma) {  // cv::getGaussianKernel(ksize, sigma, CV_32F)
  Taps t;
  t.n = ksize;
  if (sigma <= 0 && ksize == 3) {
    t.k[0] = 0.25f;
    t.k[1] = 0.5f;
    t.k[2] = 0.25f;
    return t;
  }
  const double s = sigma > 0 ? sigma : ((ksize - 1) * 0.5 - 1) * 0.3 + 0.8;
  double k[kMaxTaps], sum = 0;
  for (int i = 0; i < ksize; ++i) {
    const double x = i - (ksize - 1) * 0.5;
    k[i] = exp(-0.5 / (s * s) * x * x);
    sum += k[i];
  }
  for (int i = 0; i < ksize; ++i) t.k[i] = (float)(k[i] / sum);
  return t;
}

// 6x6 SPD system inverse (Gauss-Jordan with partial pivoting; the matrix is tiny and well conditioned)
void invert6(double a[6][6], double inv[6][6]) {
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) inv[i][j] = i == j;
  for (int c = 0; c < 6; ++c) {
    int piv = c;
    for (int r = c + 1; r < 6; ++r)
      if (fabs(a[r][c]) > fabs(a[piv][c])) piv = r;
    for (int j = 0; j < 6; ++j) {
      const double t1 = a[c][j];
      a[c][j] = a[piv][j];
      a[piv][j] = t1;
      const double t2 = inv[c][j];
      inv[c][j] = inv[piv][j];
      inv[piv][j] = t2;
    }
    const double d = 1.0 / a[c][c];
    for (int j = 0; j < 6; ++j) {
      a[c][j] *= d;
      inv[c][j] *= d;
    }
    for (int r = 0; r < 6; ++r) {
      if (r == c) continue;
      const double f = a[r][c];
      for (int j = 0; j < 6; ++j) {
        a[r][j] -= f * a[c][j];
        inv[r][j] -= f * inv[c][j];
      }
    }
  }
}

PolyTaps poly_taps(int n, double sigma) {  // FarnebackPrepareGaussian
  PolyTaps t;
  t.n = n;
  if (sigma < 1.1920929e-07) sigma = n * 0.3;
  float g[33];
  double s = 0;
  for (int x = -n; x <= n; ++x) {
    g[x + n] = (float)exp(-x * x / (2 * sigma * sigma));
    s += g[x + n];
  }
  s = 1.0 / s;
  for (int x = -n; x <= n; ++x) g[x + n] = (float)(g[x + n] * s);
  for (int k = 0; k <= n; ++k) {
    t.g[k] = g[n + k];
    t.xg[k] = (float)(k * g[n + k]);
    t.xxg[k] = (float)(k * k * g[n + k]);
  }
  double G[6][6] = {{0}}, inv[6][6];
  for (int y = -n; y <= n; ++y)
    for (int x = -n; x <= n; ++x) {
      const double w = (double)g[y + n] * g[x + n];
      G[0][0] += w;
      G[1][1] += w * x * x;
      G[3][3] += w * x * x * x * x;
      G[5][5] += w * x * x * y * y;
    }
  G[2][2] = G[0][3] = G[0][4] = G[3][0] = G[4][0] = G[1][1];
  G[4][4] = G[3][3];
  G[3][4] = G[4][3] = G[5][5];
  invert6(G, inv);
  t.ig11 = inv[1][1];
  t.ig03 = inv[0][3];
  t.ig33 = inv[3][3];
  t.ig55 = inv[5][5];
  return t;
}

int cv_round(double v) { return (int)nearbyint(v); }  // round-half-to-even under the default rounding mode

constexpr int kMaxLevels = 12;
struct Plan {  // pyramid geometry + workspace carve-up (offsets in floats)
  int levels;                       // usable pyramid depth (OpenCV: min_size = 32)
  int h[kMaxLevels + 1], w[kMaxLevels + 1];
  size_t r_off[kMaxLevels + 1];     // offset of level k inside one frame's R pyramid (5 floats per pixel)
  size_t pyr_floats;                // one frame's R pyramid
  size_t img, tmp, blur, I, row3, pyr[2], M, V, flow[2], total;
};
int make_plan(Plan& p, int H, int W, double pyr_scale, int levels) {
  int k = 0;
  double scale = 1.0;
  for (; k < levels && k < kMaxLevels; ++k) {
    scale *= pyr_scale;
    if (W * scale < 32 || H * scale < 32) break;
  }
  p.levels = k;
  size_t o = 0;
  for (k = 0; k <= p.levels; ++k) {
    scale = 1.0;
    for (int i = 0; i < k; ++i) scale *= pyr_scale;
    p.w[k] = cv_round(W * scale);
    p.h[k] = cv_round(H * scale);
    p.r_off[k] = o;
    o += ((size_t)p.h[k] * p.w[k] * 5 + 3) & ~(size_t)3;
  }
  p.pyr_floats = o;
  const size_t n = (size_t)H * W;
  o = 0;
  auto take = [&](size_t cnt) {
    const size_t at = o;
    o += (cnt + 3) & ~(size_t)3;
    return at;
  };
  p.img = take(n);
  p.tmp = take(n);
  p.blur = take(n);
  p.I = take(n);
  p.row3 = take(3 * n);
  p.pyr[0] = take(p.pyr_floats);
  p.pyr[1] = take(p.pyr_floats);
  p.M = take(5 * n);
  p.V = take(10 * n);  // 5 doubles per pixel
  p.flow[0] = take(2 * n);
  p.flow[1] = take(2 * n);
  p.total = o;
  return TVI_OK;
}

struct Params {
  int H, W, channels, winsize, iterations;
  double pyr_scale;
  PolyTaps pt;
};

// one frame: 8-bit image -> polynomial-expansion pyramid R[level] (each level blurs the FULL-resolution frame, as OpenCV)
int prepare_frame(const uint8_t* img, float* ws, const Plan& p, const Params& a, float* pyr, cudaStream_t st) {
  const int64_t npix = (int64_t)a.H * a.W;
  const unsigned nb = (unsigned)((npix + 255) / 256);
  if (a.channels == 3)
    bgr_to_gray_kernel<<<nb, 256, 0, st>>>(img, ws + p.img, npix);
  else
    u8_to_f32_kernel<<<nb, 256, 0, st>>>(img, ws + p.img, npix);
  TVI_LAUNCH_OK("farneback input conversion");
  const dim3 grid_full((a.W + 127) / 128, a.H);
  for (int k = p.levels; k >= 0; --k) {
    double scale = 1.0;
    for (int i = 0; i < k; ++i) scale *= a.pyr_scale;
    const double sigma = (1.0 / scale - 1) * 0.5;
    int smooth = cv_round(sigma * 5) | 1;
    if (smooth < 3) smooth = 3;
    TVI_CHECK_ARG(smooth <= kMaxTaps, "tvi_farneback: pyramid too deep for the blur kernel (%d taps)", smooth);
    const Taps gt = gaussian_taps(smooth, sigma);
    const int h = p.h[k], w = p.w[k];
    const dim3 grid((w + 127) / 128, h);
    gauss_kernel<false><<<grid_full, 128, 0, st>>>(ws + p.img, ws + p.tmp, a.H, a.W, gt);
    gauss_kernel<true><<<grid_full, 128, 0, st>>>(ws + p.tmp, ws + p.blur, a.H, a.W, gt);
    resize_kernel<1><<<grid, 128, 0, st>>>(ws + p.blur, ws + p.I, a.H, a.W, h, w, 1.0f);
    polyexp_v_kernel<<<grid, 128, 0, st>>>(ws + p.I, ws + p.row3, h, w, a.pt);
    polyexp_h_kernel<<<grid, 128, 0, st>>>(ws + p.row3, pyr + p.r_off[k], h, w, a.pt);
    TVI_LAUNCH_OK("farneback pyramid level");
  }
  return TVI_OK;
}

// one pair: coarse-to-fine flow from the two frames' pyramids
int solve_pair(const float* pyr0, const float* pyr1, float* ws, const Plan& p, const Params& a, float* flow, cudaStream_t st) {
  int cur = 0;
  for (int k = p.levels; k >= 0; --k) {
    const int h = p.h[k], w = p.w[k];
    const dim3 grid((w + 127) / 128, h);
    float* fl = (k == 0) ? flow : ws + p.flow[cur];
    if (k == p.levels) {
      TVI_CUDA_OK(cudaMemsetAsync(fl, 0, sizeof(float) * 2 * (size_t)h * w, st));
    } else {
      resize_kernel<2><<<grid, 128, 0, st>>>(ws + p.flow[cur ^ 1], fl, p.h[k + 1], p.w[k + 1], h, w, (float)(1.0 / a.pyr_scale));
      TVI_LAUNCH_OK("farneback resize_flow");
    }
    const float* R0 = pyr0 + p.r_off[k];
    const float* R1 = pyr1 + p.r_off[k];
    update_matrices_kernel<<<grid, 128, 0, st>>>(R0, R1, fl, ws + p.M, h, w);
    for (int it = 0; it < a.iterations; ++it) {
      box_v_kernel<<<dim3((w * 5 + 127) / 128, (h + kBoxRows - 1) / kBoxRows), 128, 0, st>>>(ws + p.M, (double*)(ws + p.V), h, w,
                                                                                            a.winsize / 2);
      box_h_solve_kernel<<<grid, 128, 0, st>>>((const double*)(ws + p.V), fl, h, w, a.winsize / 2,
                                               1.0 / ((double)a.winsize * a.winsize));
      if (it < a.iterations - 1) update_matrices_kernel<<<grid, 128, 0, st>>>(R0, R1, fl, ws + p.M, h, w);
      TVI_LAUNCH_OK("farneback iteration");
    }
    cur ^= 1;
  }
  return TVI_OK;
}

int check_params(const char* fn, int H, int W, int channels, double pyr_scale, int levels, int winsize, int iterations, int poly_n) {
  TVI_CHECK_ARG(channels == 1 || channels == 3, "%s: images must be 8-bit gray (1) or BGR (3), got %d channels", fn, channels);
  TVI_CHECK_ARG(pyr_scale > 0 && pyr_scale < 1 && levels >= 0, "%s: pyr_scale must be in (0, 1) and levels >= 0", fn);
  TVI_CHECK_ARG(winsize >= 1 && winsize % 2 == 1 && winsize <= 2 * kBoxMaxM + 1 && iterations >= 1,
                "%s: winsize must be odd and at most %d, iterations >= 1", fn, 2 * kBoxMaxM + 1);
  TVI_CHECK_ARG(poly_n >= 1 && poly_n <= 15, "%s: poly_n=%d outside [1, 15]", fn, poly_n);
  TVI_CHECK_ARG(H <= 65535, "%s: more than 65535 rows", fn);
  return TVI_OK;
}

}  // namespace
}  // namespace tvi

extern "C" size_t tvi_farneback_workspace_bytes(int H, int W, double pyr_scale, int levels) {
  if (H <= 0 || W <= 0 || !(pyr_scale > 0 && pyr_scale < 1) || levels < 0) return 0;
  tvi::Plan p;
  tvi::make_plan(p, H, W, pyr_scale, levels);
  return p.total * sizeof(float);
}

extern "C" int tvi_farneback_f32(const uint8_t* prev, const uint8_t* next, int channels, float* flow, void* workspace, int H, int W,
                                 double pyr_scale, int levels, int winsize, int iterations, int poly_n, double poly_sigma,
                                 void* stream) {
  return tvi_farneback_clip_f32(prev, next, 2, channels, flow, workspace, H, W, pyr_scale, levels, winsize, iterations, poly_n,
                                poly_sigma, stream);
}

extern "C" int tvi_farneback_clip_f32(const uint8_t* frames, const uint8_t* second, int64_t nframes, int channels, float* flows,
                                      void* workspace, int H, int W, double pyr_scale, int levels, int winsize, int iterations,
                                      int poly_n, double poly_sigma, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(H >= 0 && W >= 0 && nframes >= 0, "tvi_farneback_clip_f32: negative size");
  if (H == 0 || W == 0 || nframes < 2) return TVI_OK;
  TVI_CHECK_ARG(frames && flows && workspace, "tvi_farneback_clip_f32: null pointer");
  TVI_CHECK_ARG(second == nullptr || nframes == 2, "tvi_farneback_clip_f32: a separate second frame is only meaningful for one pair");
  int rc = check_params("tvi_farneback_clip_f32", H, W, channels, pyr_scale, levels, winsize, iterations, poly_n);
  if (rc) return rc;
  TVI_CHECK_ARG((uintptr_t)workspace % 16 == 0 && (uintptr_t)flows % 8 == 0, "tvi_farneback_clip_f32: misaligned pointer");
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = (float*)workspace;
  Plan p;
  make_plan(p, H, W, pyr_scale, levels);
  Params a;
  a.H = H;
  a.W = W;
  a.channels = channels;
  a.winsize = winsize;
  a.iterations = iterations;
  a.pyr_scale = pyr_scale;
  a.pt = poly_taps(poly_n, poly_sigma);
  const size_t frame_bytes = (size_t)H * W * channels;
  // every frame's pyramid is built ONCE: frame i + 1 is "next" of pair i and "prev" of pair i + 1 (ping-pong buffers)
  rc = prepare_frame(frames, ws, p, a, ws + p.pyr[0], st);
  if (rc) return rc;
  for (int64_t i = 0; i + 1 < nframes; ++i) {
    const uint8_t* nxt = second != nullptr ? second : frames + (size_t)(i + 1) * frame_bytes;
    rc = prepare_frame(nxt, ws, p, a, ws + p.pyr[(i + 1) & 1], st);
    if (rc) return rc;
    rc = solve_pair(ws + p.pyr[i & 1], ws + p.pyr[(i + 1) & 1], ws, p, a, flows + (size_t)i * H * W * 2, st);
    if (rc) return rc;
  }
  return TVI_OK;
}
