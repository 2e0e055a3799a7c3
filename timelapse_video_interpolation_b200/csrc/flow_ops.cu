// A6 / A7 — optical-flow checks.
//
// tvi_flow_cosine_f32 : mean per-pixel cosine similarity of two flow fields.  Follows
//   cosine_similarity(), dataset_construction/step4_flow_check.py:25-34 operation by operation
//   in fp32 (mul, add, sqrt, +1e-6, div — no FMA contraction) and accumulates the mean in fp64.
//   HBM-bound: 16 bytes per pixel, read once with 128-bit loads.
//
// tvi_backward_warp_f32 : backward warp of img2 by f_fwd plus forward/backward consistency
//   (the reference file has no warp, oracle = F.grid_sample(bilinear,
//   zeros, align_corners=True)).  One thread per 4 consecutive pixels: coalesced 128-bit loads
//   of the flow, texture-path gathers of the 4 bilinear taps (smooth flows keep them in L1/L2),
//   128-bit stores of the warped channels, one 32-bit store of four mask bytes.
#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kCosThreads = 256;

__device__ __forceinline__ float cos_px(float ax, float ay, float bx, float by) {
  const float dot = __fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by));
  const float na = __fadd_rn(__fsqrt_rn(__fadd_rn(__fmul_rn(ax, ax), __fmul_rn(ay, ay))), 1e-6f);
  const float nb = __fadd_rn(__fsqrt_rn(__fadd_rn(__fmul_rn(bx, bx), __fmul_rn(by, by))), 1e-6f);
  return __fdiv_rn(dot, __fmul_rn(na, nb));
}

__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// partial[pair * nblk + blockIdx.x] = sum over this block's pixels of cos_px
template <bool HWC>
__global__ void __launch_bounds__(kCosThreads)
flow_cosine_partial_kernel(const float* __restrict__ fa, const float* __restrict__ fb, double* __restrict__ partial,
                           int64_t npix) {
  const int64_t pair = blockIdx.y;
  const float* a = fa + pair * 2 * npix;
  const float* b = fb + pair * 2 * npix;
  double acc = 0.0;
  const int64_t tid = (int64_t)blockIdx.x * kCosThreads + threadIdx.x;
  const int64_t nthr = (int64_t)gridDim.x * kCosThreads;
  const bool aligned = (((uintptr_t)a | (uintptr_t)b) % 16 == 0) && (HWC || npix % 4 == 0);
  if (HWC) {
    // [npix][2]: one float4 = two pixels
    const int64_t nvec = aligned ? npix / 2 : 0;
    const float4* a4 = reinterpret_cast<const float4*>(a);
    const float4* b4 = reinterpret_cast<const float4*>(b);
    int64_t i = tid;
    for (; i + nthr < nvec; i += 2 * nthr) {  // two independent 128-bit loads per tensor in flight
      const float4 x0 = ld_stream(a4 + i), y0 = ld_stream(b4 + i);
      const float4 x1 = ld_stream(a4 + i + nthr), y1 = ld_stream(b4 + i + nthr);
      float s = __fadd_rn(cos_px(x0.x, x0.y, y0.x, y0.y), cos_px(x0.z, x0.w, y0.z, y0.w));
      float u = __fadd_rn(cos_px(x1.x, x1.y, y1.x, y1.y), cos_px(x1.z, x1.w, y1.z, y1.w));
      acc += (double)s + (double)u;
    }
    for (; i < nvec; i += nthr) {
      const float4 x0 = ld_stream(a4 + i), y0 = ld_stream(b4 + i);
      acc += (double)cos_px(x0.x, x0.y, y0.x, y0.y) + (double)cos_px(x0.z, x0.w, y0.z, y0.w);
    }
    for (int64_t p = nvec * 2 + tid; p < npix; p += nthr) acc += (double)cos_px(a[2 * p], a[2 * p + 1], b[2 * p], b[2 * p + 1]);
  } else {
    // [2][npix]: one float4 from each plane = four pixels
    const int64_t nvec = aligned ? npix / 4 : 0;
    const float4* ax4 = reinterpret_cast<const float4*>(a);
    const float4* ay4 = reinterpret_cast<const float4*>(a + npix);
    const float4* bx4 = reinterpret_cast<const float4*>(b);
    const float4* by4 = reinterpret_cast<const float4*>(b + npix);
    for (int64_t i = tid; i < nvec; i += nthr) {
      const float4 ax = ld_stream(ax4 + i), ay = ld_stream(ay4 + i), bx = ld_stream(bx4 + i), by = ld_stream(by4 + i);
      acc += (double)cos_px(ax.x, ay.x, bx.x, by.x) + (double)cos_px(ax.y, ay.y, bx.y, by.y) +
             (double)cos_px(ax.z, ay.z, bx.z, by.z) + (double)cos_px(ax.w, ay.w, bx.w, by.w);
    }
    for (int64_t p = nvec * 4 + tid; p < npix; p += nthr) acc += (double)cos_px(a[p], a[npix + p], b[p], b[npix + p]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ double s_w[kCosThreads / 32];
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kCosThreads / 32; ++w) t += s_w[w];
    partial[pair * gridDim.x + blockIdx.x] = t;
  }
}

// one warp per pair: fixed-order sum of the block partials, then the mean
__global__ void flow_cosine_final_kernel(const double* __restrict__ partial, float* __restrict__ out, int nblk,
                                         int64_t npairs, int64_t npix) {
  const int64_t pair = (int64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
  if (pair >= npairs) return;
  const int lane = threadIdx.x & 31;
  double acc = 0.0;
  for (int i = lane; i < nblk; i += 32) acc += partial[pair * nblk + i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) out[pair] = (float)(acc / (double)npix);
}

int cosine_blocks(int64_t npix) {
  // ~16 pixels per thread and never more than 1024 partials per pair
  int64_t nblk = (npix + (int64_t)kCosThreads * 16 - 1) / ((int64_t)kCosThreads * 16);
  if (nblk < 1) nblk = 1;
  if (nblk > 1024) nblk = 1024;
  return (int)nblk;
}

// ---------------------------------------------------------------------------------------------
// backward warp + consistency
// ---------------------------------------------------------------------------------------------
struct Tap {
  int x0, y0;      // top-left integer tap
  float wx, wy;    // fractional weights (exact: f - floor(f))
  bool in00, in01, in10, in11, valid;
};

__device__ __forceinline__ Tap make_tap(int x, int y, float fx, float fy, int W, int H) {
  Tap t;
  // p + f with p integer: floor(p + f) = p + floor(f) and frac(p + f) = f - floor(f), both exact
  // in fp32, so the bilinear weights carry no coordinate rounding at all.
  const float flx = floorf(fx), fly = floorf(fy);
  t.wx = fx - flx;
  t.wy = fy - fly;
  // clamp the integer part far outside the image so the int conversion cannot overflow
  const float cx = fminf(fmaxf(flx, -1.0e6f), 1.0e6f), cy = fminf(fmaxf(fly, -1.0e6f), 1.0e6f);
  t.x0 = x + (int)cx;
  t.y0 = y + (int)cy;
  const bool xin0 = t.x0 >= 0 && t.x0 < W, xin1 = t.x0 + 1 >= 0 && t.x0 + 1 < W;
  const bool yin0 = t.y0 >= 0 && t.y0 < H, yin1 = t.y0 + 1 >= 0 && t.y0 + 1 < H;
  t.in00 = xin0 && yin0;
  t.in01 = xin1 && yin0;
  t.in10 = xin0 && yin1;
  t.in11 = xin1 && yin1;
  // sample point inside [0, W-1] x [0, H-1] (closed); NaN flows are invalid
  const bool vx = (t.x0 >= 0) && (t.x0 < W - 1 || (t.x0 == W - 1 && t.wx == 0.0f));
  const bool vy = (t.y0 >= 0) && (t.y0 < H - 1 || (t.y0 == H - 1 && t.wy == 0.0f));
  t.valid = vx && vy && (fx == fx) && (fy == fy);
  return t;
}

__device__ __forceinline__ float bilerp(const float* __restrict__ plane, const Tap& t, int W) {
  const int64_t base = (int64_t)t.y0 * W + t.x0;
  const float v00 = t.in00 ? __ldg(plane + base) : 0.0f;
  const float v01 = t.in01 ? __ldg(plane + base + 1) : 0.0f;
  const float v10 = t.in10 ? __ldg(plane + base + W) : 0.0f;
  const float v11 = t.in11 ? __ldg(plane + base + W + 1) : 0.0f;
  // grid_sample's weights: nw = (1-wx)(1-wy), ne = wx(1-wy), sw = (1-wx)wy, se = wx wy
  const float ux = 1.0f - t.wx, uy = 1.0f - t.wy;
  return v00 * (ux * uy) + v01 * (t.wx * uy) + v10 * (ux * t.wy) + v11 * (t.wx * t.wy);
}

constexpr int kWarpThreads = 256;

// PX pixels per thread, a block apart: every load / store of a warp covers 32 consecutive pixels (one or
// two 128-byte lines for the gathers when the flow is smooth), and the PX independent pixels of a thread
// keep 20 * PX gather loads in flight.  Measured on 32 pairs of 1080p with smooth flows: 0.82 ms
// (3.3 TB/s of algorithmic traffic) against 1.18 ms for the first version (4 adjacent pixels per thread, 128-bit flow loads), whose
// warp-level gathers span four lines each.  ncu (profiles/r1_c): DRAM traffic = 1.0 x algorithmic, ~300
// SASS instructions per pixel (address arithmetic and tap predicates): the kernel is issue-bound, not HBM-bound.
// Round 2 tried to shrink the instruction stream and measured every variant against this one on the same box
// (0.818 ms for 32 pairs of 1080p): tap offsets + weight products hoisted into the tap record with clamped
// unconditional loads 2.10 ms (87 registers), the same with a 48- / 40-register cap 1.07 / 1.13 ms, predicated loads +
// hoisted weights + float-reciprocal pixel decode (no emulated 64-bit division) 1.10 ms, only the division replaced
// 1.18 ms (54 registers).  Fewer instructions made ptxas unroll the channel loop and hoist more gathers per thread; the
// kernel's speed comes from many resident warps each with few gathers in flight, and this version is kept.
template <int PX>
__global__ void __launch_bounds__(kWarpThreads)
backward_warp_kernel(const float* __restrict__ img2, const float* __restrict__ f_fwd, const float* __restrict__ f_bwd,
                             float* __restrict__ warped, uint8_t* __restrict__ mask, int C, int H, int W, float alpha,
                             float beta) {
  const int64_t b = blockIdx.y;
  const int64_t npix = (int64_t)H * W;
  const int64_t off0 = (int64_t)blockIdx.x * (kWarpThreads * PX) + threadIdx.x;
  const float* ff = f_fwd + b * 2 * npix;
  const float* fbw = f_bwd + b * 2 * npix;
  float fx[PX], fy[PX];
  bool in[PX];
#pragma unroll
  for (int i = 0; i < PX; ++i) {
    const int64_t off = off0 + (int64_t)i * kWarpThreads;
    in[i] = off < npix;
    fx[i] = in[i] ? __ldg(ff + off) : 0.f;
    fy[i] = in[i] ? __ldg(ff + npix + off) : 0.f;
  }
  Tap tap[PX];
  float bx[PX], by[PX];
#pragma unroll
  for (int i = 0; i < PX; ++i) {
    const int64_t off = in[i] ? off0 + (int64_t)i * kWarpThreads : 0;
    const int y = (int)(off / W);
    const int x = (int)(off - (int64_t)y * W);
    tap[i] = make_tap(x, y, fx[i], fy[i], W, H);
    bx[i] = bilerp(fbw, tap[i], W);
    by[i] = bilerp(fbw + npix, tap[i], W);
  }
#pragma unroll
  for (int i = 0; i < PX; ++i) {
    const float sx = fx[i] + bx[i], sy = fy[i] + by[i];
    const float lhs = sx * sx + sy * sy;
    const float rhs = alpha * (fx[i] * fx[i] + fy[i] * fy[i] + bx[i] * bx[i] + by[i] * by[i]) + beta;
    if (in[i]) mask[b * npix + off0 + (int64_t)i * kWarpThreads] = (uint8_t)((lhs > rhs ? 1u : 0u) | (tap[i].valid ? 2u : 0u));
  }
  for (int c = 0; c < C; ++c) {
    const float* plane = img2 + (b * C + c) * npix;
    float r[PX];
#pragma unroll
    for (int i = 0; i < PX; ++i) r[i] = bilerp(plane, tap[i], W);
#pragma unroll
    for (int i = 0; i < PX; ++i)
      if (in[i]) __stcs(warped + (b * C + c) * npix + off0 + (int64_t)i * kWarpThreads, r[i]);
  }
}

}  // namespace
}  // namespace tvi

extern "C" {

size_t tvi_flow_cosine_workspace_bytes(int64_t npairs, int H, int W) {
  if (npairs <= 0 || H <= 0 || W <= 0) return 0;
  return (size_t)npairs * (size_t)tvi::cosine_blocks((int64_t)H * W) * sizeof(double);
}

int tvi_flow_cosine_f32(const float* fa, const float* fb, float* out, void* workspace, int64_t npairs, int H, int W,
                        int hwc_layout, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(npairs >= 0 && H >= 0 && W >= 0, "tvi_flow_cosine_f32: negative size");
  if (npairs == 0) return TVI_OK;
  TVI_CHECK_ARG(fa && fb && out, "tvi_flow_cosine_f32: null pointer");
  TVI_CHECK_ARG(H > 0 && W > 0, "tvi_flow_cosine_f32: empty flow field (the reference's np.mean of nothing is NaN)");
  TVI_CHECK_ARG(workspace && (uintptr_t)workspace % 8 == 0, "tvi_flow_cosine_f32: workspace missing or not 8-byte aligned");
  TVI_CHECK_ARG(npairs <= 65535, "tvi_flow_cosine_f32: at most 65535 pairs per call");
  const int64_t npix = (int64_t)H * W;
  const int nblk = cosine_blocks(npix);
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((unsigned)nblk, (unsigned)npairs);
  if (hwc_layout)
    flow_cosine_partial_kernel<true><<<grid, kCosThreads, 0, st>>>(fa, fb, (double*)workspace, npix);
  else
    flow_cosine_partial_kernel<false><<<grid, kCosThreads, 0, st>>>(fa, fb, (double*)workspace, npix);
  TVI_LAUNCH_OK("flow_cosine_partial_kernel");
  const int wpb = 4;
  flow_cosine_final_kernel<<<(unsigned)((npairs + wpb - 1) / wpb), wpb * 32, 0, st>>>((const double*)workspace, out, nblk,
                                                                                      npairs, npix);
  TVI_LAUNCH_OK("flow_cosine_final_kernel");
  return TVI_OK;
}

int tvi_backward_warp_f32(const float* img2, const float* f_fwd, const float* f_bwd, float* warped, uint8_t* mask,
                          int64_t B, int C, int H, int W, float alpha, float beta, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(B >= 0 && C >= 0 && H >= 0 && W >= 0, "tvi_backward_warp_f32: negative size");
  if (B == 0 || H == 0 || W == 0) return TVI_OK;
  TVI_CHECK_ARG(f_fwd && f_bwd && mask && (C == 0 || (img2 && warped)), "tvi_backward_warp_f32: null pointer");
  TVI_CHECK_ARG(B <= 65535, "tvi_backward_warp_f32: at most 65535 image pairs per call");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t npix = (int64_t)H * W;
  constexpr int kPx = 2;
  TVI_CHECK_ARG((npix + kWarpThreads * kPx - 1) / (kWarpThreads * kPx) < ((int64_t)1 << 31), "tvi_backward_warp_f32: image too large");
  dim3 grid((unsigned)((npix + kWarpThreads * kPx - 1) / (kWarpThreads * kPx)), (unsigned)B);
  backward_warp_kernel<kPx><<<grid, kWarpThreads, 0, st>>>(img2, f_fwd, f_bwd, warped, mask, C, H, W, alpha, beta);
  TVI_LAUNCH_OK("backward_warp_kernel");
  return TVI_OK;
}

}  // extern "C"
