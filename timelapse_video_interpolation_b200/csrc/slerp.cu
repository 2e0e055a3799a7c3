// A1 — first/last-frame latent slerp (the reference file
// frame_slerp_interp.py is 0 bytes, so the arithmetic below IS the spec, mirrored by
// oracle/slerp_ref.py).
//
// One thread-block cluster per clip.  Pass 1 streams v0/v1 once from HBM and reduces
// <v0,v1>, |v0|^2, |v1|^2 in fp64 (warp shuffle -> CTA smem -> DSMEM across the cluster, summed
// in rank order so the result is run-to-run deterministic).  Every CTA then derives the 2*T
// blend coefficients in fp64 and pass 2 re-reads its slice of v0/v1 (now an L2 hit: a clip is
// 295 KB) and streams the T output frames with 128-bit evict-first stores.
// HBM traffic per clip = (2 + T) * n * 4 bytes, which is the algorithmic minimum.
#include <cooperative_groups.h>

#include "tvi_common.cuh"

namespace cg = cooperative_groups;

namespace tvi {
namespace {

constexpr int kSlerpThreads = 512;
constexpr int kMaxT = 64;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(kSlerpThreads)
slerp_kernel(const float4* __restrict__ v0, const float4* __restrict__ v1, float4* __restrict__ out, int64_t n4,
             int T, double dot_threshold) {
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned crank = cluster.block_rank();
  const unsigned csize = cluster.num_blocks();
  const int64_t clip = blockIdx.x / csize;

  __shared__ double s_warp[3][kSlerpThreads / 32];
  __shared__ double s_part[3];  // this CTA's partial sums, read by cluster peers through DSMEM
  __shared__ float s_a[kMaxT], s_b[kMaxT];

  // slice of the clip owned by this CTA (in float4 units)
  const int64_t per = (n4 + csize - 1) / csize;
  const int64_t lo = per * crank;
  const int64_t hi = (lo + per < n4) ? lo + per : n4;
  const float4* p0 = v0 + clip * n4;
  const float4* p1 = v1 + clip * n4;

  double dot = 0.0, n0 = 0.0, n1 = 0.0;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kSlerpThreads) {
    const float4 a = p0[i];
    const float4 b = p1[i];
    dot += (double)a.x * b.x + (double)a.y * b.y + (double)a.z * b.z + (double)a.w * b.w;
    n0 += (double)a.x * a.x + (double)a.y * a.y + (double)a.z * a.z + (double)a.w * a.w;
    n1 += (double)b.x * b.x + (double)b.y * b.y + (double)b.z * b.z + (double)b.w * b.w;
  }
  dot = warp_sum(dot);
  n0 = warp_sum(n0);
  n1 = warp_sum(n1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_warp[0][warp] = dot;
    s_warp[1][warp] = n0;
    s_warp[2][warp] = n1;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    double acc = 0.0;
    for (int w = 0; w < kSlerpThreads / 32; ++w) acc += s_warp[threadIdx.x][w];
    s_part[threadIdx.x] = acc;
  }
  cluster.sync();

  if (threadIdx.x == 0) {
    double tot[3] = {0.0, 0.0, 0.0};
    for (unsigned r = 0; r < csize; ++r) {
      const double* peer = cluster.map_shared_rank(s_part, r);
      tot[0] += peer[0];
      tot[1] += peer[1];
      tot[2] += peer[2];
    }
    s_warp[0][0] = tot[0];
    s_warp[1][0] = tot[1];
    s_warp[2][0] = tot[2];
  }
  __syncthreads();
  if (threadIdx.x < T) {
    const double d = s_warp[0][0], q0 = s_warp[1][0], q1 = s_warp[2][0];
    const double den = sqrt(q0) * sqrt(q1);
    const double t = (T > 1) ? (double)threadIdx.x / (double)(T - 1) : 0.0;
    double a = 1.0 - t, b = t;
    if (den > 0.0) {
      double c = d / den;
      c = c > 1.0 ? 1.0 : (c < -1.0 ? -1.0 : c);
      if (fabs(c) <= dot_threshold) {
        const double theta = acos(c);
        const double s = sin(theta);
        a = sin((1.0 - t) * theta) / s;
        b = sin(t * theta) / s;
      }
    }
    s_a[threadIdx.x] = (float)a;
    s_b[threadIdx.x] = (float)b;
  }
  // peers may still be reading s_part through DSMEM: nobody may exit before everybody has read
  cluster.sync();

  float4* po = out + clip * (int64_t)T * n4;
  for (int64_t i = lo + threadIdx.x; i < hi; i += kSlerpThreads) {
    const float4 a = p0[i];
    const float4 b = p1[i];
#pragma unroll 2
    for (int t = 0; t < T; ++t) {
      const float ca = s_a[t], cb = s_b[t];
      float4 r;
      r.x = __fadd_rn(__fmul_rn(ca, a.x), __fmul_rn(cb, b.x));
      r.y = __fadd_rn(__fmul_rn(ca, a.y), __fmul_rn(cb, b.y));
      r.z = __fadd_rn(__fmul_rn(ca, a.z), __fmul_rn(cb, b.z));
      r.w = __fadd_rn(__fmul_rn(ca, a.w), __fmul_rn(cb, b.w));
      __stcs(po + (int64_t)t * n4 + i, r);
    }
  }
}

}  // namespace
}  // namespace tvi

extern "C" int tvi_slerp_f32(const float* v0, const float* v1, float* out, int64_t nclips, int64_t n, int T,
                             float dot_threshold, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(nclips >= 0 && n >= 0, "tvi_slerp_f32: negative size");
  if (nclips == 0 || n == 0) return TVI_OK;
  TVI_CHECK_ARG(v0 && v1 && out, "tvi_slerp_f32: null pointer");
  TVI_CHECK_ARG(T >= 2 && T <= kMaxT, "tvi_slerp_f32: T=%d outside [2,%d]", T, kMaxT);
  TVI_CHECK_ARG(n % 4 == 0, "tvi_slerp_f32: n=%lld is not a multiple of 4", (long long)n);
  TVI_CHECK_ARG(((uintptr_t)v0 | (uintptr_t)v1 | (uintptr_t)out) % 16 == 0, "tvi_slerp_f32: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(nclips * 8 < (int64_t)1 << 31, "tvi_slerp_f32: too many clips for one launch");

  // Few clips: spread one clip over up to 8 CTAs so the launch still covers the chip.
  const int64_t n4 = n / 4;
  int csize = 1;
  const int sms = num_sms();
  while (csize < 8 && nclips * csize < 2 * (int64_t)sms && n4 / (csize * 2) >= kSlerpThreads) csize *= 2;

  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(nclips * csize));
  cfg.blockDim = dim3(kSlerpThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TVI_CUDA_OK(cudaLaunchKernelEx(&cfg, slerp_kernel, (const float4*)v0, (const float4*)v1, (float4*)out, n4, T,
                                 (double)dot_threshold));
  return TVI_OK;
}
