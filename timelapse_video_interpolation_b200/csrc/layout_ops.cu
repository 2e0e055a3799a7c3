// Layout glue of TransformerSpatioTemporalModel.forward (rows N1 "AlphaBlender" and N2 of SURVEY.md 8.6), all HBM-bound:
//   groupnorm_stats_kernel / groupnorm_tokens_kernel
//       self.norm (GroupNorm, 32 groups, eps 1e-6) on the NCHW feature map fused with
//       `permute(0, 2, 3, 1).reshape(BF, H*W, C)` (src/models/transformer_temporal.py:300-304): statistics per
//       (image, group) in one read, then normalise + affine + NCHW -> token-major transpose through a shared-memory tile.
//   tokens_to_nchw_add_kernel
//       `reshape(BF, H, W, C).permute(0, 3, 1, 2).contiguous()` + `hidden_states + residual` (:354-357).
//   blend_rows_kernel
//       AlphaBlender.forward (diffusers 0.24 models/resnet.py, called at :349-351):
//       out = alpha * x_spatial + (1 - alpha) * x_temporal with one alpha per image (frame).
// Every kernel moves 16 bytes per thread per access; the transposes go through a [64][72] bf16 tile so that both the
// NCHW side (64 spatial positions = 128 contiguous bytes per channel) and the token side (64 channels = 128 bytes per
// token) are read and written in full 128-byte segments.
#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kGnSlice = 16384;  // elements of one (image, group) span per statistics CTA
constexpr int kTile = 64;

__device__ __forceinline__ void unpack8f(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

// partial[(ng * nsplit + split)] = (sum, sum of squares) of one slice of the contiguous span of group ng
__global__ void __launch_bounds__(256)
groupnorm_stats_kernel(const tvi_bf16* __restrict__ x, double2* __restrict__ partial, int64_t span, int nsplit) {
  const int64_t ng = blockIdx.x / nsplit;
  const int split = (int)(blockIdx.x - ng * nsplit);
  const int64_t lo = (int64_t)split * kGnSlice;
  int64_t hi = lo + kGnSlice;
  if (hi > span) hi = span;
  const tvi_bf16* base = x + ng * span;
  float s = 0.f, ss = 0.f;
  for (int64_t i = lo + threadIdx.x * 8; i < hi; i += 256 * 8) {  // span % 8 == 0
    float f[8];
    unpack8f(__ldg(reinterpret_cast<const uint4*>(base + i)), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s += f[j];
      ss = fmaf(f[j], f[j], ss);
    }
  }
  // fixed-order block reduction in double
  __shared__ double red[2][8];
  double ds = s, dss = ss;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ds += __shfl_xor_sync(0xffffffffu, ds, o);
    dss += __shfl_xor_sync(0xffffffffu, dss, o);
  }
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = ds;
    red[1][threadIdx.x >> 5] = dss;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int w = 0; w < 8; ++w) {
      a += red[0][w];
      b += red[1][w];
    }
    partial[blockIdx.x] = make_double2(a, b);
  }
}

// y[n, s, c] = (x[n, c, s] - mean[n, g(c)]) * rstd[n, g(c)] * gamma[c] + beta[c];  kNorm = false: the bare NCHW -> token-major
// transpose (backward of tokens_to_nchw_add: d tokens = transpose(d out))
template <bool kNorm>
__global__ void __launch_bounds__(256)
groupnorm_tokens_kernel(const tvi_bf16* __restrict__ x, const double2* __restrict__ partial, const tvi_bf16* __restrict__ gamma,
                        const tvi_bf16* __restrict__ beta, tvi_bf16* __restrict__ y, int C, int HW, int groups, int nsplit,
                        int64_t ldy, float eps) {
  // [channel][spatial] as it is read, 16-byte chunk k of row r stored at k ^ (r / 8 % 8): the vector stores of a quarter
  // warp fill one 128-byte row, and the 2-byte gathers of the transposing read (rows 8 apart for the lanes of a warp) land
  // in eight different 16-byte bank groups.  (The first version transposed with scalar stores into a padded [spatial][channel]
  // tile whose rows 8 apart shared a bank: 8-way conflicts on every store.)
  __shared__ __align__(16) uint4 tile[kTile][8];
  __shared__ float s_scale[kTile], s_shift[kTile];
  const int n = blockIdx.z;
  const int c0 = blockIdx.y * kTile, s0 = blockIdx.x * kTile;
  const int cpg = C / groups;
  if (threadIdx.x < kTile) {
    const int c = c0 + threadIdx.x;
    float sc = kNorm ? 0.f : 1.f, sh = 0.f;
    if (kNorm && c < C) {
      const int g = c / cpg;
      const double2* pp = partial + ((int64_t)n * groups + g) * nsplit;
      double a = 0.0, b = 0.0;
      for (int i = 0; i < nsplit; ++i) {
        a += pp[i].x;
        b += pp[i].y;
      }
      const double cnt = (double)cpg * HW;
      const double mean = a / cnt;
      double var = b / cnt - mean * mean;
      if (var < 0.0) var = 0.0;
      const float rstd = (float)(1.0 / sqrt(var + (double)eps));
      const float ga = gamma ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(gamma + c)) : 1.f;
      const float be = beta ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(beta + c)) : 0.f;
      sc = rstd * ga;
      sh = be - (float)mean * sc;
    }
    s_scale[threadIdx.x] = sc;
    s_shift[threadIdx.x] = sh;
  }
  __syncthreads();
  const int chunk = threadIdx.x & 7;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int cl = (threadIdx.x >> 3) + 32 * j;
    const int c = c0 + cl, s = s0 + chunk * 8;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (c < C && s < HW) {  // HW % 8 == 0
      float f[8];
      unpack8f(__ldg(reinterpret_cast<const uint4*>(x + ((int64_t)n * C + c) * HW + s)), f);
      const float sc = s_scale[cl], sh = s_shift[cl];
      v = make_uint4(pack_bf16x2(fmaf(f[0], sc, sh), fmaf(f[1], sc, sh)), pack_bf16x2(fmaf(f[2], sc, sh), fmaf(f[3], sc, sh)),
                     pack_bf16x2(fmaf(f[4], sc, sh), fmaf(f[5], sc, sh)), pack_bf16x2(fmaf(f[6], sc, sh), fmaf(f[7], sc, sh)));
    }
    tile[cl][chunk ^ ((cl >> 3) & 7)] = v;
  }
  __syncthreads();
  const uint16_t* t16 = reinterpret_cast<const uint16_t*>(&tile[0][0]);
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int sl = (threadIdx.x >> 3) + 32 * j;
    const int s = s0 + sl, c = c0 + chunk * 8;
    if (s < HW && c < C) {  // C % 8 == 0
      // channels chunk * 8 + i of spatial position sl: row r = chunk * 8 + i holds it in chunk (sl / 8) ^ chunk
      const int off = (((sl >> 3) ^ chunk) << 3) + (sl & 7);
      uint32_t o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t lo = t16[(chunk * 8 + 2 * i) * 64 + off], hi = t16[(chunk * 8 + 2 * i + 1) * 64 + off];
        o[i] = lo | (hi << 16);
      }
      *reinterpret_cast<uint4*>(y + ((int64_t)n * HW + s) * ldy + c) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// out[n, c, s] = y[n, s, c] + residual[n, c, s]
__global__ void __launch_bounds__(256)
tokens_to_nchw_add_kernel(const tvi_bf16* __restrict__ y, const tvi_bf16* __restrict__ residual, tvi_bf16* __restrict__ out,
                          int C, int HW, int64_t ldy) {
  __shared__ __align__(16) uint4 tile[kTile][8];  // [spatial][channel] as read, chunk k of row r at k ^ (r / 8 % 8) (see above)
  const int n = blockIdx.z;
  const int c0 = blockIdx.y * kTile, s0 = blockIdx.x * kTile;
  const int chunk = threadIdx.x & 7;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int sl = (threadIdx.x >> 3) + 32 * j;
    const int s = s0 + sl, c = c0 + chunk * 8;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (s < HW && c < C) v = __ldg(reinterpret_cast<const uint4*>(y + ((int64_t)n * HW + s) * ldy + c));
    tile[sl][chunk ^ ((sl >> 3) & 7)] = v;
  }
  __syncthreads();
  const uint16_t* t16 = reinterpret_cast<const uint16_t*>(&tile[0][0]);
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int cl = (threadIdx.x >> 3) + 32 * j;
    const int c = c0 + cl, s = s0 + chunk * 8;
    if (c < C && s < HW) {
      const int64_t off = ((int64_t)n * C + c) * HW + s;
      // spatial positions chunk * 8 + i of channel cl: row r = chunk * 8 + i holds it in chunk (cl / 8) ^ chunk
      const int so = (((cl >> 3) ^ chunk) << 3) + (cl & 7);
      float a[8], b[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = __uint_as_float((uint32_t)t16[(chunk * 8 + i) * 64 + so] << 16);
      if (residual != nullptr) {
        unpack8f(__ldg(reinterpret_cast<const uint4*>(residual + off)), b);
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] += b[i];
      }
      *reinterpret_cast<uint4*>(out + off) =
          make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
    }
  }
}

// out[r, :] = alpha[r / rows_per_group] * xs[r, :] + (1 - alpha[...]) * xt[r, :]
__global__ void __launch_bounds__(256)
blend_rows_kernel(const tvi_bf16* __restrict__ xs, const tvi_bf16* __restrict__ xt, const float* __restrict__ alpha,
                  tvi_bf16* __restrict__ out, int64_t nrows, int64_t rows_per_group, int C8, int64_t lds, int64_t ldt, int64_t ldo) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nrows * C8) return;
  const int64_t r = idx / C8;
  const int c = (int)(idx - r * C8);
  const float a = __ldg(alpha + r / rows_per_group);
  const uint4 us = __ldg(reinterpret_cast<const uint4*>(xs + r * lds) + c);
  if (a == 1.0f) {  // image-only frames take the spatial branch untouched (also when x_temporal is not finite)
    *(reinterpret_cast<uint4*>(out + r * ldo) + c) = us;
    return;
  }
  float fs[8], ft[8];
  unpack8f(us, fs);
  unpack8f(__ldg(reinterpret_cast<const uint4*>(xt + r * ldt) + c), ft);
  const float b = 1.0f - a;
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
    o[i] = pack_bf16x2(fmaf(a, fs[2 * i], b * ft[2 * i]), fmaf(a, fs[2 * i + 1], b * ft[2 * i + 1]));
  *(reinterpret_cast<uint4*>(out + r * ldo) + c) = make_uint4(o[0], o[1], o[2], o[3]);
}

// ---------------------------------------------------------------------------------------------------------------
// Backward of the glue (training through the module-level swap, src/train.py back-propagates through
// transformer_temporal.py:300-304, :349-357).  With xhat = (x - mean) rstd per (image, group) and g = dn * gamma:
//   dx = rstd * (g - mean_grp(g) - xhat * mean_grp(g xhat)),  dgamma_c = sum_{n,s} dn xhat,  dbeta_c = sum_{n,s} dn.
// Per (image, channel) A = sum_s dn and B = sum_s dn x give every group mean and both parameter gradients, so the
// backward is: statistics of x (forward kernel), one tile pass for A / B (fp32 partials per spatial tile, summed in a
// fixed order: deterministic, no atomics), one tile pass that writes dx = k1_c dn + k2_c x + k3_c in NCHW.
// ---------------------------------------------------------------------------------------------------------------
__global__ void gn_stats_finalize_kernel(const double2* __restrict__ partial, float2* __restrict__ mr, int64_t ngroups, int nsplit,
                                         double cnt, float eps) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ngroups) return;
  double a = 0.0, b = 0.0;
  for (int k = 0; k < nsplit; ++k) {
    a += partial[i * nsplit + k].x;
    b += partial[i * nsplit + k].y;
  }
  const double mean = a / cnt;
  double var = b / cnt - mean * mean;
  if (var < 0.0) var = 0.0;
  mr[i] = make_float2((float)mean, (float)(1.0 / sqrt(var + (double)eps)));
}

// partial[((n * s_tiles + st) * C + c)] = (sum_s dn[n, s, c], sum_s dn[n, s, c] x[n, c, s]) over the 64 positions of tile st
__global__ void __launch_bounds__(256)
gn_bwd_reduce_kernel(const tvi_bf16* __restrict__ x, const tvi_bf16* __restrict__ dn, float2* __restrict__ partial, int C, int HW,
                     int64_t lddn) {
  __shared__ __align__(16) uint4 tile[kTile][8];  // dn as [spatial][channel], chunk k of row r at k ^ (r / 8 % 8) (see groupnorm_tokens_kernel)
  const int n = blockIdx.z;
  const int c0 = blockIdx.y * kTile, s0 = blockIdx.x * kTile;
  const int chunk = threadIdx.x & 7;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int sl = (threadIdx.x >> 3) + 32 * j;
    const int sp = s0 + sl, c = c0 + chunk * 8;
    uint4 u = make_uint4(0, 0, 0, 0);
    if (sp < HW && c < C) u = __ldg(reinterpret_cast<const uint4*>(dn + ((int64_t)n * HW + sp) * lddn + c));
    tile[sl][chunk ^ ((sl >> 3) & 7)] = u;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int cl = (threadIdx.x >> 3) + 32 * j;
    const int c = c0 + cl, sp = s0 + chunk * 8;
    float a = 0.f, b = 0.f;
    if (c < C && sp < HW) {
      float f[8];
      unpack8f(__ldg(reinterpret_cast<const uint4*>(x + ((int64_t)n * C + c) * HW + sp)), f);
      const uint16_t* t16 = reinterpret_cast<const uint16_t*>(&tile[0][0]);
      const int so = (((cl >> 3) ^ chunk) << 3) + (cl & 7);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float d = __uint_as_float((uint32_t)t16[(chunk * 8 + i) * 64 + so] << 16);
        a += d;
        b = fmaf(d, f[i], b);
      }
    }
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {  // the 8 chunks of a channel are 8 consecutive lanes
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    if (chunk == 0 && c < C) partial[((int64_t)n * gridDim.x + blockIdx.x) * C + c] = make_float2(a, b);
  }
}

// ab[n * C + c] = sum over the spatial tiles, in order
__global__ void gn_bwd_finalize_kernel(const float2* __restrict__ partial, float2* __restrict__ ab, int64_t N, int C, int s_tiles) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C) return;
  const int64_t n = i / C;
  const int c = (int)(i - n * C);
  float a = 0.f, b = 0.f;
  for (int t = 0; t < s_tiles; ++t) {
    const float2 v = partial[(n * s_tiles + t) * C + c];
    a += v.x;
    b += v.y;
  }
  ab[i] = make_float2(a, b);
}

__global__ void gn_bwd_params_kernel(const float2* __restrict__ ab, const float2* __restrict__ mr, tvi_bf16* __restrict__ dgamma,
                                     tvi_bf16* __restrict__ dbeta, int64_t N, int C, int groups) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const int g = c / (C / groups);
  float dg = 0.f, db = 0.f;
  for (int64_t n = 0; n < N; ++n) {
    const float2 v = ab[n * C + c];
    const float2 m = mr[n * groups + g];
    dg += m.y * (v.y - m.x * v.x);
    db += v.x;
  }
  if (dgamma != nullptr) reinterpret_cast<__nv_bfloat16*>(dgamma)[c] = __float2bfloat16_rn(dg);
  if (dbeta != nullptr) reinterpret_cast<__nv_bfloat16*>(dbeta)[c] = __float2bfloat16_rn(db);
}

// dx[n, c, s] = k1 dn[n, s, c] + k2 x[n, c, s] + k3 with k1 = rstd gamma_c, k2 = -rstd^2 m2, k3 = -rstd m1 + rstd^2 m2 mean,
// m1 = mean_grp(gamma A), m2 = mean_grp(gamma rstd (B - mean A))
__global__ void __launch_bounds__(256)
gn_bwd_apply_kernel(const tvi_bf16* __restrict__ x, const tvi_bf16* __restrict__ dn, const tvi_bf16* __restrict__ gamma,
                    const float2* __restrict__ ab, const float2* __restrict__ mr, tvi_bf16* __restrict__ dx, int C, int HW,
                    int groups, int64_t lddn) {
  __shared__ __align__(16) uint4 tile[kTile][8];  // dn as [spatial][channel], chunk k of row r at k ^ (r / 8 % 8) (see groupnorm_tokens_kernel)
  __shared__ float s_k1[kTile], s_k2[kTile], s_k3[kTile];
  const int n = blockIdx.z;
  const int c0 = blockIdx.y * kTile, s0 = blockIdx.x * kTile;
  const int cpg = C / groups;
  if (threadIdx.x < kTile) {
    const int c = c0 + threadIdx.x;
    float k1 = 0.f, k2 = 0.f, k3 = 0.f;
    if (c < C) {
      const int g = c / cpg;
      const float2 m = mr[(int64_t)n * groups + g];
      float s1 = 0.f, s2 = 0.f;
      for (int j = g * cpg; j < (g + 1) * cpg; ++j) {
        const float ga = gamma ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(gamma + j)) : 1.f;
        const float2 v = ab[(int64_t)n * C + j];
        s1 = fmaf(ga, v.x, s1);
        s2 = fmaf(ga * m.y, v.y - m.x * v.x, s2);
      }
      const float inv = 1.f / ((float)cpg * (float)HW);
      const float m1 = s1 * inv, m2 = s2 * inv;
      const float ga = gamma ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(gamma + c)) : 1.f;
      k1 = m.y * ga;
      k2 = -m.y * m.y * m2;
      k3 = -m.y * m1 - k2 * m.x;
    }
    s_k1[threadIdx.x] = k1;
    s_k2[threadIdx.x] = k2;
    s_k3[threadIdx.x] = k3;
  }
  const int chunk = threadIdx.x & 7;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int sl = (threadIdx.x >> 3) + 32 * j;
    const int sp = s0 + sl, c = c0 + chunk * 8;
    uint4 u = make_uint4(0, 0, 0, 0);
    if (sp < HW && c < C) u = __ldg(reinterpret_cast<const uint4*>(dn + ((int64_t)n * HW + sp) * lddn + c));
    tile[sl][chunk ^ ((sl >> 3) & 7)] = u;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int cl = (threadIdx.x >> 3) + 32 * j;
    const int c = c0 + cl, sp = s0 + chunk * 8;
    if (c < C && sp < HW) {
      const int64_t off = ((int64_t)n * C + c) * HW + sp;
      float f[8], o[8];
      unpack8f(__ldg(reinterpret_cast<const uint4*>(x + off)), f);
      const float k1 = s_k1[cl], k2 = s_k2[cl], k3 = s_k3[cl];
      const uint16_t* t16 = reinterpret_cast<const uint16_t*>(&tile[0][0]);
      const int so = (((cl >> 3) ^ chunk) << 3) + (cl & 7);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        o[i] = fmaf(k1, __uint_as_float((uint32_t)t16[(chunk * 8 + i) * 64 + so] << 16), fmaf(k2, f[i], k3));
      *reinterpret_cast<uint4*>(dx + off) =
          make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
    }
  }
}

// AlphaBlender backward: dxs = alpha dy, dxt = (1 - alpha) dy; dalpha[frame] = sum dy (xs - xt) as per-CTA fp32 partials
// [frames][blocks_per_frame] (summed by the caller in a fixed order) when xs / xt are given
__global__ void __launch_bounds__(256)
blend_rows_bwd_kernel(const tvi_bf16* __restrict__ dy, const tvi_bf16* __restrict__ xs, const tvi_bf16* __restrict__ xt,
                      const float* __restrict__ alpha, tvi_bf16* __restrict__ dxs, tvi_bf16* __restrict__ dxt,
                      float* __restrict__ dalpha_partial, int64_t rows_per_group, int C8, int64_t ld) {
  const int64_t frame = blockIdx.y;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // 16-byte chunk inside the frame
  const bool act = idx < rows_per_group * C8;
  const float a = __ldg(alpha + frame), b = 1.0f - a;
  float acc = 0.f;
  if (act) {
    const int64_t r = frame * rows_per_group + idx / C8;
    const int c = (int)(idx % C8);
    float g[8];
    unpack8f(__ldg(reinterpret_cast<const uint4*>(dy + r * ld) + c), g);
    *(reinterpret_cast<uint4*>(dxs + r * ld) + c) = make_uint4(pack_bf16x2(a * g[0], a * g[1]), pack_bf16x2(a * g[2], a * g[3]),
                                                               pack_bf16x2(a * g[4], a * g[5]), pack_bf16x2(a * g[6], a * g[7]));
    *(reinterpret_cast<uint4*>(dxt + r * ld) + c) = make_uint4(pack_bf16x2(b * g[0], b * g[1]), pack_bf16x2(b * g[2], b * g[3]),
                                                               pack_bf16x2(b * g[4], b * g[5]), pack_bf16x2(b * g[6], b * g[7]));
    if (dalpha_partial != nullptr) {
      float fs[8], ft[8];
      unpack8f(__ldg(reinterpret_cast<const uint4*>(xs + r * ld) + c), fs);
      unpack8f(__ldg(reinterpret_cast<const uint4*>(xt + r * ld) + c), ft);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc = fmaf(g[i], fs[i] - ft[i], acc);
    }
  }
  if (dalpha_partial == nullptr) return;
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];
    dalpha_partial[frame * gridDim.x + blockIdx.x] = t;
  }
}

int gn_splits(int64_t span) { return (int)((span + kGnSlice - 1) / kGnSlice); }

}  // namespace
}  // namespace tvi

extern "C" size_t tvi_groupnorm_tokens_workspace_bytes(int64_t N, int C, int HW, int groups) {
  if (N <= 0 || C <= 0 || HW <= 0 || groups <= 0 || C % groups != 0) return 0;
  return (size_t)N * groups * tvi::gn_splits((int64_t)(C / groups) * HW) * sizeof(double2);
}

extern "C" int tvi_groupnorm_tokens_bf16(const tvi_bf16* x, const tvi_bf16* gamma, const tvi_bf16* beta, tvi_bf16* y,
                                         void* workspace, int64_t N, int C, int HW, int groups, int64_t ldy, float eps,
                                         void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(N >= 0 && C >= 0 && HW >= 0, "tvi_groupnorm_tokens_bf16: negative size");
  if (N == 0 || C == 0 || HW == 0) return TVI_OK;
  TVI_CHECK_ARG(x && y && workspace, "tvi_groupnorm_tokens_bf16: null pointer");
  TVI_CHECK_ARG(groups >= 1 && C % groups == 0, "tvi_groupnorm_tokens_bf16: C=%d is not a multiple of groups=%d", C, groups);
  TVI_CHECK_ARG(C % 8 == 0 && HW % 8 == 0, "tvi_groupnorm_tokens_bf16: C=%d and H*W=%d must be multiples of 8", C, HW);
  TVI_CHECK_ARG(ldy % 8 == 0 && ldy >= C, "tvi_groupnorm_tokens_bf16: bad output row stride");
  TVI_CHECK_ARG(((uintptr_t)x | (uintptr_t)y | (uintptr_t)workspace) % 16 == 0, "tvi_groupnorm_tokens_bf16: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(eps > 0.f, "tvi_groupnorm_tokens_bf16: eps must be positive");
  TVI_CHECK_ARG(N <= 65535, "tvi_groupnorm_tokens_bf16: more than 65535 images in one call");
  const int64_t span = (int64_t)(C / groups) * HW;
  const int nsplit = gn_splits(span);
  TVI_CHECK_ARG(N * groups * nsplit < ((int64_t)1 << 31), "tvi_groupnorm_tokens_bf16: grid too large");
  cudaStream_t st = (cudaStream_t)stream;
  groupnorm_stats_kernel<<<(unsigned)(N * groups * nsplit), 256, 0, st>>>(x, (double2*)workspace, span, nsplit);
  TVI_LAUNCH_OK("groupnorm_stats_kernel");
  dim3 grid((HW + kTile - 1) / kTile, (C + kTile - 1) / kTile, (unsigned)N);
  groupnorm_tokens_kernel<true><<<grid, 256, 0, st>>>(x, (const double2*)workspace, gamma, beta, y, C, HW, groups, nsplit, ldy, eps);
  TVI_LAUNCH_OK("groupnorm_tokens_kernel");
  return TVI_OK;
}

extern "C" int tvi_tokens_to_nchw_add_bf16(const tvi_bf16* y, const tvi_bf16* residual, tvi_bf16* out, int64_t N, int C, int HW,
                                           int64_t ldy, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(N >= 0 && C >= 0 && HW >= 0, "tvi_tokens_to_nchw_add_bf16: negative size");
  if (N == 0 || C == 0 || HW == 0) return TVI_OK;
  TVI_CHECK_ARG(y && out, "tvi_tokens_to_nchw_add_bf16: null pointer");
  TVI_CHECK_ARG(C % 8 == 0 && HW % 8 == 0, "tvi_tokens_to_nchw_add_bf16: C=%d and H*W=%d must be multiples of 8", C, HW);
  TVI_CHECK_ARG(ldy % 8 == 0 && ldy >= C, "tvi_tokens_to_nchw_add_bf16: bad row stride");
  TVI_CHECK_ARG(((uintptr_t)y | (uintptr_t)residual | (uintptr_t)out) % 16 == 0, "tvi_tokens_to_nchw_add_bf16: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(N <= 65535, "tvi_tokens_to_nchw_add_bf16: more than 65535 images in one call");
  dim3 grid((HW + kTile - 1) / kTile, (C + kTile - 1) / kTile, (unsigned)N);
  tokens_to_nchw_add_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(y, residual, out, C, HW, ldy);
  TVI_LAUNCH_OK("tokens_to_nchw_add_kernel");
  return TVI_OK;
}

extern "C" int tvi_blend_rows_bf16(const tvi_bf16* x_spatial, const tvi_bf16* x_temporal, const float* alpha, tvi_bf16* out,
                                   int64_t nrows, int64_t rows_per_group, int C, int64_t lds, int64_t ldt, int64_t ldo,
                                   void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(nrows >= 0 && C >= 0, "tvi_blend_rows_bf16: negative size");
  if (nrows == 0 || C == 0) return TVI_OK;
  TVI_CHECK_ARG(x_spatial && x_temporal && alpha && out, "tvi_blend_rows_bf16: null pointer");
  TVI_CHECK_ARG(rows_per_group >= 1 && nrows % rows_per_group == 0, "tvi_blend_rows_bf16: nrows must be a multiple of rows_per_group");
  TVI_CHECK_ARG(C % 8 == 0 && lds % 8 == 0 && ldt % 8 == 0 && ldo % 8 == 0 && lds >= C && ldt >= C && ldo >= C,
                "tvi_blend_rows_bf16: C and the row strides must be multiples of 8");
  TVI_CHECK_ARG(((uintptr_t)x_spatial | (uintptr_t)x_temporal | (uintptr_t)out) % 16 == 0 && (uintptr_t)alpha % 4 == 0,
                "tvi_blend_rows_bf16: pointers must be 16-byte aligned");
  const int64_t total = nrows * (C / 8);
  TVI_CHECK_ARG((total + 255) / 256 < ((int64_t)1 << 31), "tvi_blend_rows_bf16: grid too large");
  blend_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x_spatial, x_temporal, alpha, out, nrows,
                                                                                       rows_per_group, C / 8, lds, ldt, ldo);
  TVI_LAUNCH_OK("blend_rows_kernel");
  return TVI_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// backward entry points of the glue
// ---------------------------------------------------------------------------------------------------------------
extern "C" int tvi_nchw_to_tokens_bf16(const tvi_bf16* x, tvi_bf16* y, int64_t N, int C, int HW, int64_t ldy, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(N >= 0 && C >= 0 && HW >= 0, "tvi_nchw_to_tokens_bf16: negative size");
  if (N == 0 || C == 0 || HW == 0) return TVI_OK;
  TVI_CHECK_ARG(x && y, "tvi_nchw_to_tokens_bf16: null pointer");
  TVI_CHECK_ARG(C % 8 == 0 && HW % 8 == 0, "tvi_nchw_to_tokens_bf16: C=%d and H*W=%d must be multiples of 8", C, HW);
  TVI_CHECK_ARG(ldy % 8 == 0 && ldy >= C, "tvi_nchw_to_tokens_bf16: bad output row stride");
  TVI_CHECK_ARG(((uintptr_t)x | (uintptr_t)y) % 16 == 0, "tvi_nchw_to_tokens_bf16: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(N <= 65535, "tvi_nchw_to_tokens_bf16: more than 65535 images in one call");
  dim3 grid((HW + kTile - 1) / kTile, (C + kTile - 1) / kTile, (unsigned)N);
  groupnorm_tokens_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(x, nullptr, nullptr, nullptr, y, C, HW, 1, 0, ldy, 1.f);
  TVI_LAUNCH_OK("groupnorm_tokens_kernel<transpose>");
  return TVI_OK;
}

namespace tvi {
namespace {
struct GnBwdLayout {
  size_t stats, mr, partial, ab, total;
  int nsplit, s_tiles;
};
GnBwdLayout gn_bwd_layout(int64_t N, int C, int HW, int groups) {
  GnBwdLayout l;
  l.nsplit = gn_splits((int64_t)(C / groups) * HW);
  l.s_tiles = (HW + kTile - 1) / kTile;
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  l.stats = 0;
  l.mr = up(l.stats + (size_t)N * groups * l.nsplit * sizeof(double2));
  l.partial = up(l.mr + (size_t)N * groups * sizeof(float2));
  l.ab = up(l.partial + (size_t)N * l.s_tiles * C * sizeof(float2));
  l.total = up(l.ab + (size_t)N * C * sizeof(float2));
  return l;
}
}  // namespace
}  // namespace tvi

extern "C" size_t tvi_groupnorm_tokens_bwd_workspace_bytes(int64_t N, int C, int HW, int groups) {
  if (N <= 0 || C <= 0 || HW <= 0 || groups <= 0 || C % groups != 0) return 0;
  return tvi::gn_bwd_layout(N, C, HW, groups).total;
}

extern "C" int tvi_groupnorm_tokens_bwd_bf16(const tvi_bf16* x, const tvi_bf16* dn, const tvi_bf16* gamma, tvi_bf16* dx,
                                             tvi_bf16* dgamma, tvi_bf16* dbeta, void* workspace, int64_t N, int C, int HW,
                                             int groups, int64_t lddn, float eps, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(N >= 0 && C >= 0 && HW >= 0, "tvi_groupnorm_tokens_bwd_bf16: negative size");
  if (N == 0 || C == 0 || HW == 0) return TVI_OK;
  TVI_CHECK_ARG(x && dn && workspace, "tvi_groupnorm_tokens_bwd_bf16: null pointer");
  TVI_CHECK_ARG(dx || dgamma || dbeta, "tvi_groupnorm_tokens_bwd_bf16: no gradient requested");
  TVI_CHECK_ARG(groups >= 1 && C % groups == 0, "tvi_groupnorm_tokens_bwd_bf16: C=%d is not a multiple of groups=%d", C, groups);
  TVI_CHECK_ARG(C % 8 == 0 && HW % 8 == 0, "tvi_groupnorm_tokens_bwd_bf16: C=%d and H*W=%d must be multiples of 8", C, HW);
  TVI_CHECK_ARG(lddn % 8 == 0 && lddn >= C, "tvi_groupnorm_tokens_bwd_bf16: bad row stride of dn");
  TVI_CHECK_ARG(((uintptr_t)x | (uintptr_t)dn | (uintptr_t)dx | (uintptr_t)workspace) % 16 == 0,
                "tvi_groupnorm_tokens_bwd_bf16: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(eps > 0.f, "tvi_groupnorm_tokens_bwd_bf16: eps must be positive");
  TVI_CHECK_ARG(N <= 65535, "tvi_groupnorm_tokens_bwd_bf16: more than 65535 images in one call");
  const GnBwdLayout l = gn_bwd_layout(N, C, HW, groups);
  TVI_CHECK_ARG(N * groups * l.nsplit < ((int64_t)1 << 31), "tvi_groupnorm_tokens_bwd_bf16: grid too large");
  uint8_t* ws = (uint8_t*)workspace;
  double2* stats = (double2*)(ws + l.stats);
  float2* mr = (float2*)(ws + l.mr);
  float2* partial = (float2*)(ws + l.partial);
  float2* ab = (float2*)(ws + l.ab);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t span = (int64_t)(C / groups) * HW;
  groupnorm_stats_kernel<<<(unsigned)(N * groups * l.nsplit), 256, 0, st>>>(x, stats, span, l.nsplit);
  TVI_LAUNCH_OK("groupnorm_stats_kernel");
  gn_stats_finalize_kernel<<<(unsigned)((N * groups + 127) / 128), 128, 0, st>>>(stats, mr, N * groups, l.nsplit, (double)span, eps);
  TVI_LAUNCH_OK("gn_stats_finalize_kernel");
  dim3 grid((unsigned)l.s_tiles, (C + kTile - 1) / kTile, (unsigned)N);
  gn_bwd_reduce_kernel<<<grid, 256, 0, st>>>(x, dn, partial, C, HW, lddn);
  TVI_LAUNCH_OK("gn_bwd_reduce_kernel");
  gn_bwd_finalize_kernel<<<(unsigned)((N * C + 255) / 256), 256, 0, st>>>(partial, ab, N, C, l.s_tiles);
  TVI_LAUNCH_OK("gn_bwd_finalize_kernel");
  if (dgamma != nullptr || dbeta != nullptr) {
    gn_bwd_params_kernel<<<(C + 127) / 128, 128, 0, st>>>(ab, mr, dgamma, dbeta, N, C, groups);
    TVI_LAUNCH_OK("gn_bwd_params_kernel");
  }
  if (dx != nullptr) {
    gn_bwd_apply_kernel<<<grid, 256, 0, st>>>(x, dn, gamma, ab, mr, dx, C, HW, groups, lddn);
    TVI_LAUNCH_OK("gn_bwd_apply_kernel");
  }
  return TVI_OK;
}

extern "C" int tvi_blend_rows_bwd_blocks(int64_t rows_per_group, int C) {
  if (rows_per_group <= 0 || C <= 0) return 0;
  return (int)((rows_per_group * (C / 8) + 255) / 256);
}

extern "C" int tvi_blend_rows_bwd_bf16(const tvi_bf16* dy, const tvi_bf16* x_spatial, const tvi_bf16* x_temporal, const float* alpha,
                                       tvi_bf16* dx_spatial, tvi_bf16* dx_temporal, float* dalpha_partial, int64_t frames,
                                       int64_t rows_per_group, int C, int64_t ld, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(frames >= 0 && rows_per_group >= 0 && C >= 0, "tvi_blend_rows_bwd_bf16: negative size");
  if (frames == 0 || rows_per_group == 0 || C == 0) return TVI_OK;
  TVI_CHECK_ARG(dy && alpha && dx_spatial && dx_temporal, "tvi_blend_rows_bwd_bf16: null pointer");
  TVI_CHECK_ARG(dalpha_partial == nullptr || (x_spatial && x_temporal), "tvi_blend_rows_bwd_bf16: dalpha needs x_spatial and x_temporal");
  TVI_CHECK_ARG(C % 8 == 0 && ld % 8 == 0 && ld >= C, "tvi_blend_rows_bwd_bf16: C and the row stride must be multiples of 8");
  TVI_CHECK_ARG(((uintptr_t)dy | (uintptr_t)x_spatial | (uintptr_t)x_temporal | (uintptr_t)dx_spatial | (uintptr_t)dx_temporal) % 16 == 0 &&
                    ((uintptr_t)alpha | (uintptr_t)dalpha_partial) % 4 == 0,
                "tvi_blend_rows_bwd_bf16: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(frames <= 65535, "tvi_blend_rows_bwd_bf16: more than 65535 frames in one call");
  dim3 grid((unsigned)tvi_blend_rows_bwd_blocks(rows_per_group, C), (unsigned)frames);
  blend_rows_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dy, x_spatial, x_temporal, alpha, dx_spatial, dx_temporal,
                                                                dalpha_partial, rows_per_group, C / 8, ld);
  TVI_LAUNCH_OK("blend_rows_bwd_kernel");
  return TVI_OK;
}
