// LayerNorm over the channel dimension of token-major bf16 activations (N1 of SURVEY.md 8.6).
// Replaces nn.LayerNorm norm1 / norm2 / norm3 / norm_in of BasicTransformerBlock and
// TemporalBasicTransformerBlock (src/models/attention.py:270, :300, :321, :441, :446, :462): per token
//   y = (x - mean) / sqrt(var + eps) * gamma + beta        (biased variance, statistics in fp32).
// HBM-bound: 2 * M * C * 2 bytes.  A warp owns R consecutive rows (R = 2 at C <= 1024) and issues all their 128-bit
// loads before the first use, so 1.3+ KB per warp are in flight; a row stays in registers between the statistics and
// the normalisation (C <= 2048: up to 8 chunks of 8 channels per lane); two-pass variance (sum of squared
// deviations) so that large means do not cancel.
#include <cstdlib>

#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kLnWarps = 8;

template <int NCH, int R>  // 16-byte chunks per lane, rows per warp (all R * NCH loads are in flight before the first use)
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_kernel(const tvi_bf16* __restrict__ x, const tvi_bf16* __restrict__ gamma, const tvi_bf16* __restrict__ beta,
                 tvi_bf16* __restrict__ y, int64_t M, int C8, int64_t ldx, int64_t ldy, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t row0 = ((int64_t)blockIdx.x * kLnWarps + (threadIdx.x >> 5)) * R;
  if (row0 >= M) return;
  uint4 raw[R][NCH];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int64_t row = row0 + r < M ? row0 + r : M - 1;  // tail rows re-read the last row (never stored)
    const uint4* xr = reinterpret_cast<const uint4*>(x + row * ldx);
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = lane + 32 * i;
      raw[r][i] = c < C8 ? __ldg(xr + c) : make_uint4(0u, 0u, 0u, 0u);
    }
  }
  // affine parameters: one 16-byte chunk of gamma / beta per (lane, i), shared by the R rows
  float g[NCH][8], b[NCH][8];
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
    uint4 ug = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);  // bf16 1.0 pairs
    uint4 ub = make_uint4(0u, 0u, 0u, 0u);
    if (c < C8) {
      if (gamma != nullptr) ug = __ldg(reinterpret_cast<const uint4*>(gamma) + c);
      if (beta != nullptr) ub = __ldg(reinterpret_cast<const uint4*>(beta) + c);
    }
    const __nv_bfloat162* hg = reinterpret_cast<const __nv_bfloat162*>(&ug);
    const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&ub);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float2 tg = __bfloat1622float2(hg[q]), tb = __bfloat1622float2(hb[q]);
      g[i][2 * q] = tg.x;
      g[i][2 * q + 1] = tg.y;
      b[i][2 * q] = tb.x;
      b[i][2 * q + 1] = tb.y;
    }
  }
  const float inv_c = 1.0f / (float)(C8 * 8);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    float v[NCH][8];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[r][i]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(h[q]);
        v[i][2 * q] = t.x;
        v[i][2 * q + 1] = t.y;
        sum += t.x + t.y;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum * inv_c;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      if (lane + 32 * i < C8) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float d = v[i][q] - mean;
          sq = fmaf(d, d, sq);
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = rsqrtf(sq * inv_c + eps);
    if (row0 + r < M) {
      uint4* yr = reinterpret_cast<uint4*>(y + (row0 + r) * ldy);
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        if (c < C8) {
          uint4 o;
          o.x = pack_bf16x2(fmaf((v[i][0] - mean) * rstd, g[i][0], b[i][0]), fmaf((v[i][1] - mean) * rstd, g[i][1], b[i][1]));
          o.y = pack_bf16x2(fmaf((v[i][2] - mean) * rstd, g[i][2], b[i][2]), fmaf((v[i][3] - mean) * rstd, g[i][3], b[i][3]));
          o.z = pack_bf16x2(fmaf((v[i][4] - mean) * rstd, g[i][4], b[i][4]), fmaf((v[i][5] - mean) * rstd, g[i][5], b[i][5]));
          o.w = pack_bf16x2(fmaf((v[i][6] - mean) * rstd, g[i][6], b[i][6]), fmaf((v[i][7] - mean) * rstd, g[i][7], b[i][7]));
          yr[c] = o;
        }
      }
    }
  }
}

// The three SVD widths (C = 320 / 640 / 1280 = 40 G columns, G = 8 / 16 / 32): G lanes share a row, a warp owns R = 32 / G
// consecutive rows, every lane the five 16-byte chunks (lane % G) + G j of its row — every lane busy in every iteration
// (the row-per-warp layout above leaves 24 of 64 lane slots idle at C = 320 and was issue-bound at 3.5 TB/s), a warp-level
// load covers whole 128-byte lines, and the two row reductions are log2(G) shuffles on one value per lane.  (The first
// version of this kernel dealt chunk lane + 32 i to row (lane + 32 i) / C8: per-row sums then needed a select per chunk
// and a full butterfly for each of the R rows — 40 shuffles per pass at C = 320 against 6 now.)  Packed f32x2 arithmetic.
// kStats: write (mean, rstd) of every row to `stats` [M][2] fp32 instead of the normalised rows (the LayerNorm is then
// applied by the epilogue of the GEMM that consumes the rows, tvi_gemm_bf16_tn_ln): a read-only pass, half the traffic.
template <int R, bool kStats>
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_flat_kernel(const tvi_bf16* __restrict__ x, const tvi_bf16* __restrict__ gamma, const tvi_bf16* __restrict__ beta,
                      tvi_bf16* __restrict__ y, float2* __restrict__ stats, int64_t M, int64_t ldx, int64_t ldy, float eps) {
  constexpr int G = 32 / R;
  constexpr int NCH = 5;
  const int lane = threadIdx.x & 31;
  const int rs = lane / G, lg = lane % G;
  const int64_t row0 = ((int64_t)blockIdx.x * kLnWarps + (threadIdx.x >> 5)) * R;
  if (row0 >= M) return;
  const int64_t row = row0 + rs;
  const bool live = row < M;
  const int64_t rr = live ? row : M - 1;  // tail slots re-read the last row (never stored)
  uint4 raw[NCH];
#pragma unroll
  for (int i = 0; i < NCH; ++i) raw[i] = __ldg(reinterpret_cast<const uint4*>(x + rr * ldx) + lg + G * i);
  uint64_t v[NCH][4];
  uint64_t acc = f32x2_pack(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const uint32_t w[4] = {raw[i].x, raw[i].y, raw[i].z, raw[i].w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      v[i][q] = f32x2_pack(__uint_as_float(w[q] << 16), __uint_as_float(w[q] & 0xffff0000u));
      acc = f32x2_add(acc, v[i][q]);
    }
  }
  float a0, a1;
  f32x2_unpack(acc, a0, a1);
  float sum = a0 + a1;
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  constexpr float inv_c = 1.0f / (float)(40 * G);
  const float mean = sum * inv_c;
  const uint64_t nm = f32x2_pack(-mean, -mean);
  acc = f32x2_pack(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < NCH; ++i)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      v[i][q] = f32x2_add(v[i][q], nm);  // deviations
      acc = f32x2_fma(v[i][q], v[i][q], acc);
    }
  f32x2_unpack(acc, a0, a1);
  float sq = a0 + a1;
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = rsqrtf(sq * inv_c + eps);
  if (kStats) {
    if (lg == 0 && live) stats[row] = make_float2(mean, rstd);
    return;
  }
  const uint64_t rs2 = f32x2_pack(rstd, rstd);
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int col = lg + G * i;
    uint4 ug = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);  // bf16 1.0 pairs
    uint4 ub = make_uint4(0u, 0u, 0u, 0u);
    if (gamma != nullptr) ug = __ldg(reinterpret_cast<const uint4*>(gamma) + col);
    if (beta != nullptr) ub = __ldg(reinterpret_cast<const uint4*>(beta) + col);
    const uint32_t wg[4] = {ug.x, ug.y, ug.z, ug.w}, wb[4] = {ub.x, ub.y, ub.z, ub.w};
    uint32_t o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint64_t g2 = f32x2_pack(__uint_as_float(wg[q] << 16), __uint_as_float(wg[q] & 0xffff0000u));
      const uint64_t b2 = f32x2_pack(__uint_as_float(wb[q] << 16), __uint_as_float(wb[q] & 0xffff0000u));
      float lo, hi;
      f32x2_unpack(f32x2_fma(f32x2_mul(v[i][q], rs2), g2, b2), lo, hi);
      o[q] = pack_bf16x2(lo, hi);
    }
    if (live) *(reinterpret_cast<uint4*>(y + row * ldy) + col) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

template <int R>
int launch_ln_flat(const tvi_bf16* x, const tvi_bf16* gamma, const tvi_bf16* beta, tvi_bf16* y, int64_t M, int64_t ldx, int64_t ldy,
                   float eps, cudaStream_t st) {
  const int64_t rows_per_block = (int64_t)kLnWarps * R;
  const int64_t blocks = (M + rows_per_block - 1) / rows_per_block;
  TVI_CHECK_ARG(blocks < ((int64_t)1 << 31), "tvi_layernorm_bf16: too many rows for one call");
  layernorm_flat_kernel<R, false><<<(unsigned)blocks, kLnWarps * 32, 0, st>>>(x, gamma, beta, y, nullptr, M, ldx, ldy, eps);
  TVI_LAUNCH_OK("layernorm_flat_kernel");
  return TVI_OK;
}

template <int R>
int launch_ln_stats(const tvi_bf16* x, float2* stats, int64_t M, int64_t ldx, float eps, cudaStream_t st) {
  const int64_t rows_per_block = (int64_t)kLnWarps * R;
  const int64_t blocks = (M + rows_per_block - 1) / rows_per_block;
  TVI_CHECK_ARG(blocks < ((int64_t)1 << 31), "tvi_layernorm_stats_f32: too many rows for one call");
  layernorm_flat_kernel<R, true><<<(unsigned)blocks, kLnWarps * 32, 0, st>>>(x, nullptr, nullptr, nullptr, stats, M, ldx, 0, eps);
  TVI_LAUNCH_OK("layernorm_flat_kernel<stats>");
  return TVI_OK;
}

template <int NCH, int R>
int launch_ln(const tvi_bf16* x, const tvi_bf16* gamma, const tvi_bf16* beta, tvi_bf16* y, int64_t M, int C8, int64_t ldx,
              int64_t ldy, float eps, cudaStream_t st) {
  const int64_t rows_per_block = (int64_t)kLnWarps * R;
  const int64_t blocks = (M + rows_per_block - 1) / rows_per_block;
  TVI_CHECK_ARG(blocks < ((int64_t)1 << 31), "tvi_layernorm_bf16: too many rows for one call");
  layernorm_kernel<NCH, R><<<(unsigned)blocks, kLnWarps * 32, 0, st>>>(x, gamma, beta, y, M, C8, ldx, ldy, eps);
  TVI_LAUNCH_OK("layernorm_kernel");
  return TVI_OK;
}

}  // namespace
}  // namespace tvi

extern "C" int tvi_layernorm_bf16(const tvi_bf16* x, const tvi_bf16* gamma, const tvi_bf16* beta, tvi_bf16* y, int64_t M, int C,
                                  int64_t ldx, int64_t ldy, float eps, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && C >= 0, "tvi_layernorm_bf16: negative size");
  if (M == 0 || C == 0) return TVI_OK;
  TVI_CHECK_ARG(x && y, "tvi_layernorm_bf16: null pointer");
  TVI_CHECK_ARG(C % 8 == 0 && C <= 2048, "tvi_layernorm_bf16: C=%d must be a multiple of 8 and at most 2048", C);
  TVI_CHECK_ARG(ldx % 8 == 0 && ldy % 8 == 0 && ldx >= C && ldy >= C, "tvi_layernorm_bf16: bad row strides");
  TVI_CHECK_ARG(((uintptr_t)x | (uintptr_t)y | (uintptr_t)gamma | (uintptr_t)beta) % 16 == 0,
                "tvi_layernorm_bf16: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(eps > 0.f, "tvi_layernorm_bf16: eps must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  const int C8 = C / 8;
#ifdef TVI_GEMM_TUNE  // tuning builds only: TVI_LN_GENERIC=1 times the generic kernel at the SVD widths
  const bool flat = getenv("TVI_LN_GENERIC") == nullptr;
#else
  const bool flat = true;
#endif
  if (flat) {
    if (C8 == 40) return launch_ln_flat<4>(x, gamma, beta, y, M, ldx, ldy, eps, st);
    if (C8 == 80) return launch_ln_flat<2>(x, gamma, beta, y, M, ldx, ldy, eps, st);
    if (C8 == 160) return launch_ln_flat<1>(x, gamma, beta, y, M, ldx, ldy, eps, st);
  }
  // rows per warp: measured at C = 320, M = 258 048: 1 row 0.106 ms, 2 rows 0.094 ms, 4 rows 0.096 ms
  if (C8 <= 64) return launch_ln<2, 2>(x, gamma, beta, y, M, C8, ldx, ldy, eps, st);
  if (C8 <= 128) return launch_ln<4, 2>(x, gamma, beta, y, M, C8, ldx, ldy, eps, st);
  return launch_ln<8, 1>(x, gamma, beta, y, M, C8, ldx, ldy, eps, st);
}

extern "C" int tvi_layernorm_stats_supported(int C) { return C == 320 || C == 640 || C == 1280; }

extern "C" int tvi_layernorm_stats_f32(const tvi_bf16* x, float* stats, int64_t M, int C, int64_t ldx, float eps, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && C >= 0, "tvi_layernorm_stats_f32: negative size");
  if (M == 0) return TVI_OK;
  TVI_CHECK_ARG(x && stats, "tvi_layernorm_stats_f32: null pointer");
  if (!tvi_layernorm_stats_supported(C))
    return fail(TVI_ERR_UNSUPPORTED, "tvi_layernorm_stats_f32: C=%d (320, 640 and 1280 — the SVD widths — are built)", C);
  TVI_CHECK_ARG(ldx % 8 == 0 && ldx >= C, "tvi_layernorm_stats_f32: bad row stride");
  TVI_CHECK_ARG((uintptr_t)x % 16 == 0 && (uintptr_t)stats % 8 == 0, "tvi_layernorm_stats_f32: x must be 16-byte, stats 8-byte aligned");
  TVI_CHECK_ARG(eps > 0.f, "tvi_layernorm_stats_f32: eps must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  float2* out = reinterpret_cast<float2*>(stats);
  if (C == 320) return launch_ln_stats<4>(x, out, M, ldx, eps, st);
  if (C == 640) return launch_ln_stats<2>(x, out, M, ldx, eps, st);
  return launch_ln_stats<1>(x, out, M, ldx, eps, st);
}

// ---------------------------------------------------------------------------------------------------------------
// LayerNorm backward (training config 5 through the block-level swap): with xhat = (x - mean) * rstd, g = dy * gamma
//   dx     = rstd * (g - mean_c(g) - xhat * mean_c(g * xhat))
//   dgamma = sum_rows dy * xhat,   dbeta = sum_rows dy
// (autograd's native_layer_norm_backward).  One warp per row recomputes the statistics from x (nothing is saved by the
// forward); the row sums of dgamma / dbeta are accumulated per block in shared memory in a fixed order and written as
// fp32 partials [blocks][2][C]; `ln_bwd_final_kernel` adds the partials in order -> deterministic, no atomics.
namespace tvi {
namespace {

constexpr int kLnBwdWarps = 8;
constexpr int kLnBwdRowsPerWarp = 8;

template <int NCH>
__global__ void __launch_bounds__(kLnBwdWarps * 32)
layernorm_bwd_kernel(const tvi_bf16* __restrict__ x, const tvi_bf16* __restrict__ dy, const tvi_bf16* __restrict__ gamma,
                     const tvi_bf16* __restrict__ dres, tvi_bf16* __restrict__ dx, float* __restrict__ partial, int64_t M, int C8,
                     int64_t ldx, int64_t lddy, int64_t lddres, int64_t lddx, float eps) {
  extern __shared__ float s_acc[];  // [warps][2][C]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int C = C8 * 8;
  float dg[NCH][8], db[NCH][8], gm[NCH][8];
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
#pragma unroll
    for (int q = 0; q < 8; ++q) dg[i][q] = db[i][q] = 0.f, gm[i][q] = 1.f;
    if (c < C8 && gamma != nullptr) {
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(gamma) + c);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(h[q]);
        gm[i][2 * q] = t.x;
        gm[i][2 * q + 1] = t.y;
      }
    }
  }
  const float inv_c = 1.0f / (float)C;
  // persistent: a warp walks row groups with a grid stride, so the number of dgamma / dbeta partials is the (small) grid size
  for (int64_t row0 = ((int64_t)blockIdx.x * kLnBwdWarps + warp) * kLnBwdRowsPerWarp; row0 < M;
       row0 += (int64_t)gridDim.x * kLnBwdWarps * kLnBwdRowsPerWarp) {
  // the next row's x / dy are requested before the current row's four dependent warp reductions (latency-bound otherwise)
  uint4 nx[NCH], nd[NCH];
  auto fetch = [&](int64_t row) {
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = lane + 32 * i;
      if (c < C8 && row < M) {
        nx[i] = __ldg(reinterpret_cast<const uint4*>(x + row * ldx) + c);
        nd[i] = __ldg(reinterpret_cast<const uint4*>(dy + row * lddy) + c);
      } else {
        nx[i] = nd[i] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  };
  fetch(row0);
  for (int r = 0; r < kLnBwdRowsPerWarp; ++r) {
    const int64_t row = row0 + r;
    if (row >= M) break;
    float xv[NCH][8], gv[NCH][8];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&nx[i]);
      const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&nd[i]);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 a = __bfloat1622float2(hx[q]), b = __bfloat1622float2(hd[q]);
        xv[i][2 * q] = a.x;
        xv[i][2 * q + 1] = a.y;
        gv[i][2 * q] = b.x;
        gv[i][2 * q + 1] = b.y;
        sum += a.x + a.y;
      }
    }
    if (r + 1 < kLnBwdRowsPerWarp) fetch(row + 1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum * inv_c;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i)
      if (lane + 32 * i < C8) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          xv[i][q] -= mean;
          sq = fmaf(xv[i][q], xv[i][q], sq);
        }
      }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = rsqrtf(sq * inv_c + eps);
    float s1 = 0.f, s2 = 0.f;  // sum g, sum g * xhat
#pragma unroll
    for (int i = 0; i < NCH; ++i)
      if (lane + 32 * i < C8) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float xhat = xv[i][q] * rstd;
          const float dyv = gv[i][q];
          dg[i][q] = fmaf(dyv, xhat, dg[i][q]);
          db[i][q] += dyv;
          const float g = dyv * gm[i][q];
          xv[i][q] = xhat;
          gv[i][q] = g;
          s1 += g;
          s2 = fmaf(g, xhat, s2);
        }
      }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    s1 *= inv_c;
    s2 *= inv_c;
    if (dx != nullptr) {
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        const int c = lane + 32 * i;
        if (c < C8) {
          float o[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) o[q] = rstd * (gv[i][q] - s1 - xv[i][q] * s2);
          if (dres != nullptr) {  // the gradient of the residual branch `x + f(LN(x))`: dx = LN'(dy) + dres in one pass
            const uint4 ur = __ldg(reinterpret_cast<const uint4*>(dres + row * lddres) + c);
            const __nv_bfloat162* hr = reinterpret_cast<const __nv_bfloat162*>(&ur);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float2 t = __bfloat1622float2(hr[q]);
              o[2 * q] += t.x;
              o[2 * q + 1] += t.y;
            }
          }
          *(reinterpret_cast<uint4*>(dx + row * lddx) + c) =
              make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
        }
      }
    }
  }
  }
  // per-block reduction over the warps in a fixed order
  float* mine = s_acc + (size_t)warp * 2 * C;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
    if (c < C8) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        mine[c * 8 + q] = dg[i][q];
        mine[C + c * 8 + q] = db[i][q];
      }
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < 2 * C; j += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < kLnBwdWarps; ++w) s += s_acc[(size_t)w * 2 * C + j];
    partial[(size_t)blockIdx.x * 2 * C + j] = s;
  }
}

// The SVD widths (C = 320 / 640 / 1280 = 40 * G columns, G = 8 / 16 / 32): G lanes share a row, 32 / G rows per warp pass,
// every lane owns the five 16-byte chunks (lane % G) + G * j of its row — all 32 lanes carry data (the kernel above runs
// its second chunk pass with 8 of 32 lanes at C = 320), a warp-level load covers whole 128-byte lines, the four row
// reductions are log2(G) shuffles on ONE value per lane instead of five per row, and the arithmetic is packed (f32x2).
// The kernel above is bound by its instruction stream (~450 per row at C = 320), not by HBM or latency: staging the rows
// through shared memory (1-D bulk copies, then a per-lane cp.async FIFO) only added instructions, 0.155 -> 0.227 / 0.180 ms.
template <int G>
__global__ void __launch_bounds__(kLnBwdWarps * 32, 1)
layernorm_bwd_flat_kernel(const tvi_bf16* __restrict__ x, const tvi_bf16* __restrict__ dy, const tvi_bf16* __restrict__ gamma,
                          const tvi_bf16* __restrict__ dres, tvi_bf16* __restrict__ dx, float* __restrict__ partial, int64_t M,
                          int64_t ldx, int64_t lddy, int64_t lddres, int64_t lddx, float eps) {
  constexpr int R = 32 / G, C8 = 5 * G, C = C8 * 8;
  extern __shared__ float s_acc[];  // [warps][R][2][C]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int rs = lane / G, lg = lane % G;
  auto unpack2 = [](uint32_t w) { return f32x2_pack(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u)); };
  uint64_t dg[5][4], db[5][4];
  uint4 ug[5];  // this lane's gamma chunks (bf16), constant over the rows
#pragma unroll
  for (int j = 0; j < 5; ++j) {
#pragma unroll
    for (int q = 0; q < 4; ++q) dg[j][q] = db[j][q] = f32x2_pack(0.f, 0.f);
    ug[j] = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);  // bf16 1.0 pairs
    if (gamma != nullptr) ug[j] = __ldg(reinterpret_cast<const uint4*>(gamma) + lg + G * j);
  }
  constexpr float inv_c = 1.0f / (float)C;
  for (int64_t row0 = ((int64_t)blockIdx.x * kLnBwdWarps + warp) * R; row0 < M; row0 += (int64_t)gridDim.x * kLnBwdWarps * R) {
    const int64_t row = row0 + rs;
    const bool live = row < M;
    const int64_t rr = live ? row : M - 1;  // tail slots re-read the last row, contribute nothing and store nothing
    uint4 rx[5], rd[5], rres[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const int c = lg + G * j;
      rx[j] = __ldg(reinterpret_cast<const uint4*>(x + rr * ldx) + c);
      rd[j] = __ldg(reinterpret_cast<const uint4*>(dy + rr * lddy) + c);
      rres[j] = make_uint4(0u, 0u, 0u, 0u);
      if (dres != nullptr) rres[j] = __ldg(reinterpret_cast<const uint4*>(dres + rr * lddres) + c);
      if (!live) rd[j] = make_uint4(0u, 0u, 0u, 0u);
    }
    uint64_t xv[5][4];
    uint64_t acc = f32x2_pack(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const uint32_t w[4] = {rx[j].x, rx[j].y, rx[j].z, rx[j].w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        xv[j][q] = unpack2(w[q]);
        acc = f32x2_add(acc, xv[j][q]);
      }
    }
    float a0, a1;
    f32x2_unpack(acc, a0, a1);
    float sum = a0 + a1;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float nmean = -sum * inv_c;
    const uint64_t nm2 = f32x2_pack(nmean, nmean);
    acc = f32x2_pack(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 5; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        xv[j][q] = f32x2_add(xv[j][q], nm2);
        acc = f32x2_fma(xv[j][q], xv[j][q], acc);
      }
    f32x2_unpack(acc, a0, a1);
    float sq = a0 + a1;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const float rstd = rsqrtf(sq * inv_c + eps);
    const uint64_t rs2 = f32x2_pack(rstd, rstd);
    uint64_t s1v = f32x2_pack(0.f, 0.f), s2v = f32x2_pack(0.f, 0.f);  // sum g, sum g * xhat (g = dy * gamma)
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const uint32_t wd[4] = {rd[j].x, rd[j].y, rd[j].z, rd[j].w}, wg[4] = {ug[j].x, ug[j].y, ug[j].z, ug[j].w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint64_t xhat = f32x2_mul(xv[j][q], rs2);
        const uint64_t d2 = unpack2(wd[q]);
        dg[j][q] = f32x2_fma(d2, xhat, dg[j][q]);
        db[j][q] = f32x2_add(db[j][q], d2);
        const uint64_t g2 = f32x2_mul(d2, unpack2(wg[q]));
        s1v = f32x2_add(s1v, g2);
        s2v = f32x2_fma(g2, xhat, s2v);
        xv[j][q] = xhat;
      }
    }
    f32x2_unpack(s1v, a0, a1);
    float s1 = a0 + a1;
    f32x2_unpack(s2v, a0, a1);
    float s2 = a0 + a1;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
      s2 += __shfl_xor_sync(0xffffffffu, s2, o);
    }
    if (dx != nullptr && live) {
      const uint64_t ns1 = f32x2_pack(-s1 * inv_c, -s1 * inv_c), ns2 = f32x2_pack(-s2 * inv_c, -s2 * inv_c);
#pragma unroll
      for (int j = 0; j < 5; ++j) {
        const uint32_t wd[4] = {rd[j].x, rd[j].y, rd[j].z, rd[j].w}, wg[4] = {ug[j].x, ug[j].y, ug[j].z, ug[j].w};
        const uint32_t wr[4] = {rres[j].x, rres[j].y, rres[j].z, rres[j].w};
        uint32_t o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          // dx = rstd * (g - mean(g) - xhat * mean(g xhat)) (+ the residual branch's gradient)
          uint64_t t = f32x2_add(f32x2_mul(unpack2(wd[q]), unpack2(wg[q])), ns1);
          t = f32x2_fma(xv[j][q], ns2, t);
          t = f32x2_fma(t, rs2, unpack2(wr[q]));
          float t0, t1;
          f32x2_unpack(t, t0, t1);
          o[q] = pack_bf16x2(t0, t1);
        }
        *(reinterpret_cast<uint4*>(dx + row * lddx) + lg + G * j) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
  // per-block reduction over (warp, row slot) in a fixed order
  float* mine = s_acc + ((size_t)warp * R + rs) * 2 * C;
#pragma unroll
  for (int j = 0; j < 5; ++j) {
    const int c = lg + G * j;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float v0, v1;
      f32x2_unpack(dg[j][q], v0, v1);
      *reinterpret_cast<float2*>(mine + c * 8 + 2 * q) = make_float2(v0, v1);
      f32x2_unpack(db[j][q], v0, v1);
      *reinterpret_cast<float2*>(mine + C + c * 8 + 2 * q) = make_float2(v0, v1);
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < 2 * C; j += blockDim.x) {
    float s = 0.f;
    for (int w = 0; w < kLnBwdWarps * R; ++w) s += s_acc[(size_t)w * 2 * C + j];
    partial[(size_t)blockIdx.x * 2 * C + j] = s;
  }
}

// one warp per output column: lane l adds blocks l, l + 32, ... in order, then a fixed shuffle tree (deterministic); one
// thread per column walked all 148+ partials serially
__global__ void __launch_bounds__(256)
ln_bwd_final_kernel(const float* __restrict__ partial, tvi_bf16* __restrict__ dgamma, tvi_bf16* __restrict__ dbeta, int nblocks,
                    int C) {
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (j >= 2 * C) return;
  float s = 0.f;
  for (int b = lane; b < nblocks; b += 32) s += partial[(size_t)b * 2 * C + j];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  tvi_bf16* dst = j < C ? dgamma : dbeta;
  if (lane == 0 && dst != nullptr) reinterpret_cast<__nv_bfloat16*>(dst)[j < C ? j : j - C] = __float2bfloat16_rn(s);
}

int64_t ln_bwd_blocks(int64_t M) {
  const int64_t rows_per_block = (int64_t)kLnBwdWarps * kLnBwdRowsPerWarp;
  const int64_t need = (M + rows_per_block - 1) / rows_per_block;
  const int64_t cap = 2 * (int64_t)num_sms();
  return need < cap ? need : cap;
}

template <int NCH>
int launch_ln_bwd(const tvi_bf16* x, const tvi_bf16* dy, const tvi_bf16* gamma, const tvi_bf16* dres, tvi_bf16* dx, float* partial,
                  int64_t M, int C8, int64_t ldx, int64_t lddy, int64_t lddres, int64_t lddx, float eps, cudaStream_t st) {
  const size_t smem = (size_t)kLnBwdWarps * 2 * C8 * 8 * sizeof(float);
  static bool configured[64] = {false};
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(layernorm_bwd_kernel<NCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * 2048 * 4));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  layernorm_bwd_kernel<NCH><<<(unsigned)ln_bwd_blocks(M), kLnBwdWarps * 32, smem, st>>>(x, dy, gamma, dres, dx, partial, M, C8, ldx,
                                                                                         lddy, lddres, lddx, eps);
  TVI_LAUNCH_OK("layernorm_bwd_kernel");
  return TVI_OK;
}

// number of fp32 partial rows tvi_layernorm_bwd_workspace_bytes provides (device-independent)
int64_t ln_bwd_ws_blocks(int64_t M) {
  const int64_t need = (M + kLnBwdWarps * kLnBwdRowsPerWarp - 1) / (kLnBwdWarps * kLnBwdRowsPerWarp);
  return need < 4096 ? need : 4096;
}

template <int G>
int launch_ln_bwd_flat(const tvi_bf16* x, const tvi_bf16* dy, const tvi_bf16* gamma, const tvi_bf16* dres, tvi_bf16* dx,
                       float* partial, int64_t M, int64_t ldx, int64_t lddy, int64_t lddres, int64_t lddx, float eps,
                       cudaStream_t st, int64_t* blocks_out) {
  constexpr int R = 32 / G, C = 40 * G;
  constexpr size_t smem = (size_t)kLnBwdWarps * R * 2 * C * sizeof(float);  // 80 KB at every G
  static bool configured[64] = {false};
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(layernorm_bwd_flat_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  // persistent: one block per SM (the 80 accumulator registers per lane leave room for one), never more partial rows
  // than the workspace holds
  int64_t blocks = (M + kLnBwdWarps * R - 1) / (kLnBwdWarps * R);
  if (blocks > num_sms()) blocks = num_sms();
  if (blocks > ln_bwd_ws_blocks(M)) blocks = ln_bwd_ws_blocks(M);
  layernorm_bwd_flat_kernel<G><<<(unsigned)blocks, kLnBwdWarps * 32, smem, st>>>(x, dy, gamma, dres, dx, partial, M, ldx, lddy,
                                                                                  lddres, lddx, eps);
  TVI_LAUNCH_OK("layernorm_bwd_flat_kernel");
  *blocks_out = blocks;
  return TVI_OK;
}

}  // namespace
}  // namespace tvi

extern "C" size_t tvi_layernorm_bwd_workspace_bytes(int64_t M, int C) {
  if (M <= 0 || C <= 0) return 0;
  // device-independent upper bound of the partial count (the launch uses min(row groups, 2 x SMs) blocks)
  return (size_t)tvi::ln_bwd_ws_blocks(M) * 2 * (size_t)C * sizeof(float);
}

extern "C" int tvi_layernorm_bwd_bf16(const tvi_bf16* x, const tvi_bf16* dy, const tvi_bf16* gamma, tvi_bf16* dx, tvi_bf16* dgamma,
                                      tvi_bf16* dbeta, void* workspace, int64_t M, int C, int64_t ldx, int64_t lddy, int64_t lddx,
                                      float eps, void* stream) {
  return tvi_layernorm_bwd_res_bf16(x, dy, gamma, nullptr, dx, dgamma, dbeta, workspace, M, C, ldx, lddy, 0, lddx, eps, stream);
}

extern "C" int tvi_layernorm_bwd_res_bf16(const tvi_bf16* x, const tvi_bf16* dy, const tvi_bf16* gamma, const tvi_bf16* dres,
                                          tvi_bf16* dx, tvi_bf16* dgamma, tvi_bf16* dbeta, void* workspace, int64_t M, int C,
                                          int64_t ldx, int64_t lddy, int64_t lddres, int64_t lddx, float eps, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(dres == nullptr || (dx != nullptr && lddres % 8 == 0 && lddres >= C && (uintptr_t)dres % 16 == 0),
                "tvi_layernorm_bwd_bf16: dres needs dx, a row stride that is a multiple of 8 and 16-byte alignment");
  TVI_CHECK_ARG(M >= 0 && C >= 0, "tvi_layernorm_bwd_bf16: negative size");
  if (M == 0 || C == 0) return TVI_OK;
  TVI_CHECK_ARG(x && dy && workspace, "tvi_layernorm_bwd_bf16: null pointer");
  TVI_CHECK_ARG(C % 8 == 0 && C <= 2048, "tvi_layernorm_bwd_bf16: C=%d must be a multiple of 8 and at most 2048", C);
  TVI_CHECK_ARG(ldx % 8 == 0 && lddy % 8 == 0 && ldx >= C && lddy >= C && (dx == nullptr || (lddx % 8 == 0 && lddx >= C)),
                "tvi_layernorm_bwd_bf16: bad row strides");
  TVI_CHECK_ARG(((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dx | (uintptr_t)gamma | (uintptr_t)workspace) % 16 == 0,
                "tvi_layernorm_bwd_bf16: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(eps > 0.f, "tvi_layernorm_bwd_bf16: eps must be positive");
  TVI_CHECK_ARG(ln_bwd_blocks(M) < ((int64_t)1 << 31), "tvi_layernorm_bwd_bf16: too many rows for one call");
  cudaStream_t st = (cudaStream_t)stream;
  const int C8 = C / 8;
  int rc;
  int64_t blocks = ln_bwd_blocks(M);
  if (C8 == 40)
    rc = launch_ln_bwd_flat<8>(x, dy, gamma, dres, dx, (float*)workspace, M, ldx, lddy, lddres, lddx, eps, st, &blocks);
  else if (C8 == 80)
    rc = launch_ln_bwd_flat<16>(x, dy, gamma, dres, dx, (float*)workspace, M, ldx, lddy, lddres, lddx, eps, st, &blocks);
  else if (C8 == 160)
    rc = launch_ln_bwd_flat<32>(x, dy, gamma, dres, dx, (float*)workspace, M, ldx, lddy, lddres, lddx, eps, st, &blocks);
  else if (C8 <= 64)
    rc = launch_ln_bwd<2>(x, dy, gamma, dres, dx, (float*)workspace, M, C8, ldx, lddy, lddres, lddx, eps, st);
  else if (C8 <= 128)
    rc = launch_ln_bwd<4>(x, dy, gamma, dres, dx, (float*)workspace, M, C8, ldx, lddy, lddres, lddx, eps, st);
  else if (C8 <= 160)
    rc = launch_ln_bwd<5>(x, dy, gamma, dres, dx, (float*)workspace, M, C8, ldx, lddy, lddres, lddx, eps, st);
  else
    rc = launch_ln_bwd<8>(x, dy, gamma, dres, dx, (float*)workspace, M, C8, ldx, lddy, lddres, lddx, eps, st);
  if (rc) return rc;
  if (dgamma != nullptr || dbeta != nullptr) {
    ln_bwd_final_kernel<<<(2 * C + 7) / 8, 256, 0, st>>>((const float*)workspace, dgamma, dbeta, (int)blocks, C);
    TVI_LAUNCH_OK("ln_bwd_final_kernel");
  }
  return TVI_OK;
}
