// LayerNorm over the channel dimension of token-major bf16 activations (N1 of SURVEY.md 8.6).
// Replaces nn.LayerNorm norm1 / norm2 / norm3 / norm_in of BasicTransformerBlock and
// TemporalBasicTransformerBlock (src/models/attention.py:270, :300, :321, :441, :446, :462): per token
//   y = (x - mean) / sqrt(var + eps) * gamma + beta        (biased variance, statistics in fp32).
// HBM-bound: 2 * M * C * 2 bytes.  One warp per row; the row stays in registers between the statistics and the
// normalisation (C <= 2048: up to 8 chunks of 8 channels per lane), 128-bit coalesced loads and stores,
// two-pass variance (sum of squared deviations) so that large means do not cancel.
#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kLnWarps = 8;

template <int NCH>  // 16-byte chunks per lane
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_kernel(const tvi_bf16* __restrict__ x, const tvi_bf16* __restrict__ gamma, const tvi_bf16* __restrict__ beta,
                 tvi_bf16* __restrict__ y, int64_t M, int C8, int64_t ldx, int64_t ldy, float eps) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * kLnWarps + (threadIdx.x >> 5);
  if (row >= M) return;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * ldx);
  float v[NCH][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
    if (c < C8) {
      const uint4 u = __ldg(xr + c);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 t = __bfloat1622float2(h[q]);
        v[i][2 * q] = t.x;
        v[i][2 * q + 1] = t.y;
        sum += t.x + t.y;
      }
    } else {
#pragma unroll
      for (int q = 0; q < 8; ++q) v[i][q] = 0.f;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float inv_c = 1.0f / (float)(C8 * 8);
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
  uint4* yr = reinterpret_cast<uint4*>(y + row * ldy);
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    const int c = lane + 32 * i;
    if (c < C8) {
      float g[8], b[8];
      if (gamma != nullptr) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(gamma) + c);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 t = __bfloat1622float2(h[q]);
          g[2 * q] = t.x;
          g[2 * q + 1] = t.y;
        }
      } else {
#pragma unroll
        for (int q = 0; q < 8; ++q) g[q] = 1.f;
      }
      if (beta != nullptr) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(beta) + c);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float2 t = __bfloat1622float2(h[q]);
          b[2 * q] = t.x;
          b[2 * q + 1] = t.y;
        }
      } else {
#pragma unroll
        for (int q = 0; q < 8; ++q) b[q] = 0.f;
      }
      uint4 o;
      o.x = pack_bf16x2(fmaf((v[i][0] - mean) * rstd, g[0], b[0]), fmaf((v[i][1] - mean) * rstd, g[1], b[1]));
      o.y = pack_bf16x2(fmaf((v[i][2] - mean) * rstd, g[2], b[2]), fmaf((v[i][3] - mean) * rstd, g[3], b[3]));
      o.z = pack_bf16x2(fmaf((v[i][4] - mean) * rstd, g[4], b[4]), fmaf((v[i][5] - mean) * rstd, g[5], b[5]));
      o.w = pack_bf16x2(fmaf((v[i][6] - mean) * rstd, g[6], b[6]), fmaf((v[i][7] - mean) * rstd, g[7], b[7]));
      yr[c] = o;
    }
  }
}

template <int NCH>
int launch_ln(const tvi_bf16* x, const tvi_bf16* gamma, const tvi_bf16* beta, tvi_bf16* y, int64_t M, int C8, int64_t ldx,
              int64_t ldy, float eps, cudaStream_t st) {
  const int64_t blocks = (M + kLnWarps - 1) / kLnWarps;
  TVI_CHECK_ARG(blocks < ((int64_t)1 << 31), "tvi_layernorm_bf16: too many rows for one call");
  layernorm_kernel<NCH><<<(unsigned)blocks, kLnWarps * 32, 0, st>>>(x, gamma, beta, y, M, C8, ldx, ldy, eps);
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
  if (C8 <= 64) return launch_ln<2>(x, gamma, beta, y, M, C8, ldx, ldy, eps, st);
  if (C8 <= 128) return launch_ln<4>(x, gamma, beta, y, M, C8, ldx, ldy, eps, st);
  return launch_ln<8>(x, gamma, beta, y, M, C8, ldx, ldy, eps, st);
}
