// Grouped column sums over token-major bf16 activations (HBM-bound, fp32 accumulation):
//   out[g, c] = sum_{r in group g} x[r, c]
// Used by the backward pass (training config 5): the bias gradient of attn.to_out[0] (one group = all
// tokens; autograd's sum over the batch for F.linear, src/models/attention_processor.py:754 / :1251)
// and the backward of the one-key cross-attention broadcast (every query row of a group receives the
// same projected value row, so its gradient is the column sum of dY over the group).
// Each thread owns one 16-byte column chunk; a warp reads 512 contiguous bytes of a row; row tiles of
// a large group go to fp32 partials that a second kernel adds in a fixed order (deterministic).
#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kCsTileRows = 1024;
constexpr int kCsTy = 8;

__global__ void __launch_bounds__(32 * kCsTy)
colsum_kernel(const tvi_bf16* __restrict__ x, tvi_bf16* __restrict__ out, float* __restrict__ partial,
              int64_t rows_per_group, int tiles_per_group, int C8, int64_t ldx, int64_t ldo, int64_t inner) {
  __shared__ float red[256][8];  // [threadIdx.y][threadIdx.x]: 32 x 8 threads, or 8 x 32 for widths that are not multiples of 256
  const int bx = blockDim.x, by = blockDim.y;
  const int cx = blockIdx.y * bx + threadIdx.x;
  const int ty = threadIdx.y;
  const int64_t group = blockIdx.x / tiles_per_group;
  const int tile = (int)(blockIdx.x - group * tiles_per_group);
  const int64_t rlo = group * rows_per_group + (int64_t)tile * kCsTileRows;
  int64_t rhi = rlo + kCsTileRows;
  const int64_t gend = (group + 1) * rows_per_group;
  if (rhi > gend) rhi = gend;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  // inner > 1: interleaved groups — logical row i of group (b, s) is physical row (b * rows_per_group + i) * inner + s
  const int64_t gbase = group * rows_per_group;
  const int64_t pbase = (group / inner) * rows_per_group * inner + group % inner;
  if (cx < C8) {
    for (int64_t r = rlo + ty; r < rhi; r += by) {
      const int64_t pr = inner == 1 ? r : pbase + (r - gbase) * inner;
      const uint4 u = __ldg(reinterpret_cast<const uint4*>(x + pr * ldx + cx * 8));
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(h[i]);
        acc[2 * i] += t.x;
        acc[2 * i + 1] += t.y;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[ty * bx + threadIdx.x][j] = acc[j];
  __syncthreads();
  if (ty == 0 && cx < C8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = 0.f;
      for (int t = 0; t < by; ++t) s += red[t * bx + threadIdx.x][j];
      acc[j] = s;
    }
    if (tiles_per_group == 1) {
      uint4 v;
      v.x = pack_bf16x2(acc[0], acc[1]);
      v.y = pack_bf16x2(acc[2], acc[3]);
      v.z = pack_bf16x2(acc[4], acc[5]);
      v.w = pack_bf16x2(acc[6], acc[7]);
      *reinterpret_cast<uint4*>(out + group * ldo + cx * 8) = v;
    } else {
      float* dst = partial + ((group * tiles_per_group + tile) * (int64_t)C8 + cx) * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
  }
}

// One warp per (group, column pair): lane l adds tiles l, l + 32, ... in order, then a fixed shuffle tree -> deterministic.
// (One THREAD per column pair walked all `tiles` partials serially: 252 dependent-latency loads for the L0 bias gradient,
// more than the streaming pass itself took.)
__global__ void __launch_bounds__(256)
colsum_final_kernel(const float* __restrict__ partial, tvi_bf16* __restrict__ out, int64_t ngroups, int tiles_per_group, int C,
                    int64_t ldo) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;  // one column pair per warp
  const int lane = threadIdx.x & 31;
  const int half = C / 2;
  if (i >= ngroups * half) return;
  const int64_t g = i / half;
  const int c = (int)(i - g * half) * 2;
  float a = 0.f, b = 0.f;
  for (int t = lane; t < tiles_per_group; t += 32) {
    const float2 v = *reinterpret_cast<const float2*>(partial + (g * tiles_per_group + t) * (int64_t)C + c);
    a += v.x;
    b += v.y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  if (lane == 0) *reinterpret_cast<uint32_t*>(out + g * ldo + c) = pack_bf16x2(a, b);
}

// out[r, :] = v[group(r), :] (+ residual[r, :]): the one-key cross-attention result (every query row of a group
// receives the group's projected value row) with the block's residual add fused (`attn_output + hidden_states`,
// src/models/attention.py:316-317, :459-460).  group(r) = (r / (inner * rows_per_group)) * inner + r % inner: inner == 1
// is the contiguous [group, row] order, inner == S addresses the temporal groups (b, s) from the [B, T, S] token order
// (the transposes at :436-438 / :474-476 folded away).  One 16-byte chunk per thread.
__global__ void __launch_bounds__(256)
bcast_rows_add_kernel(const tvi_bf16* __restrict__ v, const tvi_bf16* __restrict__ residual, tvi_bf16* __restrict__ out,
                      int64_t nrows, int64_t rows_per_group, int64_t inner, int C8, int64_t ldv, int64_t ldr, int64_t ldo) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nrows * C8) return;
  const int64_t r = idx / C8;
  const int c = (int)(idx - r * C8);
  const int64_t group = (r / (inner * rows_per_group)) * inner + r % inner;
  const uint4 a = __ldg(reinterpret_cast<const uint4*>(v + group * ldv) + c);
  if (residual == nullptr) {
    *(reinterpret_cast<uint4*>(out + r * ldo) + c) = a;
    return;
  }
  const uint4 b = __ldg(reinterpret_cast<const uint4*>(residual + r * ldr) + c);
  const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&b);
  uint32_t o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = __bfloat1622float2(ha[i]), y = __bfloat1622float2(hb[i]);
    o[i] = pack_bf16x2(x.x + y.x, x.y + y.y);
  }
  *(reinterpret_cast<uint4*>(out + r * ldo) + c) = make_uint4(o[0], o[1], o[2], o[3]);
}

int cs_tiles(int64_t rows_per_group) { return (int)((rows_per_group + kCsTileRows - 1) / kCsTileRows); }

}  // namespace
}  // namespace tvi

extern "C" size_t tvi_colsum_bf16_workspace_bytes(int64_t nrows, int64_t rows_per_group, int C) {
  if (nrows <= 0 || rows_per_group <= 0 || C <= 0) return 0;
  const int tiles = tvi::cs_tiles(rows_per_group);
  if (tiles == 1) return 0;
  return (size_t)(nrows / rows_per_group) * tiles * (size_t)C * sizeof(float);
}

extern "C" int tvi_colsum_bf16(const tvi_bf16* x, tvi_bf16* out, void* workspace, int64_t nrows, int64_t rows_per_group, int C,
                               int64_t ldx, int64_t ldo, void* stream) {
  return tvi_colsum_bf16_interleaved(x, out, workspace, nrows, rows_per_group, 1, C, ldx, ldo, stream);
}

extern "C" int tvi_colsum_bf16_interleaved(const tvi_bf16* x, tvi_bf16* out, void* workspace, int64_t nrows, int64_t rows_per_group,
                                           int64_t group_inner, int C, int64_t ldx, int64_t ldo, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(nrows >= 0 && C >= 0, "tvi_colsum_bf16: negative size");
  if (nrows == 0 || C == 0) return TVI_OK;
  TVI_CHECK_ARG(x && out, "tvi_colsum_bf16: null pointer");
  TVI_CHECK_ARG(rows_per_group >= 1 && nrows % rows_per_group == 0, "tvi_colsum_bf16: nrows must be a multiple of rows_per_group");
  TVI_CHECK_ARG(C % 8 == 0 && ldx % 8 == 0 && ldo % 8 == 0 && ldx >= C && ldo >= C, "tvi_colsum_bf16: C and the row strides must be multiples of 8");
  TVI_CHECK_ARG(((uintptr_t)x | (uintptr_t)out) % 16 == 0, "tvi_colsum_bf16: pointers must be 16-byte aligned");
  const int64_t ngroups = nrows / rows_per_group;
  TVI_CHECK_ARG(group_inner >= 1 && ngroups % group_inner == 0, "tvi_colsum_bf16: the group count must be a multiple of group_inner");
  const int tiles = cs_tiles(rows_per_group);
  TVI_CHECK_ARG(tiles == 1 || (workspace != nullptr && (uintptr_t)workspace % 16 == 0), "tvi_colsum_bf16: workspace required (16-byte aligned)");
  TVI_CHECK_ARG(ngroups * tiles < ((int64_t)1 << 31), "tvi_colsum_bf16: grid too large");
  cudaStream_t st = (cudaStream_t)stream;
  const int C8 = C / 8;
  // (A variant that deals the 256 threads of a block out as row slots x column chunks, so that C = 320 keeps every lane
  // busy, was measured slower — 0.097 vs 0.067 ms at [258 048, 320]: 252 blocks of one tile each leave the SMs under-filled.)
  // 32 column chunks x 8 row lanes per block; widths like C = 320 (40 chunks) would leave the second block of a row with
  // 8 of 32 lanes busy, so they take 8 chunks x 32 row lanes (a warp = four 128-byte row segments)
  // (only for long groups: 32 row lanes need rows to work on — the one-key dV sums have 14-row groups)
  const int bx = (C8 % 32 != 0 && C8 % 8 == 0 && rows_per_group >= 256) ? 8 : 32;
  dim3 grid((unsigned)(ngroups * tiles), (unsigned)((C8 + bx - 1) / bx));
  dim3 block((unsigned)bx, (unsigned)(256 / bx));
  colsum_kernel<<<grid, block, 0, st>>>(x, out, (float*)workspace, rows_per_group, tiles, C8, ldx, ldo, group_inner);
  TVI_LAUNCH_OK("colsum_kernel");
  if (tiles > 1) {
    const int64_t pairs = ngroups * (C / 2);  // one warp each
    colsum_final_kernel<<<(unsigned)((pairs + 7) / 8), 256, 0, st>>>((const float*)workspace, out, ngroups, tiles, C, ldo);
    TVI_LAUNCH_OK("colsum_final_kernel");
  }
  return TVI_OK;
}

extern "C" int tvi_bcast_rows_add_bf16(const tvi_bf16* v, const tvi_bf16* residual, tvi_bf16* out, int64_t nrows,
                                       int64_t rows_per_group, int64_t group_inner, int C, int64_t ldv, int64_t ldr, int64_t ldo,
                                       void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(nrows >= 0 && C >= 0, "tvi_bcast_rows_add_bf16: negative size");
  if (nrows == 0 || C == 0) return TVI_OK;
  TVI_CHECK_ARG(v && out, "tvi_bcast_rows_add_bf16: null pointer");
  TVI_CHECK_ARG(rows_per_group >= 1 && group_inner >= 1 && nrows % (rows_per_group * group_inner) == 0,
                "tvi_bcast_rows_add_bf16: nrows must be a multiple of rows_per_group * group_inner");
  TVI_CHECK_ARG(C % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0 && ldv >= C && ldo >= C && (residual == nullptr || (ldr % 8 == 0 && ldr >= C)),
                "tvi_bcast_rows_add_bf16: C and the row strides must be multiples of 8");
  TVI_CHECK_ARG(((uintptr_t)v | (uintptr_t)residual | (uintptr_t)out) % 16 == 0, "tvi_bcast_rows_add_bf16: pointers must be 16-byte aligned");
  const int64_t total = nrows * (C / 8);
  TVI_CHECK_ARG((total + 255) / 256 < ((int64_t)1 << 31), "tvi_bcast_rows_add_bf16: grid too large");
  bcast_rows_add_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(v, residual, out, nrows, rows_per_group,
                                                                                           group_inner, C / 8, ldv, ldr, ldo);
  TVI_LAUNCH_OK("bcast_rows_add_kernel");
  return TVI_OK;
}
