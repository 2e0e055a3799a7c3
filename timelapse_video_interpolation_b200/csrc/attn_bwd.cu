// Backward of the three attention cores (training config 5: src/train_decouples.py back-propagates
// through F.scaled_dot_product_attention / xformers / the eager baddbmm-softmax-bmm of
// src/models/attention_processor.py:598-615, :750, :1243 at the call sites of rows A3a-A3d).
//
// With P = softmax(scale . Q K^T), O = P V and D_i = sum_d dO_id O_id = sum_j P_ij dP_ij:
//     dV = P^T dO        dP = dO V^T        dS = P o (dP - D)        dQ = scale . dS K      dK = scale . dS^T Q
//
// tvi_attn_spatial_bwd  (long sequences, flash-style recomputation from Q, K, V, LSE; nothing S x Skv is
//   ever stored).  Two passes of ONE mma.sync kernel template, so no atomics and a fixed summation order:
//     rows = keys   : a CTA owns 128 key rows, streams the queries in blocks of 64, computes S^T = K Q^T
//                     and dP^T = V dO^T, accumulates dV += P^T dO and dK += dS^T Q in registers;
//     rows = queries: a CTA owns 128 query rows, streams the keys, accumulates dQ += dS K.
//   P^T / dS^T (resp. dS) go from the accumulator registers straight into the next product's A operand.
// tvi_attn_temporal_bwd (T <= 64): one warp per (sequence, head); P and dS of the whole T x T problem
//   are staged in shared memory as bf16 and re-read transposed with ldmatrix.trans for dV and dK.
//   Needs neither O nor LSE (D is recomputed from P and dP).  HBM bytes: 7 * rows * heads*d * 2.
// tvi_attn_fewkeys_bwd  (1..8 context keys shared by rows_per_group query rows): one 16-byte channel
//   chunk per thread; dK / dV of a group are accumulated in registers, reduced across the CTA's rows in
//   shared memory in a fixed order, and across row tiles through fp32 partials.
#include "mma_sync.cuh"

namespace tvi {
namespace {

constexpr float kLog2e = 1.4426950408889634f;

// ---------------------------------------------------------------------------------------------
// D[row, head] = sum_d dO . O
// ---------------------------------------------------------------------------------------------
template <int DH>
__global__ void attn_bwd_dot_kernel(const tvi_bf16* __restrict__ o, const tvi_bf16* __restrict__ dout, float* __restrict__ D,
                                    int64_t nrows, int heads, int64_t ldo, int64_t lddo) {
  // 8 * DH lanes per (row, head): one 16-byte chunk each
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t item = idx / (8 * DH);
  const int c = (int)(idx % (8 * DH));
  const bool in = item < nrows * heads;
  const int64_t row = in ? item / heads : 0;
  const int h = in ? (int)(item - row * heads) : 0;
  const uint4 a = __ldg(reinterpret_cast<const uint4*>(o + row * ldo + h * 64 * DH + c * 8));
  const uint4 b = __ldg(reinterpret_cast<const uint4*>(dout + row * lddo + h * 64 * DH + c * 8));
  const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&b);
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = __bfloat1622float2(ha[i]), y = __bfloat1622float2(hb[i]);
    s = fmaf(x.x, y.x, s);
    s = fmaf(x.y, y.y, s);
  }
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  if (DH == 2) s += __shfl_xor_sync(0xffffffffu, s, 8);
  if (in && c == 0) D[item] = s;
}

// ---------------------------------------------------------------------------------------------
// flash backward (spatial)
// ---------------------------------------------------------------------------------------------
constexpr int kFbBr = 128;  // rows per CTA
constexpr int kFbBc = 64;   // streamed columns per stage
constexpr int kFbStages = 3;
constexpr int kFbThreads = 256;
constexpr int kFbRowHalf = kFbBr * 128;  // one 64-column half of a row tile
constexpr int kFbColHalf = kFbBc * 128;  // one 64-column half of a streamed tile
// DH = d / 64 swizzled halves per tile (d = 128: the class default heads of unet_spatio_temporal_condition.py:94)
constexpr int fb_stage_bytes(int DH) { return 2 * DH * kFbColHalf + 2 * kFbBc * 4; }  // two tiles + lse[64] + D[64]
constexpr int fb_smem(int DH) { return 2 * DH * kFbRowHalf + kFbStages * fb_stage_bytes(DH); }

struct FbParams {
  const tvi_bf16 *r1, *r2;  // row operands      rows = keys: K, V     rows = queries: Q, dO
  const tvi_bf16 *c1, *c2;  // streamed operands rows = keys: Q, dO    rows = queries: K, V
  int64_t ldr1, ldr2, ldc1, ldc2;
  const float *lse, *D;     // [batch * Sq, heads]
  tvi_bf16 *out1, *out2;    // rows = keys: dK, dV     rows = queries: dQ, -
  int64_t ldo1, ldo2;
  int heads, nrows, ncols;  // per sequence
  float scale;
};

template <bool kRowsAreKeys, int DH>
__global__ void __launch_bounds__(kFbThreads, 1) attn_bwd_flash_kernel(const FbParams p) {
  constexpr int kFbRowBytes = DH * kFbRowHalf, kFbColBytes = DH * kFbColHalf, kFbStageBytes = fb_stage_bytes(DH);
  constexpr int ND = 8 * DH;  // 8-column tiles across d
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, qd = lane & 3;
  const int bh = blockIdx.y;
  const int b = bh / p.heads, h = bh - b * p.heads;
  const int r0 = blockIdx.x * kFbBr;

  const uint32_t sR1 = smem_u32(smem_raw);
  const uint32_t sR2 = sR1 + kFbRowBytes;
  const uint32_t sSt = sR2 + kFbRowBytes;
  const tvi_bf16* r1b = p.r1 + ((int64_t)b * p.nrows) * p.ldr1 + h * 64 * DH;
  const tvi_bf16* r2b = p.r2 + ((int64_t)b * p.nrows) * p.ldr2 + h * 64 * DH;
  const tvi_bf16* c1b = p.c1 + ((int64_t)b * p.ncols) * p.ldc1 + h * 64 * DH;
  const tvi_bf16* c2b = p.c2 + ((int64_t)b * p.ncols) * p.ldc2 + h * 64 * DH;
  // lse / D are indexed by the query: a column when rows are keys, a row otherwise
  const int nq = kRowsAreKeys ? p.ncols : p.nrows;
  const float* lse_b = p.lse + ((int64_t)b * nq) * p.heads + h;
  const float* D_b = p.D + ((int64_t)b * nq) * p.heads + h;

  auto load_stage = [&](int st, int j0) {
    const uint32_t dst = sSt + st * kFbStageBytes;
#pragma unroll
    for (int i = 0; i < (kFbBc * ND) / kFbThreads; ++i) {
      const int idx = tid + i * kFbThreads;
      const int r = idx / ND, c = idx % ND;
      const bool in = (j0 + r) < p.ncols;
      const uint32_t off = (c >> 3) * kFbColHalf + tile_off(r, c & 7);
      const int64_t row = in ? (j0 + r) : 0;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + off), "l"(c1b + row * p.ldc1 + c * 8),
                   "r"(in ? 16 : 0)
                   : "memory");
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + kFbColBytes + off),
                   "l"(c2b + row * p.ldc2 + c * 8), "r"(in ? 16 : 0)
                   : "memory");
    }
    if (kRowsAreKeys && tid < 2 * kFbBc) {
      const int r = tid & (kFbBc - 1);
      const bool in = (j0 + r) < p.ncols;
      const float* src = (tid < kFbBc ? lse_b : D_b) + (int64_t)(in ? j0 + r : 0) * p.heads;
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst + 2 * kFbColBytes + tid * 4), "l"(src),
                   "r"(in ? 4 : 0)
                   : "memory");
    }
  };

  // ---- prologue: the two row tiles + first two column stages -------------------------------
#pragma unroll
  for (int i = 0; i < (kFbBr * ND) / kFbThreads; ++i) {
    const int idx = tid + i * kFbThreads;
    const int r = idx / ND, c = idx % ND;
    const bool in = (r0 + r) < p.nrows;
    const int64_t row = in ? (r0 + r) : 0;
    const uint32_t roff = (c >> 3) * kFbRowHalf + tile_off(r, c & 7);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sR1 + roff), "l"(r1b + row * p.ldr1 + c * 8),
                 "r"(in ? 16 : 0)
                 : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sR2 + roff), "l"(r2b + row * p.ldr2 + c * 8),
                 "r"(in ? 16 : 0)
                 : "memory");
  }
  const int nblk = (p.ncols + kFbBc - 1) / kFbBc;
  load_stage(0, 0);
  cp_async_commit();
  if (nblk > 1) load_stage(1, kFbBc);
  cp_async_commit();

  const float sl2 = p.scale * kLog2e;
  // rows = queries: this thread's two rows' lse / D stay in registers
  float lr0 = 0.f, lr1 = 0.f, dr0 = 0.f, dr1 = 0.f;
  if (!kRowsAreKeys) {
    const int q0 = r0 + warp * 16 + g, q1 = q0 + 8;
    if (q0 < p.nrows) {
      lr0 = __ldg(lse_b + (int64_t)q0 * p.heads) * kLog2e;
      dr0 = __ldg(D_b + (int64_t)q0 * p.heads);
    }
    if (q1 < p.nrows) {
      lr1 = __ldg(lse_b + (int64_t)q1 * p.heads) * kLog2e;
      dr1 = __ldg(D_b + (int64_t)q1 * p.heads);
    }
  }

  float acc1[ND][4];                         // dS . C1  (dK or dQ)
  float acc2[kRowsAreKeys ? ND : 1][4];      // P . C2   (dV)
#pragma unroll
  for (int nt = 0; nt < ND; ++nt) acc1[nt][0] = acc1[nt][1] = acc1[nt][2] = acc1[nt][3] = 0.f;
#pragma unroll
  for (int nt = 0; nt < (kRowsAreKeys ? ND : 1); ++nt) acc2[nt][0] = acc2[nt][1] = acc2[nt][2] = acc2[nt][3] = 0.f;

  const int arow = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;  // ldmatrix row of this lane in the row tiles

  for (int j = 0; j < nblk; ++j) {
    cp_async_wait<1>();
    __syncthreads();
    if (j + 2 < nblk) load_stage((j + 2) % kFbStages, (j + 2) * kFbBc);
    cp_async_commit();
    const uint32_t sC1 = sSt + (j % kFbStages) * kFbStageBytes;
    const uint32_t sC2 = sC1 + kFbColBytes;
    const float* sLse = reinterpret_cast<const float*>(smem_raw + 2 * kFbRowBytes + (j % kFbStages) * kFbStageBytes +
                                                       2 * kFbColBytes);
    const float* sD = sLse + kFbBc;

    // ---- S = R1 C1^T and dP = R2 C2^T (16 x 64 per warp) -------------------------------------
    float s[8][4], dp[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      dp[nt][0] = dp[nt][1] = dp[nt][2] = dp[nt][3] = 0.f;
    }
#pragma unroll
    for (int ks = 0; ks < 4 * DH; ++ks) {
      uint32_t a1[4], a2[4];
      const uint32_t aoff = (ks >> 2) * kFbRowHalf + tile_off(arow, 2 * (ks & 3) + (lane >> 4));
      ldsm_x4(sR1 + aoff, a1[0], a1[1], a1[2], a1[3]);
      ldsm_x4(sR2 + aoff, a2[0], a2[1], a2[2], a2[3]);
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        uint32_t b0, b1, b2, b3;
        const uint32_t boff = (ks >> 2) * kFbColHalf + tile_off(np * 16 + (lane & 7) + (lane >> 4) * 8, 2 * (ks & 3) + ((lane >> 3) & 1));
        ldsm_x4(sC1 + boff, b0, b1, b2, b3);
        mma_bf16_16816(s[2 * np], a1, b0, b1);
        mma_bf16_16816(s[2 * np + 1], a1, b2, b3);
        ldsm_x4(sC2 + boff, b0, b1, b2, b3);
        mma_bf16_16816(dp[2 * np], a2, b0, b1);
        mma_bf16_16816(dp[2 * np + 1], a2, b2, b3);
      }
    }
    // ---- P = exp(scale S - lse), dS = P o (dP - D) -------------------------------------------
    uint32_t pfr[8][2], dfr[8][2];
    const int j0 = j * kFbBc;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int col = nt * 8 + qd * 2;
      float l0a, l0b, l1a, l1b, d0a, d0b, d1a, d1b;  // (row g | g+8) x (col | col+1)
      if (kRowsAreKeys) {
        const float2 lv = *reinterpret_cast<const float2*>(sLse + col);
        const float2 dv = *reinterpret_cast<const float2*>(sD + col);
        l0a = l1a = lv.x * kLog2e;
        l0b = l1b = lv.y * kLog2e;
        d0a = d1a = dv.x;
        d0b = d1b = dv.y;
      } else {
        l0a = l0b = lr0;
        l1a = l1b = lr1;
        d0a = d0b = dr0;
        d1a = d1b = dr1;
      }
      float p0 = fast_exp2(fmaf(s[nt][0], sl2, -l0a)), p1 = fast_exp2(fmaf(s[nt][1], sl2, -l0b));
      float p2 = fast_exp2(fmaf(s[nt][2], sl2, -l1a)), p3 = fast_exp2(fmaf(s[nt][3], sl2, -l1b));
      if (j0 + col >= p.ncols) p0 = p2 = 0.f;  // streamed columns past the sequence
      if (j0 + col + 1 >= p.ncols) p1 = p3 = 0.f;
      pfr[nt][0] = pack_bf16x2(p0, p1);
      pfr[nt][1] = pack_bf16x2(p2, p3);
      dfr[nt][0] = pack_bf16x2(p0 * (dp[nt][0] - d0a), p1 * (dp[nt][1] - d0b));
      dfr[nt][1] = pack_bf16x2(p2 * (dp[nt][2] - d1a), p3 * (dp[nt][3] - d1b));
    }
    // ---- acc1 += dS C1, acc2 += P C2 -----------------------------------------------------------
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const uint32_t ad[4] = {dfr[2 * kk][0], dfr[2 * kk][1], dfr[2 * kk + 1][0], dfr[2 * kk + 1][1]};
      const uint32_t ap[4] = {pfr[2 * kk][0], pfr[2 * kk][1], pfr[2 * kk + 1][0], pfr[2 * kk + 1][1]};
#pragma unroll
      for (int nc = 0; nc < ND; nc += 2) {
        uint32_t b0, b1, b2, b3;
        const uint32_t boff = (nc >> 3) * kFbColHalf + tile_off(kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, (nc & 7) + (lane >> 4));
        ldsm_x4_t(sC1 + boff, b0, b1, b2, b3);
        mma_bf16_16816(acc1[nc], ad, b0, b1);
        mma_bf16_16816(acc1[nc + 1], ad, b2, b3);
        if (kRowsAreKeys) {
          ldsm_x4_t(sC2 + boff, b0, b1, b2, b3);
          mma_bf16_16816(acc2[nc], ap, b0, b1);
          mma_bf16_16816(acc2[nc + 1], ap, b2, b3);
        }
      }
    }
  }
  cp_async_wait<0>();
  __syncwarp();

  // ---- epilogue: each warp stages its 16 rows over its own rows of the row tiles ----------------
  uint8_t* pR1 = smem_raw;
  uint8_t* pR2 = smem_raw + kFbRowBytes;
  const int ra = warp * 16 + g, rb = ra + 8;
#pragma unroll
  for (int nt = 0; nt < ND; ++nt) {
    const int hoff = (nt >> 3) * kFbRowHalf;
    *reinterpret_cast<uint32_t*>(pR1 + hoff + tile_off(ra, nt & 7) + qd * 4) = pack_bf16x2(acc1[nt][0] * p.scale, acc1[nt][1] * p.scale);
    *reinterpret_cast<uint32_t*>(pR1 + hoff + tile_off(rb, nt & 7) + qd * 4) = pack_bf16x2(acc1[nt][2] * p.scale, acc1[nt][3] * p.scale);
    if (kRowsAreKeys) {
      *reinterpret_cast<uint32_t*>(pR2 + hoff + tile_off(ra, nt & 7) + qd * 4) = pack_bf16x2(acc2[nt][0], acc2[nt][1]);
      *reinterpret_cast<uint32_t*>(pR2 + hoff + tile_off(rb, nt & 7) + qd * 4) = pack_bf16x2(acc2[nt][2], acc2[nt][3]);
    }
  }
  __syncwarp();
  tvi_bf16* o1b = p.out1 + ((int64_t)b * p.nrows) * p.ldo1 + h * 64 * DH;
  tvi_bf16* o2b = kRowsAreKeys ? p.out2 + ((int64_t)b * p.nrows) * p.ldo2 + h * 64 * DH : nullptr;
#pragma unroll
  for (int i = 0; i < 4 * DH; ++i) {
    const int idx = lane + 32 * i;
    const int r = warp * 16 + idx / ND, c = idx % ND;
    if (r0 + r < p.nrows) {
      const int soff = (c >> 3) * kFbRowHalf + tile_off(r, c & 7);
      *reinterpret_cast<uint4*>(o1b + (int64_t)(r0 + r) * p.ldo1 + c * 8) = *reinterpret_cast<const uint4*>(pR1 + soff);
      if (kRowsAreKeys)
        *reinterpret_cast<uint4*>(o2b + (int64_t)(r0 + r) * p.ldo2 + c * 8) = *reinterpret_cast<const uint4*>(pR2 + soff);
    }
  }
}

template <bool kRowsAreKeys, int DH>
int launch_flash_bwd(const FbParams& p, int batch, cudaStream_t st) {
  static bool configured[64] = {false};
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(attn_bwd_flash_kernel<kRowsAreKeys, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, fb_smem(DH)));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  dim3 grid((unsigned)((p.nrows + kFbBr - 1) / kFbBr), (unsigned)(batch * p.heads));
  attn_bwd_flash_kernel<kRowsAreKeys, DH><<<grid, kFbThreads, fb_smem(DH), st>>>(p);
  TVI_LAUNCH_OK("attn_bwd_flash_kernel");
  return TVI_OK;
}

// ---------------------------------------------------------------------------------------------
// temporal (short-sequence) backward
// ---------------------------------------------------------------------------------------------
// DH = d / 64: the d-wide tiles (Q, K, V, dO, the output staging) are DH swizzled 64-column halves; P and dS are
// [queries][keys <= 64] and do not depend on d
template <int MT, int DH>
struct TemporalBwdCfg {
  static constexpr int kRows = 16 * MT;
  static constexpr int kHalfBytes = kRows * 128;
  static constexpr int kTileBytes = DH * kHalfBytes;
  static constexpr int kStageBytes = DH * 16 * 128;
  static constexpr int kWarpBytes = 4 * kTileBytes + 2 * kHalfBytes + kStageBytes;  // Q K V dO | P dS | 16-row output staging
  static constexpr int kWarps = (MT == 1) ? 8 : (MT == 4 && DH == 2) ? 2 : 4;
  static constexpr int kSmemBytes = kWarps * kWarpBytes;
};

struct TbParams {
  const tvi_bf16 *q, *k, *v, *dout;
  tvi_bf16 *dq, *dk, *dv;
  int64_t ldq, ldk, ldv, lddo, lddq, lddk, lddv;
  int64_t nitems;
  int heads, Tq, Tk;
  float scale;
  int64_t inner;  // > 1: interleaved sequences — token t of sequence s is row (s / inner) * inner * T + t * inner + s % inner
};

template <int MT, int DH>
__global__ void __launch_bounds__(TemporalBwdCfg<MT, DH>::kWarps * 32) attn_temporal_bwd_kernel(const TbParams p) {
  using Cfg = TemporalBwdCfg<MT, DH>;
  constexpr int kRows = Cfg::kRows;
  constexpr int kHalf = Cfg::kHalfBytes;
  constexpr int ND = 8 * DH;  // 8-column tiles across d
  constexpr int NT = 2 * MT;
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, qd = lane & 3;
  uint8_t* wptr = smem_raw + warp * Cfg::kWarpBytes;
  const uint32_t sQ = smem_u32(wptr);
  const uint32_t sK = sQ + Cfg::kTileBytes;
  const uint32_t sV = sK + Cfg::kTileBytes;
  const uint32_t sDO = sV + Cfg::kTileBytes;
  const uint32_t sP = sDO + Cfg::kTileBytes;
  const uint32_t sDS = sP + kHalf;
  uint8_t* pP = wptr + 4 * Cfg::kTileBytes;
  uint8_t* pDS = pP + kHalf;
  uint8_t* pStage = pDS + kHalf;
  const float sl2 = p.scale * kLog2e;

  // 16 staged rows -> coalesced 128-bit stores (rows past `limit` are dropped)
  auto flush_rows = [&](tvi_bf16* dst, int64_t ld, int row0, int limit) {
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 4 * DH; ++i) {
      const int idx = lane + 32 * i;
      const int r = idx / ND, c = idx % ND;
      if (row0 + r < limit)
        *reinterpret_cast<uint4*>(dst + (int64_t)(row0 + r) * ld + c * 8) =
            *reinterpret_cast<const uint4*>(pStage + (c >> 3) * 2048 + tile_off(r, c & 7));
    }
    __syncwarp();
  };
  auto stage_acc = [&](const float (*acc)[4], float mul) {
#pragma unroll
    for (int nt = 0; nt < ND; ++nt) {
      uint8_t* ph = pStage + (nt >> 3) * 2048;
      *reinterpret_cast<uint32_t*>(ph + tile_off(g, nt & 7) + qd * 4) = pack_bf16x2(acc[nt][0] * mul, acc[nt][1] * mul);
      *reinterpret_cast<uint32_t*>(ph + tile_off(g + 8, nt & 7) + qd * 4) = pack_bf16x2(acc[nt][2] * mul, acc[nt][3] * mul);
    }
  };

  for (int64_t item = (int64_t)blockIdx.x * Cfg::kWarps + warp; item < p.nitems; item += (int64_t)gridDim.x * Cfg::kWarps) {
    const int64_t seq = item / p.heads;
    const int head = (int)(item - seq * p.heads);
    // first row of the sequence and the row stride between its tokens (as the forward kernel, attn_short.cu)
    const int64_t so = seq / p.inner, si = seq - so * p.inner;
    const int64_t q0 = so * p.inner * p.Tq + si, k0 = so * p.inner * p.Tk + si;
    const int64_t ldq = p.ldq * p.inner, ldk = p.ldk * p.inner, ldv = p.ldv * p.inner, lddo = p.lddo * p.inner;
    const tvi_bf16* qb = p.q + q0 * p.ldq + head * 64 * DH;
    const tvi_bf16* kb = p.k + k0 * p.ldk + head * 64 * DH;
    const tvi_bf16* vb = p.v + k0 * p.ldv + head * 64 * DH;
    const tvi_bf16* dob = p.dout + q0 * p.lddo + head * 64 * DH;
    {
#pragma unroll
      for (int i = 0; i < kRows * DH / 4; ++i) {
        const int idx = lane + 32 * i;
        const int r = idx / ND, c = idx % ND;
        const uint32_t off = (c >> 3) * kHalf + tile_off(r, c & 7);
        const bool pq = r < p.Tq, pk = r < p.Tk;
        cp_async_16(wptr + off, pq ? qb + r * ldq + c * 8 : qb, pq);
        cp_async_16(wptr + Cfg::kTileBytes + off, pk ? kb + r * ldk + c * 8 : kb, pk);
        cp_async_16(wptr + 2 * Cfg::kTileBytes + off, pk ? vb + r * ldv + c * 8 : vb, pk);
        cp_async_16(wptr + 3 * Cfg::kTileBytes + off, pq ? dob + r * lddo + c * 8 : dob, pq);
      }
      cp_async_commit();
      cp_async_wait<0>();
      __syncwarp();
    }

    // ---- phase 1: per query tile: P, dP, dS (staged as bf16) and dQ ------------------------------
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      if (mt * 16 >= p.Tq) break;
      float s[NT][4], dp[NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        dp[nt][0] = dp[nt][1] = dp[nt][2] = dp[nt][3] = 0.f;
      }
#pragma unroll
      for (int ks = 0; ks < 4 * DH; ++ks) {
        uint32_t a1[4], a2[4];
        const uint32_t aoff = (ks >> 2) * kHalf + tile_off(mt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, 2 * (ks & 3) + (lane >> 4));
        ldsm_x4(sQ + aoff, a1[0], a1[1], a1[2], a1[3]);
        ldsm_x4(sDO + aoff, a2[0], a2[1], a2[2], a2[3]);
#pragma unroll
        for (int np = 0; np < MT; ++np) {
          uint32_t b0, b1, b2, b3;
          const uint32_t boff = (ks >> 2) * kHalf + tile_off(np * 16 + (lane & 7) + (lane >> 4) * 8, 2 * (ks & 3) + ((lane >> 3) & 1));
          ldsm_x4(sK + boff, b0, b1, b2, b3);
          mma_bf16_16816(s[2 * np], a1, b0, b1);
          mma_bf16_16816(s[2 * np + 1], a1, b2, b3);
          ldsm_x4(sV + boff, b0, b1, b2, b3);
          mma_bf16_16816(dp[2 * np], a2, b0, b1);
          mma_bf16_16816(dp[2 * np + 1], a2, b2, b3);
        }
      }
      float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int col = nt * 8 + qd * 2;
        if (col >= p.Tk) s[nt][0] = s[nt][2] = -INFINITY;
        if (col + 1 >= p.Tk) s[nt][1] = s[nt][3] = -INFINITY;
        m0 = fmaxf(m0, fmaxf(s[nt][0], s[nt][1]));
        m1 = fmaxf(m1, fmaxf(s[nt][2], s[nt][3]));
      }
      m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
      m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
      m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
      m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
      const float ms0 = m0 * sl2, ms1 = m1 * sl2;
      float l0 = 0.f, l1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        s[nt][0] = fast_exp2(fmaf(s[nt][0], sl2, -ms0));
        s[nt][1] = fast_exp2(fmaf(s[nt][1], sl2, -ms0));
        s[nt][2] = fast_exp2(fmaf(s[nt][2], sl2, -ms1));
        s[nt][3] = fast_exp2(fmaf(s[nt][3], sl2, -ms1));
        l0 += s[nt][0] + s[nt][1];
        l1 += s[nt][2] + s[nt][3];
      }
      l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
      l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
      const float inv0 = 1.f / l0, inv1 = 1.f / l1;
      float D0 = 0.f, D1 = 0.f;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        s[nt][0] *= inv0;
        s[nt][1] *= inv0;
        s[nt][2] *= inv1;
        s[nt][3] *= inv1;
        D0 += s[nt][0] * dp[nt][0] + s[nt][1] * dp[nt][1];
        D1 += s[nt][2] * dp[nt][2] + s[nt][3] * dp[nt][3];
      }
      D0 += __shfl_xor_sync(0xffffffffu, D0, 1);
      D0 += __shfl_xor_sync(0xffffffffu, D0, 2);
      D1 += __shfl_xor_sync(0xffffffffu, D1, 1);
      D1 += __shfl_xor_sync(0xffffffffu, D1, 2);
      uint32_t dfr[NT][2];
      const int ra = mt * 16 + g, rb = ra + 8;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        dfr[nt][0] = pack_bf16x2(s[nt][0] * (dp[nt][0] - D0) * p.scale, s[nt][1] * (dp[nt][1] - D0) * p.scale);
        dfr[nt][1] = pack_bf16x2(s[nt][2] * (dp[nt][2] - D1) * p.scale, s[nt][3] * (dp[nt][3] - D1) * p.scale);
        *reinterpret_cast<uint32_t*>(pP + tile_off(ra, nt) + qd * 4) = pack_bf16x2(s[nt][0], s[nt][1]);
        *reinterpret_cast<uint32_t*>(pP + tile_off(rb, nt) + qd * 4) = pack_bf16x2(s[nt][2], s[nt][3]);
        *reinterpret_cast<uint32_t*>(pDS + tile_off(ra, nt) + qd * 4) = dfr[nt][0];
        *reinterpret_cast<uint32_t*>(pDS + tile_off(rb, nt) + qd * 4) = dfr[nt][1];
      }
      // dQ tile = dS K
      float acc[ND][4];
#pragma unroll
      for (int nt = 0; nt < ND; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
      for (int kk = 0; kk < MT; ++kk) {
        if (kk * 16 >= p.Tk) break;
        const uint32_t a[4] = {dfr[2 * kk][0], dfr[2 * kk][1], dfr[2 * kk + 1][0], dfr[2 * kk + 1][1]};
#pragma unroll
        for (int nc = 0; nc < ND; nc += 2) {
          uint32_t b0, b1, b2, b3;
          ldsm_x4_t(sK + (nc >> 3) * kHalf + tile_off(kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, (nc & 7) + (lane >> 4)), b0, b1, b2, b3);
          mma_bf16_16816(acc[nc], a, b0, b1);
          mma_bf16_16816(acc[nc + 1], a, b2, b3);
        }
      }
      stage_acc(acc, 1.f);
      flush_rows(p.dq + q0 * p.lddq + head * 64 * DH, p.lddq * p.inner, mt * 16, p.Tq);
    }
    __syncwarp();

    // ---- phase 2: per key tile: dV = P^T dO, dK = dS^T Q (A operands re-read transposed) ---------
#pragma unroll
    for (int kt = 0; kt < MT; ++kt) {
      if (kt * 16 >= p.Tk) break;
      float av[ND][4], ak[ND][4];
#pragma unroll
      for (int nt = 0; nt < ND; ++nt) {
        av[nt][0] = av[nt][1] = av[nt][2] = av[nt][3] = 0.f;
        ak[nt][0] = ak[nt][1] = ak[nt][2] = ak[nt][3] = 0.f;
      }
#pragma unroll
      for (int qt = 0; qt < MT; ++qt) {
        if (qt * 16 >= p.Tq) break;
        // A[m = key][k = query] is stored as [query][key]: ldmatrix.trans; matrix i of the x4 load =
        // (queries 8*(i>>1).., keys 8*(i&1)..)
        const uint32_t aoff = tile_off(qt * 16 + ((lane >> 4) & 1) * 8 + (lane & 7), 2 * kt + ((lane >> 3) & 1));
        uint32_t ap[4], ad[4];
        ldsm_x4_t(sP + aoff, ap[0], ap[1], ap[2], ap[3]);
        ldsm_x4_t(sDS + aoff, ad[0], ad[1], ad[2], ad[3]);
#pragma unroll
        for (int nc = 0; nc < ND; nc += 2) {
          uint32_t b0, b1, b2, b3;
          const uint32_t boff = (nc >> 3) * kHalf + tile_off(qt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, (nc & 7) + (lane >> 4));
          ldsm_x4_t(sDO + boff, b0, b1, b2, b3);
          mma_bf16_16816(av[nc], ap, b0, b1);
          mma_bf16_16816(av[nc + 1], ap, b2, b3);
          ldsm_x4_t(sQ + boff, b0, b1, b2, b3);
          mma_bf16_16816(ak[nc], ad, b0, b1);
          mma_bf16_16816(ak[nc + 1], ad, b2, b3);
        }
      }
      stage_acc(av, 1.f);
      flush_rows(p.dv + k0 * p.lddv + head * 64 * DH, p.lddv * p.inner, kt * 16, p.Tk);
      stage_acc(ak, 1.f);
      flush_rows(p.dk + k0 * p.lddk + head * 64 * DH, p.lddk * p.inner, kt * 16, p.Tk);
    }
    __syncwarp();  // the tiles are refilled by the next item
  }
}

template <int MT, int DH>
int launch_temporal_bwd(const TbParams& p, cudaStream_t st) {
  using Cfg = TemporalBwdCfg<MT, DH>;
  static bool configured[64] = {false};  // per device: the attribute belongs to the device's context
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(attn_temporal_bwd_kernel<MT, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  const int per_sm = (227 * 1024) / (Cfg::kSmemBytes + 1024);
  int64_t grid = (p.nitems + Cfg::kWarps - 1) / Cfg::kWarps;
  const int64_t cap = (int64_t)num_sms() * (per_sm > 0 ? per_sm : 1) * 4;
  if (grid > cap) grid = cap;
  attn_temporal_bwd_kernel<MT, DH><<<(unsigned)grid, Cfg::kWarps * 32, Cfg::kSmemBytes, st>>>(p);
  TVI_LAUNCH_OK("attn_temporal_bwd_kernel");
  return TVI_OK;
}

// ---------------------------------------------------------------------------------------------
// few-key cross attention backward
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void unpack8b(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

constexpr int kFewBwdRowsPerThread = 32;      // default rows per thread and tile (the launcher picks 16 .. 128, see launch_fewkeys_bwd)
constexpr int kFewBwdMinRowsPerThread = 16;
constexpr int kFewBwdMaxRowsPerThread = 128;
constexpr int kFkDepth = 6;  // two-key kernel: rows in flight per thread (16-byte q and dO chunks in a shared-memory FIFO)

struct FkParams {
  const tvi_bf16 *q, *k, *v, *dout;
  tvi_bf16 *dq, *dk, *dv;
  float* partial;  // [groups][tiles][2][L][C] fp32 when tiles_per_group > 1
  int64_t ldq, ldk, ldv, lddo, lddq, lddk, lddv;
  int64_t rows_per_group;
  int tiles_per_group, L, Tq;
  int rows_per_thread;  // rows of a tile = row lanes (blockDim.y) * rows_per_thread
  float scale;
  const float* bias;
  int64_t inner;  // > 1: interleaved groups (logical row i of group (b, s) is physical row (b * rows_per_group + i) * inner + s)
};

template <int LT, int LPH>
__global__ void __launch_bounds__(256) attn_fewkeys_bwd_kernel(const FkParams p) {
  extern __shared__ float s_red[];  // [2][LT][C]
  const int cx = threadIdx.x;
  const int ry = threadIdx.y, RY = blockDim.y;
  const int C = blockDim.x * 8;
  const int64_t group = blockIdx.x / p.tiles_per_group;
  const int tile = blockIdx.x - (int)(group * p.tiles_per_group);
  const int64_t rows_per_tile = (int64_t)RY * p.rows_per_thread;
  const int64_t rlo = group * p.rows_per_group + (int64_t)tile * rows_per_tile;
  int64_t rhi = rlo + rows_per_tile;
  const int64_t gend = (group + 1) * p.rows_per_group;
  if (rhi > gend) rhi = gend;
  const int L = p.L;
  const int64_t gbase = group * p.rows_per_group;
  const int64_t pbase = (group / p.inner) * p.rows_per_group * p.inner + group % p.inner;
  auto phys = [&](int64_t r) -> int64_t { return p.inner == 1 ? r : pbase + (r - gbase) * p.inner; };

  float kf[LT][8], vf[LT][8], dk[LT][8], dv[LT][8];
#pragma unroll
  for (int l = 0; l < LT; ++l) {
    if (l < L) {
      unpack8b(__ldg(reinterpret_cast<const uint4*>(p.k + (group * L + l) * p.ldk + cx * 8)), kf[l]);
      unpack8b(__ldg(reinterpret_cast<const uint4*>(p.v + (group * L + l) * p.ldv + cx * 8)), vf[l]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) kf[l][j] = vf[l][j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) dk[l][j] = dv[l][j] = 0.f;
  }

  if constexpr (LT == 2) {
    // Two keys (L == 2): P = [w, 1 - w], w = sigma(z), z = scale * q . (k0 - k1) + (b0 - b1).  With g = dO . (v0 - v1):
    // dS0 = -dS1 = w (1 - w) g, so dQ = scale dS0 (k0 - k1), dK0 = -dK1 = scale sum dS0 q: two dot products and 24
    // accumulators per row instead of four and 32.
    // Packed fp32 (FFMA2 / FMUL2) throughout: the kernel is bound by its dependent instruction chain per row, not by bytes.
    uint64_t kd2[4], vd2[4], dk2[4], dv0[4], dv1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      kd2[i] = f32x2_pack(kf[0][2 * i] - kf[1][2 * i], kf[0][2 * i + 1] - kf[1][2 * i + 1]);
      vd2[i] = f32x2_pack(vf[0][2 * i] - vf[1][2 * i], vf[0][2 * i + 1] - vf[1][2 * i + 1]);
      dk2[i] = dv0[i] = dv1[i] = f32x2_pack(0.f, 0.f);
    }
    auto unpack_pairs = [](const uint4& u, uint64_t* o) {  // bf16 pair -> fp32 pair: low half << 16, high half masked
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) o[i] = f32x2_pack(__uint_as_float(w[i] << 16), __uint_as_float(w[i] & 0xffff0000u));
    };
    // Rows stream through a per-thread FIFO in shared memory (cp.async, kFkDepth rows ahead of the arithmetic; a thread
    // only reads back the 16-byte chunks it copied itself, so no barrier is involved).  The register ping-pong it replaces
    // kept ONE group of two rows in flight per thread — ~900 cycles of arithmetic against a longer DRAM latency under load:
    // the ncu capture had a third of all stall samples on the first use of the loaded registers.
    // Row addressing without per-row 64-bit arithmetic: this thread's rows are rlo + ry + it * RY, a constant pointer
    // step apart (the first version spent ~45 % of its instructions on phys(), the tail clamp and the bias index).
    const int nrows_tile = (int)(rhi - rlo);                       // <= RY * kFewBwdRowsPerThread
    const int nit = (nrows_tile + RY - 1) / RY;                    // block-uniform
    const int ry0 = ry < nrows_tile ? ry : 0;
    const int64_t prow0 = phys(rlo + ry0);
    const int64_t pstep = (int64_t)RY * (p.inner == 1 ? 1 : p.inner);  // physical rows between two rows of this thread
    const tvi_bf16* q_ptr = p.q + prow0 * p.ldq + cx * 8;          // row of the next FIFO request
    const tvi_bf16* g_ptr = p.dout + prow0 * p.lddo + cx * 8;
    tvi_bf16* dq_ptr = p.dq != nullptr ? p.dq + prow0 * p.lddq + cx * 8 : nullptr;  // row being computed
    const int64_t q_step = pstep * p.ldq, g_step = pstep * p.lddo, dq_step = pstep * p.lddq;
    const int tq = p.Tq;
    int t_row = p.bias != nullptr ? (int)((rlo + ry) % tq) : 0;
    const int t_step = RY % tq;
    const int nthreads = blockDim.x * RY, tid = ry * blockDim.x + cx;
    // [kFkDepth][2][nthreads], in the space the row lanes' partial sums take after the loop
    uint4* fifo = reinterpret_cast<uint4*>(s_red + (size_t)2 * LT * C) + tid;
    auto request = [&](int it) {  // always one commit group, so the wait below counts rows
      if (it * RY + ry < nrows_tile) {
        uint4* dst = fifo + (size_t)(it % kFkDepth) * 2 * nthreads;
        cp_async_16(dst, q_ptr, true);
        cp_async_16(dst + nthreads, g_ptr, true);
        q_ptr += q_step;
        g_ptr += g_step;
      }
      cp_async_commit();
    };
#pragma unroll
    for (int it = 0; it < kFkDepth; ++it) request(it);
    for (int it = 0; it < nit; ++it) {
      cp_async_wait<kFkDepth - 1>();
      const bool act = it * RY + ry < nrows_tile;
      const uint4* src = fifo + (size_t)(it % kFkDepth) * 2 * nthreads;
      uint4 qv = make_uint4(0u, 0u, 0u, 0u), gv = qv;  // a row lane past the tile computes on zeros: contributes nothing
      if (act) {
        qv = src[0];
        gv = src[nthreads];
      }
      uint64_t qp[4], gp[4];
      unpack_pairs(qv, qp);
      unpack_pairs(gv, gp);
      request(it + kFkDepth);  // the slot's values are in registers
      uint64_t z2 = f32x2_mul(qp[0], kd2[0]), g2 = f32x2_mul(gp[0], vd2[0]);
#pragma unroll
      for (int i = 1; i < 4; ++i) {
        z2 = f32x2_fma(qp[i], kd2[i], z2);
        g2 = f32x2_fma(gp[i], vd2[i], g2);
      }
      float z, zb, g, gb;
      f32x2_unpack(z2, z, zb);
      f32x2_unpack(g2, g, gb);
      z += zb;
      g += gb;
#pragma unroll
      for (int off = LPH / 2; off > 0; off >>= 1) {
        z += __shfl_xor_sync(0xffffffffu, z, off);
        g += __shfl_xor_sync(0xffffffffu, g, off);
      }
      z *= p.scale;
      if (p.bias != nullptr) {
        z += __ldg(p.bias + t_row * 2) - __ldg(p.bias + t_row * 2 + 1);
        t_row += t_step;  // (rlo + ry + it * RY) mod Tq without a division
        if (t_row >= tq) t_row -= tq;
      }
      float w = __fdividef(1.f, 1.f + __expf(-z));
      float w1 = 1.f - w;
      if (!act) w = w1 = 0.f;
      const float ds = w * w1 * g * p.scale;
      const uint64_t ds2 = f32x2_pack(ds, ds), w2 = f32x2_pack(w, w), w12 = f32x2_pack(w1, w1);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        dk2[i] = f32x2_fma(ds2, qp[i], dk2[i]);
        dv0[i] = f32x2_fma(w2, gp[i], dv0[i]);
        dv1[i] = f32x2_fma(w12, gp[i], dv1[i]);
      }
      if (act && dq_ptr != nullptr) {
        uint32_t o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float a, b;
          f32x2_unpack(f32x2_mul(ds2, kd2[i]), a, b);
          o[i] = pack_bf16x2(a, b);
        }
        *reinterpret_cast<uint4*>(dq_ptr) = make_uint4(o[0], o[1], o[2], o[3]);
        dq_ptr += dq_step;
      }
    }
    cp_async_wait<0>();
    if (RY > 1) __syncthreads();  // every thread is done with its FIFO slots: the space becomes s_part below
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f32x2_unpack(dk2[i], dk[0][2 * i], dk[0][2 * i + 1]);
      f32x2_unpack(dv0[i], dv[0][2 * i], dv[0][2 * i + 1]);
      f32x2_unpack(dv1[i], dv[1][2 * i], dv[1][2 * i + 1]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) dk[1][j] = -dk[0][j];
  } else {
  // the Q / dO chunks of kUnroll rows are requested before the first of them is used (memory-level parallelism: the
  // arithmetic of a row is a dependent chain of ~150 instructions)
  constexpr int kUnroll = 2;
  for (int it0 = 0; it0 < p.rows_per_thread; it0 += kUnroll) {  // rows_per_thread is even
    if (rlo + (int64_t)it0 * RY >= rhi) break;  // block-uniform
    uint4 qv[kUnroll], gv[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t rr = rlo + (int64_t)(it0 + u) * RY + ry;
      const int64_t r = rr < rhi ? rr : rlo;
      qv[u] = p.q != nullptr ? __ldg(reinterpret_cast<const uint4*>(p.q + phys(r) * p.ldq + cx * 8)) : make_uint4(0, 0, 0, 0);
      gv[u] = __ldg(reinterpret_cast<const uint4*>(p.dout + phys(r) * p.lddo + cx * 8));
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
    const int it = it0 + u;
    if (rlo + (int64_t)it * RY >= rhi) break;  // block-uniform
    const int64_t rr = rlo + (int64_t)it * RY + ry;
    const bool act = rr < rhi;
    const int64_t r = act ? rr : rlo;
    float qf[8], gof[8];
    unpack8b(qv[u], qf);
    unpack8b(gv[u], gof);
    float logit[LT], dpv[LT];
#pragma unroll
    for (int l = 0; l < LT; ++l) {
      float a = 0.f, b2 = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        a = fmaf(qf[j], kf[l][j], a);
        b2 = fmaf(gof[j], vf[l][j], b2);
      }
#pragma unroll
      for (int off = LPH / 2; off > 0; off >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, off);
        b2 += __shfl_xor_sync(0xffffffffu, b2, off);
      }
      logit[l] = a * p.scale;
      dpv[l] = b2;
    }
    if (p.bias != nullptr) {
      const int t = (int)(r % p.Tq);
#pragma unroll
      for (int l = 0; l < LT; ++l)
        if (l < L) logit[l] += __ldg(p.bias + t * L + l);
    }
    float m = -INFINITY;
#pragma unroll
    for (int l = 0; l < LT; ++l) {
      if (l >= L) logit[l] = -INFINITY;
      m = fmaxf(m, logit[l]);
    }
    float sum = 0.f, pr[LT];
#pragma unroll
    for (int l = 0; l < LT; ++l) {
      pr[l] = __expf(logit[l] - m);
      sum += pr[l];
    }
    const float inv = act ? 1.f / sum : 0.f;  // inactive rows contribute nothing
    float Dd = 0.f;
#pragma unroll
    for (int l = 0; l < LT; ++l) {
      pr[l] *= inv;
      Dd = fmaf(pr[l], dpv[l], Dd);
    }
    float dqf[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) dqf[j] = 0.f;
#pragma unroll
    for (int l = 0; l < LT; ++l) {
      const float ds = pr[l] * (dpv[l] - Dd) * p.scale;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dqf[j] = fmaf(ds, kf[l][j], dqf[j]);
        dk[l][j] = fmaf(ds, qf[j], dk[l][j]);
        dv[l][j] = fmaf(pr[l], gof[j], dv[l][j]);
      }
    }
    if (act && p.dq != nullptr) {
      uint4 o;
      o.x = pack_bf16x2(dqf[0], dqf[1]);
      o.y = pack_bf16x2(dqf[2], dqf[3]);
      o.z = pack_bf16x2(dqf[4], dqf[5]);
      o.w = pack_bf16x2(dqf[6], dqf[7]);
      *reinterpret_cast<uint4*>(p.dq + phys(r) * p.lddq + cx * 8) = o;
    }
    }
  }

  }  // LT != 2

  // ---- fixed-order reduction over the CTA's row lanes: every lane writes its sums, ONE barrier, lane 0 adds them in
  // order (a chain of RY barriers cost 14 % of the stall cycles) ------------------------------------------------
  float* s_part = RY > 1 ? s_red + 2 * LT * C : s_red;  // [RY][2][LT][C]; a single row lane writes the sums directly
#pragma unroll
  for (int l = 0; l < LT; ++l) {
#pragma unroll
    for (int j = 0; j < 8; j += 4) {
      *reinterpret_cast<float4*>(&s_part[((ry * 2 + 0) * LT + l) * C + cx * 8 + j]) = make_float4(dk[l][j], dk[l][j + 1], dk[l][j + 2], dk[l][j + 3]);
      *reinterpret_cast<float4*>(&s_part[((ry * 2 + 1) * LT + l) * C + cx * 8 + j]) = make_float4(dv[l][j], dv[l][j + 1], dv[l][j + 2], dv[l][j + 3]);
    }
  }
  __syncthreads();
  if (ry == 0 && RY > 1) {
#pragma unroll
    for (int l = 0; l < LT; ++l) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float a = 0.f, b2 = 0.f;
        for (int t = 0; t < RY; ++t) {
          a += s_part[((t * 2 + 0) * LT + l) * C + cx * 8 + j];
          b2 += s_part[((t * 2 + 1) * LT + l) * C + cx * 8 + j];
        }
        s_red[(0 * LT + l) * C + cx * 8 + j] = a;
        s_red[(1 * LT + l) * C + cx * 8 + j] = b2;
      }
    }
  }
  if (ry == 0) {
#pragma unroll
    for (int l = 0; l < LT; ++l) {
      if (l >= L) continue;
      const float* a = &s_red[(0 * LT + l) * C + cx * 8];
      const float* b2 = &s_red[(1 * LT + l) * C + cx * 8];
      if (p.tiles_per_group == 1) {
        uint4 o;
        o.x = pack_bf16x2(a[0], a[1]);
        o.y = pack_bf16x2(a[2], a[3]);
        o.z = pack_bf16x2(a[4], a[5]);
        o.w = pack_bf16x2(a[6], a[7]);
        *reinterpret_cast<uint4*>(p.dk + (group * L + l) * p.lddk + cx * 8) = o;
        o.x = pack_bf16x2(b2[0], b2[1]);
        o.y = pack_bf16x2(b2[2], b2[3]);
        o.z = pack_bf16x2(b2[4], b2[5]);
        o.w = pack_bf16x2(b2[6], b2[7]);
        *reinterpret_cast<uint4*>(p.dv + (group * L + l) * p.lddv + cx * 8) = o;
      } else {
        float* dst = p.partial + (((group * p.tiles_per_group + tile) * 2 + 0) * (int64_t)L + l) * C + cx * 8;
        *reinterpret_cast<float4*>(dst) = make_float4(a[0], a[1], a[2], a[3]);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(a[4], a[5], a[6], a[7]);
        dst = p.partial + (((group * p.tiles_per_group + tile) * 2 + 1) * (int64_t)L + l) * C + cx * 8;
        *reinterpret_cast<float4*>(dst) = make_float4(b2[0], b2[1], b2[2], b2[3]);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(b2[4], b2[5], b2[6], b2[7]);
      }
    }
  }
}

// one warp per (group, dK | dV, key, column pair): lane l adds tiles l, l + 32, ... in order, then a fixed shuffle tree
__global__ void __launch_bounds__(256)
fewkeys_bwd_final_kernel(const float* __restrict__ partial, tvi_bf16* __restrict__ dk, tvi_bf16* __restrict__ dv, int64_t ngroups,
                         int tiles, int L, int C, int64_t lddk, int64_t lddv) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int half = C / 2;
  const int64_t total = ngroups * 2 * L * half;
  if (i >= total) return;
  const int cp = (int)(i % half);
  const int l = (int)((i / half) % L);
  const int which = (int)((i / ((int64_t)half * L)) % 2);
  const int64_t g = i / ((int64_t)half * L * 2);
  float a = 0.f, b = 0.f;
  for (int t = lane; t < tiles; t += 32) {
    const float2 v = *reinterpret_cast<const float2*>(partial + (((g * tiles + t) * 2 + which) * (int64_t)L + l) * C + cp * 2);
    a += v.x;
    b += v.y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  tvi_bf16* dst = which == 0 ? dk + (g * L + l) * lddk : dv + (g * L + l) * lddv;
  if (lane == 0) *reinterpret_cast<uint32_t*>(dst + cp * 2) = pack_bf16x2(a, b);
}

int fk_block_rows(int C8) {
  const int unit = (C8 % 32 == 0) ? 1 : (C8 % 16 == 0) ? 2 : 4;
  int RY = unit * (256 / (C8 * unit) > 0 ? 256 / (C8 * unit) : 1);
  return RY;
}

int fk_tiles(int64_t rows_per_group, int C8, int rows_per_thread) {
  const int64_t rows_per_tile = (int64_t)fk_block_rows(C8) * rows_per_thread;
  return (int)((rows_per_group + rows_per_tile - 1) / rows_per_tile);
}

template <int LT, int LPH>
int launch_fewkeys_bwd(const FkParams& p_in, int64_t ngroups, int C8, cudaStream_t st, int* tiles_out) {
  FkParams p = p_in;
  const int RY = fk_block_rows(C8);
  TVI_CHECK_ARG(C8 * RY <= 256, "tvi_attn_fewkeys_bwd: heads*d = %d is too wide for this kernel", C8 * 8);
  const size_t sums = (size_t)2 * LT * C8 * 8 * sizeof(float);
  size_t sm = sums * (RY > 1 ? 1 + RY : 1);  // the sums + every row lane's partials
  if (LT == 2) {                              // the two-key kernel's row FIFO shares the partials' space
    const size_t fifo = (size_t)kFkDepth * 2 * C8 * RY * 16;
    if (sums + fifo > sm) sm = sums + fifo;
  }
  TVI_CHECK_ARG(sm <= 200 * 1024, "tvi_attn_fewkeys_bwd: L*heads*d too large for the reduction buffer");
  static size_t configured[64] = {0};  // per device: largest dynamic shared memory size opted into so far
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (sm > 48 * 1024 && (dev < 0 || dev >= 64 || sm > configured[dev])) {
    TVI_CUDA_OK(cudaFuncSetAttribute(attn_fewkeys_bwd_kernel<LT, LPH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    if (dev >= 0 && dev < 64) configured[dev] = sm;
  }
  // Tile height.  Every tile costs the same and pays a K / V prologue and a partial-sum epilogue, so the grid should be ONE
  // wave that fills the resident CTA slots, or else waves whose last one is full: among 16 .. 128 rows per thread take the
  // height with the best fill (near-ties: the taller tile).  L0 spatial cross-attention (28 groups x 9 216 rows, 5 CTAs x
  // 148 SMs resident): 32 rows per thread = 2 016 tiles = 2.72 waves; 88 rows = 27 x 28 = 756 ... the search decides.
  int per_sm = 1;
  TVI_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, attn_fewkeys_bwd_kernel<LT, LPH>, C8 * RY, sm));
  const int64_t resident = (int64_t)num_sms() * (per_sm > 0 ? per_sm : 1);
  int best_rpt = kFewBwdRowsPerThread;
  double best_eff = -1.0;
  for (int rpt = kFewBwdMaxRowsPerThread; rpt >= kFewBwdMinRowsPerThread; rpt -= 2) {
    const int64_t blocks = ngroups * fk_tiles(p.rows_per_group, C8, rpt);
    const int64_t waves = (blocks + resident - 1) / resident;
    const double eff = (double)blocks / (double)(waves * resident);
    if (eff > best_eff + 0.02) {
      best_eff = eff;
      best_rpt = rpt;
    }
  }
  p.rows_per_thread = best_rpt;
  p.tiles_per_group = fk_tiles(p.rows_per_group, C8, best_rpt);
  *tiles_out = p.tiles_per_group;
  TVI_CHECK_ARG((int64_t)p.tiles_per_group * ngroups < ((int64_t)1 << 31), "tvi_attn_fewkeys_bwd: grid too large");
  dim3 block((unsigned)C8, (unsigned)RY);
  dim3 grid((unsigned)(p.tiles_per_group * ngroups));
  attn_fewkeys_bwd_kernel<LT, LPH><<<grid, block, sm, st>>>(p);
  TVI_LAUNCH_OK("attn_fewkeys_bwd_kernel");
  return TVI_OK;
}

}  // namespace
}  // namespace tvi

namespace tvi {
size_t spatial_bwd_tc_workspace_bytes(int batch, int heads, int S);
int launch_spatial_bwd_tc(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* o, const tvi_bf16* dout, const float* lse,
                          tvi_bf16* dq, tvi_bf16* dk, tvi_bf16* dv, void* workspace, int batch, int heads, int S, int Skv, int64_t ldq,
                          int64_t ldk, int64_t ldv, int64_t ldo, int64_t lddo, int64_t lddq, int64_t lddk, int64_t lddv, float scale,
                          cudaStream_t st);
}  // namespace tvi

extern "C" {

size_t tvi_attn_spatial_bwd_workspace_bytes(int batch, int heads, int S) {
  if (batch <= 0 || heads <= 0 || S <= 0) return 0;
  // the tcgen05 passes keep lse2 and -D as [batch*heads][S padded to 128] planes; the mma.sync passes need D only
  return tvi::spatial_bwd_tc_workspace_bytes(batch, heads, S);
}

int tvi_attn_spatial_bwd_algo(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* o, const tvi_bf16* dout,
                              const float* lse, tvi_bf16* dq, tvi_bf16* dk, tvi_bf16* dv, void* workspace, int batch, int heads,
                              int S, int Skv, int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int64_t lddo,
                              int64_t lddq, int64_t lddk, int64_t lddv, float scale, int algo, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(batch >= 0 && heads >= 0 && S >= 0 && Skv >= 0, "tvi_attn_spatial_bwd: negative size");
  if (batch == 0 || heads == 0 || S == 0) return TVI_OK;
  TVI_CHECK_ARG(Skv >= 1, "tvi_attn_spatial_bwd: softmax over zero keys is undefined");
  TVI_CHECK_ARG(q && k && v && o && dout && lse && dq && dk && dv && workspace, "tvi_attn_spatial_bwd: null pointer");
  if (d != 64 && d != 128) return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_spatial_bwd: head dim %d (64 and 128 are built)", d);
  TVI_CHECK_ARG((ldq | ldk | ldv | ldo | lddo | lddq | lddk | lddv) % 8 == 0, "tvi_attn_spatial_bwd: row strides must be multiples of 8 elements");
  const int64_t C = (int64_t)heads * d;
  TVI_CHECK_ARG(ldq >= C && ldk >= C && ldv >= C && ldo >= C && lddo >= C && lddq >= C && lddk >= C && lddv >= C,
                "tvi_attn_spatial_bwd: row stride smaller than heads*d");
  TVI_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o | (uintptr_t)dout | (uintptr_t)dq | (uintptr_t)dk |
                 (uintptr_t)dv) % 16 == 0, "tvi_attn_spatial_bwd: pointers must be 16-byte aligned");
  TVI_CHECK_ARG((int64_t)batch * heads <= 65535, "tvi_attn_spatial_bwd: batch*heads exceeds 65535");
  TVI_CHECK_ARG(algo >= TVI_SPATIAL_AUTO && algo <= TVI_SPATIAL_TCGEN05, "tvi_attn_spatial_bwd_algo: unknown algo %d", algo);
  cudaStream_t st = (cudaStream_t)stream;
  if (algo == TVI_SPATIAL_AUTO) {
    // a CTA of the tcgen05 passes owns 128 rows and streams the other side in steps of 128; short sequences (L2 / mid
    // level: 576, 144 tokens) stay on the mma.sync passes, like the forward
    // (and d = 128: the tcgen05 passes are built for 64-wide heads)
    algo = (S >= 1024 && Skv >= 1024 && d == 64) ? TVI_SPATIAL_TCGEN05 : TVI_SPATIAL_MMA_SYNC;
  }
  if (algo == TVI_SPATIAL_TCGEN05 && d != 64)
    return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_spatial_bwd: the tcgen05 passes are built for head dim 64 (got %d)", d);
  if (algo == TVI_SPATIAL_TCGEN05)
    return launch_spatial_bwd_tc(q, k, v, o, dout, lse, dq, dk, dv, workspace, batch, heads, S, Skv, ldq, ldk, ldv, ldo, lddo, lddq, lddk,
                                 lddv, scale, st);
  float* D = (float*)workspace;
  {
    const int64_t nrows = (int64_t)batch * S;
    const int64_t threads = nrows * heads * (d / 8);
    if (d == 64) attn_bwd_dot_kernel<1><<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(o, dout, D, nrows, heads, ldo, lddo);
    else attn_bwd_dot_kernel<2><<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(o, dout, D, nrows, heads, ldo, lddo);
    TVI_LAUNCH_OK("attn_bwd_dot_kernel");
  }
  FbParams p;
  p.lse = lse;
  p.D = D;
  p.heads = heads;
  p.scale = scale;
  // pass 1: rows = keys -> dK, dV
  p.r1 = k; p.ldr1 = ldk; p.r2 = v; p.ldr2 = ldv;
  p.c1 = q; p.ldc1 = ldq; p.c2 = dout; p.ldc2 = lddo;
  p.out1 = dk; p.ldo1 = lddk; p.out2 = dv; p.ldo2 = lddv;
  p.nrows = Skv; p.ncols = S;
  int rc = d == 64 ? launch_flash_bwd<true, 1>(p, batch, st) : launch_flash_bwd<true, 2>(p, batch, st);
  if (rc) return rc;
  // pass 2: rows = queries -> dQ
  p.r1 = q; p.ldr1 = ldq; p.r2 = dout; p.ldr2 = lddo;
  p.c1 = k; p.ldc1 = ldk; p.c2 = v; p.ldc2 = ldv;
  p.out1 = dq; p.ldo1 = lddq; p.out2 = nullptr; p.ldo2 = 0;
  p.nrows = S; p.ncols = Skv;
  return d == 64 ? launch_flash_bwd<false, 1>(p, batch, st) : launch_flash_bwd<false, 2>(p, batch, st);
}

int tvi_attn_spatial_bwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* o, const tvi_bf16* dout,
                         const float* lse, tvi_bf16* dq, tvi_bf16* dk, tvi_bf16* dv, void* workspace, int batch, int heads,
                         int S, int Skv, int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int64_t lddo,
                         int64_t lddq, int64_t lddk, int64_t lddv, float scale, void* stream) {
  return tvi_attn_spatial_bwd_algo(q, k, v, o, dout, lse, dq, dk, dv, workspace, batch, heads, S, Skv, d, ldq, ldk, ldv, ldo, lddo, lddq,
                                   lddk, lddv, scale, TVI_SPATIAL_AUTO, stream);
}

int tvi_attn_temporal_bwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* dout, tvi_bf16* dq,
                          tvi_bf16* dk, tvi_bf16* dv, int64_t nseq, int heads, int Tq, int Tk, int d, int64_t ldq, int64_t ldk,
                          int64_t ldv, int64_t lddo, int64_t lddq, int64_t lddk, int64_t lddv, float scale, void* stream) {
  return tvi_attn_temporal_bwd_interleaved(q, k, v, dout, dq, dk, dv, nseq, heads, Tq, Tk, d, 1, ldq, ldk, ldv, lddo, lddq, lddk, lddv,
                                           scale, stream);
}

int tvi_attn_temporal_bwd_interleaved(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* dout, tvi_bf16* dq,
                                      tvi_bf16* dk, tvi_bf16* dv, int64_t nseq, int heads, int Tq, int Tk, int d, int64_t seq_inner,
                                      int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo, int64_t lddq, int64_t lddk, int64_t lddv,
                                      float scale, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(seq_inner >= 1 && (nseq == 0 || nseq % seq_inner == 0), "tvi_attn_temporal_bwd: nseq must be a multiple of seq_inner");
  TVI_CHECK_ARG(nseq >= 0 && heads >= 0, "tvi_attn_temporal_bwd: negative size");
  if (nseq == 0 || heads == 0) return TVI_OK;
  TVI_CHECK_ARG(q && k && v && dout && dq && dk && dv, "tvi_attn_temporal_bwd: null pointer");
  if (d != 64 && d != 128) return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_temporal_bwd: head dim %d (64 and 128 are built)", d);
  TVI_CHECK_ARG(Tq >= 1 && Tq <= 64 && Tk >= 1 && Tk <= 64, "tvi_attn_temporal_bwd: Tq=%d Tk=%d outside [1,64]", Tq, Tk);
  TVI_CHECK_ARG((ldq | ldk | ldv | lddo | lddq | lddk | lddv) % 8 == 0, "tvi_attn_temporal_bwd: row strides must be multiples of 8 elements");
  const int64_t C = (int64_t)heads * d;
  TVI_CHECK_ARG(ldq >= C && ldk >= C && ldv >= C && lddo >= C && lddq >= C && lddk >= C && lddv >= C,
                "tvi_attn_temporal_bwd: row stride smaller than heads*d");
  TVI_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)dout | (uintptr_t)dq | (uintptr_t)dk | (uintptr_t)dv) % 16 == 0,
                "tvi_attn_temporal_bwd: pointers must be 16-byte aligned");
  TbParams p;
  p.q = q; p.k = k; p.v = v; p.dout = dout; p.dq = dq; p.dk = dk; p.dv = dv;
  p.ldq = ldq; p.ldk = ldk; p.ldv = ldv; p.lddo = lddo; p.lddq = lddq; p.lddk = lddk; p.lddv = lddv;
  p.nitems = nseq * heads;
  p.heads = heads; p.Tq = Tq; p.Tk = Tk; p.scale = scale;
  p.inner = seq_inner;
  const int T = Tq > Tk ? Tq : Tk;
  cudaStream_t st = (cudaStream_t)stream;
#define TVI_TBWD(MT) (d == 64 ? launch_temporal_bwd<MT, 1>(p, st) : launch_temporal_bwd<MT, 2>(p, st))
  if (T <= 16) return TVI_TBWD(1);
  if (T <= 32) return TVI_TBWD(2);
  return TVI_TBWD(4);
#undef TVI_TBWD
}

size_t tvi_attn_fewkeys_bwd_workspace_bytes(int64_t nrows, int64_t rows_per_group, int heads, int L, int d) {
  if (nrows <= 0 || rows_per_group <= 0 || heads <= 0 || L <= 0 || d <= 0) return 0;
  const int C = heads * d;
  const int tiles = tvi::fk_tiles(rows_per_group, C / 8, tvi::kFewBwdMinRowsPerThread);  // the most the launcher may use
  if (tiles == 1) return 0;
  return (size_t)(nrows / rows_per_group) * tiles * 2 * L * (size_t)C * sizeof(float);
}

int tvi_attn_fewkeys_bwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* dout, tvi_bf16* dq,
                         tvi_bf16* dk, tvi_bf16* dv, void* workspace, int64_t nrows, int64_t rows_per_group, int heads, int L,
                         int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo, int64_t lddq, int64_t lddk,
                         int64_t lddv, float scale, int Tq, const float* bias, void* stream) {
  return tvi_attn_fewkeys_bwd_interleaved(q, k, v, dout, dq, dk, dv, workspace, nrows, rows_per_group, heads, L, d, 1, ldq, ldk, ldv,
                                          lddo, lddq, lddk, lddv, scale, Tq, bias, stream);
}

int tvi_attn_fewkeys_bwd_interleaved(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* dout, tvi_bf16* dq,
                                     tvi_bf16* dk, tvi_bf16* dv, void* workspace, int64_t nrows, int64_t rows_per_group, int heads,
                                     int L, int d, int64_t group_inner, int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo,
                                     int64_t lddq, int64_t lddk, int64_t lddv, float scale, int Tq, const float* bias, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(nrows >= 0 && heads >= 0, "tvi_attn_fewkeys_bwd: negative size");
  if (nrows == 0 || heads == 0) return TVI_OK;
  TVI_CHECK_ARG(k && v && dout && dk && dv, "tvi_attn_fewkeys_bwd: null pointer");
  TVI_CHECK_ARG((q == nullptr) == (dq == nullptr), "tvi_attn_fewkeys_bwd: q and dq must both be given or both be null");
  TVI_CHECK_ARG(q != nullptr || L == 1, "tvi_attn_fewkeys_bwd: q may only be null when L == 1");
  if (d != 64 && d != 128) return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_fewkeys_bwd: head dim %d (64 and 128 are built)", d);
  TVI_CHECK_ARG(L >= 1 && L <= 8, "tvi_attn_fewkeys_bwd: L=%d outside [1,8]", L);
  TVI_CHECK_ARG(rows_per_group >= 1 && nrows % rows_per_group == 0, "tvi_attn_fewkeys_bwd: nrows must be a multiple of rows_per_group");
  TVI_CHECK_ARG(Tq >= 1, "tvi_attn_fewkeys_bwd: Tq must be >= 1");
  const int C = heads * d;
  TVI_CHECK_ARG((ldk | ldv | lddo | lddk | lddv) % 8 == 0 && (q == nullptr || (ldq | lddq) % 8 == 0),
                "tvi_attn_fewkeys_bwd: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)dout | (uintptr_t)dq | (uintptr_t)dk | (uintptr_t)dv) % 16 == 0,
                "tvi_attn_fewkeys_bwd: pointers must be 16-byte aligned");
  const int64_t ngroups = nrows / rows_per_group;
  TVI_CHECK_ARG(group_inner >= 1 && ngroups % group_inner == 0, "tvi_attn_fewkeys_bwd: the group count must be a multiple of group_inner");
  FkParams p;
  p.inner = group_inner;
  p.q = q; p.k = k; p.v = v; p.dout = dout; p.dq = dq; p.dk = dk; p.dv = dv;
  p.partial = (float*)workspace;
  p.ldq = ldq; p.ldk = ldk; p.ldv = ldv; p.lddo = lddo; p.lddq = lddq; p.lddk = lddk; p.lddv = lddv;
  p.rows_per_group = rows_per_group;
  p.tiles_per_group = fk_tiles(rows_per_group, C / 8, kFewBwdMinRowsPerThread);  // upper bound; the launcher sets the real one
  p.rows_per_thread = kFewBwdRowsPerThread;
  p.L = L; p.Tq = Tq; p.scale = scale; p.bias = bias;
  TVI_CHECK_ARG(p.tiles_per_group == 1 || (workspace != nullptr && (uintptr_t)workspace % 16 == 0),
                "tvi_attn_fewkeys_bwd: workspace required (16-byte aligned)");
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  int tiles = 1;
#define TVI_FEWB(LT) (d == 64 ? launch_fewkeys_bwd<LT, 8>(p, ngroups, C / 8, st, &tiles) : launch_fewkeys_bwd<LT, 16>(p, ngroups, C / 8, st, &tiles))
  if (L == 1) rc = TVI_FEWB(1);
  else if (L == 2) rc = TVI_FEWB(2);
  else if (L <= 4) rc = TVI_FEWB(4);
  else rc = TVI_FEWB(8);
#undef TVI_FEWB
  if (rc != TVI_OK) return rc;
  if (tiles > 1) {
    const int64_t total = ngroups * 2 * L * (C / 2);
    fewkeys_bwd_final_kernel<<<(unsigned)((total + 7) / 8), 256, 0, st>>>((const float*)workspace, dk, dv, ngroups, tiles, L, C,
                                                                              lddk, lddv);
    TVI_LAUNCH_OK("fewkeys_bwd_final_kernel");
  }
  return TVI_OK;
}

}  // extern "C"
