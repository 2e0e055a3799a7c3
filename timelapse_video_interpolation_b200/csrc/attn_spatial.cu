// A3a — spatial self-attention core, O = softmax(scale . Q K^T) V over `batch` independent
// sequences (frames) of S query rows / Skv key rows, every head, d = 64 or 128, no mask, non-causal.
// Replaces F.scaled_dot_product_attention / xformers.ops.memory_efficient_attention on
// [BF, h, S, 64] (src/models/attention_processor.py:1243, :1156; caller src/models/attention.py:276-281).
//
// Flash-style single pass with online fp32 softmax; the S x Skv score matrix never leaves the SM.
// One CTA = 128 query rows of one (frame, head): 8 warps x 16 rows.  Q fragments live in
// registers for the whole pass; K/V tiles of 64 keys stream through a 3-deep cp.async ring of
// XOR-swizzled 128-byte rows (conflict-free ldmatrix); QK^T and PV run on mma.sync m16n8k16 with
// P re-used straight from the accumulator registers; O is staged through shared memory so every
// global store is a coalesced 128-bit write.  Heads are read in place from the packed
// [tokens, heads*d] projection output (row stride ld*), so no head-split copy is ever made.
// d = 128 (the class default `num_attention_heads=(5, 10, 10, 20)` at C = 1280,
// src/models/unet_spatio_temporal_condition.py:94) is the same kernel over DH = 2 swizzled 64-column halves per tile,
// one CTA per SM (Q fragments + the 128-wide accumulator need the whole register file).
// FLOPs: 4 * batch * heads * S * Skv * d.
#include "mma_sync.cuh"

namespace tvi {
namespace {

constexpr int kSpBr = 128;  // query rows per CTA
constexpr int kSpBc = 64;   // keys per pipeline stage
constexpr int kSpStages = 3;
constexpr int kSpThreads = 256;
constexpr int kSpQBytes = kSpBr * 128;   // one 64-column half of the Q tile
constexpr int kSpKVBytes = kSpBc * 128;  // one 64-column half of a K (or V) stage
constexpr int sp_smem(int DH) { return DH * (kSpQBytes + kSpStages * 2 * kSpKVBytes); }  // DH * (16 KB + 48 KB)

// K and V of one stage: [DH halves][64 keys][64 channels], K first
template <int DH>
__device__ __forceinline__ void sp_load_kv(uint32_t dstK, uint32_t dstV, const tvi_bf16* kb, const tvi_bf16* vb, int j0,
                                           int Skv, int64_t ldk, int64_t ldv, int tid) {
#pragma unroll
  for (int i = 0; i < (kSpBc * 8 * DH) / kSpThreads; ++i) {
    const int idx = tid + i * kSpThreads;
    const int r = idx / (8 * DH), c = idx % (8 * DH);
    const bool in = (j0 + r) < Skv;
    const uint32_t off = (c >> 3) * kSpKVBytes + tile_off(r, c & 7);
    const int64_t row = in ? (j0 + r) : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dstK + off), "l"(kb + row * ldk + c * 8),
                 "r"(in ? 16 : 0)
                 : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dstV + off), "l"(vb + row * ldv + c * 8),
                 "r"(in ? 16 : 0)
                 : "memory");
  }
}

template <int DH>
__global__ void __launch_bounds__(kSpThreads, DH == 1 ? 2 : 1)
attn_spatial_kernel(const tvi_bf16* __restrict__ q, const tvi_bf16* __restrict__ k, const tvi_bf16* __restrict__ v,
                    tvi_bf16* __restrict__ o, float* __restrict__ lse, int heads, int S, int Skv, int64_t ldq, int64_t ldk,
                    int64_t ldv, int64_t ldo, float scale) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, qd = lane & 3;
  const int bh = blockIdx.y;
  const int b = bh / heads, h = bh - b * heads;
  const int q0 = blockIdx.x * kSpBr;

  const uint32_t sQ = smem_u32(smem_raw);
  const uint32_t sKV = sQ + DH * kSpQBytes;
  constexpr int kStage = 2 * DH * kSpKVBytes;  // K halves then V halves
  const tvi_bf16* qb = q + ((int64_t)b * S) * ldq + h * 64 * DH;
  const tvi_bf16* kb = k + ((int64_t)b * Skv) * ldk + h * 64 * DH;
  const tvi_bf16* vb = v + ((int64_t)b * Skv) * ldv + h * 64 * DH;

  // ---- prologue: Q tile + first K/V stages ----------------------------------------------
#pragma unroll
  for (int i = 0; i < (kSpBr * 8 * DH) / kSpThreads; ++i) {
    const int idx = tid + i * kSpThreads;
    const int r = idx / (8 * DH), c = idx % (8 * DH);
    const bool in = (q0 + r) < S;
    const int64_t row = in ? (q0 + r) : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sQ + (c >> 3) * kSpQBytes + tile_off(r, c & 7)), "l"(qb + row * ldq + c * 8),
                 "r"(in ? 16 : 0)
                 : "memory");
  }
  const int nblk = (Skv + kSpBc - 1) / kSpBc;
  sp_load_kv<DH>(sKV, sKV + DH * kSpKVBytes, kb, vb, 0, Skv, ldk, ldv, tid);
  cp_async_commit();  // group 0: Q + stage 0
  if (nblk > 1) sp_load_kv<DH>(sKV + kStage, sKV + kStage + DH * kSpKVBytes, kb, vb, kSpBc, Skv, ldk, ldv, tid);
  cp_async_commit();  // group 1: stage 1 (possibly empty)

  cp_async_wait<1>();
  __syncthreads();
  uint32_t qf[4 * DH][4];
#pragma unroll
  for (int ks = 0; ks < 4 * DH; ++ks)
    ldsm_x4(sQ + (ks >> 2) * kSpQBytes + tile_off(warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, 2 * (ks & 3) + (lane >> 4)),
            qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);

  const float sl2 = scale * 1.4426950408889634f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  float acc[8 * DH][4];
#pragma unroll
  for (int nt = 0; nt < 8 * DH; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;

  for (int j = 0; j < nblk; ++j) {
    // stage j has landed for this thread (at most the newest group may be pending) ...
    cp_async_wait<1>();
    // ... and for everybody; also: everybody is done reading stage j-1, which is refilled next
    __syncthreads();
    if (j + 2 < nblk) {
      const int st = (j + 2) % kSpStages;
      sp_load_kv<DH>(sKV + st * kStage, sKV + st * kStage + DH * kSpKVBytes, kb, vb, (j + 2) * kSpBc, Skv, ldk, ldv, tid);
    }
    cp_async_commit();
    const uint32_t sK = sKV + (j % kSpStages) * kStage;
    const uint32_t sV = sK + DH * kSpKVBytes;

    // ---- S = Q K^T (16 x 64 per warp) ---------------------------------------------------
    float s[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 4 * DH; ++ks) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(sK + (ks >> 2) * kSpKVBytes + tile_off(np * 16 + (lane & 7) + (lane >> 4) * 8, 2 * (ks & 3) + ((lane >> 3) & 1)),
                b0, b1, b2, b3);
        mma_bf16_16816(s[2 * np], qf[ks], b0, b1);
        mma_bf16_16816(s[2 * np + 1], qf[ks], b2, b3);
      }
    }
    if ((j + 1) * kSpBc > Skv) {  // ragged last block: mask keys past the sequence
      const int j0 = j * kSpBc;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const int col = j0 + nt * 8 + qd * 2;
        if (col >= Skv) s[nt][0] = s[nt][2] = -INFINITY;
        if (col + 1 >= Skv) s[nt][1] = s[nt][3] = -INFINITY;
      }
    }
    // ---- online softmax ------------------------------------------------------------------
    float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      bm0 = fmaxf(bm0, fmaxf(s[nt][0], s[nt][1]));
      bm1 = fmaxf(bm1, fmaxf(s[nt][2], s[nt][3]));
    }
    bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 1));
    bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 2));
    bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 1));
    bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 2));
    const float mn0 = fmaxf(m0, bm0), mn1 = fmaxf(m1, bm1);  // finite: every block has >= 1 real key
    const float a0 = fast_exp2((m0 - mn0) * sl2), a1 = fast_exp2((m1 - mn1) * sl2);
    m0 = mn0;
    m1 = mn1;
    const float ms0 = mn0 * sl2, ms1 = mn1 * sl2;
    float rs0 = 0.f, rs1 = 0.f;
    uint32_t pfr[8][2];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const float e0 = fast_exp2(fmaf(s[nt][0], sl2, -ms0)), e1 = fast_exp2(fmaf(s[nt][1], sl2, -ms0));
      const float e2 = fast_exp2(fmaf(s[nt][2], sl2, -ms1)), e3 = fast_exp2(fmaf(s[nt][3], sl2, -ms1));
      rs0 += e0 + e1;
      rs1 += e2 + e3;
      pfr[nt][0] = pack_bf16x2(e0, e1);
      pfr[nt][1] = pack_bf16x2(e2, e3);
    }
    l0 = l0 * a0 + rs0;  // per-thread partial row sums; reduced across the quad at the end
    l1 = l1 * a1 + rs1;
#pragma unroll
    for (int nt = 0; nt < 8 * DH; ++nt) {
      acc[nt][0] *= a0;
      acc[nt][1] *= a0;
      acc[nt][2] *= a1;
      acc[nt][3] *= a1;
    }
    // ---- O += P V ---------------------------------------------------------------------------
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const uint32_t a[4] = {pfr[2 * kk][0], pfr[2 * kk][1], pfr[2 * kk + 1][0], pfr[2 * kk + 1][1]};
#pragma unroll
      for (int nc = 0; nc < 8 * DH; nc += 2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(sV + (nc >> 3) * kSpKVBytes + tile_off(kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, (nc & 7) + (lane >> 4)),
                  b0, b1, b2, b3);
        mma_bf16_16816(acc[nc], a, b0, b1);
        mma_bf16_16816(acc[nc + 1], a, b2, b3);
      }
    }
  }
  cp_async_wait<0>();

  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = 1.f / l0, inv1 = 1.f / l1;

  // Q tile is dead (fragments are in registers since before the loop): stage O over this warp's 16 rows.
  uint8_t* pQ = smem_raw;
  const int r0 = warp * 16 + g, r1 = r0 + 8;
#pragma unroll
  for (int nt = 0; nt < 8 * DH; ++nt) {
    uint8_t* ph = pQ + (nt >> 3) * kSpQBytes;
    *reinterpret_cast<uint32_t*>(ph + tile_off(r0, nt & 7) + qd * 4) = pack_bf16x2(acc[nt][0] * inv0, acc[nt][1] * inv0);
    *reinterpret_cast<uint32_t*>(ph + tile_off(r1, nt & 7) + qd * 4) = pack_bf16x2(acc[nt][2] * inv1, acc[nt][3] * inv1);
  }
  if (lse != nullptr && qd == 0) {
    if (q0 + r0 < S) lse[((int64_t)b * S + q0 + r0) * heads + h] = m0 * scale + __logf(l0);
    if (q0 + r1 < S) lse[((int64_t)b * S + q0 + r1) * heads + h] = m1 * scale + __logf(l1);
  }
  __syncwarp();
  tvi_bf16* ob = o + ((int64_t)b * S) * ldo + h * 64 * DH;
#pragma unroll
  for (int i = 0; i < 4 * DH; ++i) {
    const int idx = i * 32 + lane;
    const int r = warp * 16 + idx / (8 * DH), c = idx % (8 * DH);
    if (q0 + r < S) {
      const uint4 val = *reinterpret_cast<const uint4*>(pQ + (c >> 3) * kSpQBytes + tile_off(r, c & 7));
      *reinterpret_cast<uint4*>(ob + (int64_t)(q0 + r) * ldo + c * 8) = val;
    }
  }
}

}  // namespace
}  // namespace tvi

namespace tvi {
int launch_spatial_tc(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse, int batch, int heads,
                       int S, int Skv, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, float scale, int poly, cudaStream_t st);
}

extern "C" int tvi_attn_spatial_fwd_algo(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse,
                                         int batch, int heads, int S, int Skv, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                                         int64_t ldo, float scale, int algo, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(batch >= 0 && heads >= 0 && S >= 0 && Skv >= 0, "tvi_attn_spatial_fwd: negative size");
  if (batch == 0 || heads == 0 || S == 0) return TVI_OK;
  TVI_CHECK_ARG(Skv >= 1, "tvi_attn_spatial_fwd: softmax over zero keys is undefined");
  TVI_CHECK_ARG(q && k && v && o, "tvi_attn_spatial_fwd: null pointer");
  if (d != 64 && d != 128) return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_spatial_fwd: head dim %d (64 and 128 are built)", d);
  TVI_CHECK_ARG(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0, "tvi_attn_spatial_fwd: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(ldq >= heads * d && ldk >= heads * d && ldv >= heads * d && ldo >= heads * d, "tvi_attn_spatial_fwd: row stride smaller than heads*d");
  TVI_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o) % 16 == 0, "tvi_attn_spatial_fwd: pointers must be 16-byte aligned");
  TVI_CHECK_ARG((int64_t)batch * heads <= 65535, "tvi_attn_spatial_fwd: batch*heads exceeds 65535");
  TVI_CHECK_ARG(algo >= TVI_SPATIAL_AUTO && algo <= TVI_SPATIAL_TCGEN05, "tvi_attn_spatial_fwd_algo: unknown algo %d", algo);
  if (algo == TVI_SPATIAL_AUTO) {
    // a CTA of the tcgen05 kernel owns 256 query rows and needs a few key tiles to amortise its
    // prologue; short sequences (L2 / mid level: 576, 144 tokens) stay on the mma.sync kernel
    // (and d = 128 heads: the tcgen05 kernel's TMEM / shared-memory plan is built for 64-wide heads)
    algo = (S >= 1024 && Skv >= 1024 && d == 64) ? TVI_SPATIAL_TCGEN05 : TVI_SPATIAL_MMA_SYNC;
  }
  if (algo == TVI_SPATIAL_TCGEN05) {
    if (d != 64) return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_spatial_fwd: the tcgen05 kernel is built for head dim 64 (got %d)", d);
    int poly = 4;
#ifdef TVI_TC_POLY_SWEEP  // tuning builds only (tools/micro): share of exponentials on the FMA pipe from the environment
    static int env_poly = -1;
    if (env_poly < 0) {
      const char* e = getenv("TVI_TC_POLY");
      env_poly = e ? atoi(e) : 4;
    }
    poly = env_poly;
#endif
    return launch_spatial_tc(q, k, v, o, lse, batch, heads, S, Skv, ldq, ldk, ldv, ldo, scale, poly, (cudaStream_t)stream);
  }
  static bool configured[64] = {false};
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(attn_spatial_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, sp_smem(1)));
    TVI_CUDA_OK(cudaFuncSetAttribute(attn_spatial_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sp_smem(2)));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  dim3 grid((unsigned)((S + kSpBr - 1) / kSpBr), (unsigned)(batch * heads));
  if (d == 64)
    attn_spatial_kernel<1><<<grid, kSpThreads, sp_smem(1), (cudaStream_t)stream>>>(q, k, v, o, lse, heads, S, Skv, ldq, ldk, ldv,
                                                                                 ldo, scale);
  else
    attn_spatial_kernel<2><<<grid, kSpThreads, sp_smem(2), (cudaStream_t)stream>>>(q, k, v, o, lse, heads, S, Skv, ldq, ldk, ldv,
                                                                                 ldo, scale);
  TVI_LAUNCH_OK("attn_spatial_kernel");
  return TVI_OK;
}

extern "C" int tvi_attn_spatial_fwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse, int batch,
                                    int heads, int S, int Skv, int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo,
                                    float scale, void* stream) {
  return tvi_attn_spatial_fwd_algo(q, k, v, o, lse, batch, heads, S, Skv, d, ldq, ldk, ldv, ldo, scale, TVI_SPATIAL_AUTO, stream);
}
