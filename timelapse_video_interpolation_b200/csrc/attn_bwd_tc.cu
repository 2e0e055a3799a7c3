// Backward of the spatial self-attention core on the 5th-generation tensor cores (tcgen05 / TMEM / TMA): what autograd
// derives for F.scaled_dot_product_attention / xformers at src/models/attention_processor.py:1243, :1156 when
// src/train_decouples.py:1068 (`accelerator.backward`) runs; dX through the frozen spatial blocks is needed on every
// training step (:707-720).  Flash-style recomputation from Q, K, V, the forward's LSE and D = rowsum(dO o O):
//     P = exp(scale S - LSE)      dV = P^T dO      dP = dO V^T      dS = P o (dP - D)      dQ = scale dS K      dK = scale dS^T Q
// in TWO passes of one kernel template, so there are no atomics and the summation order is fixed:
//     pass 1 (rows = keys)   : a CTA owns 128 keys  (R1 = K, R2 = V),  streams the queries (C1 = Q, C2 = dO) -> dV, dK
//     pass 2 (rows = queries): a CTA owns 128 queries (R1 = Q, R2 = dO), streams the keys (C1 = K, C2 = V)   -> dQ
// Per step of 128 streamed columns, with rows on the TMEM lanes:
//     S'  = R1 C1^T    (M=128 N=128 K=64, SS)          pass 1: S^T          pass 2: S
//     dP' = R2 C2^T    (M=128 N=128 K=64, SS)          pass 1: dP^T         pass 2: dP
//     P'  = exp2(S' scale log2e - lse2[query])  -> bf16 over the first half of the S' buffer       (X warps)
//     dS' = P' o (dP' - D[query])                -> bf16 over the first half of the dP' buffer      (Y warps)
//     accP += P'  C2   (M=128 N=64 K=128, TS, C2 MN-major)      pass 1 only: dV
//     accD += dS' C1   (M=128 N=64 K=128, TS, C1 MN-major)      pass 1: dK   pass 2: dQ          (scale in the epilogue)
// 640 threads, warp-specialised: warp 0 TMA producer (resident R tiles, a 4-stage ring of C tiles and, in pass 1, the
// per-query lse2 / -D vectors of the step), warp 1 MMA issuer (warp-uniform loop), warp 2 owns the tensor memory,
// warps 4..11 "X" (two threads per row, 64 of the step's 128 columns each: the exponentials; the 16-per-clock MUFU unit is
// this role's resource, 1 of 4 score pairs takes the FMA-pipe polynomial), warps 12..19 "Y" (two threads per row: dS on the
// FMA / ALU pipes).  Every sub-partition holds two X and two Y warps: a single warp per role could not issue its ~480
// dependent instructions per step faster than one per 4 cycles (2150 cycles per step against 1024 of tensor work), and
// the exponentials of step i+1 overlap the dS arithmetic of step i.  A thread's P' / dS' go over the first 32 columns of
// its OWN 64 score columns (behind its read front), so the bf16 A operands sit at columns [0,32) and [64,96) of a buffer.
// Tensor memory: three rotating 128-column buffers B0 B1 B2 at [0,384), accP [384,448), accD [448,512).  Step i keeps S'_i
// (then P'_i) in B[2i mod 3] and dP'_i (then dS'_i) in B[(2i+1) mod 3]; the third buffer already holds S'_{i+1}.  dP'_{i+1}
// goes into the buffer of P'_i as soon as the Y warps have copied P'_i into registers (p_read) and accP_i is issued, i.e.
// while dS'_i is still being computed, and S'_{i+2} into the buffer of dS'_i after accD_i: no product waits on the step
// before it, the X warps run one step ahead of the Y warps, and Y_{i+1} starts as soon as Y_i ends.
// Nothing of size S x Skv reaches memory.  7 GEMM units of 2 B h S Skv 64 FLOPs per launch pair (S' and dP' twice).
#include "tvi_common.cuh"

#ifndef TVI_BWD_POLY
#define TVI_BWD_POLY 0x11u
#endif

namespace tvi {
namespace {

constexpr int kBtThreads = 640;
constexpr int kBtRows = 128;      // rows per CTA (UMMA M)
constexpr int kBtCols = 128;      // streamed columns per step (UMMA N of S' / dP', K of the accumulating products)
constexpr int kBtStages = 4;
constexpr int kBtTile = kBtRows * 128;                 // one [128][64] bf16 tile, 128-byte rows, SWIZZLE_128B
constexpr int kBtVec = kBtCols * 4;                    // one fp32 vector of the step's columns
constexpr int kBtStageBytes = 2 * kBtTile + 2 * kBtVec;  // C1 | C2 | lse2 | -D
constexpr int kBtSmem = 1024 + 2 * kBtTile + kBtStages * kBtStageBytes;

constexpr uint32_t kBtColAccP = 384, kBtColAccD = 448;  // rotating buffers: column 128 * k

struct BtParams {
  int nrows, ncols;     // per sequence: pass 1 (Skv, S), pass 2 (S, Skv)
  int n_steps;          // ceil(ncols / 128)
  int heads;
  int q_pad;            // padded query count of the lse2 / -D planes (multiple of 128)
  float scale;
  const float* lse2;    // [batch*heads][q_pad]: LSE * log2(e)   (+inf in the padding: P = 0 there)
  const float* negd;    // [batch*heads][q_pad]: -rowsum(dO o O) (0 in the padding)
  tvi_bf16* outP;       // pass 1: dV          (rows = keys)
  tvi_bf16* outD;       // pass 1: dK          pass 2: dQ
  int64_t ldoP, ldoD;
};

struct BtBarriers {
  uint64_t r_full;
  uint64_t c_full[kBtStages];
  uint64_t c_empty[kBtStages];
  // by step parity (one completion per two steps): a parity wait is only sound for a waiter at most one phase behind,
  // and the X warps run a step ahead of the Y warps
  uint64_t s_full[2];    // S' of a step is in tensor memory
  uint64_t p_full[2];    // P' of a step is stored (4 X warps)
  uint64_t dp_full[2];   // dP' of a step is in tensor memory
  // one completion per step; the issuing warp, which takes them in program order, is the only waiter
  uint64_t p_read;       // the Y warps hold P' of a step in registers: its buffer may take dP' of the next step
  uint64_t ds_full;      // dS' of a step is stored (4 Y warps)
  uint64_t acc_done;
};

template <bool kKeysOnRows, uint32_t kPolyMask>
__global__ void __launch_bounds__(kBtThreads, 1)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmR1, const __grid_constant__ CUtensorMap tmR2,
                   const __grid_constant__ CUtensorMap tmC1, const __grid_constant__ CUtensorMap tmC2, const BtParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) BtBarriers bars;
  __shared__ uint32_t s_tmem_base;

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  // [R1 | R2 | stage 0: C1 C2 lse2 negD | stage 1 ...]
  const uint32_t sR1 = base, sR2 = base + kBtTile, sStage = base + 2 * kBtTile;

  const int bh = blockIdx.y;
  const int b = bh / p.heads, h = bh - b * p.heads;
  const int row0 = blockIdx.x * kBtRows;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmR1);
    tma_prefetch_desc(&tmR2);
    tma_prefetch_desc(&tmC1);
    tma_prefetch_desc(&tmC2);
    mbar_init(&bars.r_full, 1);
    for (int s = 0; s < kBtStages; ++s) {
      mbar_init(&bars.c_full[s], 1);
      mbar_init(&bars.c_empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars.s_full[i], 1);
      mbar_init(&bars.p_full[i], 8);
      mbar_init(&bars.dp_full[i], 1);
    }
    mbar_init(&bars.p_read, 8);
    mbar_init(&bars.ds_full, 8);
    mbar_init(&bars.acc_done, 1);
    mbar_fence_init();
  }
  if (warp == 2) {
    tmem_alloc(&s_tmem_base, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      mbar_arrive_expect_tx(&bars.r_full, 2 * kBtTile);
      tma_load_3d(base_ptr, &tmR1, &bars.r_full, h * 64, row0, b);
      tma_load_3d(base_ptr + kBtTile, &tmR2, &bars.r_full, h * 64, row0, b);
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < p.n_steps; ++i) {
        TVI_WAIT(&bars.c_empty[s], ph ^ 1, 1, i);
        uint8_t* dst = base_ptr + 2 * kBtTile + s * kBtStageBytes;
        mbar_arrive_expect_tx(&bars.c_full[s], kKeysOnRows ? kBtStageBytes : 2 * kBtTile);
        tma_load_3d(dst, &tmC1, &bars.c_full[s], h * 64, i * kBtCols, b);
        tma_load_3d(dst + kBtTile, &tmC2, &bars.c_full[s], h * 64, i * kBtCols, b);
        if (kKeysOnRows) {  // the step's columns are queries: their lse2 / -D
          const int64_t off = (int64_t)bh * p.q_pad + (int64_t)i * kBtCols;
          tma_load_1d(dst + 2 * kBtTile, p.lse2 + off, kBtVec, &bars.c_full[s]);
          tma_load_1d(dst + 2 * kBtTile + kBtVec, p.negd + off, kBtVec, &bars.c_full[s]);
        }
        if (++s == kBtStages) {
          s = 0;
          ph ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t leader = elect_one() ? 1u : 0u;
    if (tmem_base != 0u) __trap();  // 512 columns = the whole tensor memory of the SM: the allocation starts at column 0
    const uint32_t idesc_ss = umma_idesc_bf16(kBtRows, kBtCols, 0, 0);
    const uint32_t idesc_ts = umma_idesc_bf16(kBtRows, 64, 0, 1);  // B MN-major ([streamed rows][d])
    const uint32_t hi_desc = (1024u >> 4) | (1u << 14) | (kUmmaLayoutSw128 << 29);
    auto desc64 = [&](uint32_t lo) {
      uint64_t d;
      asm("mov.b64 %0, {%1,%2};" : "=l"(d) : "r"(lo), "r"(hi_desc));
      return d;
    };
    const uint32_t lo_r1 = ((sR1 >> 4) & 0x3FFFu) | (1u << 16), lo_r2 = ((sR2 >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t lo_c0_k = ((sStage >> 4) & 0x3FFFu) | (1u << 16);    // C tiles as K-major B operand (LBO 16)
    const uint32_t lo_c0_mn = ((sStage >> 4) & 0x3FFFu) | (64u << 16);  // C tiles as MN-major B operand (LBO 1024)
    constexpr uint32_t kStage16 = kBtStageBytes >> 4, kTile16 = kBtTile >> 4;
    auto issue_s = [&](int st, uint32_t col, int par) {       // S' = R1 C1^T
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        umma_bf16_ss_pred(col, desc64(lo_r1 + kk * 2), desc64(lo_c0_k + (uint32_t)st * kStage16 + kk * 2), idesc_ss, kk != 0 ? 1u : 0u,
                          leader);
      umma_commit_pred(&bars.s_full[par], leader);
    };
    auto issue_dp = [&](int st, uint32_t col, int par) {      // dP' = R2 C2^T
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        umma_bf16_ss_pred(col, desc64(lo_r2 + kk * 2), desc64(lo_c0_k + (uint32_t)st * kStage16 + kTile16 + kk * 2), idesc_ss,
                          kk != 0 ? 1u : 0u, leader);
      umma_commit_pred(&bars.dp_full[par], leader);
    };
    const int n = p.n_steps;
    mbar_wait_uniform(&bars.r_full, 0);
    mbar_wait_uniform(&bars.c_full[0], 0);
    tc_fence_after();
    issue_s(0, 0, 0);        // S'_0 -> B0
    issue_dp(0, 128, 0);     // dP'_0 -> B1
    if (n > 1) {
      mbar_wait_uniform(&bars.c_full[1], 0);
      tc_fence_after();
      issue_s(1, 256, 1);    // S'_1 -> B2
    }
    int st = 0;              // ring stage of step i
    uint32_t cs = 0;         // column of the buffer of S'_i / P'_i; dP'_i / dS'_i sit in the next buffer (mod 3)
    for (int i = 0; i < n; ++i) {
      const int par = i & 1;
      const uint32_t cd = cs == 256 ? 0u : cs + 128;
      const uint32_t acc = i != 0 ? 1u : 0u;
      const uint32_t lo_mn = lo_c0_mn + (uint32_t)st * kStage16;
      const int st1 = (st + 1) & (kBtStages - 1);
      if (kKeysOnRows) {
        // accP += P'_i C2_i
        mbar_wait_uniform(&bars.p_full[par], (uint32_t)(i >> 1) & 1u);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)
          umma_bf16_ts_pred(kBtColAccP, cs + (kk >> 2) * 64 + (kk & 3) * 8, desc64(lo_mn + kTile16 + kk * 128), idesc_ts,
                            kk != 0 ? 1u : acc, leader);
      }
      // dP'_{i+1} into the buffer of P'_i: the Y warps have it in registers, accP_i (just issued) reads it first (in order)
      mbar_wait_uniform(&bars.p_read, (uint32_t)i & 1u);
      tc_fence_after();
      if (i + 1 < n) issue_dp(st1, cs, par ^ 1);
      // accD += dS'_i C1_i
      mbar_wait_uniform(&bars.ds_full, (uint32_t)i & 1u);
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < 8; ++kk)
        umma_bf16_ts_pred(kBtColAccD, cd + (kk >> 2) * 64 + (kk & 3) * 8, desc64(lo_mn + kk * 128), idesc_ts, kk != 0 ? 1u : acc, leader);
      umma_commit_pred(&bars.c_empty[st], leader);  // the stage's tiles have been read once these products complete
      if (i + 2 < n) {
        const int st2 = (st + 2) & (kBtStages - 1);
        mbar_wait_uniform(&bars.c_full[st2], (uint32_t)((i + 2) / kBtStages) & 1u);
        tc_fence_after();
        issue_s(st2, cd, par);                      // S'_{i+2} into the buffer of dS'_i (read by accD_i, in order)
      }
      st = st1;
      cs = cd == 256 ? 0u : cd + 128;               // S'_{i+1} sits two buffers on
    }
    umma_commit_pred(&bars.acc_done, leader);
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== X (exponentials) / Y (dS) warps, then the epilogue =====================
    const bool is_x = warp < 12;
    const int quad = warp & 3;  // the TMEM lane quadrant a warp may touch is fixed by its index
    const int ch = ((warp - 4) >> 2) & 1;  // which 64 of the step's 128 columns
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const uint32_t tm = tmem_base + lane_base;
    const int rr = quad * 32 + lane;
    const int row = row0 + rr;
    const float sl2 = p.scale * 1.4426950408889634f;
    const uint64_t sl2x2 = f32x2_pack(sl2, sl2);
    // rows = queries: this thread's own lse2 / -D
    float row_lse2 = 0.f, row_negd = 0.f;
    if (!kKeysOnRows && row < p.nrows) {
      row_lse2 = p.lse2[(int64_t)bh * p.q_pad + row];
      row_negd = p.negd[(int64_t)bh * p.q_pad + row];
    }
    int st = 0;
    uint32_t cs = 0;  // column of the buffer of S'_i / P'_i
    if (is_x) {
      for (int i = 0; i < p.n_steps; ++i) {
        const int par = i & 1;
        const float* vec = reinterpret_cast<const float*>(base_ptr + 2 * kBtTile + st * kBtStageBytes + 2 * kBtTile);  // lse2 of the columns
        // the lse2 vector of the step arrived with the stage's tiles (observe that barrier itself before reading it)
        if (kKeysOnRows) TVI_WAIT(&bars.c_full[st], (uint32_t)(i / kBtStages) & 1u, 6, i);
        TVI_WAIT(&bars.s_full[par], (uint32_t)(i >> 1) & 1u, 2, i);
        tc_fence_after();
        const uint32_t tS = tm + cs + ch * 64;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t ra[32];
          tmem_ld_x32(tS + c * 32, ra);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int k = 0; k < 32; k += 4) {
            float n0, n1, n2, n3;
            if (kKeysOnRows) {
              const float4 l4 = *reinterpret_cast<const float4*>(vec + ch * 64 + c * 32 + k);  // same address in every lane: broadcast
              n0 = -l4.x, n1 = -l4.y, n2 = -l4.z, n3 = -l4.w;
            } else {
              n0 = n1 = n2 = n3 = -row_lse2;
            }
            uint64_t xa = f32x2_fma(f32x2_pack(__uint_as_float(ra[k]), __uint_as_float(ra[k + 1])), sl2x2, f32x2_pack(n0, n1));
            uint64_t xb = f32x2_fma(f32x2_pack(__uint_as_float(ra[k + 2]), __uint_as_float(ra[k + 3])), sl2x2, f32x2_pack(n2, n3));
            float a0, a1, b0, b1;
            if ((kPolyMask >> ((k >> 1) & 7)) & 1u) {
              xa = exp2_poly_x2(xa);
              f32x2_unpack(xa, a0, a1);
            } else {
              f32x2_unpack(xa, a0, a1);
              a0 = fast_exp2(a0);
              a1 = fast_exp2(a1);
            }
            if ((kPolyMask >> (((k >> 1) + 1) & 7)) & 1u) {
              xb = exp2_poly_x2(xb);
              f32x2_unpack(xb, b0, b1);
            } else {
              f32x2_unpack(xb, b0, b1);
              b0 = fast_exp2(b0);
              b1 = fast_exp2(b1);
            }
            pk[k >> 1] = pack_bf16x2(a0, a1);
            pk[(k >> 1) + 1] = pack_bf16x2(b0, b1);
          }
          // P' chunk c (32 bf16 = 16 columns) over columns [16c, 16c+16) of this thread's own 64 score columns: behind its read front
          tmem_st_x16(tS + c * 16, pk);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars.p_full[par]);
        if (++st == kBtStages) st = 0;
        cs = cs >= 128 ? cs - 128 : 256;  // + 2 buffers (mod 3)
      }
    } else {
      for (int i = 0; i < p.n_steps; ++i) {
        const int par = i & 1;
        const float* vec = reinterpret_cast<const float*>(base_ptr + 2 * kBtTile + st * kBtStageBytes + 2 * kBtTile + kBtVec);  // -D of the columns
        if (kKeysOnRows) TVI_WAIT(&bars.c_full[st], (uint32_t)(i / kBtStages) & 1u, 7, i);
        TVI_WAIT(&bars.p_full[par], (uint32_t)(i >> 1) & 1u, 3, i);
        tc_fence_after();
        // this thread's P'_i (64 bf16 = 32 columns) into registers at once, so that the buffer can take dP' of the next step
        uint32_t pp[32];
        tmem_ld_x32(tm + cs + ch * 64, pp);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars.p_read);
        TVI_WAIT(&bars.dp_full[par], (uint32_t)(i >> 1) & 1u, 4, i);
        tc_fence_after();
        const uint32_t tD = tm + (cs == 256 ? 0u : cs + 128) + ch * 64;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t dp[32];
          tmem_ld_x32(tD + c * 32, dp);
          tmem_ld_wait();
          uint32_t dk[16];
#pragma unroll
          for (int k = 0; k < 32; k += 4) {
            float n0, n1, n2, n3;
            if (kKeysOnRows) {
              const float4 d4 = *reinterpret_cast<const float4*>(vec + ch * 64 + c * 32 + k);
              n0 = d4.x, n1 = d4.y, n2 = d4.z, n3 = d4.w;
            } else {
              n0 = n1 = n2 = n3 = row_negd;
            }
            const uint32_t w0 = pp[c * 16 + (k >> 1)], w1 = pp[c * 16 + (k >> 1) + 1];
            const uint64_t p01 = f32x2_pack(__uint_as_float(w0 << 16), __uint_as_float(w0 & 0xffff0000u));
            const uint64_t p23 = f32x2_pack(__uint_as_float(w1 << 16), __uint_as_float(w1 & 0xffff0000u));
            const uint64_t t01 = f32x2_add(f32x2_pack(__uint_as_float(dp[k]), __uint_as_float(dp[k + 1])), f32x2_pack(n0, n1));
            const uint64_t t23 = f32x2_add(f32x2_pack(__uint_as_float(dp[k + 2]), __uint_as_float(dp[k + 3])), f32x2_pack(n2, n3));
            float a0, a1, b0, b1;
            f32x2_unpack(f32x2_mul(t01, p01), a0, a1);
            f32x2_unpack(f32x2_mul(t23, p23), b0, b1);
            dk[k >> 1] = pack_bf16x2(a0, a1);
            dk[(k >> 1) + 1] = pack_bf16x2(b0, b1);
          }
          tmem_st_x16(tD + c * 16, dk);  // dS' chunk c over columns [16c, 16c+16) of this thread's own 64 dP' columns
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars.ds_full);
        if (++st == kBtStages) st = 0;
        cs = cs >= 128 ? cs - 128 : 256;  // + 2 buffers (mod 3)
      }
    }
    // ---- epilogue: accumulators -> bf16 -> global (one full 128-byte row segment per thread) ----
    TVI_WAIT(&bars.acc_done, 0, 5, 0);
    tc_fence_after();
    // one role (X: accP = dV in pass 1; Y: accD = dK resp. dQ, times scale) per accumulator, 32 of its 64 columns per warp
    {
      const bool mine = kKeysOnRows || !is_x;   // pass 2 has only accD: the Y warps store it
      const uint32_t tA = tm + ((kKeysOnRows && is_x) ? kBtColAccP : kBtColAccD) + ch * 32;
      tvi_bf16* out = (kKeysOnRows && is_x) ? p.outP : p.outD;
      const int64_t ld = (kKeysOnRows && is_x) ? p.ldoP : p.ldoD;
      const float f = (kKeysOnRows && is_x) ? 1.f : p.scale;
      if (mine) {
        uint4* dst = reinterpret_cast<uint4*>(out + ((int64_t)b * p.nrows + row) * ld + h * 64 + ch * 32);
        uint32_t a[32];
        tmem_ld_x32(tA, a);
        tmem_ld_wait();
        if (row < p.nrows) {
#pragma unroll
          for (int k = 0; k < 32; k += 8)
            dst[k >> 3] = make_uint4(pack_bf16x2(__uint_as_float(a[k]) * f, __uint_as_float(a[k + 1]) * f),
                                     pack_bf16x2(__uint_as_float(a[k + 2]) * f, __uint_as_float(a[k + 3]) * f),
                                     pack_bf16x2(__uint_as_float(a[k + 4]) * f, __uint_as_float(a[k + 5]) * f),
                                     pack_bf16x2(__uint_as_float(a[k + 6]) * f, __uint_as_float(a[k + 7]) * f));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// lse2 = LSE log2(e) and -D = -rowsum(dO o O), transposed to [batch*heads][q_pad] planes (contiguous per (frame, head), so
// that pass 1 can fetch the 128 values of a step with one bulk copy each); the padding reads as P = 0, D = 0.
__global__ void attn_bwd_tc_prep_kernel(const tvi_bf16* __restrict__ o, const tvi_bf16* __restrict__ dout, const float* __restrict__ lse,
                                        float* __restrict__ lse2, float* __restrict__ negd, int batch, int S, int heads, int q_pad,
                                        int64_t ldo, int64_t lddo) {
  // 8 lanes per (frame, head, query): one 16-byte chunk each
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t item = idx >> 3;
  const int c = (int)(idx & 7);
  const int64_t total = (int64_t)batch * heads * q_pad;
  const bool in = item < total;
  const int64_t bh = in ? item / q_pad : 0;
  const int s = in ? (int)(item - bh * q_pad) : 0;
  const int b = (int)(bh / heads), h = (int)(bh - (int64_t)b * heads);
  const bool real = in && s < S;
  float acc = 0.f;
  if (real) {
    const int64_t r = (int64_t)b * S + s;
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(o + r * ldo + h * 64 + c * 8));
    const uint4 d = __ldg(reinterpret_cast<const uint4*>(dout + r * lddo + h * 64 + c * 8));
    const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&d);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 x = __bfloat1622float2(ha[i]), y = __bfloat1622float2(hd[i]);
      acc = fmaf(x.x, y.x, acc);
      acc = fmaf(x.y, y.y, acc);
    }
  }
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  if (in && c == 0) {
    lse2[item] = real ? lse[((int64_t)b * S + s) * heads + h] * 1.4426950408889634f : INFINITY;
    negd[item] = real ? -acc : 0.f;
  }
}

}  // namespace

size_t spatial_bwd_tc_workspace_bytes(int batch, int heads, int S) {
  const int64_t q_pad = ((int64_t)S + kBtCols - 1) / kBtCols * kBtCols;
  return (size_t)batch * heads * q_pad * 2 * sizeof(float);
}

// Arguments are validated by tvi_attn_spatial_bwd (attn_bwd.cu).
int launch_spatial_bwd_tc(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* o, const tvi_bf16* dout, const float* lse,
                          tvi_bf16* dq, tvi_bf16* dk, tvi_bf16* dv, void* workspace, int batch, int heads, int S, int Skv, int64_t ldq,
                          int64_t ldk, int64_t ldv, int64_t ldo, int64_t lddo, int64_t lddq, int64_t lddk, int64_t lddv, float scale,
                          cudaStream_t st) {
  const int q_pad = (S + kBtCols - 1) / kBtCols * kBtCols;
  float* lse2 = (float*)workspace;
  float* negd = lse2 + (size_t)batch * heads * q_pad;
  {
    const int64_t threads = (int64_t)batch * heads * q_pad * 8;
    attn_bwd_tc_prep_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(o, dout, lse, lse2, negd, batch, S, heads, q_pad, ldo, lddo);
    TVI_LAUNCH_OK("attn_bwd_tc_prep_kernel");
  }
  auto make = [&](CUtensorMap* m, const tvi_bf16* ptr, int rows, int64_t ld) {
    // [heads*64 (contiguous), rows of one sequence, batch]; rows past the sequence read as zero
    const uint64_t dims[3] = {(uint64_t)heads * 64, (uint64_t)rows, (uint64_t)batch};
    const uint64_t strides[2] = {(uint64_t)ld * 2, (uint64_t)rows * (uint64_t)ld * 2};
    const uint32_t box[3] = {64u, 128u, 1u};
    return make_tensor_map(m, ptr, 2, 3, dims, strides, box, Swizzle::B128);
  };
  CUtensorMap tmQ, tmK, tmV, tmDo;
  int rc;
  if ((rc = make(&tmQ, q, S, ldq))) return rc;
  if ((rc = make(&tmK, k, Skv, ldk))) return rc;
  if ((rc = make(&tmV, v, Skv, ldv))) return rc;
  if ((rc = make(&tmDo, dout, S, lddo))) return rc;
  constexpr uint32_t kPoly = TVI_BWD_POLY;  // score pairs (of every 8) on the FMA-pipe polynomial
  static bool configured[64] = {};
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(attn_bwd_tc_kernel<true, kPoly>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBtSmem));
    TVI_CUDA_OK(cudaFuncSetAttribute(attn_bwd_tc_kernel<false, kPoly>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBtSmem));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  BtParams p;
  p.heads = heads;
  p.q_pad = q_pad;
  p.scale = scale;
  p.lse2 = lse2;
  p.negd = negd;
  // pass 1: rows = keys -> dV, dK
  p.nrows = Skv;
  p.ncols = S;
  p.n_steps = (S + kBtCols - 1) / kBtCols;
  p.outP = dv;
  p.ldoP = lddv;
  p.outD = dk;
  p.ldoD = lddk;
  {
    dim3 grid((unsigned)((Skv + kBtRows - 1) / kBtRows), (unsigned)(batch * heads));
    attn_bwd_tc_kernel<true, kPoly><<<grid, kBtThreads, kBtSmem, st>>>(tmK, tmV, tmQ, tmDo, p);
    TVI_LAUNCH_OK("attn_bwd_tc_kernel<keys>");
  }
  // pass 2: rows = queries -> dQ
  p.nrows = S;
  p.ncols = Skv;
  p.n_steps = (Skv + kBtCols - 1) / kBtCols;
  p.outP = nullptr;
  p.ldoP = 0;
  p.outD = dq;
  p.ldoD = lddq;
  {
    dim3 grid((unsigned)((S + kBtRows - 1) / kBtRows), (unsigned)(batch * heads));
    attn_bwd_tc_kernel<false, kPoly><<<grid, kBtThreads, kBtSmem, st>>>(tmQ, tmDo, tmK, tmV, p);
    TVI_LAUNCH_OK("attn_bwd_tc_kernel<queries>");
  }
  return TVI_OK;
}

}  // namespace tvi
