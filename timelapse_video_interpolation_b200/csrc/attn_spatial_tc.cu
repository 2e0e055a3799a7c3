// A3a — spatial self-attention core on the 5th-generation tensor cores, persistent version
// (tcgen05 / TMEM / TMA).  O = softmax(scale . Q K^T) V over `batch` independent sequences (frames) of S
// query rows and Skv key rows, every head, d = 64, no mask, non-causal.  Replaces
// F.scaled_dot_product_attention / xformers.ops.memory_efficient_attention on [BF, h, S, 64]
// (src/models/attention_processor.py:1243, :1156; caller src/models/attention.py:276-281).
//
// One CTA per SM walks a static queue of work items (256 query rows = two 128-row tiles of one
// (frame, head)), streaming the keys in steps of 64 through ONE K/V ring that keeps running across
// items.  640 threads, warp-specialised:
//   warp 0        TMA producer: K_j / V_j ([64 keys x 64] each) into an 8-stage ring, and the Q tiles
//                 ([128 rows x 64], SWIZZLE_128B) of the NEXT item into the other half of a double buffer.
//   warps 1, 2    tcgen05.mma issuers, one per tile; convergent warp-uniform loops (all waits end in a vote),
//                 so every descriptor lives in uniform registers:
//                     S_t[g&1]  = Q_t K_j^T            (M=128 N=64 K=64, SS: A = Q_t in smem, B = K_j K-major)
//                     Oa_t     += P_t[:, 0:32] V_j[0:32]   (TS: A = P in TMEM, B = V_j MN-major, K = 32)
//                     Ob_t     += P_t[:,32:64] V_j[32:64]
//                 The chain that has to fit under one softmax step is  P_g stored -> PV_g -> QK_{g+2}  (S is
//                 double-buffered and P_g lives inside S[g&1]), so the loop body is exactly that chain: descriptors
//                 are built before the wait for P_g, the K/V and Q waits of the next QK are taken after it.  The
//                 chain crosses item boundaries (QK_0 / QK_1 of the next item run under the epilogue of this one).
//   warp 3        owns the tensor-memory allocation.
//   warps 4..11   softmax of tile 0, warps 12..19 of tile 1: two threads per query row (TMEM lane), 32 of the
//                 step's 64 key columns each.  The two halves of a row are INDEPENDENT streams: each keeps its own
//                 reference maximum and row sum and owns its own output accumulator (Oa / Ob), so there is no
//                 per-step exchange between them; they meet once per item in the epilogue
//                 (O = (Oa 2^(ma-m) + Ob 2^(mb-m)) / (la 2^(ma-m) + lb 2^(mb-m))).
//                 Per step: tcgen05.ld of 32 scores, exponentials against the CURRENT reference without looking
//                 for the row maximum first (speculative; the half-row sum tells whether that was sound, else
//                 the exact pass runs: row maximum, lazy reference update, rescale of the own accumulator),
//                 P as bf16 over the first 16 columns of the thread's own score columns, tcgen05.st.
//   epilogue      combine, normalise -> bf16 -> swizzled smem (the item's dead Q buffer) -> TMA store (clips S tails).
// TMEM columns of tile t (base 256 t): S[0] [0,64)  S[1] [64,128)  Oa [128,192)  Ob [192,256).
// FLOPs: 4 * batch * heads * S * Skv * 64.  With d = 64 the exponential unit (16 ex2 / clk / SM: 1024 clk per 256 x 64
// step against 640 clk of MMA) is the ceiling, so 4 of a thread's 16 score pairs per step take a polynomial exp2 on
// the FMA pipe (tools/micro/softmax_step_bench.cu: 1032 -> 804 clk per step for the register math alone).
// What the round-2 timelines / ncu captures showed (profiles/r2_spatial_fwd.md): the round-1 kernel (two threads per
// row sharing one reference maximum through a named barrier every step, Q parked in TMEM, one-shot CTAs) ran at
// 1298 clk per step with the exponential unit 81 % busy, and moving exponentials to the FMA pipe made it SLOWER,
// because the step period was E + L (exponential phase + per-step latency of the pair exchange / TMEM round trips /
// the MMA warp's 1000-cycle issue loop), not the unit's throughput.  This kernel removes the exchange (independent
// halves, two accumulators; Q moves to shared memory to make room), shortens the MMA chain, and keeps running across
// items: 1076 clk per step under ncu.
#include <stdlib.h>

#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kP2Threads = 640;
constexpr int kP2BM = 128;     // query rows per tile (UMMA M)
constexpr int kP2BN = 64;      // keys per step
constexpr int kP2Stages = 8;
constexpr int kP2KBytes = kP2BN * 128;         // [64 keys][64 bf16]
constexpr int kP2StageBytes = 2 * kP2KBytes;   // K | V
constexpr int kP2QBytes = kP2BM * 128;         // one Q tile / one O staging tile
// [Q tile t, buffer b] x 4 | (K,V) x stages
constexpr int kP2Smem = 1024 + 4 * kP2QBytes + kP2Stages * kP2StageBytes;

constexpr uint32_t kP2ColTile = 256, kP2ColS = 0, kP2ColOa = 128, kP2ColOb = 192;

struct P2Params {
  int S, Skv, heads;
  int n_kv;        // key steps per item
  int n_qblk;      // 256-row blocks per (frame, head)
  int n_items;     // batch * heads * n_qblk
  float scale;
  float inv_qblk, inv_heads;  // 1 / n_qblk, 1 / heads (division-free item decoding)
  float* lse;      // nullable, [batch*S, heads]
};

struct P2Barriers {
  uint64_t q_full[2][2];      // [tile][buffer]: Q tile landed
  uint64_t buf_free[2][2];    // [tile][buffer]: the O store that staged through this buffer has been read out
  uint64_t kv_full[kP2Stages];
  uint64_t kv_empty[kP2Stages];
  uint64_t s_full[2][2];      // [tile][S buffer]
  uint64_t p_full[2][2];      // [tile][S buffer]: P of both halves stored (8 warps)
  uint64_t o_full[2];         // [tile]: one completion per step, after the PV products of that step
  uint64_t o_done[2];         // [tile]: one completion per ITEM, after its last PV products (a parity wait is only sound
                              // for a waiter at most one phase behind: the other half may trail by two steps)
};

#ifdef TVI_TC_TIMELINE
// debug build only (tools/micro/tc_timeline.cu): cycle stamps of 16 steps of one CTA
__device__ long long* g_tc_timeline = nullptr;
#define P2_TL(e)                                                                                   \
  do {                                                                                             \
    if (g_tc_timeline && blockIdx.x == 77 && lane == 0 && g >= 200 && g < 216)                    \
      g_tc_timeline[(((threadIdx.x >> 5) - 4) * 16 + (g - 200)) * 8 + (e)] = clock64();           \
  } while (0)
#define P2_TLM(e)                                                                                  \
  do {                                                                                             \
    if (g_tc_timeline && blockIdx.x == 77 && lane == 0 && g >= 200 && g < 216)                    \
      g_tc_timeline[2048 + ((warp - 1) * 16 + (g - 200)) * 8 + (e)] = clock64();                  \
  } while (0)
#else
#define P2_TL(e)
#define P2_TLM(e)
#endif

template <bool B>
struct P2Tag {
  static constexpr bool value = B;
};

// a / b for 0 <= a < 2^24, b >= 1 without the integer-division subroutine (inv_b = 1.0f / b from the host)
__device__ __forceinline__ int p2_div(int a, int b, float inv_b) {
  int q = (int)((float)a * inv_b);
  const int r = a - q * b;
  if (r < 0) --q;
  else if (r >= b) ++q;
  return q;
}

// kPolyMask: bit i set = pair i of a thread's 16 score pairs takes the polynomial path.
template <uint32_t kPolyMask>
__global__ void __launch_bounds__(kP2Threads, 1)
attn_spatial_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const P2Params p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) P2Barriers bars;
  __shared__ uint32_t s_tmem_base;
  __shared__ float2 s_xchg[2][2][kP2BM];  // [tile][half][row]: (reference maximum, row sum) for the epilogue

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);  // warp-uniform by construction
  const int lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t sQ = base;                       // + (t * 2 + b) * kP2QBytes
  const uint32_t sKV = base + 4 * kP2QBytes;      // + s * kP2StageBytes (+ kP2KBytes for V)

  // static queue: item w = blockIdx.x + i * gridDim.x; consecutive CTAs share a (frame, head) -> K/V hits in L2
  const int n_my = ((int)blockIdx.x < p.n_items) ? (p.n_items - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;  // (once)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
    for (int t = 0; t < 2; ++t)
      for (int i = 0; i < 2; ++i) {
        mbar_init(&bars.q_full[t][i], 1);
        mbar_init(&bars.buf_free[t][i], 4);
        mbar_init(&bars.s_full[t][i], 1);
        mbar_init(&bars.p_full[t][i], 8);
      }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&bars.o_full[t], 1);
      mbar_init(&bars.o_done[t], 1);
    }
    for (int s = 0; s < kP2Stages; ++s) {
      mbar_init(&bars.kv_full[s], 1);
      mbar_init(&bars.kv_empty[s], 2);
    }
    mbar_fence_init();
  }
  if (warp == 3) {
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
      auto load_q = [&](int i) {  // Q tiles of local item i into buffer i & 1
        const int w = (int)blockIdx.x + i * (int)gridDim.x;
        const int bh = p2_div(w, p.n_qblk, p.inv_qblk), qb = w - bh * p.n_qblk;
        const int b = p2_div(bh, p.heads, p.inv_heads), h = bh - b * p.heads;
        const int buf = i & 1;
        for (int t = 0; t < 2; ++t) {
          // the O store of item i-2 staged through this buffer: wait until it has been read out
          TVI_WAIT(&bars.buf_free[t][buf], (uint32_t)((i >> 1) & 1) ^ 1u, 1, i);
          mbar_arrive_expect_tx(&bars.q_full[t][buf], kP2QBytes);
          tma_load_3d(base_ptr + (t * 2 + buf) * kP2QBytes, &tmQ, &bars.q_full[t][buf], h * 64, qb * 2 * kP2BM + t * kP2BM, b);
        }
      };
      // Q of item i+1 goes into the buffer that staged the output of item i-1.  The producer runs up to a ring ahead
      // of the consumers, so it asks for that buffer only after K/V step kP2Stages+1 of item i: by then the softmax
      // warps are provably inside item i (its stage kP2Stages-ish was consumed), i.e. past the epilogue of item i-1,
      // and the wait cannot hold up the K/V stream (which the last PV of item i-1 transitively depends on).
      constexpr int kQAt = kP2Stages + 1;
      if (n_my > 0) load_q(0);
      int s = 0;
      uint32_t ph = 0;
      for (int i = 0; i < n_my; ++i) {
        const int w = (int)blockIdx.x + i * (int)gridDim.x;
        const int bh = p2_div(w, p.n_qblk, p.inv_qblk);
        const int b = p2_div(bh, p.heads, p.inv_heads), h = bh - b * p.heads;
        for (int j = 0; j < p.n_kv; ++j) {
          if (j == kQAt && i + 1 < n_my) load_q(i + 1);
          TVI_WAIT(&bars.kv_empty[s], ph ^ 1, 2, j);
          uint8_t* kdst = base_ptr + 4 * kP2QBytes + s * kP2StageBytes;
          mbar_arrive_expect_tx(&bars.kv_full[s], kP2StageBytes);
          tma_load_3d(kdst, &tmK, &bars.kv_full[s], h * 64, j * kP2BN, b);
          tma_load_3d(kdst + kP2KBytes, &tmV, &bars.kv_full[s], h * 64, j * kP2BN, b);
          if (++s == kP2Stages) {
            s = 0;
            ph ^= 1;
          }
        }
        if (p.n_kv <= kQAt && i + 1 < n_my) load_q(i + 1);
      }
    }
  } else if (warp == 1 || warp == 2) {
    // ===================== MMA issuers =====================
    const uint32_t leader = elect_one() ? 1u : 0u;
    // 512 columns = the whole tensor memory of the SM, so the allocation starts at column 0; the literal keeps
    // the TMEM operands compile-time uniform.
    if (tmem_base != 0u) __trap();
    const int t = warp - 1;
    const uint32_t tile = (uint32_t)t * kP2ColTile;
    const uint32_t idesc_qk = umma_idesc_bf16(kP2BM, kP2BN, 0, 0);
    const uint32_t idesc_pv = umma_idesc_bf16(kP2BM, 64, 0, 1);  // B = V is MN-major ([keys][d])
    const int total = n_my * p.n_kv;  // global steps of this CTA
    // The critical path of a tile is  P_g stored -> PV_g -> QK_{g+2} (S[g&1] is free once PV_g has read P_g) -> softmax
    // of step g+2, which has to fit under the softmax of step g+1.  So the loop body is exactly that chain, and the
    // waits for the inputs of the NEXT QK (K/V stage, Q tile) are taken in advance, off the chain.
    struct Cursor {  // position of a global step: ring stage + phase, step inside its item, local item
      int s, j, item;
      uint32_t ph;
    };
    auto advance = [&](Cursor& c) {
      if (++c.s == kP2Stages) {
        c.s = 0;
        c.ph ^= 1;
      }
      if (++c.j == p.n_kv) {
        c.j = 0;
        ++c.item;
      }
    };
    auto wait_inputs = [&](const Cursor& c) {
      if (c.j == 0) mbar_wait_uniform(&bars.q_full[t][c.item & 1], (uint32_t)(c.item >> 1) & 1u);
      mbar_wait_uniform(&bars.kv_full[c.s], c.ph);
    };
    // descriptors are computed ahead of the wait they follow, so that only tcgen05 instructions sit on the
    // P_g -> PV_g -> QK_{g+2} chain
    struct QkDesc {
      uint64_t a[4], b[4];
    };
    // Shared-memory descriptors differ only in their 14-bit start-address field (every operand sits below 256 KB and is
    // 16-byte aligned, so adding byte_offset >> 4 to the low word never carries out of it): one add per descriptor
    // instead of re-encoding it - the issuing warps share their sub-partitions' issue slots with four softmax warps.
    const uint32_t hi_desc = (1024u >> 4) | (1u << 14) | (kUmmaLayoutSw128 << 29);  // SBO 1024 B, version 1, 128-byte swizzle
    const uint32_t lo_q0 = (((sQ + (uint32_t)(t * 2) * kP2QBytes) >> 4) & 0x3FFFu) | ((16u >> 4) << 16);   // K-major: LBO 16
    const uint32_t lo_k0 = ((sKV >> 4) & 0x3FFFu) | ((16u >> 4) << 16);
    const uint32_t lo_v0 = (((sKV + kP2KBytes) >> 4) & 0x3FFFu) | ((1024u >> 4) << 16);                    // MN-major: LBO 1024
    auto desc64 = [&](uint32_t lo) {
      uint64_t d;
      asm("mov.b64 %0, {%1,%2};" : "=l"(d) : "r"(lo), "r"(hi_desc));
      return d;
    };
    auto make_qk = [&](const Cursor& c, QkDesc& d) {
      const uint32_t lo_q = lo_q0 + (uint32_t)(c.item & 1) * (kP2QBytes >> 4);
      const uint32_t lo_k = lo_k0 + (uint32_t)c.s * (kP2StageBytes >> 4);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        d.a[kk] = desc64(lo_q + kk * 2);   // 32 bytes (16 bf16 of the reduction) per MMA
        d.b[kk] = desc64(lo_k + kk * 2);
      }
    };
    auto issue_qk = [&](const QkDesc& d, int g) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        umma_bf16_ss_pred(tile + kP2ColS + (uint32_t)(g & 1) * 64, d.a[kk], d.b[kk], idesc_qk, kk != 0 ? 1u : 0u, leader);
      umma_commit_pred(&bars.s_full[t][g & 1], leader);
    };
    Cursor cq = {0, 0, 0, 0u};  // step whose QK is issued next
    Cursor cp = {0, 0, 0, 0u};  // step whose PV is issued next
    QkDesc qd;
    for (int g = 0; g < 2 && g < total; ++g) {
      wait_inputs(cq);
      tc_fence_after();
      make_qk(cq, qd);
      issue_qk(qd, g);
      advance(cq);
    }
    if (total > 2) wait_inputs(cq);
    for (int g = 0; g < total; ++g) {
      // everything the two products of this iteration need, before the wait
      const uint32_t lo_v = lo_v0 + (uint32_t)cp.s * (kP2StageBytes >> 4);
      const uint32_t pcol = tile + kP2ColS + (uint32_t)(g & 1) * 64;
      uint64_t vd[4];
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)  // 16 keys per MMA = two 8-row groups of 128-byte rows (SBO = 1024 B); one 64-wide N atom
        vd[kk] = desc64(lo_v + kk * (2048 >> 4));
      const uint32_t acc0 = cp.j != 0 ? 1u : 0u;
      const bool last_of_item = cp.j == p.n_kv - 1;
      uint64_t* bar_kv_empty = &bars.kv_empty[cp.s];
      const bool more_qk = g + 2 < total;
      if (more_qk) make_qk(cq, qd);

      mbar_wait_uniform(&bars.p_full[t][g & 1], (uint32_t)(g >> 1) & 1u);
      tc_fence_after();
      P2_TLM(0);
      umma_bf16_ts_pred(tile + kP2ColOa, pcol + 0, vd[0], idesc_pv, acc0, leader);
      umma_bf16_ts_pred(tile + kP2ColOa, pcol + 8, vd[1], idesc_pv, 1u, leader);
      umma_bf16_ts_pred(tile + kP2ColOb, pcol + 32, vd[2], idesc_pv, acc0, leader);
      umma_bf16_ts_pred(tile + kP2ColOb, pcol + 40, vd[3], idesc_pv, 1u, leader);
      // S[g&1] held P_g, read by the PV products just issued: MMAs of one thread complete in order
      if (more_qk) issue_qk(qd, g + 2);
      umma_commit_pred(&bars.o_full[t], leader);
      umma_commit_pred(bar_kv_empty, leader);  // this tile is done with K / V of step g (2 arrivals free the stage)
      if (last_of_item) umma_commit_pred(&bars.o_done[t], leader);
      P2_TLM(1);
      advance(cp);
      if (more_qk) {
        advance(cq);
        if (g + 3 < total) {
          wait_inputs(cq);
          tc_fence_after();
        }
        P2_TLM(2);
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== softmax + epilogue (warps 4..19) =====================
    const int sw = warp - 4;
    const int t = sw >> 3, half = (sw >> 2) & 1;
    const int quad = warp & 3;  // the TMEM lane quadrant a warp may touch is fixed by its index
    const uint32_t pair_bar = 1u + t * 4 + quad;  // 64 threads: the two halves of 32 rows
    const uint32_t tile_bar = 9u + t;             // 256 threads: all softmax warps of the tile
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const uint32_t tile_tmem = tmem_base + t * kP2ColTile + lane_base;
    const uint32_t tS = tile_tmem + kP2ColS + half * 32;   // + buf*64: this thread's 32 scores; P over the first 16 columns
    const uint32_t tOwn = tile_tmem + (half ? kP2ColOb : kP2ColOa);
    const int rr = quad * 32 + lane;  // row inside the tile
    uint64_t* bar_s = bars.s_full[t];
    uint64_t* bar_p = bars.p_full[t];
    uint64_t* bar_o = &bars.o_full[t];

    const float sl2 = p.scale * 1.4426950408889634f;
    const uint64_t sl2x2 = f32x2_pack(sl2, sl2);
    const bool last_ragged = (p.Skv & (kP2BN - 1)) != 0;
    uint32_t ra[32];  // this thread's 32 scores of the current step
    // global step g of this CTA: S buffer = g & 1, barrier phase = (g >> 1) & 1, kept as two toggled registers
    uint32_t buf = 0, ph = 0;
#ifdef TVI_TC_TIMELINE
    int g = 0;
#endif

    for (int i = 0; i < n_my; ++i) {
      float m_ref = -INFINITY;      // reference maximum of this half-row, raw (unscaled) logit units
      float l = 0.f;                // running sum of exp2((s - m_ref) * sl2) over this half's key columns
      // first scores of the item (QK_0 of this item ran under the previous item's epilogue)
      TVI_WAIT(&bar_s[buf], ph, 5, i);
      tc_fence_after();
      tmem_ld_x32(tS + buf * 64, ra);

      // One key step.  kRagged: last step of an item that runs past Skv.  kFirst: first step of an item (no
      // reference yet: the exact pass is taken unconditionally).
      auto step = [&](int j, auto ragged_tag) {
        constexpr bool kRagged = decltype(ragged_tag)::value;
        tmem_ld_wait();
        P2_TL(0);
        uint32_t pk[16];
        float hs;  // sum of the step's exponentials
        auto mask_tail = [&]() {
          if (kRagged) {
            const int valid = p.Skv - j * kP2BN - half * 32;
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (c >= valid) ra[c] = 0xff800000u;  // -inf
          }
        };
        auto row_max = [&]() {
          float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            mx0 = fmax3(mx0, __uint_as_float(ra[c]), __uint_as_float(ra[c + 1]));
            mx1 = fmax3(mx1, __uint_as_float(ra[c + 2]), __uint_as_float(ra[c + 3]));
          }
          return fmaxf(mx0, mx1);
        };
        // Exponentials IN PLACE over the score registers (the exact pass re-reads the scores from tensor memory, where
        // they stay until P is stored), in two sweeps: all the ex2 first, sums and bf16 packing afterwards.  With the
        // scores dead early and the consumers a whole sweep behind the MUFU results, one warp alone can keep the
        // 16-per-clock exponential unit busy, so the unit idles only when every warp of the sub-partition is between steps.
        auto exp_pass = [&]() {
          // a half-row that has not seen a finite score yet keeps m_ref = -inf: evaluate against 0 (all its scores
          // are -inf -> P = 0)
          const float neg = (m_ref == -INFINITY) ? 0.f : -m_ref * sl2;
          const uint64_t negx2 = f32x2_pack(neg, neg);
#pragma unroll
          for (int c = 0; c < 32; c += 2) {
            uint64_t x = f32x2_fma(f32x2_pack(__uint_as_float(ra[c]), __uint_as_float(ra[c + 1])), sl2x2, negx2);
            float e0, e1;
            if ((kPolyMask >> (c >> 1)) & 1u) {
              x = exp2_poly_x2(x);
              f32x2_unpack(x, e0, e1);
            } else {
              f32x2_unpack(x, e0, e1);
              e0 = fast_exp2(e0);
              e1 = fast_exp2(e1);
            }
            ra[c] = __float_as_uint(e0);
            ra[c + 1] = __float_as_uint(e1);
          }
          uint64_t sum_a = f32x2_pack(0.f, 0.f), sum_b = sum_a;
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            sum_a = f32x2_add(sum_a, f32x2_pack(__uint_as_float(ra[c]), __uint_as_float(ra[c + 1])));
            sum_b = f32x2_add(sum_b, f32x2_pack(__uint_as_float(ra[c + 2]), __uint_as_float(ra[c + 3])));
            pk[c >> 1] = pack_bf16x2(__uint_as_float(ra[c]), __uint_as_float(ra[c + 1]));
            pk[(c >> 1) + 1] = pack_bf16x2(__uint_as_float(ra[c + 2]), __uint_as_float(ra[c + 3]));
          }
          float s0, s1, s2, s3;
          f32x2_unpack(sum_a, s0, s1);
          f32x2_unpack(sum_b, s2, s3);
          hs = (s0 + s1) + (s2 + s3);
        };
        mask_tail();
        if (j == 0) {
          // first step of an item: no reference yet, take the row maximum of the step (l = 0, and the first PV
          // product overwrites the accumulator: nothing to rescale)
          const float mx = row_max();
          if (mx > m_ref) m_ref = mx;
          exp_pass();
        } else {
          // Speculative pass against the CURRENT reference without looking for the row maximum first; P <= hs <= 2^12
          // keeps every value far inside the bf16 / fp32 range and the reference does not have to be the exact maximum
          // (O and l carry the same factor).  A sum that is larger, infinite or NaN sends the warp through the exact pass.
          exp_pass();
          if (__any_sync(0xffffffffu, !(hs <= 4096.f))) {
            tmem_ld_x32(tS + buf * 64, ra);  // the scores are still there: P is stored below
            tmem_ld_wait();
            mask_tail();
            const float mx = row_max();
            // advance the reference only when the row maximum outgrew it by more than 2^8 (always from -inf)
            const bool need = (mx > m_ref) && ((m_ref == -INFINITY) || (mx - m_ref) * sl2 > 8.f);
            float alpha = 1.f;
            if (need) {
              alpha = (m_ref == -INFINITY) ? 0.f : fast_exp2((m_ref - mx) * sl2);
              m_ref = mx;
              l *= alpha;
            }
            if (__any_sync(0xffffffffu, need)) {
              // rescale this half's accumulator: the PV products of step g-1 must have landed; those of step g are
              // not issued before this warp arrives on bar_p below
              TVI_WAIT(bar_o, buf ^ 1u, 6, j);  // one completion per step: parity of step g-1
              tc_fence_after();
#pragma unroll 1
              for (int c0 = 0; c0 < 64; c0 += 8) {  // small chunks: the 32 scores stay in registers for the second pass
                uint32_t o[8];
                tmem_ld_x8(tOwn + c0, o);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 8; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
                tmem_st_x8(tOwn + c0, o);
              }
            }
            exp_pass();
          }
        }
        P2_TL(1);
        l += hs;
        // P (32 bf16 = 16 columns) overwrites the first half of this thread's OWN 32 score columns
        tmem_st_x16(tS + buf * 64, pk);
        P2_TL(2);
        // the next step's scores (other S buffer, normally long finished) travel to registers while the P store drains
        if (j + 1 < p.n_kv) {
          TVI_WAIT(&bar_s[buf ^ 1], ph ^ buf, 7, j);  // phase of step g+1 (flips when the buffer index wraps)
          tc_fence_after();
          tmem_ld_x32(tS + (buf ^ 1) * 64, ra);
        }
        P2_TL(3);
        tmem_st_wait();
        tc_fence_before();
        P2_TL(4);
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_p[buf]);
        ph ^= buf;
        buf ^= 1u;
#ifdef TVI_TC_TIMELINE
        ++g;
#endif
      };
      for (int j = 0; j < p.n_kv - 1; ++j) step(j, P2Tag<false>{});
      if (last_ragged) step(p.n_kv - 1, P2Tag<true>{});
      else step(p.n_kv - 1, P2Tag<false>{});

      // ---- epilogue: combine the halves, normalise -> bf16 -> swizzled staging -> TMA store ----------------
      const int w = (int)blockIdx.x + i * (int)gridDim.x;
      const int bh = p2_div(w, p.n_qblk, p.inv_qblk), qb = w - bh * p.n_qblk;
      const int b = p2_div(bh, p.heads, p.inv_heads), h = bh - b * p.heads;
      const int q_row0 = qb * 2 * kP2BM + t * kP2BM;
      const int row = q_row0 + rr;  // row inside the sequence
      s_xchg[t][half][rr] = make_float2(m_ref, l);
      named_bar_sync(pair_bar, 64);
      const float2 other = s_xchg[t][half ^ 1][rr];
      const float m = fmaxf(m_ref, other.x);  // finite: Skv >= 1 puts a valid key in half 0's first column
      const float f_own = (m_ref == -INFINITY) ? 0.f : fast_exp2((m_ref - m) * sl2);
      const float f_oth = (other.x == -INFINITY) ? 0.f : fast_exp2((other.x - m) * sl2);
      const float lsum = l * f_own + other.y * f_oth;
      const float inv = 1.f / lsum;
      const float fa = (half == 0 ? f_own : f_oth) * inv, fb = (half == 0 ? f_oth : f_own) * inv;
      // every PV product of this item has landed in both accumulators
      TVI_WAIT(&bars.o_done[t], (uint32_t)i & 1u, 8, 0);
      tc_fence_after();
      // this thread writes output columns [32 half, 32 half + 32) of its row
      uint4 outv[4];
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        uint32_t oa[16], ob[16];
        tmem_ld_x16(tile_tmem + kP2ColOa + half * 32 + ch * 16, oa);
        tmem_ld_x16(tile_tmem + kP2ColOb + half * 32 + ch * 16, ob);
        tmem_ld_wait();
        uint32_t pkd[8];
#pragma unroll
        for (int c = 0; c < 16; c += 2)
          pkd[c >> 1] = pack_bf16x2(__uint_as_float(oa[c]) * fa + __uint_as_float(ob[c]) * fb,
                                    __uint_as_float(oa[c + 1]) * fa + __uint_as_float(ob[c + 1]) * fb);
        outv[ch * 2 + 0] = make_uint4(pkd[0], pkd[1], pkd[2], pkd[3]);
        outv[ch * 2 + 1] = make_uint4(pkd[4], pkd[5], pkd[6], pkd[7]);
      }
      // both accumulators are in registers in all 8 warps before any of them lets the next item's first PV start
      tc_fence_before();
      named_bar_sync(tile_bar, 256);
      // staging = this item's Q buffer: every QK product of the item completed before the last PV did
      uint8_t* stage_ptr = base_ptr + (t * 2 + (i & 1)) * kP2QBytes;
      uint8_t* row_ptr = stage_ptr + rr * 128;
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        const int chunk = half * 4 + ch;
        *reinterpret_cast<uint4*>(row_ptr + ((chunk ^ (rr & 7)) << 4)) = outv[ch];
      }
      if (half == 0 && p.lse != nullptr && row < p.S) p.lse[((int64_t)b * p.S + row) * p.heads + h] = m * p.scale + __logf(lsum);
      fence_proxy_async_smem();
      named_bar_sync(pair_bar, 64);  // both halves of the 32 rows are staged
      if (half == 0 && lane == 0) {
        if (q_row0 + quad * 32 < p.S) {
          tma_store_3d(&tmO, stage_ptr + quad * 32 * 128, h * 64, q_row0 + quad * 32, b);
          tma_store_commit();
          tma_store_wait_read<0>();
        }
        mbar_arrive(&bars.buf_free[t][i & 1]);
      }
      __syncwarp();
    }
    if (half == 0 && lane == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 3) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

// Arguments are already validated by tvi_attn_spatial_fwd_algo (attn_spatial.cu).
int launch_spatial_tc(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse, int batch, int heads,
                       int S, int Skv, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, float scale, int poly,
                       cudaStream_t st) {
  CUtensorMap tmQ, tmK, tmV, tmO;
  struct Spec {
    CUtensorMap* m;
    const void* ptr;
    int rows;
    int64_t ld;
    uint32_t box_rows;
  } specs[4] = {{&tmQ, q, S, ldq, (uint32_t)kP2BM}, {&tmK, k, Skv, ldk, (uint32_t)kP2BN}, {&tmV, v, Skv, ldv, (uint32_t)kP2BN},
                {&tmO, o, S, ldo, 32}};
  for (const Spec& sp : specs) {
    // [heads*64 (contiguous), rows of one sequence, batch]; rows past the sequence read as zero / are not written
    const uint64_t dims[3] = {(uint64_t)heads * 64, (uint64_t)sp.rows, (uint64_t)batch};
    const uint64_t strides[2] = {(uint64_t)sp.ld * 2, (uint64_t)sp.rows * (uint64_t)sp.ld * 2};
    const uint32_t box[3] = {64u, sp.box_rows, 1u};
    const int rc = make_tensor_map(sp.m, sp.ptr, 2, 3, dims, strides, box, Swizzle::B128);
    if (rc) return rc;
  }
  using Kern = void (*)(CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, P2Params);
  // share of the exponentials evaluated on the FMA pipe: 0, 4 (default), 5 or 6 of a thread's 16 score pairs per step.
  // Measured at L0 (28 x 5 x 9216 tokens, 100 back-to-back launches, power-limited clocks): 0 -> 3.50 ms, 4 -> 3.30,
  // 5 -> 3.33, 6 -> 3.36; single launch under ncu: 0 -> 1280 cycles per key step, 4 -> 1076.
  Kern kern;
  int slot;
  switch (poly) {
    case 0: kern = attn_spatial_tc_kernel<0x0000u>; slot = 0; break;
    case 2: kern = attn_spatial_tc_kernel<0x0101u>; slot = 4; break;
    case 3: kern = attn_spatial_tc_kernel<0x0421u>; slot = 5; break;
    case 5: kern = attn_spatial_tc_kernel<0x1249u>; slot = 2; break;
    case 6: kern = attn_spatial_tc_kernel<0x2525u>; slot = 3; break;
    default: kern = attn_spatial_tc_kernel<0x1111u>; slot = 1; break;
  }
  static bool configured[64][6] = {};
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev][slot]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kP2Smem));
    if (dev >= 0 && dev < 64) configured[dev][slot] = true;
  }
  P2Params p;
  p.S = S;
  p.Skv = Skv;
  p.heads = heads;
  p.n_kv = (Skv + kP2BN - 1) / kP2BN;
  p.n_qblk = (S + 2 * kP2BM - 1) / (2 * kP2BM);
  p.n_items = batch * heads * p.n_qblk;
  if ((int64_t)batch * heads * p.n_qblk >= (1 << 24)) return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_spatial_fwd: more than 2^24 work items");
  p.scale = scale;
  p.inv_qblk = 1.0f / (float)p.n_qblk;
  p.inv_heads = 1.0f / (float)heads;
  p.lse = lse;
  const int grid = p.n_items < num_sms() ? p.n_items : num_sms();
  kern<<<grid, kP2Threads, kP2Smem, st>>>(tmQ, tmK, tmV, tmO, p);
  TVI_LAUNCH_OK("attn_spatial_tc_kernel");
  return TVI_OK;
}

#ifdef TVI_TC_TIMELINE
extern "C" int tvi_debug_tc_timeline(long long* dev_buf) {
  return cudaMemcpyToSymbol(g_tc_timeline, &dev_buf, sizeof(dev_buf)) == cudaSuccess ? 0 : -1;
}
#endif

}  // namespace tvi
