// A3a — spatial self-attention core on the 5th-generation tensor cores (tcgen05 / TMEM / TMA).
// O = softmax(scale . Q K^T) V over `batch` independent sequences (frames) of S query rows and
// Skv key rows, every head, d = 64, no mask, non-causal.  Replaces
// F.scaled_dot_product_attention / xformers.ops.memory_efficient_attention on [BF, h, S, 64]
// (src/models/attention_processor.py:1243, :1156; caller src/models/attention.py:276-281).
//
// One CTA owns 256 query rows (two 128-row tiles) of one (frame, head) and streams the keys in
// steps of 64.  608 threads, warp-specialised:
//   warp 0        TMA producer: K_j / V_j ([64 keys x 64] each) into an 8-stage smem ring
//                 (3-D tensor maps [heads*64, Skv, batch]: sequence tails are zero-filled).
//   warps 1, 2    tcgen05.mma issuers, one per tile (the whole warp walks the loop so descriptors stay
//                 in uniform registers; one elected lane's instructions are enabled).  Both products
//                 take their A operand from tensor memory, so shared memory only feeds the 64-wide B
//                 operands:
//                     S_t[j&1] = Q_t K_j^T      (M=128 N=64 K=64, A = Q_t in TMEM, B = K_j K-major)
//                     O_t     += P_t[j&1] V_j   (M=128 N=64 K=64, A = P in TMEM, B = V_j MN-major)
//                 issued per tile as  QK_j PV_{j-1}: the scores of step j are produced while the
//                 softmax warps are still working on step j-1 (S is double-buffered per tile), and
//                 the two tiles only meet at the K/V ring (a stage is freed by both tiles' commits).
//   warps 3..10   softmax of tile 0, warps 11..18 softmax of tile 1: TWO threads per query row
//                 (TMEM lane), 32 of the step's 64 key columns each, so four softmax warps share every
//                 SM sub-partition.  The threads first park their Q half-rows in TMEM, then per step read
//                 their 32 scores with tcgen05.ld, exchange the row maximum through shared memory, do
//                 the fp32 online softmax (FMNMX3 / FFMA2 / FADD2; ex2 on the MUFU unit for 3 of 4
//                 score pairs, a cubic polynomial on the FMA pipe for the rest) and write P as bf16
//                 over the first half of the same S buffer with tcgen05.st.
//                 The reference maximum only moves when the row maximum outgrows it by 2^8 (P stays
//                 <= 256, exact in O / l), so the O accumulator in TMEM is rescaled rarely, by the
//                 same threads.
//   epilogue      O_t / l -> bf16 -> swizzled smem -> TMA store, one 32-row box per warp (clips S tails).
// TMEM columns of tile t (base 256 t): S[0] [0,64) S[1] [64,128) O [128,192) Q [192,224).
// FLOPs: 4 * batch * heads * S * Skv * 64.  With d = 64 the exp throughput (16 / clk / SM), not the
// tensor pipe, is the per-SM ceiling: 2 x 128 x 64 exps against 4 x 128 MMA cycles per step.
#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kTcThreads = 608;
constexpr int kTcBM = 128;      // query rows per tile (UMMA M)
constexpr int kTcBN = 64;       // keys per step (UMMA N of QK^T, K of PV)
constexpr int kTcStages = 8;
constexpr int kTcKBytes = kTcBN * 128;      // [64 keys][64 bf16]
constexpr int kTcStageBytes = 2 * kTcKBytes;
constexpr int kTcOutBytes = kTcBM * 128;    // output staging of one tile
constexpr int kTcSmem = 1024 + 2 * kTcOutBytes + kTcStages * kTcStageBytes;

constexpr uint32_t kColTile = 256, kColS = 0, kColO = 128, kColQ = 192;

struct TcParams {
  const tvi_bf16* q;
  int64_t ldq;
  int S, Skv, heads;
  int n_kv;       // key steps
  float scale;
  float* lse;     // nullable, [batch*S, heads]
};

struct TcBarriers {
  uint64_t q_full[2];
  uint64_t kv_full[kTcStages];
  uint64_t kv_empty[kTcStages];
  uint64_t s_full[2][2];
  uint64_t p_full[2][2];
  uint64_t o_full[2];
  uint64_t o_done[2];  // completes once per tile, after the last PV product
};

#include "attn_spatial_tc_softmax.cuh"

template <uint32_t kPolyMask>
__global__ void __launch_bounds__(kTcThreads, 1)
attn_spatial_tc_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                       const __grid_constant__ CUtensorMap tmO, const TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) TcBarriers bars;
  __shared__ uint32_t s_tmem_base;
  __shared__ float s_xchg[2][4 * 256];  // per tile: {row sums, row maxima} x [step parity][half][row] exchange slots

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  // [O staging tile 0 | O staging tile 1 | (K,V) x stages]
  const uint32_t sKV = base + 2 * kTcOutBytes;

  const int bh = blockIdx.y;
  const int b = bh / p.heads, h = bh - b * p.heads;
  const int q0 = blockIdx.x * 2 * kTcBM;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
    for (int t = 0; t < 2; ++t) {
      mbar_init(&bars.q_full[t], 8);
      mbar_init(&bars.o_full[t], 1);
      mbar_init(&bars.o_done[t], 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&bars.s_full[t][i], 1);
        mbar_init(&bars.p_full[t][i], 8);
      }
    }
    for (int s = 0; s < kTcStages; ++s) {
      mbar_init(&bars.kv_full[s], 1);
      mbar_init(&bars.kv_empty[s], 2);
    }
    mbar_fence_init();
  }
  if (warp == 1) {
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
      int s = 0;
      uint32_t ph = 0;
      for (int j = 0; j < p.n_kv; ++j) {
        TVI_WAIT(&bars.kv_empty[s], ph ^ 1, 1, j);
        uint8_t* kdst = base_ptr + 2 * kTcOutBytes + s * kTcStageBytes;
        mbar_arrive_expect_tx(&bars.kv_full[s], kTcStageBytes);
        tma_load_3d(kdst, &tmK, &bars.kv_full[s], h * 64, j * kTcBN, b);
        tma_load_3d(kdst + kTcKBytes, &tmV, &bars.kv_full[s], h * 64, j * kTcBN, b);
        if (++s == kTcStages) {
          s = 0;
          ph ^= 1;
        }
      }
    }
  } else if (warp == 1 || warp == 2) {
    // ===================== MMA issuers =====================
    // The whole warp walks the loop (convergent control flow keeps every descriptor in uniform
    // registers); only the elected lane's tcgen05 instructions are enabled.
    const uint32_t leader = elect_one() ? 1u : 0u;
    // 512 columns = the whole tensor memory of the SM, so the allocation can only start at column 0;
    // using the literal keeps the TMEM operands compile-time uniform.
    if (tmem_base != 0u) __trap();  // (both issuing warps check)
    const uint32_t idesc_qk = umma_idesc_bf16(kTcBM, kTcBN, 0, 0);
    const uint32_t idesc_pv = umma_idesc_bf16(kTcBM, 64, 0, 1);  // B = V is MN-major ([keys][d])
    auto issue_qk = [&](int t, int s, int j) {
      const uint32_t k_addr = sKV + s * kTcStageBytes;
      const uint32_t tile = t * kColTile;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const uint64_t bd = umma_smem_desc(k_addr + kk * 32, 16, 1024, kUmmaLayoutSw128);
        umma_bf16_ts_pred(tile + kColS + (j & 1) * 64, tile + kColQ + kk * 8, bd, idesc_qk, kk != 0 ? 1u : 0u, leader);
      }
      umma_commit_pred(&bars.s_full[t][j & 1], leader);
    };
    auto issue_pv = [&](int t, int s, int j) {
      const uint32_t v_addr = sKV + s * kTcStageBytes + kTcKBytes;
      const uint32_t tile = t * kColTile;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        // 16 keys per MMA = two 8-row groups of 128-byte rows (SBO = 1024 B); one 64-wide N atom
        const uint64_t bd = umma_smem_desc(v_addr + kk * 2048, 1024, 1024, kUmmaLayoutSw128);
        umma_bf16_ts_pred(tile + kColO, tile + kColS + (j & 1) * 64 + kk * 8, bd, idesc_pv, (j | kk) != 0 ? 1u : 0u,
                          leader);
      }
      umma_commit_pred(&bars.o_full[t], leader);
    };
    // One issuing warp per tile (warp 1 -> tile 0, warp 2 -> tile 1), each blocking only on its own
    // tile's barriers: neither tile's softmax is gated by the other tile's progress until the 8-stage
    // K/V ring runs out, so their exp phases can drift apart instead of colliding on the MUFU unit.
    const int t = (warp == 1) ? 0 : 1;
    TVI_WAIT(&bars.q_full[t], 0, 2, 0);
    tc_fence_after();
    int s = 0, sp = 0;  // stage of step j, stage of step j-1
    uint32_t ph = 0;
    for (int j = 0; j <= p.n_kv; ++j) {
      if (j < p.n_kv) {
        TVI_WAIT(&bars.kv_full[s], ph, 3, j);
        tc_fence_after();
        // S[j&1] last held P_{j-2}, read by PV_t(j-2): issued earlier by this thread, MMAs complete in order
        issue_qk(t, s, j);
      }
      if (j >= 1) {
        const int jj = j - 1;
        TVI_WAIT(&bars.p_full[t][jj & 1], (uint32_t)(jj >> 1) & 1u, 4, jj);
        tc_fence_after();
        issue_pv(t, sp, jj);
        umma_commit_pred(&bars.kv_empty[sp], leader);  // this tile is done with K_jj / V_jj (2 arrivals free the stage)
        // the epilogue waits on a barrier of its own: a parity wait on o_full is only sound for a waiter
        // that is at most one phase behind, and a softmax warp can reach the epilogue after BOTH of the
        // last two PV products have completed
        if (jj == p.n_kv - 1) umma_commit_pred(&bars.o_done[t], leader);
      }
      sp = s;
      if (++s == kTcStages) {
        s = 0;
        ph ^= 1;
      }
    }
    __syncwarp();
  } else {
    // ===================== softmax + epilogue (warps 3..18) =====================
    const int sw = warp - 3;
    const int t = sw >> 3, half = (sw >> 2) & 1;
    const int quad = warp & 3;  // the TMEM lane quadrant a warp may touch is fixed by its index
    tc_softmax_role<kPolyMask>(p, tmem_base + t * kColTile, quad, half, 1u + t * 4 + quad, s_xchg[t], lane, b, h,
                               q0 + t * kTcBM, base_ptr + t * kTcOutBytes, &bars.q_full[t], bars.s_full[t], bars.p_full[t],
                               &bars.o_full[t], &bars.o_done[t], &tmO);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace
}  // namespace tvi

#include "attn_spatial_tc_host.inc"
