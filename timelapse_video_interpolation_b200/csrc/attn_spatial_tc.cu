// A3a — spatial self-attention core on the 5th-generation tensor cores (tcgen05 / TMEM / TMA).
// O = softmax(scale . Q K^T) V over `batch` independent sequences (frames) of S query rows and
// Skv key rows, every head, d = 64, no mask, non-causal.  Replaces
// F.scaled_dot_product_attention / xformers.ops.memory_efficient_attention on [BF, h, S, 64]
// (src/models/attention_processor.py:1243, :1156; caller src/models/attention.py:276-281).
//
// One CTA owns 256 query rows (two 128-row tiles) of one (frame, head) and streams the keys in
// tiles of 128.  320 threads, warp-specialised:
//   warp 0        TMA producer: the two Q tiles once, then K_j / V_j into a 4-stage smem ring
//                 (3-D tensor maps [heads*64, S, batch] so sequence tails are zero-filled / clipped).
//   warp 1        tcgen05.mma issuer (one lane):  S_t = Q_t K_j^T   (SS, M=128 N=128 K=64)
//                                                O_t += P_t V_j    (TS: P read from TMEM, V_j is the
//                                                                   MN-major smem operand, N=64)
//                 ordered  PV0_j, QK0_{j+1}, PV1_j, QK1_{j+1}  so that one tile's MMAs run under the
//                 other tile's softmax.
//   warps 2..5    softmax of tile 0, warps 6..9 softmax of tile 1: one thread per query row
//                 (TMEM lane), S read with tcgen05.ld, fp32 online softmax with FMNMX3 / FFMA2 /
//                 ex2, P written back to TMEM as bf16 with tcgen05.st.  The running maximum is only
//                 advanced when it grows by more than 2^8 (P stays <= 256, exact in the final
//                 O / l), so the O accumulator in TMEM is rescaled rarely, by the same threads.
//   epilogue      O_t / l -> bf16 -> swizzled smem (the dead Q tile) -> TMA store, one 32-row box per warp.
// TMEM columns: S0 [0,128) S1 [128,256) O0 [256,320) O1 [320,384) P0 [384,448) P1 [448,512).
// FLOPs: 4 * batch * heads * S * Skv * 64.  With d = 64 the exp throughput (16 / clk / SM), not the
// tensor pipe, is the per-SM ceiling: 128x128 exps per 2 x 256 MMA cycles.
#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kTcThreads = 320;
constexpr int kTcBM = 128;       // query rows per tile (UMMA M)
constexpr int kTcBN = 128;       // keys per stage (UMMA N of QK^T, K of PV)
constexpr int kTcStages = 4;
constexpr int kTcTileBytes = 128 * 128;  // 128 rows x 64 bf16
constexpr int kTcSmem = 1024 + 2 * kTcTileBytes + kTcStages * 2 * kTcTileBytes;

constexpr uint32_t kColS = 0, kColO = 256, kColP = 384;

struct TcParams {
  int S, Skv, heads;
  int n_kv;       // key tiles
  float scale;
  float* lse;     // nullable, [batch*S, heads]
};

#include "attn_spatial_tc_softmax.cuh"

__global__ void __launch_bounds__(kTcThreads, 1)
attn_spatial_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO,
                       const TcParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_q[2];
  __shared__ __align__(8) uint64_t bar_k_full[kTcStages];
  __shared__ __align__(8) uint64_t bar_v_full[kTcStages];
  __shared__ __align__(8) uint64_t bar_kv_empty[kTcStages];
  __shared__ __align__(8) uint64_t bar_s_full[2];
  __shared__ __align__(8) uint64_t bar_p_full[2];
  __shared__ __align__(8) uint64_t bar_o_full[2];
  __shared__ uint32_t s_tmem_base;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  // [Q0 | Q1 | (K,V) x stages]
  const uint32_t sQ = base;
  const uint32_t sKV = base + 2 * kTcTileBytes;

  const int bh = blockIdx.y;
  const int b = bh / p.heads, h = bh - b * p.heads;
  const int q0 = blockIdx.x * 2 * kTcBM;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
    for (int t = 0; t < 2; ++t) {
      mbar_init(&bar_q[t], 1);
      mbar_init(&bar_s_full[t], 1);
      mbar_init(&bar_p_full[t], 4);
      mbar_init(&bar_o_full[t], 1);
    }
    for (int s = 0; s < kTcStages; ++s) {
      mbar_init(&bar_k_full[s], 1);
      mbar_init(&bar_v_full[s], 1);
      mbar_init(&bar_kv_empty[s], 1);
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
      for (int t = 0; t < 2; ++t) {
        mbar_arrive_expect_tx(&bar_q[t], kTcTileBytes);
        tma_load_3d(base_ptr + t * kTcTileBytes, &tmQ, &bar_q[t], h * 64, q0 + t * kTcBM, b);
      }
      int s = 0;
      uint32_t ph = 0;
      for (int j = 0; j < p.n_kv; ++j) {
        mbar_wait(&bar_kv_empty[s], ph ^ 1);
        uint8_t* kdst = base_ptr + 2 * kTcTileBytes + s * 2 * kTcTileBytes;
        mbar_arrive_expect_tx(&bar_k_full[s], kTcTileBytes);
        tma_load_3d(kdst, &tmK, &bar_k_full[s], h * 64, j * kTcBN, b);
        mbar_arrive_expect_tx(&bar_v_full[s], kTcTileBytes);
        tma_load_3d(kdst + kTcTileBytes, &tmV, &bar_v_full[s], h * 64, j * kTcBN, b);
        if (++s == kTcStages) {
          s = 0;
          ph ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc_qk = umma_idesc_bf16(kTcBM, kTcBN, 0, 0);
      const uint32_t idesc_pv = umma_idesc_bf16(kTcBM, 64, 0, 1);  // B = V is MN-major ([keys][d])
      auto issue_qk = [&](int t, int s) {
        const uint32_t a_addr = sQ + t * kTcTileBytes;
        const uint32_t b_addr = sKV + s * 2 * kTcTileBytes;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t ad = umma_smem_desc(a_addr + kk * 32, 16, 1024, kUmmaLayoutSw128);
          const uint64_t bd = umma_smem_desc(b_addr + kk * 32, 16, 1024, kUmmaLayoutSw128);
          umma_bf16_ss(tmem_base + kColS + t * 128, ad, bd, idesc_qk, kk != 0 ? 1u : 0u);
        }
        umma_commit(&bar_s_full[t]);
      };
      auto issue_pv = [&](int t, int s, int j) {
        const uint32_t v_addr = sKV + s * 2 * kTcTileBytes + kTcTileBytes;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          // 16 keys per MMA = two 8-row groups of 128-byte rows (SBO = 1024 B); one 64-wide N atom
          const uint64_t bd = umma_smem_desc(v_addr + kk * 2048, 1024, 1024, kUmmaLayoutSw128);
          umma_bf16_ts(tmem_base + kColO + t * 64, tmem_base + kColP + t * 64 + kk * 8, bd, idesc_pv,
                       (j | kk) != 0 ? 1u : 0u);
        }
        umma_commit(&bar_o_full[t]);
      };
      mbar_wait(&bar_k_full[0], 0);
      mbar_wait(&bar_q[0], 0);
      tc_fence_after();
      issue_qk(0, 0);
      mbar_wait(&bar_q[1], 0);
      tc_fence_after();
      issue_qk(1, 0);
      int s = 0;
      uint32_t ph = 0;
      for (int j = 0; j < p.n_kv; ++j) {
        int sn = s + 1;
        uint32_t phn = ph;
        if (sn == kTcStages) {
          sn = 0;
          phn ^= 1;
        }
        const bool more = (j + 1) < p.n_kv;
        mbar_wait(&bar_v_full[s], ph);
        mbar_wait(&bar_p_full[0], j & 1);
        tc_fence_after();
        issue_pv(0, s, j);
        if (more) {
          mbar_wait(&bar_k_full[sn], phn);
          tc_fence_after();
          issue_qk(0, sn);
        }
        mbar_wait(&bar_p_full[1], j & 1);
        tc_fence_after();
        issue_pv(1, s, j);
        umma_commit(&bar_kv_empty[s]);  // K_j and V_j fully consumed by both tiles
        if (more) issue_qk(1, sn);
        s = sn;
        ph = phn;
      }
    }
  } else {
    // ===================== softmax + epilogue (warps 2..9) =====================
    const int t = (warp - 2) >> 2;
    tc_softmax_role(p, tmem_base, t, warp, lane, b, h, q0 + t * kTcBM, base_ptr + t * kTcTileBytes, &bar_s_full[t],
                    &bar_p_full[t], &bar_o_full[t], &tmO);
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
