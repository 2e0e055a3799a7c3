// A2/A3 — the attention projections: Y[M,N] = X[M,K] . W[N,K]^T (+ bias), bf16 in/out, fp32 accumulate.
// Replaces attn.to_q / to_k / to_v / to_out[0] (nn.Linear -> cuBLAS) of
// src/models/attention_processor.py:729,736-737,754 / :1223,1230-1231,1251.
//
// Persistent, warp-specialised tcgen05 kernel, one CTA per SM (192 threads):
//   warp 0      TMA producer: X tile [128 x 64] and W tile [BN x 64] per k-block, SWIZZLE_128B,
//               `stages`-deep mbarrier ring (full/empty).
//   warp 1      TMEM owner + MMA issuer: one lane issues tcgen05.mma (M=128, N=BN, K=16) four
//               times per k-block and commits the stage back to the producer; accumulators are
//               double-buffered in TMEM (2 x 256 columns) so the epilogue of tile i overlaps the
//               MMAs of tile i+1.
//   warps 2..5  epilogue: tcgen05.ld 32 lanes x 32 columns, + bias, bf16 pack, SWIZZLE_64B
//               staging in shared memory, TMA store (which also clips the M / N tails).
// Tiles are walked N-fastest so the CTAs resident at one time share X rows through L2.
// At the SVD shapes K is only 320..1280, so per tile the 128 x K read and 128 x BN write dominate:
// the projections sit at or below the HBM/tensor ridge and one-CTA MMA is enough.
//
// The same kernel, with MN-major operand descriptors, is the backward of nn.Linear (training
// config 5, what autograd derives from F.linear at the same call sites):
//   kBMn          dX[M,K] = dY[M,N] . W[N,K]      W is consumed in place as an MN-major B operand
//                 (TMA boxes of [64 reduction rows][64 contiguous columns], SWIZZLE_128B).
//   kAMn + kBMn   dW[N,K] = dY[M,N]^T . X[M,K]    both operands MN-major, the token reduction is split
//   + kSplit      across CTAs, fp32 partial tiles go to a workspace and wgrad_reduce_kernel sums them
//                 in a fixed order (deterministic) into bf16.
#include <stdlib.h>

#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kBM = 128;
constexpr int kBK = 64;
#ifndef TVI_GEMM_EPI_WARPS
#define TVI_GEMM_EPI_WARPS 8
#endif
constexpr int kEpiWarps = TVI_GEMM_EPI_WARPS;                        // two per TMEM lane quadrant, alternating 32-column chunks (12 / 16 warps squeeze the chain into 128 / 96 registers and are slower: profiles/r2_gemm_epi_warps.txt)
constexpr int kGemmThreads = 64 + 32 * kEpiWarps;
constexpr int kAccStride = 256;                    // TMEM columns per accumulator stage
constexpr int kABytes = kBM * kBK * 2;             // 16 KB
// One staging buffer per warp: a lone warp per scheduler cannot hide its own tcgen05.ld -> FADD -> F2FP -> STS -> TMA-store
// chain (~640 clk per 32-column chunk measured, linear in BN: the epilogue, not the MMAs, set the tile time), so the
// latency is hidden ACROSS warps instead: 8 warps x 1 buffer use the shared memory of 4 warps x 2 buffers and took the
// N = 2560 / K = 320 GEMM from 0.52 to 0.45 ms, out-projection + residual from 0.175 to 0.105 ms, QKV unchanged (0.158).
#ifndef TVI_GEMM_STAGE_BUFS
#define TVI_GEMM_STAGE_BUFS 1
#endif
constexpr int kStageBufs = TVI_GEMM_STAGE_BUFS;
#ifndef TVI_GEMM_RES_BUFS
#define TVI_GEMM_RES_BUFS 4
#endif
constexpr int kResStageBufs = TVI_GEMM_RES_BUFS;  // with a residual: the buffers also receive the residual chunk ahead of its use
constexpr int kResAhead = kResStageBufs >= 4 ? 2 : 1;  // residual chunks requested ahead of the one being processed
// warps x bufs x [32 rows][64 B] + warps x 256 fp32 bias
constexpr int kBiasBytes = 4 * 256 * 4;            // one fp32 bias strip (+ one LayerNorm column-sum strip) per accumulator stage, shared by the epilogue warps
constexpr int staging_bytes(int bufs) { return kEpiWarps * bufs * 2048 + kBiasBytes; }
constexpr int kMaxStages = 8;

// gelu(g) = g * Phi(g), Phi(g) = 0.5 * (1 + erf(g / sqrt 2)) (the erf form F.gelu computes).  Branch-free:
// E = 0.5 * erfc(|g| / sqrt 2) = exp2(-1 + u * P(u)), u = min(|g|, 6), P a degree-5 minimax fit of
// (log2(0.5 erfc(u / sqrt 2)) + 1) / u on [0, 6] (|log2 error| <= 7.1e-5); Phi = g < 0 ? E : 1 - E.  No cancellation on
// the negative side; |gelu error| <= 7.4e-6 absolute and 4.9e-5 relative over the whole line -- 80x below the bf16
// rounding of the stored product.  6 FMA + 1 MUFU.EX2 instead of erff's two divergent polynomial branches.
// `geglu_x2` evaluates it on a PAIR of gates with packed fp32 arithmetic (FFMA2 / FADD2 / FMUL2): v * g * Phi(g) for two columns.
// (the six polynomial FMAs, the bias adds and the products cost one issue slot per pair: 0.448 -> 0.436 ms for ff_in at L0;
// the epilogue chain is latency-bound, profiles/r2_gemm_ffin.md).
// Phi = 0.5 + copysign(0.5 - E, g) (E <= 0.5) replaces the compare + select: one LOP3 per element.
__device__ __forceinline__ uint64_t geglu_x2(uint64_t v2, uint64_t g2) {
  float g0, g1;
  f32x2_unpack(g2, g0, g1);
  const uint64_t u2 = f32x2_pack(fminf(fabsf(g0), 6.0f), fminf(fabsf(g1), 6.0f));
  uint64_t q = f32x2_fma(f32x2_pack(2.3433244e-05f, 2.3433244e-05f), u2, f32x2_pack(-0.00061973924f, -0.00061973924f));
  q = f32x2_fma(q, u2, f32x2_pack(0.007260352f, 0.007260352f));
  q = f32x2_fma(q, u2, f32x2_pack(-0.051418606f, -0.051418606f));
  q = f32x2_fma(q, u2, f32x2_pack(-0.46086353f, -0.46086353f));
  q = f32x2_fma(q, u2, f32x2_pack(-1.1504804f, -1.1504804f));
  float t0, t1;
  f32x2_unpack(f32x2_fma(q, u2, f32x2_pack(-1.0f, -1.0f)), t0, t1);
  const uint64_t h2 = f32x2_fma(f32x2_pack(fast_exp2(t0), fast_exp2(t1)), f32x2_pack(-1.0f, -1.0f), f32x2_pack(0.5f, 0.5f));
  float h0, h1;
  f32x2_unpack(h2, h0, h1);
  const float s0 = __uint_as_float(__float_as_uint(h0) | (__float_as_uint(g0) & 0x80000000u));
  const float s1 = __uint_as_float(__float_as_uint(h1) | (__float_as_uint(g1) & 0x80000000u));
  const uint64_t phi2 = f32x2_add(f32x2_pack(s0, s1), f32x2_pack(0.5f, 0.5f));
  return f32x2_mul(f32x2_mul(g2, phi2), v2);
}

// Backward of the same product on a pair of columns: h = v * gelu(g), d(value) = dh * gelu(g),
// d(gate) = dh * v * (Phi(g) + g * phi(g)), phi(g) = exp(-g^2 / 2) / sqrt(2 pi).  Phi as above; one more MUFU.EX2 for phi.
__device__ __forceinline__ void geglu_bwd_x2(uint64_t v2, uint64_t g2, uint64_t d2, uint64_t& h2o, uint64_t& dv2o, uint64_t& dg2o) {
  float g0, g1;
  f32x2_unpack(g2, g0, g1);
  const uint64_t u2 = f32x2_pack(fminf(fabsf(g0), 6.0f), fminf(fabsf(g1), 6.0f));
  uint64_t q = f32x2_fma(f32x2_pack(2.3433244e-05f, 2.3433244e-05f), u2, f32x2_pack(-0.00061973924f, -0.00061973924f));
  q = f32x2_fma(q, u2, f32x2_pack(0.007260352f, 0.007260352f));
  q = f32x2_fma(q, u2, f32x2_pack(-0.051418606f, -0.051418606f));
  q = f32x2_fma(q, u2, f32x2_pack(-0.46086353f, -0.46086353f));
  q = f32x2_fma(q, u2, f32x2_pack(-1.1504804f, -1.1504804f));
  float t0, t1;
  f32x2_unpack(f32x2_fma(q, u2, f32x2_pack(-1.0f, -1.0f)), t0, t1);
  const uint64_t hm2 = f32x2_fma(f32x2_pack(fast_exp2(t0), fast_exp2(t1)), f32x2_pack(-1.0f, -1.0f), f32x2_pack(0.5f, 0.5f));
  float h0, h1;
  f32x2_unpack(hm2, h0, h1);
  const float s0 = __uint_as_float(__float_as_uint(h0) | (__float_as_uint(g0) & 0x80000000u));
  const float s1 = __uint_as_float(__float_as_uint(h1) | (__float_as_uint(g1) & 0x80000000u));
  const uint64_t phi2 = f32x2_add(f32x2_pack(s0, s1), f32x2_pack(0.5f, 0.5f));
  float e0, e1;
  f32x2_unpack(f32x2_mul(f32x2_mul(g2, g2), f32x2_pack(-0.72134752f, -0.72134752f)), e0, e1);  // -g^2 / 2 * log2(e)
  const uint64_t pdf2 = f32x2_mul(f32x2_pack(fast_exp2(e0), fast_exp2(e1)), f32x2_pack(0.3989422804f, 0.3989422804f));
  const uint64_t gelu2 = f32x2_mul(g2, phi2);
  h2o = f32x2_mul(gelu2, v2);
  dv2o = f32x2_mul(gelu2, d2);
  dg2o = f32x2_mul(f32x2_mul(d2, v2), f32x2_fma(g2, pdf2, phi2));
}

struct GemmParams {
  int64_t m_tiles;
  int n_tiles;
  int k_blocks;        // k-blocks per tile (per split when kSplit)
  int k_blocks_total;  // k-blocks of the whole reduction
  int BN;
  int N;
  int stages;
  const tvi_bf16* bias;
  int geglu;           // epilogue: columns [0, BN/2) of a tile are values, [BN/2, BN) their gates -> value * gelu(gate)
  int N_store;         // columns of Y (N, or N/2 with the GEGLU epilogue)
  const tvi_bf16* residual;  // nullable [M, N_store] (ldr): added after bias / activation
  int64_t ldr;
  int stage_bufs;      // staging buffers per epilogue warp (kStageBufs, or kResStageBufs with a residual)
  int res_b;           // the CTA's B slice [BN x K] stays resident in shared memory (grid is a multiple of n_tiles)
  float* partial;      // kSplit: [splits][M_out][ldp] fp32
  // LayerNorm folded into the GEMM (A = the raw rows x, B = gamma (.) W): y = rstd * (acc - mean * colsum_n) + bias_n with
  // colsum_n = sum_k B[n, k] and bias_n = W_n . beta + b_n prepared by the host; ln_stats = [M][2] fp32 (mean, rstd)
  const float2* ln_stats;
  const float* ln_colsum;
  int64_t M_out;
  int64_t ldp;
  tvi_bf16* dpre;      // kGbwd: [M, 2 * N_store] = [d(value) | d(gate)], row stride lddpre
  int64_t lddpre;
  tvi_bf16* h_out;     // kGbwd: [M, N_store] = value * gelu(gate), row stride ldh
  int64_t ldh;
  int wide_stores;     // kGbwd: all three outputs are 32-byte aligned per row -> 256-bit stores
};

// kAMn / kBMn: the A / B operand is stored reduction-row-major ("MN-major": the output dimension is the
// contiguous one).  kSplit: tiles are (split, m, n); fp32 partials instead of the bf16 TMA-store epilogue.
// kPair (forward TN GEMM only): two CTAs of a cluster (one TPC) work on ONE 256 x BN tile with `tcgen05.mma.cta_group::2`:
// each CTA loads its own 128 rows of A and HALF of the B tile, the leader CTA issues the MMAs for both, every CTA runs the
// epilogue of its own 128 accumulator rows.  Per CTA and k-block the L2 -> SM operand stream is A + B/2 instead of A + B;
// with the weight slice resident (half of it per CTA: 80 KB at BN = 256, K = 320) only A streams at all.
// kGbwd (forward TN GEMM with the GEGLU tile layout only): the backward of the fused GEGLU projection.  The accumulator
// is the recomputed pre-activation [value | gate]; the "residual" stream is dh = d(loss)/d(h); the epilogue writes
// h = value * gelu(gate) and d(pre) = [d value | d gate] straight from registers (64 contiguous bytes per thread and
// output).  Neither the 2 * inner-wide pre-activation nor a second pass over dh touches HBM.
template <bool kAMn, bool kBMn, bool kSplit, bool kPair = false, bool kGbwd = false>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmY, const __grid_constant__ CUtensorMap tmR, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[kMaxStages];
  __shared__ __align__(8) uint64_t bar_empty[kMaxStages];
  __shared__ __align__(8) uint64_t bar_acc_full[2];
  __shared__ __align__(8) uint64_t bar_acc_empty[2];
  __shared__ __align__(8) uint64_t bar_w;
  __shared__ __align__(8) uint64_t bar_res[kEpiWarps][kResStageBufs];
  __shared__ uint32_t s_tmem_base;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // 1024-byte aligned carve-up (SWIZZLE_128B atoms are 1024 B)
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  static_assert(!kPair || (!kAMn && !kBMn && !kSplit), "the CTA-pair kernel is the forward (TN) GEMM");
  static_assert(!kGbwd || (!kAMn && !kBMn && !kSplit && !kPair), "the GEGLU-backward epilogue belongs to the forward (TN) GEMM");
  const uint32_t cta_rank = kPair ? cluster_ctarank() : 0u;
  const int64_t worker = kPair ? (int64_t)(blockIdx.x >> 1) : (int64_t)blockIdx.x;      // tile walker: a CTA, or a CTA pair
  const int64_t nworkers = kPair ? (int64_t)(gridDim.x >> 1) : (int64_t)gridDim.x;
  constexpr int kMStep = kPair ? 2 * kBM : kBM;                                          // rows of a (pair) tile
  const int bn_local = kPair ? p.BN / 2 : p.BN;                                          // B rows held by this CTA
  const uint32_t b_bytes = (uint32_t)bn_local * kBK * 2;
  // Resident-B mode (forward projections at K = 320): with gridDim.x a multiple of n_tiles the walk
  // tile = blockIdx.x + i * gridDim.x keeps nt = blockIdx.x % n_tiles, so the CTA's [BN x K] weight slice is
  // loaded ONCE and only X streams through the ring.  Re-fetching the slice per tile made the L2 -> SM
  // traffic 2.5x the X traffic (123 KB of W per 82 KB of X at N = 960) and capped the kernel at 3.8 TB/s.
  const bool res_b = p.res_b != 0;
  const uint32_t stage_bytes = res_b ? (uint32_t)kABytes : kABytes + b_bytes;
  const uint32_t res_bytes = res_b ? (uint32_t)p.k_blocks_total * b_bytes : 0u;
  const uint32_t wres = base + p.stages * stage_bytes;
  uint8_t* wres_ptr = base_ptr + p.stages * stage_bytes;
  const uint32_t staging = wres + res_bytes;
  uint8_t* staging_ptr = wres_ptr + res_bytes;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmY);
    tma_prefetch_desc(&tmR);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&bar_full[s], 1);
      mbar_init(&bar_empty[s], 1);
    }
    mbar_init(&bar_w, 1);
    for (int w = 0; w < kEpiWarps; ++w)
      for (int b = 0; b < kResStageBufs; ++b) mbar_init(&bar_res[w][b], 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bar_acc_full[s], 1);
      mbar_init(&bar_acc_empty[s], kPair ? 2 * kEpiWarps : kEpiWarps);  // pair: both CTAs' epilogue warps release the leader
    }
    mbar_fence_init();
  }
  if (warp == 1) {
    if (kPair) {
      tmem_alloc_pair(&s_tmem_base, 512);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(&s_tmem_base, 512);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (kPair) cluster_sync_all();  // the peer's barriers are initialised before any remote arrive / complete_tx
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;

  const int64_t mn_tiles = p.m_tiles * p.n_tiles;
  const int64_t n_tiles_total = kSplit ? mn_tiles * ((p.k_blocks_total + p.k_blocks - 1) / p.k_blocks) : mn_tiles;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      // pair: every load is counted on the LEADER's barrier (which expects both CTAs' bytes); each CTA waits on its own
      // `empty` barrier, released for both by the leader's multicast commit
      const uint32_t bar_w_leader = kPair ? mapa_u32(&bar_w, 0) : 0u;
      if (res_b && worker < n_tiles_total) {
        const int n0 = (int)(worker % p.n_tiles) * p.BN + (int)cta_rank * bn_local;
        if (kPair) {
          if (cta_rank == 0) mbar_arrive_expect_tx(&bar_w, 2 * res_bytes);
          for (int kb = 0; kb < p.k_blocks_total; ++kb) tma_load_2d_pair(wres_ptr + kb * b_bytes, &tmW, bar_w_leader, kb * kBK, n0);
        } else {
          mbar_arrive_expect_tx(&bar_w, res_bytes);
          for (int kb = 0; kb < p.k_blocks_total; ++kb) tma_load_2d(wres_ptr + kb * b_bytes, &tmW, &bar_w, kb * kBK, n0);
        }
      }
      for (int64_t tile = worker; tile < n_tiles_total; tile += nworkers) {
        const int64_t split = kSplit ? tile / mn_tiles : 0;
        const int64_t mn = tile - split * mn_tiles;
        const int64_t mt = mn / p.n_tiles;
        const int nt = (int)(mn - mt * p.n_tiles);
        const int m0 = (int)(mt * kMStep) + (int)cta_rank * kBM, n0 = nt * p.BN + (int)cta_rank * (kPair ? bn_local : 0);
        const int kb_lo = (int)split * p.k_blocks;
        const int kb_hi = min(kb_lo + p.k_blocks, p.k_blocks_total);
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(&bar_empty[s], ph ^ 1);
          uint8_t* a_dst = base_ptr + s * stage_bytes;
          if (kPair) {
            if (cta_rank == 0) mbar_arrive_expect_tx(&bar_full[s], 2 * stage_bytes);
            const uint32_t full_leader = mapa_u32(&bar_full[s], 0);
            tma_load_2d_pair(a_dst, &tmX, full_leader, kb * kBK, m0);
            if (!res_b) tma_load_2d_pair(a_dst + kABytes, &tmW, full_leader, kb * kBK, n0);
            if (++s == p.stages) {
              s = 0;
              ph ^= 1;
            }
            continue;
          }
          mbar_arrive_expect_tx(&bar_full[s], stage_bytes);
          if (!kAMn) {
            tma_load_2d(a_dst, &tmX, &bar_full[s], kb * kBK, m0);
          } else {
            // [64 reduction rows][64 output columns] boxes, one per 64-wide MN atom
            tma_load_2d(a_dst, &tmX, &bar_full[s], m0, kb * kBK);
            tma_load_2d(a_dst + 8192, &tmX, &bar_full[s], m0 + 64, kb * kBK);
          }
          if (res_b) {
            // B is resident
          } else if (!kBMn) {
            tma_load_2d(a_dst + kABytes, &tmW, &bar_full[s], kb * kBK, n0);
          } else {
            for (int hb = 0; hb < p.BN / 64; ++hb)
              tma_load_2d(a_dst + kABytes + hb * 8192, &tmW, &bar_full[s], n0 + hb * 64, kb * kBK);
          }
          if (++s == p.stages) {
            s = 0;
            ph ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp walks the loop (convergent control flow keeps the descriptors in uniform
    // registers); only the elected lane's tcgen05 instructions are enabled.  Issuing from inside
    // `if (lane == 0)` makes nvcc wrap every MMA in an ELECT / BRA.U.ANY loop (~75 cycles each).
    const uint32_t leader = (elect_one() && cta_rank == 0) ? 1u : 0u;  // pair: only the leader CTA issues
    const uint32_t idesc = umma_idesc_bf16(kMStep, (uint32_t)p.BN, kAMn ? 1 : 0, kBMn ? 1 : 0);
    int s = 0;
    uint32_t ph = 0;
    int as = 0;
    uint32_t aph = 0;
    if (res_b && worker < n_tiles_total && cta_rank == 0) {
      mbar_wait(&bar_w, 0);
      tc_fence_after();
    }
    for (int64_t tile = worker; tile < n_tiles_total && cta_rank == 0; tile += nworkers) {
      mbar_wait(&bar_acc_empty[as], aph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * kAccStride;
      const int64_t split = kSplit ? tile / mn_tiles : 0;
      const int kb_lo = (int)split * p.k_blocks;
      const int kb_n = min(kb_lo + p.k_blocks, p.k_blocks_total) - kb_lo;
      for (int kb = 0; kb < kb_n; ++kb) {
        mbar_wait(&bar_full[s], ph);
        tc_fence_after();
        const uint32_t a_addr = base + s * stage_bytes;
        const uint32_t b_addr = res_b ? wres + (uint32_t)(kb_lo + kb) * b_bytes : a_addr + kABytes;
#pragma unroll
        for (int kk = 0; kk < kBK / 16; ++kk) {
          // K-major: 16 reduction elements = 32 bytes along the swizzled 128-byte row.
          // MN-major: 16 reduction rows = two 8-row groups of 1024 B; 64-wide MN atoms are 8192 B apart.
          const uint64_t ad = kAMn ? umma_smem_desc(a_addr + kk * 2048, 8192, 1024, kUmmaLayoutSw128)
                                   : umma_smem_desc(a_addr + kk * 32, 16, 1024, kUmmaLayoutSw128);
          const uint64_t bd = kBMn ? umma_smem_desc(b_addr + kk * 2048, 8192, 1024, kUmmaLayoutSw128)
                                   : umma_smem_desc(b_addr + kk * 32, 16, 1024, kUmmaLayoutSw128);
          if (kPair) umma_bf16_ss_pair_pred(d_tmem, ad, bd, idesc, (kb | kk) != 0 ? 1u : 0u, leader);
          else umma_bf16_ss_pred(d_tmem, ad, bd, idesc, (kb | kk) != 0 ? 1u : 0u, leader);
        }
        // frees the smem stage (in both CTAs of a pair) once these MMAs have read it
        if (kPair) umma_commit_pair_pred(&bar_empty[s], leader);
        else umma_commit_pred(&bar_empty[s], leader);
        if (++s == p.stages) {
          s = 0;
          ph ^= 1;
        }
      }
      if (kPair) umma_commit_pair_pred(&bar_acc_full[as], leader);  // accumulator complete -> both CTAs' epilogues
      else umma_commit_pred(&bar_acc_full[as], leader);             // accumulator complete -> epilogue
      if (++as == 2) {
        as = 0;
        aph ^= 1;
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue (warps 2 .. 2 + kEpiWarps) =====================
    const int quad = warp & 3;  // TMEM lane quadrant this warp may read
    const int chalf = (warp - 2) >> 2;
    constexpr int kChunkStep = kEpiWarps / 4;
    const int nbufs = p.stage_bufs;
    uint8_t* my_staging_ptr = staging_ptr + (warp - 2) * nbufs * 2048;
    float* bias_strips = reinterpret_cast<float*>(staging_ptr + kEpiWarps * nbufs * 2048);
    float* cs_strips = bias_strips + 2 * 256;
    const bool has_ln = !kSplit && p.ln_stats != nullptr;
    const int epi_tid = threadIdx.x - 64;
    int buf = 0;
    // Residual (Y = acc + bias + R): a per-chunk global read after the accumulator wait would expose one DRAM latency
    // per 32 columns (out-projection at K = 320: 0.076 -> 0.180 ms).  Instead lane 0 TMA-loads the [32 x 32] residual
    // chunk into the chunk's own staging buffer TWO chunks ahead (4 buffers per warp; one ahead with 3 buffers: 0.125 vs 0.105 ms); the thread adds its row in
    // place and the same buffer goes out through the TMA store.
    const bool has_res = !kSplit && p.residual != nullptr;
    uint64_t* my_res_bar = bar_res[warp - 2];
    const uint32_t acc_empty_leader[2] = {kPair ? mapa_u32(&bar_acc_empty[0], 0) : 0u, kPair ? mapa_u32(&bar_acc_empty[1], 0) : 0u};
    auto release_acc = [&](int stage) {  // lane 0: this warp has read everything it needs of accumulator stage `stage`
      if (kPair) mbar_arrive_cluster(acc_empty_leader[stage]);
      else mbar_arrive(&bar_acc_empty[stage]);
    };
    int64_t ptile = worker;  // prefetch cursor: (tile, chunk) of the next residual chunk to request
    int pc = chalf;
    uint32_t pq = 0, qn = 0;     // chunks requested / consumed by this warp
    auto valid_chunks = [&](int64_t tile) {  // chunks of this tile that start inside [0, N_store)
      const int nt = (int)(tile % p.n_tiles);
      const int n0s = p.geglu ? (nt * p.BN) >> 1 : nt * p.BN;
      const int oc = p.geglu ? p.BN / 64 : p.BN / 32;
      const int v = (p.N_store - n0s + 31) / 32;
      return v < oc ? v : oc;
    };
    auto issue_residual = [&]() {  // lane 0
      while (ptile < n_tiles_total && pc >= valid_chunks(ptile)) {
        ptile += nworkers;
        pc = chalf;
      }
      if (ptile >= n_tiles_total) return;
      const int64_t mt = ptile / p.n_tiles;
      const int nt = (int)(ptile - mt * p.n_tiles);
      const int n0s = p.geglu ? (nt * p.BN) >> 1 : nt * p.BN;
      const uint32_t b = pq % kResStageBufs;
      mbar_arrive_expect_tx(&my_res_bar[b], 2048);
      tma_load_2d(my_staging_ptr + b * 2048, &tmR, &my_res_bar[b], n0s + pc * 32, (int)(mt * kMStep) + (int)cta_rank * kBM + quad * 32);
      ++pq;
      pc += kChunkStep;
    };
    if (has_res && lane == 0) {
#pragma unroll
      for (int i = 0; i < kResAhead; ++i) issue_residual();
    }
    int as = 0;
    uint32_t aph = 0;
    const int chunks = p.BN / 32;
    for (int64_t tile = worker; tile < n_tiles_total; tile += nworkers) {
      const int64_t split = kSplit ? tile / mn_tiles : 0;
      const int64_t mn = tile - split * mn_tiles;
      const int64_t mt = mn / p.n_tiles;
      const int nt = (int)(mn - mt * p.n_tiles);
      const int m0 = (int)(mt * kMStep) + (int)cta_rank * kBM, n0 = nt * p.BN;
      const float* my_bias = bias_strips + as * 256;
      const float* my_cs = cs_strips + as * 256;
      float ln_r = 1.f, ln_t = 0.f;  // y = ln_r * acc + ln_t * colsum + bias, ln_t = -mean * rstd
      if (has_ln) {
        const int64_t row = (int64_t)m0 + quad * 32 + lane;
        if (row < p.M_out) {
          const float2 st = __ldg(p.ln_stats + row);
          ln_r = st.y;
          ln_t = -st.x * st.y;
        }
      }
      if (!kSplit && p.bias != nullptr) {
        // This tile's bias slice as fp32 in the strip of accumulator stage `as` (read back as broadcast LDS.128).  One
        // barrier per tile among the epilogue warps: a warp that passes it knows every warp has finished tile i - 1's
        // predecessor, so the strip of tile i + 1 (same stage as i - 1) is free once the NEXT barrier is passed.
        if (epi_tid < p.BN)
          bias_strips[as * 256 + epi_tid] =
              (n0 + epi_tid < p.N) ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(p.bias + n0 + epi_tid)) : 0.f;
        if (has_ln && epi_tid < p.BN) cs_strips[as * 256 + epi_tid] = (n0 + epi_tid < p.N) ? __ldg(p.ln_colsum + n0 + epi_tid) : 0.f;
        named_bar_sync(1, 32 * kEpiWarps);
      }
      mbar_wait(&bar_acc_full[as], aph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + as * kAccStride;
      const int out_chunks = p.geglu ? chunks / 2 : chunks;
      if (chalf >= out_chunks) {
        // A tile with fewer OUTPUT chunks than warps per quadrant (BN = 32, or BN = 64 with GEGLU) leaves this warp without
        // work; it must still release the accumulator stage.  (Testing `chunks` here hung GEGLU tiles of BN = 64: the loop
        // below runs over out_chunks.)
        tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(as);
      }
      for (int c = chalf; c < out_chunks; c += kChunkStep) {
        uint32_t r[32], rg[32];
        tmem_ld_x32(t_row + c * 32, r);
        if (p.geglu) tmem_ld_x32(t_row + (out_chunks + c) * 32, rg);
        tmem_ld_wait();
        if (c + kChunkStep >= out_chunks) {
          // every column this warp needs of the accumulator stage is now in registers
          tc_fence_before();
          __syncwarp();
          if (lane == 0) release_acc(as);
        }
        const int ncol0 = (p.geglu ? (n0 >> 1) : n0) + c * 32;
        if (ncol0 >= p.N_store) continue;  // tile column tail: nothing to store (warp-uniform)
        if (kSplit) {
          // fp32 partial tile: each thread owns 32 consecutive columns of one output row
          const int64_t row = (int64_t)m0 + quad * 32 + lane;
          if (row < p.M_out) {
            float* dst = p.partial + ((int64_t)split * p.M_out + row) * p.ldp + ncol0;
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              if (ncol0 + j < p.N)  // N % 8 == 0: a float4 never straddles the end of the row
                *reinterpret_cast<uint4*>(dst + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
          }
          continue;
        }
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(r[j]);
        if (has_ln) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(my_bias + c * 32 + j);
            const float4 s4 = *reinterpret_cast<const float4*>(my_cs + c * 32 + j);
            f[j] = fmaf(ln_r, f[j], fmaf(ln_t, s4.x, b4.x));
            f[j + 1] = fmaf(ln_r, f[j + 1], fmaf(ln_t, s4.y, b4.y));
            f[j + 2] = fmaf(ln_r, f[j + 2], fmaf(ln_t, s4.z, b4.z));
            f[j + 3] = fmaf(ln_r, f[j + 3], fmaf(ln_t, s4.w, b4.w));
          }
        } else if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const float4 b4 = *reinterpret_cast<const float4*>(my_bias + c * 32 + j);
            f32x2_unpack(f32x2_add(f32x2_pack(f[j], f[j + 1]), f32x2_pack(b4.x, b4.y)), f[j], f[j + 1]);
            f32x2_unpack(f32x2_add(f32x2_pack(f[j + 2], f[j + 3]), f32x2_pack(b4.z, b4.w)), f[j + 2], f[j + 3]);
          }
        }
        if (!kGbwd && p.geglu) {
          // GEGLU (diffusers GEGLU.forward: hidden * F.gelu(gate), erf form): gate = the partner column
          const float* gate_bias = my_bias + (out_chunks + c) * 32;
          if (has_ln) {  // the gate pre-activation gets the same affine map of its accumulator
            const float* gate_cs = my_cs + (out_chunks + c) * 32;
#pragma unroll
            for (int j = 0; j < 32; ++j) rg[j] = __float_as_uint(fmaf(ln_r, __uint_as_float(rg[j]), fmaf(ln_t, gate_cs[j], gate_bias[j])));
          }
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p.bias != nullptr && !has_ln) b4 = *reinterpret_cast<const float4*>(gate_bias + j);
            const uint64_t ga = f32x2_add(f32x2_pack(__uint_as_float(rg[j]), __uint_as_float(rg[j + 1])), f32x2_pack(b4.x, b4.y));
            const uint64_t gb = f32x2_add(f32x2_pack(__uint_as_float(rg[j + 2]), __uint_as_float(rg[j + 3])), f32x2_pack(b4.z, b4.w));
            f32x2_unpack(geglu_x2(f32x2_pack(f[j], f[j + 1]), ga), f[j], f[j + 1]);
            f32x2_unpack(geglu_x2(f32x2_pack(f[j + 2], f[j + 3]), gb), f[j + 2], f[j + 3]);
          }
        }
        uint8_t* row_ptr = my_staging_ptr + buf * 2048 + lane * 64;
        const int sw = (lane >> 1) & 3;
        if constexpr (kGbwd) {
          // dh chunk (requested kResAhead chunks ago into this chunk's staging buffer) -> h, d(value), d(gate), all three
          // stored straight from registers: 64 contiguous bytes per thread and output (two 32-byte stores when the rows
          // are 32-byte aligned, else four 16-byte ones).  The staging buffers only ever receive dh, so the epilogue has
          // no shared-memory stores, no proxy fence and no bulk-store bookkeeping.
          if (lane == 0) issue_residual();  // into the buffer read two chunks ago (every lane passed a __syncwarp since)
          mbar_wait(&my_res_bar[buf], (qn / kResStageBufs) & 1);
          ++qn;
          const float* gate_bias = my_bias + (out_chunks + c) * 32;
          const int64_t row = (int64_t)m0 + quad * 32 + lane;
          tvi_bf16* h_dst = p.h_out + row * p.ldh + ncol0;
          tvi_bf16* dv_dst = p.dpre + row * p.lddpre + ncol0;
          tvi_bf16* dg_dst = dv_dst + p.N_store;
          uint4 oh[4], ov[4], og[4];
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            const uint4 u = *reinterpret_cast<const uint4*>(row_ptr + ((ch ^ sw) << 4));
            const uint32_t wd[4] = {u.x, u.y, u.z, u.w};
            uint32_t th[4], tv[4], tg[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const int j = ch * 8 + 2 * q;
              const uint64_t d2 = f32x2_pack(__uint_as_float(wd[q] << 16), __uint_as_float(wd[q] & 0xffff0000u));
              float2 gb = make_float2(0.f, 0.f);
              if (p.bias != nullptr) gb = *reinterpret_cast<const float2*>(gate_bias + j);
              const uint64_t g2 = f32x2_add(f32x2_pack(__uint_as_float(rg[j]), __uint_as_float(rg[j + 1])), f32x2_pack(gb.x, gb.y));
              uint64_t hh, dv, dg;
              geglu_bwd_x2(f32x2_pack(f[j], f[j + 1]), g2, d2, hh, dv, dg);
              float a, b;
              f32x2_unpack(hh, a, b);
              th[q] = pack_bf16x2(a, b);
              f32x2_unpack(dv, a, b);
              tv[q] = pack_bf16x2(a, b);
              f32x2_unpack(dg, a, b);
              tg[q] = pack_bf16x2(a, b);
            }
            oh[ch] = make_uint4(th[0], th[1], th[2], th[3]);
            ov[ch] = make_uint4(tv[0], tv[1], tv[2], tv[3]);
            og[ch] = make_uint4(tg[0], tg[1], tg[2], tg[3]);
          }
          if (row < p.M_out) {
            if (p.wide_stores) {
              if (p.h_out != nullptr) {
                st_global_256(h_dst, oh[0], oh[1]);
                st_global_256(h_dst + 16, oh[2], oh[3]);
              }
              st_global_256(dv_dst, ov[0], ov[1]);
              st_global_256(dv_dst + 16, ov[2], ov[3]);
              st_global_256(dg_dst, og[0], og[1]);
              st_global_256(dg_dst + 16, og[2], og[3]);
            } else {
#pragma unroll
              for (int ch = 0; ch < 4; ++ch) {
                if (p.h_out != nullptr) *reinterpret_cast<uint4*>(h_dst + ch * 8) = oh[ch];
                *reinterpret_cast<uint4*>(dv_dst + ch * 8) = ov[ch];
                *reinterpret_cast<uint4*>(dg_dst + ch * 8) = og[ch];
              }
            }
          }
          __syncwarp();  // every lane has read its dh row: the buffer may be re-armed (two chunks from now)
          buf = (buf + 1 == nbufs) ? 0 : buf + 1;
          continue;
        } else if (has_res) {
          // buffer (qn + 2) % 4 was last read by the store of chunk qn - 2: all but the latest store have drained
          if (lane == 0) {
            tma_store_wait_read<kResStageBufs - kResAhead - 1>();
            issue_residual();
          }
          mbar_wait(&my_res_bar[buf], (qn / kResStageBufs) & 1);
          ++qn;
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            const uint4 u = *reinterpret_cast<const uint4*>(row_ptr + ((ch ^ sw) << 4));
            const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float2 t2 = __bfloat1622float2(h2[q]);
              f[ch * 8 + 2 * q] += t2.x;
              f[ch * 8 + 2 * q + 1] += t2.y;
            }
          }
        } else {
          // staging buffer `buf` was last read by the TMA store issued kStageBufs chunks ago
          if (lane == 0) tma_store_wait_read<kStageBufs - 1>();
        }
        // (Measured alternatives of this store path, none kept: (a) a probe build WITHOUT the bulk stores runs the plain
        // N = 2 560 GEMM in 0.362 instead of 0.470 ms and QKV in 0.118 instead of 0.160 ms — the stores cost 14-26 %, the
        // rest is the chain above; (b) 64 columns per iteration (two tcgen05.ld in flight, one [32 x 128 B] SWIZZLE_128B tile
        // and one bulk store of full 128-byte rows): plain N = 2 560 0.467 -> 0.447 ms, QKV 0.159 -> 0.172 ms, dgrad unchanged.)
        // (Storing the 32 columns straight from registers with two 256-bit st.global per thread instead of this staging +
        // TMA-store path was measured slower on every forward shape: ff_in + GEGLU 0.453 -> 0.491 ms, plain N = 2 560
        // 0.469 -> 0.505 ms, out-projection 0.070 -> 0.076 ms — a warp-level row-scattered store costs 32 LSU transactions.)
        __syncwarp();
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          uint4 v;
          v.x = pack_bf16x2(f[ch * 8 + 0], f[ch * 8 + 1]);
          v.y = pack_bf16x2(f[ch * 8 + 2], f[ch * 8 + 3]);
          v.z = pack_bf16x2(f[ch * 8 + 4], f[ch * 8 + 5]);
          v.w = pack_bf16x2(f[ch * 8 + 6], f[ch * 8 + 7]);
          *reinterpret_cast<uint4*>(row_ptr + ((ch ^ sw) << 4)) = v;
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmY, reinterpret_cast<const void*>(my_staging_ptr + buf * 2048), ncol0, m0 + quad * 32);
          tma_store_commit();
        }
        buf = (buf + 1 == nbufs) ? 0 : buf + 1;
      }
      if (++as == 2) {
        as = 0;
        aph ^= 1;
      }
    }
    if (lane == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  if (kPair) cluster_sync_all();  // neither CTA leaves (or frees tensor memory) while the other may still touch it
  else __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    if (kPair) tmem_dealloc_pair(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
  }
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, tvi_bf16* __restrict__ dw, int splits, int64_t rows,
                                    int cols, int64_t ldp, int64_t lddw) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // one pair of columns per thread
  const int half = cols / 2;
  if (i >= rows * half) return;
  const int64_t r = i / half;
  const int c = (int)(i - r * half) * 2;
  float a = 0.f, b = 0.f;
  for (int s = 0; s < splits; ++s) {
    const float2 v = *reinterpret_cast<const float2*>(partial + ((int64_t)s * rows + r) * ldp + c);
    a += v.x;
    b += v.y;
  }
  *reinterpret_cast<uint32_t*>(dw + r * lddw + c) = pack_bf16x2(a, b);
}

int pick_bn(int N, int step = 32) {
  // largest tile width (multiple of `step`, <= 256) with the least padded columns
  int best = step, best_waste = 1 << 30;
  for (int bn = 256; bn >= step; bn -= step) {
    const int tiles = (N + bn - 1) / bn;
    const int waste = tiles * bn - N;
    if (waste < best_waste) {
      best_waste = waste;
      best = bn;
    }
  }
  return best;
}


// Tile width of the MN-major-B kernels (dgrad, wgrad): multiples of the 64-wide swizzle atom.  `pick_bn` (least padding)
// chose 64 for an output width of 320 — a 128 x 64 tile streams 24 KB of operands per 128 cycles of MMA.  What matters
// is the operand stream per unit of output, tiles * (A tile + B tile), against the padded MMA work; measured at the L0
// shapes (profiles/r2_gemm_bn_mn.txt): dgrad to 320 columns 0.690 -> 0.451 ms at 192 (2 560 -> 320), 0.275 -> 0.182 ms
// (960 -> 320); to 1 280 columns 256 stays best.  wgrad shares its tiles among the token splits and does best with the
// finer 128 at 320 output columns (0.818 -> 0.645 ms for dW[2560, 320]), 256 at 1 280.
int pick_bn_mn(int N, bool wgrad) {
  if (wgrad && N <= 384) return N <= 64 ? 64 : 128;
  int best = 64;
  double best_cost = 1e30;
  for (int bn = 256; bn >= 64; bn -= 64) {
    const int tiles = (N + bn - 1) / bn;
    const double waste = (double)(tiles * bn - N) / (double)(tiles * bn);
    const double cost = (double)tiles * (kABytes + bn * 128.0) * (1.0 + 0.5 * waste);
    if (cost < best_cost) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

// Operand tensor map.  K-major: dims {reduction, rows}, box {64, rows_box}.  MN-major: dims {out, reduction},
// box {64, 64}.
int operand_map(CUtensorMap* tm, const tvi_bf16* base, bool mn_major, int64_t out_extent, int64_t red_extent, int64_t ld,
                int rows_box) {
  if (!mn_major) {
    const uint64_t dims[2] = {(uint64_t)red_extent, (uint64_t)out_extent};
    const uint64_t strides[1] = {(uint64_t)ld * 2};
    const uint32_t box[2] = {(uint32_t)kBK, (uint32_t)rows_box};
    return make_tensor_map(tm, base, 2, 2, dims, strides, box, Swizzle::B128);
  }
  const uint64_t dims[2] = {(uint64_t)out_extent, (uint64_t)red_extent};
  const uint64_t strides[1] = {(uint64_t)ld * 2};
  const uint32_t box[2] = {64u, (uint32_t)kBK};
  return make_tensor_map(tm, base, 2, 2, dims, strides, box, Swizzle::B128);
}

template <bool kAMn, bool kBMn, bool kSplit, bool kPair, bool kGbwd>
int configure_once() {
  static bool configured[64] = {false};
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(gemm_bf16_kernel<kAMn, kBMn, kSplit, kPair, kGbwd>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     227 * 1024 - 1024));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  return TVI_OK;
}

// D[Mo, No] = A . B over a reduction of length R.  A is [Mo, R] (K-major) or [R, Mo] (MN-major) with row stride
// lda; B is [No, R] or [R, No] with row stride ldb.  Output: bf16 Y (ldy) through TMA, or fp32 partials when kSplit.
template <bool kAMn, bool kBMn, bool kSplit, bool kPair = false, bool kGbwd = false>
int launch_gemm(const char* name, const tvi_bf16* A, const tvi_bf16* B, const tvi_bf16* bias, tvi_bf16* Y, float* partial,
                int splits, int64_t Mo, int No, int64_t R, int64_t lda, int64_t ldb, int64_t ldy, cudaStream_t st,
                int geglu_block = 0, const tvi_bf16* residual = nullptr, int64_t ldr = 0, const float* ln_stats = nullptr,
                const float* ln_colsum = nullptr, tvi_bf16* dpre = nullptr, int64_t lddpre = 0) {
  GemmParams p;
  p.dpre = dpre;
  p.lddpre = lddpre;
  p.h_out = Y;
  p.ldh = ldy;
  p.wide_stores = (kGbwd && ((uintptr_t)Y | (uintptr_t)dpre) % 32 == 0 && (Y == nullptr || ldy % 16 == 0) && lddpre % 16 == 0 && geglu_block > 0 &&
                   (No / 2) % 16 == 0) ? 1 : 0;
  p.ln_stats = reinterpret_cast<const float2*>(ln_stats);
  p.ln_colsum = ln_colsum;
  p.BN = kBMn ? pick_bn_mn(No, kAMn) : pick_bn(No, 32);
#ifdef TVI_GEMM_TUNE  // tuning builds only (tools/micro/build_tune.sh): tile-width sweeps
  if (const char* e = getenv("TVI_GEMM_BN")) {
    const int bn = atoi(e);
    if (bn >= 32 && bn <= 256 && bn % (kBMn ? 64 : 32) == 0) p.BN = bn;
  }
#endif
  if (geglu_block > 0) p.BN = 2 * geglu_block;  // a tile = geglu_block value columns + their gate columns
  p.geglu = geglu_block > 0 ? 1 : 0;
  p.N_store = geglu_block > 0 ? No / 2 : No;
  p.residual = residual;
  p.ldr = ldr;
  p.N = No;
  constexpr int kMStep = kPair ? 2 * kBM : kBM;  // a CTA pair owns 256 rows
  p.m_tiles = (Mo + kMStep - 1) / kMStep;
  p.n_tiles = (No + p.BN - 1) / p.BN;
  p.k_blocks_total = (int)((R + kBK - 1) / kBK);
  p.k_blocks = kSplit ? (p.k_blocks_total + splits - 1) / splits : p.k_blocks_total;
  p.bias = bias;
  p.partial = partial;
  p.M_out = Mo;
  p.ldp = No;
  const int b_bytes = (kPair ? p.BN / 2 : p.BN) * kBK * 2;  // per CTA: a pair holds half of the B tile each
  const int res_bytes = p.k_blocks_total * b_bytes;
  const int sms = num_sms();
  const int workers = kPair ? sms / 2 : sms;               // tile walkers: CTAs, or CTA pairs
  // resident weight slice: plain forward GEMM, slice <= 128 KB, and enough row tiles per CTA to amortise it
  p.res_b = (!kAMn && !kBMn && !kSplit && res_bytes <= 128 * 1024 && p.n_tiles <= workers &&
             p.m_tiles * p.n_tiles >= 4 * (int64_t)workers) ? 1 : 0;
  const int stage_bytes = p.res_b ? kABytes : kABytes + b_bytes;
  p.stage_bufs = (!kSplit && residual != nullptr) ? kResStageBufs : kStageBufs;
  const int kStagingBytes = staging_bytes(p.stage_bufs);
  const int smem_cap = 227 * 1024 - 1024 /*alignment slack*/ - kStagingBytes - 1024 /*static*/ - (p.res_b ? res_bytes : 0);
  int stages = smem_cap / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  TVI_CHECK_ARG(stages >= 2, "%s: tile does not fit shared memory", name);
  p.stages = stages;
  const int smem_bytes = 1024 + stages * stage_bytes + kStagingBytes + (p.res_b ? res_bytes : 0);

  CUtensorMap tmA, tmB, tmY, tmR;
  int rc = operand_map(&tmA, A, kAMn, Mo, R, lda, kBM);
  if (rc) return rc;
  rc = operand_map(&tmB, B, kBMn, No, R, ldb, kPair ? p.BN / 2 : p.BN);
  if (rc) return rc;
  if (!kSplit) {
    const uint64_t dims[2] = {(uint64_t)p.N_store, (uint64_t)Mo};
    // kGbwd stores from registers and never uses the output map: without an H output it is built over dH instead
    const bool no_y = kGbwd && Y == nullptr;
    const uint64_t strides[1] = {(uint64_t)(no_y ? ldr : ldy) * 2};
    const uint32_t box[2] = {32u, 32u};
    rc = make_tensor_map(&tmY, no_y ? residual : Y, 2, 2, dims, strides, box, Swizzle::B64);
    if (rc) return rc;
    tmR = tmY;
    if (residual != nullptr) {
      const uint64_t rstrides[1] = {(uint64_t)ldr * 2};
      rc = make_tensor_map(&tmR, residual, 2, 2, dims, rstrides, box, Swizzle::B64);
      if (rc) return rc;
    }
  } else {
    tmY = tmA;  // unused by the split epilogue
    tmR = tmA;
  }
  rc = configure_once<kAMn, kBMn, kSplit, kPair, kGbwd>();
  if (rc) return rc;
  int64_t grid = p.m_tiles * p.n_tiles;
  if (kSplit) grid *= (p.k_blocks_total + p.k_blocks - 1) / p.k_blocks;
  if (grid > workers) grid = p.res_b ? (workers / p.n_tiles) * p.n_tiles : workers;
  if (kPair) {
    // two CTAs per tile walker, launched as clusters of 2 (the pair shares a TPC)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * grid));
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = (size_t)smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    TVI_CUDA_OK(cudaLaunchKernelEx(&cfg, gemm_bf16_kernel<kAMn, kBMn, kSplit, kPair>, tmA, tmB, tmY, tmR, p));
    return TVI_OK;
  }
  gemm_bf16_kernel<kAMn, kBMn, kSplit, kPair, kGbwd><<<(unsigned)grid, kGemmThreads, smem_bytes, st>>>(tmA, tmB, tmY, tmR, p);
  TVI_LAUNCH_OK(name);
  return TVI_OK;
}

// CTA-pair (cta_group::2) forward GEMM.  MEASURED on B200 at the L0 shapes (tools/micro/pair_gemm_check.py,
// profiles/r2_gemm_pair.txt) and OFF by default: ff_in + GEGLU 0.448 -> 0.537 ms, QKV 0.159 -> 0.220 ms, out-projection
// 0.109 -> 0.142 ms, ff_out (K = 1 280) 0.223 -> 0.221 ms.  Halving the L2 -> SM operand stream buys nothing: these GEMMs
// are bound by their epilogue (GEGLU: issue slots) and by the HBM write stream, and the pair adds a cross-CTA handshake
// per 5-k-block tile.  TVI_GEMM_PAIR=1 selects it wherever the shape allows (parity: tests/test_gpu_gemm_pair.py).
bool use_pair(int64_t M, int N, int K) {
  (void)K;
  const char* e = getenv("TVI_GEMM_PAIR");
  return e != nullptr && atoi(e) == 1 && M >= 2 * kBM && N >= 32;
}

int wgrad_splits(int64_t M, int N, int K) {
  const int64_t tiles = ((N + kBM - 1) / kBM) * (int64_t)((K + pick_bn_mn(K, true) - 1) / pick_bn_mn(K, true));
  const int64_t kblocks = (M + kBK - 1) / kBK;
  const int64_t sms = num_sms();
  // The persistent grid works through tiles * splits equal items in rounds of `sms`: pick the split count whose last
  // round is fullest.  (ceil(sms / tiles) alone lands just past a round boundary for every SVD shape — 40 tiles x 4,
  // 16 x 10, 6 x 25, 15 x 10 items on 148 SMs — and ran a second, almost empty round: ~52 % of the machine.)
  // Up to ~4 rounds are considered; a split keeps at least 16 k-blocks so the fp32 partial tile stays a small
  // fraction of what it reads.
  int64_t smax = (4 * sms + tiles - 1) / tiles;
  if (smax > kblocks / 16) smax = kblocks / 16;
  if (smax < 1) smax = 1;
  int64_t best = 1;
  double best_eff = 0.0;
  for (int64_t s = 1; s <= smax; ++s) {
    const int64_t items = tiles * s, rounds = (items + sms - 1) / sms;
    const double eff = (double)items / (double)(rounds * sms);
    if (eff > best_eff + 0.01) {  // ties go to the smaller workspace
      best_eff = eff;
      best = s;
    }
  }
  // every split must own at least one k-block
  const int64_t per = (kblocks + best - 1) / best;
  return (int)((kblocks + per - 1) / per);
}

}  // namespace
}  // namespace tvi

extern "C" int tvi_gemm_bf16_tn(const tvi_bf16* X, const tvi_bf16* W, const tvi_bf16* bias, tvi_bf16* Y, int64_t M, int N,
                                int K, int64_t ldx, int64_t ldw, int64_t ldy, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "tvi_gemm_bf16_tn: negative size");
  if (M == 0 || N == 0) return TVI_OK;
  TVI_CHECK_ARG(X && W && Y, "tvi_gemm_bf16_tn: null pointer");
  TVI_CHECK_ARG(K > 0, "tvi_gemm_bf16_tn: K must be positive");
  TVI_CHECK_ARG(K % 8 == 0 && N % 8 == 0, "tvi_gemm_bf16_tn: N=%d and K=%d must be multiples of 8", N, K);
  TVI_CHECK_ARG(ldx % 8 == 0 && ldw % 8 == 0 && ldy % 8 == 0, "tvi_gemm_bf16_tn: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(ldx >= K && ldw >= K && ldy >= N, "tvi_gemm_bf16_tn: row stride smaller than the row");
  TVI_CHECK_ARG(((uintptr_t)X | (uintptr_t)W | (uintptr_t)Y) % 16 == 0, "tvi_gemm_bf16_tn: X, W, Y must be 16-byte aligned");
  TVI_CHECK_ARG(bias == nullptr || (uintptr_t)bias % 4 == 0, "tvi_gemm_bf16_tn: bias must be 4-byte aligned");
  TVI_CHECK_ARG(M < ((int64_t)1 << 31) - 256, "tvi_gemm_bf16_tn: M too large for one call");
  if (use_pair(M, N, K))
    return launch_gemm<false, false, false, true>("tvi_gemm_bf16_tn", X, W, bias, Y, nullptr, 1, M, N, K, ldx, ldw, ldy,
                                                  (cudaStream_t)stream);
  return launch_gemm<false, false, false>("tvi_gemm_bf16_tn", X, W, bias, Y, nullptr, 1, M, N, K, ldx, ldw, ldy,
                                          (cudaStream_t)stream);
}

extern "C" int tvi_geglu_block(int inner) {
  // width of the value / gate column blocks the fused GEGLU projection expects in its re-ordered weight
#ifdef TVI_GEMM_TUNE  // tuning builds only
  if (const char* e = getenv("TVI_GEGLU_BLOCK")) {
    const int h = atoi(e);
    if (h >= 32 && h <= 128 && h % 32 == 0 && inner > 0 && inner % h == 0) return h;
  }
#endif
  for (int h = 128; h >= 32; h -= 32)
    if (inner > 0 && inner % h == 0) return h;
  return 0;
}

extern "C" int tvi_gemm_bf16_tn_ex(const tvi_bf16* X, const tvi_bf16* W, const tvi_bf16* bias, const tvi_bf16* residual,
                                   tvi_bf16* Y, int64_t M, int N, int K, int64_t ldx, int64_t ldw, int64_t ldr, int64_t ldy,
                                   int epilogue, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "tvi_gemm_bf16_tn_ex: negative size");
  if (M == 0 || N == 0) return TVI_OK;
  TVI_CHECK_ARG(X && W && Y, "tvi_gemm_bf16_tn_ex: null pointer");
  TVI_CHECK_ARG(K > 0 && K % 8 == 0 && N % 8 == 0, "tvi_gemm_bf16_tn_ex: N=%d and K=%d must be positive multiples of 8", N, K);
  TVI_CHECK_ARG(epilogue == TVI_EPILOGUE_NONE || epilogue == TVI_EPILOGUE_GEGLU, "tvi_gemm_bf16_tn_ex: unknown epilogue %d", epilogue);
  int block = 0;
  int n_store = N;
  if (epilogue == TVI_EPILOGUE_GEGLU) {
    TVI_CHECK_ARG(N % 2 == 0, "tvi_gemm_bf16_tn_ex: GEGLU needs an even number of projection columns");
    n_store = N / 2;
    block = tvi_geglu_block(n_store);
    TVI_CHECK_ARG(block > 0, "tvi_gemm_bf16_tn_ex: GEGLU inner width %d must be a multiple of 32", n_store);
  }
  TVI_CHECK_ARG(ldx % 8 == 0 && ldw % 8 == 0 && ldy % 8 == 0 && (residual == nullptr || ldr % 8 == 0),
                "tvi_gemm_bf16_tn_ex: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(ldx >= K && ldw >= K && ldy >= n_store && (residual == nullptr || ldr >= n_store),
                "tvi_gemm_bf16_tn_ex: row stride smaller than the row");
  TVI_CHECK_ARG(((uintptr_t)X | (uintptr_t)W | (uintptr_t)Y | (uintptr_t)residual) % 16 == 0,
                "tvi_gemm_bf16_tn_ex: X, W, Y, residual must be 16-byte aligned");
  TVI_CHECK_ARG(bias == nullptr || (uintptr_t)bias % 2 == 0, "tvi_gemm_bf16_tn_ex: bias must be 2-byte aligned");
  TVI_CHECK_ARG(M < ((int64_t)1 << 31) - 256, "tvi_gemm_bf16_tn_ex: M too large for one call");
  if (use_pair(M, N, K))
    return launch_gemm<false, false, false, true>("tvi_gemm_bf16_tn_ex", X, W, bias, Y, nullptr, 1, M, N, K, ldx, ldw, ldy,
                                                  (cudaStream_t)stream, block, residual, ldr);
  return launch_gemm<false, false, false>("tvi_gemm_bf16_tn_ex", X, W, bias, Y, nullptr, 1, M, N, K, ldx, ldw, ldy,
                                          (cudaStream_t)stream, block, residual, ldr);
}

extern "C" int tvi_gemm_bf16_tn_ln(const tvi_bf16* X, const tvi_bf16* Wg, const tvi_bf16* bias, const float* colsum,
                                   const float* stats, const tvi_bf16* residual, tvi_bf16* Y, int64_t M, int N, int K, int64_t ldx,
                                   int64_t ldw, int64_t ldr, int64_t ldy, int epilogue, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "tvi_gemm_bf16_tn_ln: negative size");
  if (M == 0 || N == 0) return TVI_OK;
  TVI_CHECK_ARG(X && Wg && Y && bias && colsum && stats, "tvi_gemm_bf16_tn_ln: null pointer (bias, colsum and stats are required)");
  TVI_CHECK_ARG(K > 0 && K % 8 == 0 && N % 8 == 0, "tvi_gemm_bf16_tn_ln: N=%d and K=%d must be positive multiples of 8", N, K);
  TVI_CHECK_ARG(epilogue == TVI_EPILOGUE_NONE || epilogue == TVI_EPILOGUE_GEGLU, "tvi_gemm_bf16_tn_ln: unknown epilogue %d", epilogue);
  int block = 0;
  int n_store = N;
  if (epilogue == TVI_EPILOGUE_GEGLU) {
    TVI_CHECK_ARG(N % 2 == 0, "tvi_gemm_bf16_tn_ln: GEGLU needs an even number of projection columns");
    n_store = N / 2;
    block = tvi_geglu_block(n_store);
    TVI_CHECK_ARG(block > 0, "tvi_gemm_bf16_tn_ln: GEGLU inner width %d must be a multiple of 32", n_store);
  }
  TVI_CHECK_ARG(ldx % 8 == 0 && ldw % 8 == 0 && ldy % 8 == 0 && (residual == nullptr || ldr % 8 == 0),
                "tvi_gemm_bf16_tn_ln: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(ldx >= K && ldw >= K && ldy >= n_store && (residual == nullptr || ldr >= n_store),
                "tvi_gemm_bf16_tn_ln: row stride smaller than the row");
  TVI_CHECK_ARG(((uintptr_t)X | (uintptr_t)Wg | (uintptr_t)Y | (uintptr_t)residual) % 16 == 0,
                "tvi_gemm_bf16_tn_ln: X, Wg, Y, residual must be 16-byte aligned");
  TVI_CHECK_ARG((uintptr_t)bias % 2 == 0 && (uintptr_t)colsum % 4 == 0 && (uintptr_t)stats % 8 == 0,
                "tvi_gemm_bf16_tn_ln: bias / colsum / stats are misaligned");
  TVI_CHECK_ARG(M < ((int64_t)1 << 31) - 256, "tvi_gemm_bf16_tn_ln: M too large for one call");
  return launch_gemm<false, false, false>("tvi_gemm_bf16_tn_ln", X, Wg, bias, Y, nullptr, 1, M, N, K, ldx, ldw, ldy,
                                          (cudaStream_t)stream, block, residual, ldr, stats, colsum);
}

extern "C" int tvi_gemm_bf16_geglu_bwd(const tvi_bf16* X, const tvi_bf16* Wi, const tvi_bf16* bias_i, const tvi_bf16* dH, tvi_bf16* H,
                                       tvi_bf16* dPre, int64_t M, int inner, int K, int64_t ldx, int64_t ldw, int64_t lddh,
                                       int64_t ldh, int64_t lddpre, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && inner >= 0 && K >= 0, "tvi_gemm_bf16_geglu_bwd: negative size");
  if (M == 0 || inner == 0) return TVI_OK;
  TVI_CHECK_ARG(X && Wi && dH && dPre, "tvi_gemm_bf16_geglu_bwd: null pointer");
  TVI_CHECK_ARG(K > 0 && K % 8 == 0, "tvi_gemm_bf16_geglu_bwd: K=%d must be a positive multiple of 8", K);
  const int block = tvi_geglu_block(inner);
  TVI_CHECK_ARG(block > 0, "tvi_gemm_bf16_geglu_bwd: GEGLU inner width %d must be a multiple of 32", inner);
  TVI_CHECK_ARG(ldx % 8 == 0 && ldw % 8 == 0 && lddh % 8 == 0 && (H == nullptr || ldh % 8 == 0) && lddpre % 8 == 0,
                "tvi_gemm_bf16_geglu_bwd: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(ldx >= K && ldw >= K && lddh >= inner && (H == nullptr || ldh >= inner) && lddpre >= 2 * (int64_t)inner,
                "tvi_gemm_bf16_geglu_bwd: row stride smaller than the row");
  TVI_CHECK_ARG(((uintptr_t)X | (uintptr_t)Wi | (uintptr_t)dH | (uintptr_t)H | (uintptr_t)dPre) % 16 == 0,
                "tvi_gemm_bf16_geglu_bwd: X, Wi, dH, H, dPre must be 16-byte aligned");
  TVI_CHECK_ARG(bias_i == nullptr || (uintptr_t)bias_i % 2 == 0, "tvi_gemm_bf16_geglu_bwd: bias must be 2-byte aligned");
  TVI_CHECK_ARG(M < ((int64_t)1 << 31) - 256, "tvi_gemm_bf16_geglu_bwd: M too large for one call");
  return launch_gemm<false, false, false, false, true>("tvi_gemm_bf16_geglu_bwd", X, Wi, bias_i, H, nullptr, 1, M, 2 * inner, K, ldx,
                                                       ldw, ldh, (cudaStream_t)stream, block, dH, lddh, nullptr, nullptr, dPre,
                                                       lddpre);
}

extern "C" int tvi_gemm_bf16_dgrad(const tvi_bf16* dY, const tvi_bf16* W, tvi_bf16* dX, int64_t M, int N, int K,
                                   int64_t lddy, int64_t ldw, int64_t lddx, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "tvi_gemm_bf16_dgrad: negative size");
  if (M == 0 || K == 0) return TVI_OK;
  TVI_CHECK_ARG(dY && W && dX, "tvi_gemm_bf16_dgrad: null pointer");
  TVI_CHECK_ARG(N > 0, "tvi_gemm_bf16_dgrad: N must be positive");
  TVI_CHECK_ARG(K % 8 == 0 && N % 8 == 0, "tvi_gemm_bf16_dgrad: N=%d and K=%d must be multiples of 8", N, K);
  TVI_CHECK_ARG(lddy % 8 == 0 && ldw % 8 == 0 && lddx % 8 == 0, "tvi_gemm_bf16_dgrad: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(lddy >= N && ldw >= K && lddx >= K, "tvi_gemm_bf16_dgrad: row stride smaller than the row");
  TVI_CHECK_ARG(((uintptr_t)dY | (uintptr_t)W | (uintptr_t)dX) % 16 == 0, "tvi_gemm_bf16_dgrad: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(M < ((int64_t)1 << 31) - 256, "tvi_gemm_bf16_dgrad: M too large for one call");
  // D[M, K] = A[M, N] . B, B = W stored [N (reduction), K (output, contiguous)] -> MN-major B operand
  return launch_gemm<false, true, false>("tvi_gemm_bf16_dgrad", dY, W, nullptr, dX, nullptr, 1, M, K, N, lddy, ldw, lddx,
                                         (cudaStream_t)stream);
}

extern "C" size_t tvi_gemm_bf16_wgrad_workspace_bytes(int64_t M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  return (size_t)tvi::wgrad_splits(M, N, K) * (size_t)N * (size_t)K * sizeof(float);
}

extern "C" int tvi_gemm_bf16_wgrad(const tvi_bf16* dY, const tvi_bf16* X, tvi_bf16* dW, void* workspace, int64_t M, int N,
                                   int K, int64_t lddy, int64_t ldx, int64_t lddw, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "tvi_gemm_bf16_wgrad: negative size");
  if (N == 0 || K == 0) return TVI_OK;
  TVI_CHECK_ARG(dY && X && dW && workspace, "tvi_gemm_bf16_wgrad: null pointer");
  TVI_CHECK_ARG(M > 0, "tvi_gemm_bf16_wgrad: M must be positive");
  TVI_CHECK_ARG(K % 8 == 0 && N % 8 == 0, "tvi_gemm_bf16_wgrad: N=%d and K=%d must be multiples of 8", N, K);
  TVI_CHECK_ARG(lddy % 8 == 0 && ldx % 8 == 0 && lddw % 2 == 0, "tvi_gemm_bf16_wgrad: bad row strides");
  TVI_CHECK_ARG(lddy >= N && ldx >= K && lddw >= K, "tvi_gemm_bf16_wgrad: row stride smaller than the row");
  TVI_CHECK_ARG(((uintptr_t)dY | (uintptr_t)X | (uintptr_t)workspace) % 16 == 0 && (uintptr_t)dW % 4 == 0,
                "tvi_gemm_bf16_wgrad: pointers must be 16-byte aligned (dW: 4)");
  TVI_CHECK_ARG(M < ((int64_t)1 << 31) - 256, "tvi_gemm_bf16_wgrad: M too large for one call");
  const int splits = wgrad_splits(M, N, K);
  // D[N, K] = A . B with A = dY stored [M (reduction), N (output rows, contiguous)], B = X stored [M, K]
  int rc = launch_gemm<true, true, true>("tvi_gemm_bf16_wgrad", dY, X, nullptr, nullptr, (float*)workspace, splits, N, K, M,
                                         lddy, ldx, 0, (cudaStream_t)stream);
  if (rc) return rc;
  const int64_t pairs = (int64_t)N * (K / 2);
  wgrad_reduce_kernel<<<(unsigned)((pairs + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const float*)workspace, dW, splits,
                                                                                          N, K, K, lddw);
  TVI_LAUNCH_OK("wgrad_reduce_kernel");
  return TVI_OK;
}
