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
#include "tvi_common.cuh"

namespace tvi {
namespace {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kGemmThreads = 192;
constexpr int kAccStride = 256;                    // TMEM columns per accumulator stage
constexpr int kABytes = kBM * kBK * 2;             // 16 KB
constexpr int kStagingBytes = 4 * 2 * 32 * 64;     // 4 epilogue warps x 2 buffers x [32 rows][64 B]
constexpr int kMaxStages = 8;

struct GemmParams {
  int64_t m_tiles;
  int n_tiles;
  int k_blocks;
  int BN;
  int N;
  int stages;
  const tvi_bf16* bias;
};

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                    const __grid_constant__ CUtensorMap tmY, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[kMaxStages];
  __shared__ __align__(8) uint64_t bar_empty[kMaxStages];
  __shared__ __align__(8) uint64_t bar_acc_full[2];
  __shared__ __align__(8) uint64_t bar_acc_empty[2];
  __shared__ uint32_t s_tmem_base;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // 1024-byte aligned carve-up (SWIZZLE_128B atoms are 1024 B)
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t b_bytes = (uint32_t)p.BN * kBK * 2;
  const uint32_t stage_bytes = kABytes + b_bytes;
  const uint32_t staging = base + p.stages * stage_bytes;
  uint8_t* staging_ptr = base_ptr + p.stages * stage_bytes;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmY);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&bar_full[s], 1);
      mbar_init(&bar_empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bar_acc_full[s], 1);
      mbar_init(&bar_acc_empty[s], 4);
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

  const int64_t n_tiles_total = p.m_tiles * p.n_tiles;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int64_t tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
        const int64_t mt = tile / p.n_tiles;
        const int nt = (int)(tile - mt * p.n_tiles);
        const int m0 = (int)(mt * kBM), n0 = nt * p.BN;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&bar_empty[s], ph ^ 1);
          mbar_arrive_expect_tx(&bar_full[s], stage_bytes);
          uint8_t* a_dst = base_ptr + s * stage_bytes;
          tma_load_2d(a_dst, &tmX, &bar_full[s], kb * kBK, m0);
          tma_load_2d(a_dst + kABytes, &tmW, &bar_full[s], kb * kBK, n0);
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
    const uint32_t leader = elect_one() ? 1u : 0u;
    const uint32_t idesc = umma_idesc_bf16(kBM, (uint32_t)p.BN, 0, 0);
    int s = 0;
    uint32_t ph = 0;
    int as = 0;
    uint32_t aph = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
      mbar_wait(&bar_acc_empty[as], aph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * kAccStride;
      for (int kb = 0; kb < p.k_blocks; ++kb) {
        mbar_wait(&bar_full[s], ph);
        tc_fence_after();
        const uint32_t a_addr = base + s * stage_bytes;
        const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
        for (int kk = 0; kk < kBK / 16; ++kk) {
          const uint64_t ad = umma_smem_desc(a_addr + kk * 32, 16, 1024, kUmmaLayoutSw128);
          const uint64_t bd = umma_smem_desc(b_addr + kk * 32, 16, 1024, kUmmaLayoutSw128);
          umma_bf16_ss_pred(d_tmem, ad, bd, idesc, (kb | kk) != 0 ? 1u : 0u, leader);
        }
        umma_commit_pred(&bar_empty[s], leader);  // frees the smem stage once these MMAs have read it
        if (++s == p.stages) {
          s = 0;
          ph ^= 1;
        }
      }
      umma_commit_pred(&bar_acc_full[as], leader);  // accumulator complete -> epilogue
      if (++as == 2) {
        as = 0;
        aph ^= 1;
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quad = warp & 3;  // TMEM lane quadrant this warp may read
    const uint32_t my_staging = staging + (warp - 2) * 2 * 2048;
    uint8_t* my_staging_ptr = staging_ptr + (warp - 2) * 2 * 2048;
    int buf = 0;
    int as = 0;
    uint32_t aph = 0;
    const int chunks = p.BN / 32;
    for (int64_t tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
      const int64_t mt = tile / p.n_tiles;
      const int nt = (int)(tile - mt * p.n_tiles);
      const int m0 = (int)(mt * kBM), n0 = nt * p.BN;
      mbar_wait(&bar_acc_full[as], aph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(quad * 32) << 16) + as * kAccStride;
      for (int c = 0; c < chunks; ++c) {
        uint32_t r[32];
        tmem_ld_x32(t_row + c * 32, r);
        tmem_ld_wait();
        if (c == chunks - 1) {
          // every column of this accumulator stage is now in registers
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_acc_empty[as]);
        }
        const int ncol0 = n0 + c * 32;
        if (ncol0 >= p.N) continue;  // tile column tail: nothing to store (warp-uniform)
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(r[j]);
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            // N % 8 == 0, so a pair never straddles the end of the bias vector
            if (ncol0 + j < p.N) {
              const __nv_bfloat162 b2 = *reinterpret_cast<const __nv_bfloat162*>(p.bias + ncol0 + j);
              const float2 bf = __bfloat1622float2(b2);
              f[j] += bf.x;
              f[j + 1] += bf.y;
            }
          }
        }
        // staging buffer `buf` was last read by the TMA store issued two chunks ago
        if (lane == 0) tma_store_wait_read<1>();
        __syncwarp();
        uint8_t* row_ptr = my_staging_ptr + buf * 2048 + lane * 64;
        const int sw = (lane >> 1) & 3;
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
        buf ^= 1;
      }
      if (++as == 2) {
        as = 0;
        aph ^= 1;
      }
    }
    if (lane == 0) tma_store_wait_all<0>();
    (void)my_staging;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

int pick_bn(int N) {
  // largest tile width (multiple of 32, <= 256) with the least padded columns
  int best = 32, best_waste = 1 << 30;
  for (int bn = 256; bn >= 32; bn -= 32) {
    const int tiles = (N + bn - 1) / bn;
    const int waste = tiles * bn - N;
    if (waste < best_waste) {
      best_waste = waste;
      best = bn;
    }
  }
  return best;
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

  GemmParams p;
  p.BN = pick_bn(N);
  p.N = N;
  p.m_tiles = (M + kBM - 1) / kBM;
  p.n_tiles = (N + p.BN - 1) / p.BN;
  p.k_blocks = (K + kBK - 1) / kBK;
  p.bias = bias;
  const int stage_bytes = kABytes + p.BN * kBK * 2;
  const int smem_cap = 227 * 1024 - 1024 /*alignment slack*/ - kStagingBytes - 256 /*static*/;
  int stages = smem_cap / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  TVI_CHECK_ARG(stages >= 2, "tvi_gemm_bf16_tn: tile does not fit shared memory");
  p.stages = stages;
  const int smem_bytes = 1024 + stages * stage_bytes + kStagingBytes;

  CUtensorMap tmX, tmW, tmY;
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};
    const uint64_t strides[1] = {(uint64_t)ldx * 2};
    const uint32_t box[2] = {(uint32_t)kBK, (uint32_t)kBM};
    int rc = make_tensor_map(&tmX, X, 2, 2, dims, strides, box, Swizzle::B128);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N};
    const uint64_t strides[1] = {(uint64_t)ldw * 2};
    const uint32_t box[2] = {(uint32_t)kBK, (uint32_t)p.BN};
    int rc = make_tensor_map(&tmW, W, 2, 2, dims, strides, box, Swizzle::B128);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
    const uint64_t strides[1] = {(uint64_t)ldy * 2};
    const uint32_t box[2] = {32u, 32u};
    int rc = make_tensor_map(&tmY, Y, 2, 2, dims, strides, box, Swizzle::B64);
    if (rc) return rc;
  }

  static bool configured[64] = {false};
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(gemm_bf16_tn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 256));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  int64_t grid = p.m_tiles * p.n_tiles;
  const int sms = num_sms();
  if (grid > sms) grid = sms;
  gemm_bf16_tn_kernel<<<(unsigned)grid, kGemmThreads, smem_bytes, (cudaStream_t)stream>>>(tmX, tmW, tmY, p);
  TVI_LAUNCH_OK("gemm_bf16_tn_kernel");
  return TVI_OK;
}
