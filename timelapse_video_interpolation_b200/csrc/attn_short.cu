// A3c / A3b / A3d — the two HBM-bound attention cores.
//
// tvi_attn_temporal_fwd : nseq x heads independent problems of T<=64 tokens, d=64 or 128.  One warp
//   per (sequence, head): the three T x 64 bf16 tiles are staged in shared memory with 128-bit
//   cp.async copies (each tile row is one full 128-byte line; rows live in XOR-swizzled 16-byte
//   chunks so ldmatrix is conflict-free), double-buffered per warp so the next problem's loads
//   are in flight while this one computes.  QK^T and PV run on mma.sync m16n8k16 (a 14x14
//   problem cannot fill a 128-row tcgen05 tile; CUDA-core FMAs would be 1.4x short of the HBM
//   roofline), softmax is fp32 in registers with quad shuffles, and O goes back through shared
//   memory so every global store is a coalesced 128-bit write.
//   Algorithmic HBM bytes: 4 * rows * heads*d * 2 (Q, K, V read once, O written once).
//
// tvi_attn_fewkeys_fwd : cross-attention against 1..8 context tokens that are shared by
//   rows_per_group query rows.  K/V of the group stay in registers; each thread owns one 16-byte
//   channel chunk and walks down the rows, the 64-wide dot products are reduced across the 8
//   (d=64) or 16 (d=128) lanes of a head with shuffles.  With one key the softmax is exactly 1,
//   so Q is never read.  Optional additive [Tq,L] bias and fused [Tq,L] mean-logit statistics
//   replace the reference's 2*T .item() syncs (src/train_decouples.py:437-471).
//   Algorithmic HBM bytes: 2 * rows * heads*d * 2 (Q read, O written); L == 1: rows*heads*d*2.
#include "mma_sync.cuh"

namespace tvi {
namespace {

// ---------------------------------------------------------------------------------------------
// temporal (short-sequence) attention
// ---------------------------------------------------------------------------------------------
// DH = d / 64: a tile is DH swizzled [rows][64] halves (d = 128 is the class default at C = 1280 / 10 heads,
// src/models/unet_spatio_temporal_condition.py:94)
template <int MT, int DH>
struct TemporalCfg {
  static constexpr int kRows = 16 * MT;
  static constexpr int kHalfBytes = kRows * 128;
  static constexpr int kTileBytes = DH * kHalfBytes;
  static constexpr int kStageBytes = 3 * kTileBytes;
  static constexpr int kWarps = DH == 1 ? ((MT == 4) ? 4 : 8) : (MT == 1 ? 8 : MT == 2 ? 4 : 2);
  static constexpr int kSmemBytes = kWarps * 2 * kStageBytes;
};

template <int MT, int DH>
__device__ __forceinline__ void temporal_issue_loads(uint32_t stage_smem, const tvi_bf16* q, const tvi_bf16* k,
                                                     const tvi_bf16* v, int64_t item, int heads, int Tq, int Tk,
                                                     int64_t ldq, int64_t ldk, int64_t ldv, int64_t inner, int lane) {
  constexpr int kRows = 16 * MT;
  constexpr int kTile = TemporalCfg<MT, DH>::kTileBytes;
  const int64_t seq = item / heads;
  const int head = (int)(item - seq * heads);
  // token t of sequence `seq` is row (seq / inner) * inner * T + t * inner + seq % inner: inner == 1 is the
  // contiguous [nseq, T] layout, inner == S reads the temporal sequences straight out of the [B, T, S] token order
  const int64_t so = seq / inner, si = seq - so * inner;
  const tvi_bf16* qb = q + (so * inner * Tq + si) * ldq + head * 64 * DH;
  const tvi_bf16* kb = k + (so * inner * Tk + si) * ldk + head * 64 * DH;
  const tvi_bf16* vb = v + (so * inner * Tk + si) * ldv + head * 64 * DH;
  ldq *= inner;  // consecutive tokens of a sequence are `inner` rows apart
  ldk *= inner;
  ldv *= inner;
#pragma unroll
  for (int i = 0; i < kRows * DH / 4; ++i) {
    const int idx = lane + 32 * i;
    const int r = idx / (8 * DH), c = idx % (8 * DH);
    const uint32_t off = (c >> 3) * (kRows * 128) + tile_off(r, c & 7);
    const bool pq = r < Tq, pk = r < Tk;
    // rows past the sequence are zero-filled (src-size 0): padded K rows are masked in the
    // softmax, padded V rows must be finite because they are multiplied by P = 0.
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(stage_smem + off),
                 "l"(pq ? qb + r * ldq + c * 8 : qb), "r"(pq ? 16 : 0)
                 : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(stage_smem + kTile + off),
                 "l"(pk ? kb + r * ldk + c * 8 : kb), "r"(pk ? 16 : 0)
                 : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(stage_smem + 2 * kTile + off),
                 "l"(pk ? vb + r * ldv + c * 8 : vb), "r"(pk ? 16 : 0)
                 : "memory");
  }
}

template <int MT, int DH>
__global__ void __launch_bounds__(TemporalCfg<MT, DH>::kWarps * 32)
attn_temporal_kernel(const tvi_bf16* __restrict__ q, const tvi_bf16* __restrict__ k, const tvi_bf16* __restrict__ v,
                     tvi_bf16* __restrict__ o, float* __restrict__ lse, int64_t nitems, int heads, int Tq, int Tk,
                     int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, float scale, int64_t inner) {
  using Cfg = TemporalCfg<MT, DH>;
  constexpr int kRows = Cfg::kRows;
  constexpr int kHalf = Cfg::kHalfBytes;
  constexpr int NT = 2 * MT;  // 8-key column tiles of S
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t warp_smem = smem_u32(smem_raw) + warp * 2 * Cfg::kStageBytes;
  uint8_t* warp_ptr = smem_raw + warp * 2 * Cfg::kStageBytes;

  const int64_t gw = (int64_t)blockIdx.x * Cfg::kWarps + warp;
  const int64_t stride = (int64_t)gridDim.x * Cfg::kWarps;
  const float sl2 = scale * 1.4426950408889634f;
  const int g = lane >> 2, qd = lane & 3;

  int64_t item = gw;
  int stage = 0;
  if (item < nitems) temporal_issue_loads<MT, DH>(warp_smem, q, k, v, item, heads, Tq, Tk, ldq, ldk, ldv, inner, lane);
  cp_async_commit();

  for (; item < nitems; item += stride, stage ^= 1) {
    const int64_t next = item + stride;
    if (next < nitems)
      temporal_issue_loads<MT, DH>(warp_smem + (stage ^ 1) * Cfg::kStageBytes, q, k, v, next, heads, Tq, Tk, ldq, ldk, ldv,
                                   inner, lane);
    cp_async_commit();
    cp_async_wait<1>();
    __syncwarp();

    const uint32_t sQ = warp_smem + stage * Cfg::kStageBytes;
    const uint32_t sK = sQ + Cfg::kTileBytes;
    const uint32_t sV = sK + Cfg::kTileBytes;
    uint8_t* pQ = warp_ptr + stage * Cfg::kStageBytes;
    const int64_t seq = item / heads;
    const int head = (int)(item - seq * heads);
    const int64_t so = seq / inner;
    const int64_t row0 = so * inner * Tq + (seq - so * inner);  // first query row of this sequence; rows are `inner` apart

#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      if (mt * 16 >= Tq) break;
      // ---- S = Q K^T ---------------------------------------------------------------
      float s[NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < 4 * DH; ++ks) {
        uint32_t a[4];
        ldsm_x4(sQ + (ks >> 2) * kHalf + tile_off(mt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, 2 * (ks & 3) + (lane >> 4)),
                a[0], a[1], a[2], a[3]);
#pragma unroll
        for (int np = 0; np < MT; ++np) {
          uint32_t b0, b1, b2, b3;
          ldsm_x4(sK + (ks >> 2) * kHalf + tile_off(np * 16 + (lane & 7) + (lane >> 4) * 8, 2 * (ks & 3) + ((lane >> 3) & 1)),
                  b0, b1, b2, b3);
          mma_bf16_16816(s[2 * np], a, b0, b1);
          mma_bf16_16816(s[2 * np + 1], a, b2, b3);
        }
      }
      // ---- softmax over keys (rows g and g+8 of this m-tile) ---------------------------
      float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int col = nt * 8 + qd * 2;
        if (col >= Tk) s[nt][0] = s[nt][2] = -INFINITY;
        if (col + 1 >= Tk) s[nt][1] = s[nt][3] = -INFINITY;
        m0 = fmaxf(m0, fmaxf(s[nt][0], s[nt][1]));
        m1 = fmaxf(m1, fmaxf(s[nt][2], s[nt][3]));
      }
      m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
      m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
      m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
      m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
      const float ms0 = m0 * sl2, ms1 = m1 * sl2;
      float l0 = 0.f, l1 = 0.f;
      uint32_t p[NT][2];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const float e0 = fast_exp2(fmaf(s[nt][0], sl2, -ms0)), e1 = fast_exp2(fmaf(s[nt][1], sl2, -ms0));
        const float e2 = fast_exp2(fmaf(s[nt][2], sl2, -ms1)), e3 = fast_exp2(fmaf(s[nt][3], sl2, -ms1));
        l0 += e0 + e1;
        l1 += e2 + e3;
        p[nt][0] = pack_bf16x2(e0, e1);
        p[nt][1] = pack_bf16x2(e2, e3);
      }
      l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
      l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
      l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
      // ---- O = P V ------------------------------------------------------------------
      float acc[8 * DH][4];
#pragma unroll
      for (int nt = 0; nt < 8 * DH; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
      for (int kk = 0; kk < MT; ++kk) {
        if (kk * 16 >= Tk) break;
        const uint32_t a[4] = {p[2 * kk][0], p[2 * kk][1], p[2 * kk + 1][0], p[2 * kk + 1][1]};
#pragma unroll
        for (int nc = 0; nc < 8 * DH; nc += 2) {
          uint32_t b0, b1, b2, b3;
          ldsm_x4_t(sV + (nc >> 3) * kHalf + tile_off(kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, (nc & 7) + (lane >> 4)),
                    b0, b1, b2, b3);
          mma_bf16_16816(acc[nc], a, b0, b1);
          mma_bf16_16816(acc[nc + 1], a, b2, b3);
        }
      }
      const float inv0 = 1.f / l0, inv1 = 1.f / l1;
      // every lane of the warp has consumed this m-tile's Q rows (ldmatrix above is warp-wide):
      // stage O over them.
      __syncwarp();
      const int r0 = mt * 16 + g, r1 = r0 + 8;
#pragma unroll
      for (int nt = 0; nt < 8 * DH; ++nt) {
        uint8_t* ph = pQ + (nt >> 3) * kHalf;
        *reinterpret_cast<uint32_t*>(ph + tile_off(r0, nt & 7) + qd * 4) = pack_bf16x2(acc[nt][0] * inv0, acc[nt][1] * inv0);
        *reinterpret_cast<uint32_t*>(ph + tile_off(r1, nt & 7) + qd * 4) = pack_bf16x2(acc[nt][2] * inv1, acc[nt][3] * inv1);
      }
      if (lse != nullptr && qd == 0) {
        // natural-log-sum-exp of the scaled logits
        if (r0 < Tq) lse[(row0 + r0 * inner) * heads + head] = m0 * scale + __logf(l0);
        if (r1 < Tq) lse[(row0 + r1 * inner) * heads + head] = m1 * scale + __logf(l1);
      }
    }
    __syncwarp();
    // ---- coalesced 128-bit stores of the O tile --------------------------------------
    tvi_bf16* ob = o + row0 * ldo + head * 64 * DH;
#pragma unroll
    for (int i = 0; i < kRows * DH / 4; ++i) {
      const int idx = lane + 32 * i;
      const int r = idx / (8 * DH), c = idx % (8 * DH);
      if (r < Tq) {
        const uint4 val = *reinterpret_cast<const uint4*>(pQ + (c >> 3) * kHalf + tile_off(r, c & 7));
        *reinterpret_cast<uint4*>(ob + r * inner * ldo + c * 8) = val;
      }
    }
    __syncwarp();  // the stage may be refilled by the next iteration's cp.async
  }
  cp_async_wait<0>();
}

template <int MT, int DH>
int launch_temporal(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse, int64_t nseq,
                    int heads, int Tq, int Tk, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, float scale,
                    int64_t inner, cudaStream_t st) {
  using Cfg = TemporalCfg<MT, DH>;
  static bool configured[64] = {false};  // per device: the attribute belongs to the device's context
  int dev = 0;
  TVI_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    TVI_CUDA_OK(cudaFuncSetAttribute(attn_temporal_kernel<MT, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  const int64_t nitems = nseq * heads;
  const int per_sm = (MT == 1 && DH == 1) ? 2 : 1;
  int64_t grid = (nitems + Cfg::kWarps - 1) / Cfg::kWarps;
  const int64_t cap = (int64_t)num_sms() * per_sm;
  if (grid > cap) grid = cap;
  attn_temporal_kernel<MT, DH><<<(unsigned)grid, Cfg::kWarps * 32, Cfg::kSmemBytes, st>>>(q, k, v, o, lse, nitems, heads, Tq, Tk,
                                                                                     ldq, ldk, ldv, ldo, scale, inner);
  TVI_LAUNCH_OK("attn_temporal_kernel");
  return TVI_OK;
}

// ---------------------------------------------------------------------------------------------
// few-key cross attention
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

constexpr int kFewRowsPerThread = 16;

template <int LT, int LPH, bool STATS>
__global__ void __launch_bounds__(512)
attn_fewkeys_kernel(const tvi_bf16* __restrict__ q, const tvi_bf16* __restrict__ k, const tvi_bf16* __restrict__ v,
                    tvi_bf16* __restrict__ o, int64_t rows_per_group, int tiles_per_group, int L, int64_t ldq,
                    int64_t ldk, int64_t ldv, int64_t ldo, float scale, int Tq, const float* __restrict__ bias,
                    double* __restrict__ stats_acc, int64_t inner) {
  const int cx = threadIdx.x;               // 16-byte channel chunk of the row
  const int ry = threadIdx.y, RY = blockDim.y;
  const int64_t group = blockIdx.x / tiles_per_group;
  const int tile = blockIdx.x - (int)(group * tiles_per_group);
  const int64_t rows_per_tile = (int64_t)RY * kFewRowsPerThread;
  const int64_t rlo = group * rows_per_group + (int64_t)tile * rows_per_tile;
  int64_t rhi = rlo + rows_per_tile;
  const int64_t gend = (group + 1) * rows_per_group;
  if (rhi > gend) rhi = gend;
  // inner > 1: the groups are interleaved — logical row i of group (b, s) (s < inner) is physical row
  // (b * rows_per_group + i) * inner + s, i.e. the [B*T, S, C] token order read as [B*S, T, C] (T = rows_per_group)
  const int64_t gbase = group * rows_per_group;
  const int64_t pbase = (group / inner) * rows_per_group * inner + group % inner;
  auto phys = [&](int64_t r) -> int64_t { return inner == 1 ? r : pbase + (r - gbase) * inner; };

  extern __shared__ float s_stats[];  // [Tq * LT] when STATS
  if (STATS) {
    for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < Tq * LT; i += blockDim.x * blockDim.y) s_stats[i] = 0.f;
    __syncthreads();
  }

  // K / V chunks of this group stay in registers (packed bf16)
  uint4 kr[LT], vr[LT];
#pragma unroll
  for (int l = 0; l < LT; ++l) {
    if (l < L) {
      kr[l] = __ldg(reinterpret_cast<const uint4*>(k + (group * L + l) * ldk + cx * 8));
      vr[l] = __ldg(reinterpret_cast<const uint4*>(v + (group * L + l) * ldv + cx * 8));
    } else {
      kr[l] = make_uint4(0, 0, 0, 0);
      vr[l] = make_uint4(0, 0, 0, 0);
    }
  }

  if (LT == 1 && !STATS) {
    // softmax over one key is exactly 1: O = V broadcast, Q is never read
    for (int64_t r = rlo + ry; r < rhi; r += RY) *reinterpret_cast<uint4*>(o + phys(r) * ldo + cx * 8) = vr[0];
    return;
  }

  if constexpr (LT == 2 && !STATS) {
    // Two keys: softmax([l0, l1]) = [sigma(z), 1 - sigma(z)] with z = l0 - l1 = scale * q . (k0 - k1) + (b0 - b1), and
    // O = v1 + sigma(z) * (v0 - v1): ONE dot product and one shuffle chain per row instead of two, no max / sum.
    float kd[8], vd[8], v1f[8];
    {
      float a[8], b[8];
      unpack8(kr[0], a);
      unpack8(kr[1], b);
#pragma unroll
      for (int j = 0; j < 8; ++j) kd[j] = (a[j] - b[j]) * (scale * 1.4426950408889634f);   // z in log2 units
      unpack8(vr[0], a);
      unpack8(vr[1], v1f);
#pragma unroll
      for (int j = 0; j < 8; ++j) vd[j] = a[j] - v1f[j];
    }
    constexpr int kU = 4;
    for (int it0 = 0; it0 < kFewRowsPerThread; it0 += kU) {
      if (rlo + (int64_t)it0 * RY >= rhi) break;  // block-uniform
      uint4 qv[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int64_t rr = rlo + (int64_t)(it0 + u) * RY + ry;
        qv[u] = __ldg(reinterpret_cast<const uint4*>(q + phys(rr < rhi ? rr : rlo) * ldq + cx * 8));
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (rlo + (int64_t)(it0 + u) * RY >= rhi) break;  // block-uniform: every lane reaches every shuffle
        const int64_t rr = rlo + (int64_t)(it0 + u) * RY + ry;
        const bool act = rr < rhi;
        const int64_t r = act ? rr : rlo;
        float qf[8];
        unpack8(qv[u], qf);
        float z = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) z = fmaf(qf[j], kd[j], z);
#pragma unroll
        for (int off = LPH / 2; off > 0; off >>= 1) z += __shfl_xor_sync(0xffffffffu, z, off);
        if (bias != nullptr) {
          const int t = (int)(r % Tq);
          z = fmaf(__ldg(bias + t * 2) - __ldg(bias + t * 2 + 1), 1.4426950408889634f, z);
        }
        const float w = __fdividef(1.f, 1.f + fast_exp2(-z));   // sigma; exp2(+inf) = inf -> w = 0
        uint4 outv;
        outv.x = pack_bf16x2(fmaf(w, vd[0], v1f[0]), fmaf(w, vd[1], v1f[1]));
        outv.y = pack_bf16x2(fmaf(w, vd[2], v1f[2]), fmaf(w, vd[3], v1f[3]));
        outv.z = pack_bf16x2(fmaf(w, vd[4], v1f[4]), fmaf(w, vd[5], v1f[5]));
        outv.w = pack_bf16x2(fmaf(w, vd[6], v1f[6]), fmaf(w, vd[7], v1f[7]));
        if (act) *reinterpret_cast<uint4*>(o + phys(r) * ldo + cx * 8) = outv;
      }
    }
    return;
  }

  // A warp can straddle two rows (blockDim.x = 40 at C = 320), so the trip count is made
  // block-uniform and rows past the tile are predicated: every lane reaches every shuffle.
  // The Q chunks of kFewUnroll rows are requested before the first of them is used: a streaming kernel with ~80
  // instructions per 32 bytes needs the loads of several rows in flight per thread to cover the HBM latency.
  constexpr int kFewUnroll = 4;
  static_assert(kFewRowsPerThread % kFewUnroll == 0, "row loop is unrolled in groups");
  for (int it0 = 0; it0 < kFewRowsPerThread; it0 += kFewUnroll) {
    if (rlo + (int64_t)it0 * RY >= rhi) break;
    uint4 qv[kFewUnroll];
#pragma unroll
    for (int u = 0; u < kFewUnroll; ++u) {
      const int64_t rr = rlo + (int64_t)(it0 + u) * RY + ry;
      qv[u] = __ldg(reinterpret_cast<const uint4*>(q + phys(rr < rhi ? rr : rlo) * ldq + cx * 8));
    }
#pragma unroll
    for (int u = 0; u < kFewUnroll; ++u) {
    const int it = it0 + u;
    if (rlo + (int64_t)it * RY >= rhi) break;
    const int64_t rr = rlo + (int64_t)it * RY + ry;
    const bool act = rr < rhi;
    const int64_t r = act ? rr : rlo;
    float qf[8];
    unpack8(qv[u], qf);
    float logit[LT];
#pragma unroll
    for (int l = 0; l < LT; ++l) {
      float kf[8];
      unpack8(kr[l], kf);
      float d = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) d = fmaf(qf[j], kf[j], d);
#pragma unroll
      for (int off = LPH / 2; off > 0; off >>= 1) d += __shfl_xor_sync(0xffffffffu, d, off);
      logit[l] = d * scale;
    }
    const int t = (int)(r % Tq);
    if (bias != nullptr) {
#pragma unroll
      for (int l = 0; l < LT; ++l)
        if (l < L) logit[l] += __ldg(bias + t * L + l);
    }
    if (STATS) {
      if (act && (cx & (LPH - 1)) == 0) {
#pragma unroll
        for (int l = 0; l < LT; ++l)
          if (l < L) atomicAdd(&s_stats[t * LT + l], logit[l]);
      }
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
    const float inv = 1.f / sum;
    float of[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) of[j] = 0.f;
#pragma unroll
    for (int l = 0; l < LT; ++l) {
      float vf[8];
      unpack8(vr[l], vf);
      const float w = pr[l] * inv;
#pragma unroll
      for (int j = 0; j < 8; ++j) of[j] = fmaf(w, vf[j], of[j]);
    }
    uint4 outv;
    outv.x = pack_bf16x2(of[0], of[1]);
    outv.y = pack_bf16x2(of[2], of[3]);
    outv.z = pack_bf16x2(of[4], of[5]);
    outv.w = pack_bf16x2(of[6], of[7]);
    if (act) *reinterpret_cast<uint4*>(o + phys(r) * ldo + cx * 8) = outv;
    }
  }

  if (STATS) {
    __syncthreads();
    for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < Tq * LT; i += blockDim.x * blockDim.y) {
      const int t = i / LT, l = i - t * LT;
      if (l < L) atomicAdd(&stats_acc[t * L + l], (double)s_stats[i]);
    }
  }
}

__global__ void fewkeys_stats_final_kernel(const double* __restrict__ acc, float* __restrict__ stats, int n, double inv_count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) stats[i] = (float)(acc[i] * inv_count);
}

template <int LT, int LPH>
int launch_fewkeys(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, int64_t ngroups,
                   int64_t rows_per_group, int C8, int L, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo,
                   float scale, int Tq, const float* bias, double* stats_acc, int64_t inner, cudaStream_t st) {
  // whole warps only: blockDim.x * blockDim.y must be a multiple of 32 (C8 is a multiple of 8)
  const int unit = (C8 % 32 == 0) ? 1 : (C8 % 16 == 0) ? 2 : 4;
  int RY = unit * (256 / (C8 * unit) > 0 ? 256 / (C8 * unit) : 1);
  TVI_CHECK_ARG(C8 * RY <= 512, "tvi_attn_fewkeys_fwd: heads*d = %d is too wide for this kernel", C8 * 8);
  const int64_t rows_per_tile = (int64_t)RY * kFewRowsPerThread;
  const int64_t tiles = (rows_per_group + rows_per_tile - 1) / rows_per_tile;
  TVI_CHECK_ARG(tiles * ngroups < ((int64_t)1 << 31), "tvi_attn_fewkeys_fwd: grid too large");
  dim3 block((unsigned)C8, (unsigned)RY);
  dim3 grid((unsigned)(tiles * ngroups));
  if (stats_acc != nullptr) {
    const size_t sm = (size_t)Tq * LT * sizeof(float);
    TVI_CHECK_ARG(sm <= 48 * 1024, "tvi_attn_fewkeys_fwd: Tq*L too large for the statistics buffer");
    attn_fewkeys_kernel<LT, LPH, true><<<grid, block, sm, st>>>(q, k, v, o, rows_per_group, (int)tiles, L, ldq, ldk, ldv, ldo,
                                                                scale, Tq, bias, stats_acc, inner);
  } else {
    attn_fewkeys_kernel<LT, LPH, false><<<grid, block, 0, st>>>(q, k, v, o, rows_per_group, (int)tiles, L, ldq, ldk, ldv, ldo,
                                                                scale, Tq, bias, nullptr, inner);
  }
  TVI_LAUNCH_OK("attn_fewkeys_kernel");
  return TVI_OK;
}

}  // namespace
}  // namespace tvi

extern "C" {

int tvi_attn_temporal_fwd_interleaved(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse,
                                      int64_t nseq, int heads, int Tq, int Tk, int d, int64_t seq_inner, int64_t ldq,
                                      int64_t ldk, int64_t ldv, int64_t ldo, float scale, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(nseq >= 0 && heads >= 0, "tvi_attn_temporal_fwd_interleaved: negative size");
  if (nseq == 0 || heads == 0) return TVI_OK;
  TVI_CHECK_ARG(q && k && v && o, "tvi_attn_temporal_fwd_interleaved: null pointer");
  if (d != 64 && d != 128) return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_temporal_fwd_interleaved: head dim %d (64 and 128 are built)", d);
  TVI_CHECK_ARG(Tq >= 1 && Tq <= 64 && Tk >= 1 && Tk <= 64, "tvi_attn_temporal_fwd_interleaved: Tq=%d Tk=%d outside [1,64]", Tq, Tk);
  TVI_CHECK_ARG(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0, "tvi_attn_temporal_fwd_interleaved: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(ldq >= heads * d && ldk >= heads * d && ldv >= heads * d && ldo >= heads * d, "tvi_attn_temporal_fwd_interleaved: row stride smaller than heads*d");
  TVI_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o) % 16 == 0, "tvi_attn_temporal_fwd_interleaved: pointers must be 16-byte aligned");
  TVI_CHECK_ARG(seq_inner >= 1 && nseq % seq_inner == 0, "tvi_attn_temporal_fwd_interleaved: nseq must be a multiple of seq_inner");
  cudaStream_t st = (cudaStream_t)stream;
  const int T = Tq > Tk ? Tq : Tk;
#define TVI_TEMPORAL(MT)                                                                                                  \
  (d == 64 ? launch_temporal<MT, 1>(q, k, v, o, lse, nseq, heads, Tq, Tk, ldq, ldk, ldv, ldo, scale, seq_inner, st)       \
           : launch_temporal<MT, 2>(q, k, v, o, lse, nseq, heads, Tq, Tk, ldq, ldk, ldv, ldo, scale, seq_inner, st))
  if (T <= 16) return TVI_TEMPORAL(1);
  if (T <= 32) return TVI_TEMPORAL(2);
  return TVI_TEMPORAL(4);
#undef TVI_TEMPORAL
}

int tvi_attn_temporal_fwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse, int64_t nseq,
                          int heads, int Tq, int Tk, int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo,
                          float scale, void* stream) {
  return tvi_attn_temporal_fwd_interleaved(q, k, v, o, lse, nseq, heads, Tq, Tk, d, 1, ldq, ldk, ldv, ldo, scale, stream);
}

int tvi_attn_fewkeys_fwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, int64_t nrows,
                         int64_t rows_per_group, int heads, int L, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                         int64_t ldo, float scale, int Tq, const float* bias, float* stats, void* stats_ws, void* stream) {
  return tvi_attn_fewkeys_fwd_interleaved(q, k, v, o, nrows, rows_per_group, heads, L, d, 1, ldq, ldk, ldv, ldo, scale, Tq, bias,
                                          stats, stats_ws, stream);
}

int tvi_attn_fewkeys_fwd_interleaved(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, int64_t nrows,
                                     int64_t rows_per_group, int heads, int L, int d, int64_t group_inner, int64_t ldq,
                                     int64_t ldk, int64_t ldv, int64_t ldo, float scale, int Tq, const float* bias,
                                     float* stats, void* stats_ws, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(nrows >= 0 && heads >= 0, "tvi_attn_fewkeys_fwd: negative size");
  if (nrows == 0 || heads == 0) return TVI_OK;
  TVI_CHECK_ARG(k && v && o, "tvi_attn_fewkeys_fwd: null pointer");
  if (d != 64 && d != 128) return fail(TVI_ERR_UNSUPPORTED, "tvi_attn_fewkeys_fwd: head dim %d (64 and 128 are built)", d);
  TVI_CHECK_ARG(L >= 1 && L <= 8, "tvi_attn_fewkeys_fwd: L=%d outside [1,8]", L);
  TVI_CHECK_ARG(rows_per_group >= 1 && nrows % rows_per_group == 0, "tvi_attn_fewkeys_fwd: nrows must be a multiple of rows_per_group");
  TVI_CHECK_ARG(q != nullptr || (L == 1 && stats == nullptr), "tvi_attn_fewkeys_fwd: q may only be null when L == 1 and no statistics are requested");
  TVI_CHECK_ARG(Tq >= 1, "tvi_attn_fewkeys_fwd: Tq must be >= 1");
  TVI_CHECK_ARG(stats == nullptr || stats_ws != nullptr, "tvi_attn_fewkeys_fwd: stats requires stats_ws");
  const int C = heads * d;
  TVI_CHECK_ARG(C / 8 <= 512, "tvi_attn_fewkeys_fwd: heads*d too large");
  TVI_CHECK_ARG(ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0 && (q == nullptr || ldq % 8 == 0), "tvi_attn_fewkeys_fwd: row strides must be multiples of 8 elements");
  TVI_CHECK_ARG(((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)o) % 16 == 0, "tvi_attn_fewkeys_fwd: pointers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t ngroups = nrows / rows_per_group;
  TVI_CHECK_ARG(group_inner >= 1 && ngroups % group_inner == 0, "tvi_attn_fewkeys_fwd: the group count must be a multiple of group_inner");
  double* acc = nullptr;
  if (stats != nullptr) {
    acc = (double*)stats_ws;
    TVI_CUDA_OK(cudaMemsetAsync(acc, 0, sizeof(double) * Tq * L, st));
  }
  int rc;
#define TVI_FEW(LT)                                                                                              \
  (d == 64 ? launch_fewkeys<LT, 8>(q, k, v, o, ngroups, rows_per_group, C / 8, L, ldq, ldk, ldv, ldo, scale, Tq, bias, acc, group_inner, st) \
           : launch_fewkeys<LT, 16>(q, k, v, o, ngroups, rows_per_group, C / 8, L, ldq, ldk, ldv, ldo, scale, Tq, bias, acc, group_inner, st))
  if (L == 1) rc = TVI_FEW(1);
  else if (L == 2) rc = TVI_FEW(2);
  else if (L <= 4) rc = TVI_FEW(4);
  else rc = TVI_FEW(8);
#undef TVI_FEW
  if (rc != TVI_OK) return rc;
  if (stats != nullptr) {
    const int n = Tq * L;
    const double count = (double)(nrows / Tq) * heads;
    fewkeys_stats_final_kernel<<<(n + 127) / 128, 128, 0, st>>>(acc, stats, n, count > 0 ? 1.0 / count : 0.0);
    TVI_LAUNCH_OK("fewkeys_stats_final_kernel");
  }
  return TVI_OK;
}

// ---------------------------------------------------------------------------------------------
// No-allocation variant of the few-keys core with statistics (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops).
// The Tq*L fp64 accumulators live in a file-scope device array instead of a caller-provided workspace.  In host code
// the array's NAME is only a host-side stand-in, never a device address: the address is taken once per device with
// cudaGetSymbolAddress (that device current) and handed to the unchanged entry point above as `stats_ws`.
// Every call on a device shares the array and it keeps the last call's contents: the call clears the Tq*L entries it
// uses (cudaMemsetAsync on `stream`, as the entry point above does) — calls that overlap on the GPU (different
// streams) race on it.  32 KB per GPU that loads the library, held until exit.
// ---------------------------------------------------------------------------------------------
}  // extern "C"

static __device__ __align__(256) double g_tvi_fewkeys_stats_scratch[TVI_FEWKEYS_STATIC_MAX_STATS];

namespace tvi {
namespace {
constexpr int kMaxDevices = 64;
// device address of g_tvi_fewkeys_stats_scratch on the CURRENT device (resolved once per device); nullptr + *rc on error
double* fewkeys_static_scratch(int* rc) {
  static double* ptrs[kMaxDevices] = {nullptr};
  *rc = TVI_OK;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    *rc = fail(TVI_ERR_CUDA, "tvi_attn_fewkeys_fwd_static: cudaGetDevice failed or device %d outside [0, %d)", dev, kMaxDevices);
    return nullptr;
  }
  double* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {  // racing first calls resolve the same address twice: harmless
    e = cudaGetSymbolAddress((void**)&p, g_tvi_fewkeys_stats_scratch);
    if (e != cudaSuccess || p == nullptr) {
      *rc = fail(TVI_ERR_CUDA, "tvi_attn_fewkeys_fwd_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return nullptr;
    }
    __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);
  }
  return p;
}
}  // namespace
}  // namespace tvi

extern "C" {

int tvi_attn_fewkeys_fwd_static(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, int64_t nrows,
                                int64_t rows_per_group, int heads, int L, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                                int64_t ldo, float scale, int Tq, const float* bias, float* stats, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(Tq >= 1 && L >= 1 && L <= 8, "tvi_attn_fewkeys_fwd_static: Tq=%d must be >= 1 and L=%d in [1, 8]", Tq, L);
  TVI_CHECK_ARG((int64_t)Tq * L <= TVI_FEWKEYS_STATIC_MAX_STATS,
                "tvi_attn_fewkeys_fwd_static: Tq*L = %lld statistics exceed the static scratch of %d entries "
                "(use tvi_attn_fewkeys_fwd with a caller-provided stats_ws)", (long long)Tq * L, TVI_FEWKEYS_STATIC_MAX_STATS);
  double* ws = nullptr;
  if (stats != nullptr) {
    int rc = TVI_OK;
    ws = fewkeys_static_scratch(&rc);
    if (ws == nullptr) return rc;
  }
  return tvi_attn_fewkeys_fwd_interleaved(q, k, v, o, nrows, rows_per_group, heads, L, d, 1, ldq, ldk, ldv, ldo, scale, Tq, bias,
                                          stats, ws, stream);
}

}  // extern "C"
