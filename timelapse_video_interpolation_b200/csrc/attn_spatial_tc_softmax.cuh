// Softmax / correction / epilogue role of attn_spatial_tc_kernel (included inside namespace tvi::{anon}).
// TWO threads per query row of a 128-row tile (same TMEM lane, i.e. two warps of the same lane quadrant):
// each owns 32 of the 64 key columns of a step, so every SM sub-partition holds four softmax warps and the
// MUFU / FMA phases of one warp hide the TMEM round trips and barrier waits of the others.  The pair
// shares the row maximum through shared memory (one 64-thread named barrier per step) and therefore the
// same reference maximum; the row sum is kept as two partial sums that meet in the epilogue.
// `tile_tmem` is the tile's TMEM column base.  This is the softmax of
// src/models/attention_processor.py:598-615 (softmax over keys of scale * QK^T) evaluated 64 keys at a
// time against a per-row reference maximum.

// 2^x for a pair of fp32 values on the FMA / ALU pipes instead of the 16-per-clock MUFU unit:
// n = round(x) through the 1.5*2^23 magic constant, f = x - n in [-0.5, 0.5], a cubic minimax
// polynomial for 2^f (relative error 7.6e-5, far below the bf16 rounding of P) and n added
// straight into the exponent field.  x is clamped at -125 (2^-125 instead of an underflowed 0).
__device__ __forceinline__ uint64_t exp2_poly_x2(uint64_t x) {
  float x0, x1;
  f32x2_unpack(x, x0, x1);
  const uint64_t xc = f32x2_pack(fmaxf(x0, -125.f), fmaxf(x1, -125.f));
  const uint64_t t = f32x2_add(xc, f32x2_pack(12582912.f, 12582912.f));
  const uint64_t n = f32x2_add(t, f32x2_pack(-12582912.f, -12582912.f));
  const uint64_t f = f32x2_fma(n, f32x2_pack(-1.f, -1.f), xc);
  uint64_t pl = f32x2_fma(f32x2_pack(0.05520550534f, 0.05520550534f), f, f32x2_pack(0.24261397123f, 0.24261397123f));
  pl = f32x2_fma(pl, f, f32x2_pack(0.69325476885f, 0.69325476885f));
  pl = f32x2_fma(pl, f, f32x2_pack(0.99992769957f, 0.99992769957f));
  float t0, t1, p0, p1;
  f32x2_unpack(t, t0, t1);
  f32x2_unpack(pl, p0, p1);
  const float r0 = __int_as_float((__float_as_int(t0) << 23) + __float_as_int(p0));
  const float r1 = __int_as_float((__float_as_int(t1) << 23) + __float_as_int(p1));
  return f32x2_pack(r0, r1);
}

template <bool B>
struct TcTag {
  static constexpr bool value = B;
};

// kPolyMask: bit i set = pair i (mod 8) of every 16 consecutive scores takes the polynomial path.
template <uint32_t kPolyMask>
__device__ __forceinline__ void tc_softmax_role(const TcParams& p, uint32_t tile_tmem, int quad, int half, uint32_t pair_bar,
                                                float* xchg, int lane, int b, int h, int q_row0, uint8_t* stage_ptr,
                                                uint64_t* bar_q, uint64_t* bar_s, uint64_t* bar_p, uint64_t* bar_o,
                                                uint64_t* bar_done, const CUtensorMap* tmO) {
  const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
  const uint32_t tS = tile_tmem + lane_base + kColS;              // + buf*64: scores; this thread reads + half*32
  const uint32_t tO = tile_tmem + lane_base + kColO + half * 32;  // this thread's 32 of the 64 output columns
  const uint32_t tQ = tile_tmem + lane_base + kColQ + half * 16;
  const int rr = quad * 32 + lane;  // row inside the tile
  const int row = q_row0 + rr;      // row inside the sequence

  // ---- park this thread's half of the Q row (32 bf16 = 16 packed columns) in tensor memory ----------
  {
    uint32_t qr[16];
    if (row < p.S) {
      const uint4* src = reinterpret_cast<const uint4*>(p.q + ((int64_t)b * p.S + row) * p.ldq + h * 64 + half * 32);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint4 v = __ldg(src + i);
        qr[4 * i + 0] = v.x;
        qr[4 * i + 1] = v.y;
        qr[4 * i + 2] = v.z;
        qr[4 * i + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) qr[i] = 0u;
    }
    tmem_st_x16(tQ, qr);
    tmem_st_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_q);
  }

  const float sl2 = p.scale * 1.4426950408889634f;
  const uint64_t sl2x2 = f32x2_pack(sl2, sl2);
  float m_ref = -INFINITY;  // reference maximum of this row, in raw (unscaled) logit units (identical in both halves)
  float l = 0.f;            // this half's running sum of exp2((s - m_ref) * sl2)

  uint32_t ra[32];  // this thread's 32 scores of the current step
#ifdef TVI_DEBUG_WAIT
  int n_slow = 0;
#endif
  TVI_WAIT(&bar_s[0], 0, 5, 0);
  tc_fence_after();
  tmem_ld_x32(tS + half * 32, ra);

  // One key step.  kRagged is only true for a last step that runs past Skv (peeled so that the
  // compare/selects of the mask stay out of the steady-state loop).
  auto step = [&](int j, auto ragged_tag) {
    constexpr bool kRagged = decltype(ragged_tag)::value;
    const int buf = j & 1;
    tmem_ld_wait();
    if (kRagged) {
      const int valid = p.Skv - j * kTcBN - half * 32;
#pragma unroll
      for (int c = 0; c < 32; ++c)
        if (c >= valid) ra[c] = 0xff800000u;  // -inf
    }
    // Speculative pass: evaluate the exponentials against the CURRENT reference maximum without looking
    // for the row maximum first (32 FMNMX3 per step on a half-rate pipe).  The pair's row sum of the step
    // tells whether that was sound: P <= sum <= 2^13 keeps every value far inside the bf16 / fp32 range,
    // and the reference does not have to be the exact maximum (O and l carry the same factor).  A sum that
    // is larger, infinite or NaN (first step: m_ref = -inf) sends the whole warp through the exact pass.
    uint32_t pk[16];
    float hs;  // this half's sum of the step's exponentials
    auto exp_pass = [&]() {
      const float neg = -m_ref * sl2;
      const uint64_t negx2 = f32x2_pack(neg, neg);
      uint64_t sum_a = f32x2_pack(0.f, 0.f), sum_b = sum_a;
#pragma unroll
      for (int c = 0; c < 32; c += 4) {
        uint64_t xa = f32x2_fma(f32x2_pack(__uint_as_float(ra[c]), __uint_as_float(ra[c + 1])), sl2x2, negx2);
        uint64_t xb = f32x2_fma(f32x2_pack(__uint_as_float(ra[c + 2]), __uint_as_float(ra[c + 3])), sl2x2, negx2);
        float a0, a1, b0, b1;
        if ((kPolyMask >> ((c >> 1) & 7)) & 1u) {
          xa = exp2_poly_x2(xa);
          f32x2_unpack(xa, a0, a1);
        } else {
          f32x2_unpack(xa, a0, a1);
          a0 = fast_exp2(a0);
          a1 = fast_exp2(a1);
          xa = f32x2_pack(a0, a1);
        }
        if ((kPolyMask >> (((c >> 1) + 1) & 7)) & 1u) {
          xb = exp2_poly_x2(xb);
          f32x2_unpack(xb, b0, b1);
        } else {
          f32x2_unpack(xb, b0, b1);
          b0 = fast_exp2(b0);
          b1 = fast_exp2(b1);
          xb = f32x2_pack(b0, b1);
        }
        sum_a = f32x2_add(sum_a, xa);
        sum_b = f32x2_add(sum_b, xb);
        pk[c >> 1] = pack_bf16x2(a0, a1);
        pk[(c >> 1) + 1] = pack_bf16x2(b0, b1);
      }
      float s0, s1, s2, s3;
      f32x2_unpack(sum_a, s0, s1);
      f32x2_unpack(sum_b, s2, s3);
      hs = (s0 + s1) + (s2 + s3);
    };
    exp_pass();
    // both halves see the same total (parity-double-buffered exchange slots), hence take the same branch
    float* slot = xchg + buf * 256;
    slot[half * 128 + rr] = hs;
    named_bar_sync(pair_bar, 64);
    const float total = hs + slot[(half ^ 1) * 128 + rr];
    if (__any_sync(0xffffffffu, !(total <= 8192.f))) {
#ifdef TVI_DEBUG_WAIT
      ++n_slow;
#endif
      // ---- exact pass (rare): row maximum of the step over both halves, lazy reference update ----------
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int c = 0; c < 32; c += 4) {
        mx0 = fmax3(mx0, __uint_as_float(ra[c]), __uint_as_float(ra[c + 1]));
        mx1 = fmax3(mx1, __uint_as_float(ra[c + 2]), __uint_as_float(ra[c + 3]));
      }
      float* mslot = xchg + 512 + buf * 256;
      const float mxl = fmaxf(mx0, mx1);
      mslot[half * 128 + rr] = mxl;
      named_bar_sync(pair_bar, 64);
      const float mx = fmaxf(mxl, mslot[(half ^ 1) * 128 + rr]);
      // advance the reference only when the row maximum outgrew it by more than 2^8
      const bool need = (mx - m_ref) * sl2 > 8.f;
      float alpha = 1.f;
      if (need) {
        alpha = fast_exp2((m_ref - mx) * sl2);  // 0 on the first step (m_ref = -inf)
        m_ref = mx;
        l *= alpha;
      }
      if (j > 0 && __any_sync(0xffffffffu, need)) {
        // rescale this thread's half of the O accumulator: PV_{j-1} must have landed; PV_j is not issued
        // before bar_p below
        TVI_WAIT(bar_o, (uint32_t)(j - 1) & 1u, 6, j);
        tc_fence_after();
        uint32_t o[32];
        tmem_ld_x32(tO, o);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c) o[c] = __float_as_uint(__uint_as_float(o[c]) * alpha);
        tmem_st_x32(tO, o);
      }
      exp_pass();
    }
    l += hs;
    // P (64 bf16 = 32 columns) overwrites the first half of this S buffer; this thread writes the 16
    // columns of its 32 keys.  For half 1 those are score columns 16..31, which belong to half 0: its
    // tcgen05.ld of them completed before it arrived on the pair barrier above.
    tmem_st_x16(tS + buf * 64 + half * 16, pk);
    // the next step's scores (other S buffer, normally long finished) travel to registers while the
    // P store drains
    if (j + 1 < p.n_kv) {
      TVI_WAIT(&bar_s[buf ^ 1], (uint32_t)((j + 1) >> 1) & 1u, 7, j);
      tc_fence_after();
      tmem_ld_x32(tS + (buf ^ 1) * 64 + half * 32, ra);
    }
    tmem_st_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&bar_p[buf]);
  };
  const bool last_ragged = (p.Skv & (kTcBN - 1)) != 0;
  for (int j = 0; j < p.n_kv - 1; ++j) step(j, TcTag<false>{});
  if (last_ragged) step(p.n_kv - 1, TcTag<true>{});
  else step(p.n_kv - 1, TcTag<false>{});

#ifdef TVI_DEBUG_WAIT
  if (blockIdx.x < 2 && blockIdx.y == 0 && lane == 0) printf("block %d quad %d half %d: %d exact passes of %d steps\n", blockIdx.x, quad, half, n_slow, p.n_kv);
#endif
  // ---- epilogue: O / l -> bf16 -> swizzled staging -> TMA store ------------------------------------
  // the two partial row sums meet (slots of the parity the last step did not use)
  float* slot = xchg + (p.n_kv & 1) * 256;
  slot[half * 128 + rr] = l;
  named_bar_sync(pair_bar, 64);
  l += slot[(half ^ 1) * 128 + rr];
  // every PV product of this tile has landed in the O accumulator
  TVI_WAIT(bar_done, 0, 8, 0);
  tc_fence_after();
  const float inv = 1.f / l;
  uint8_t* row_ptr = stage_ptr + rr * 128;
  {
    uint32_t o[32];
    tmem_ld_x32(tO, o);
    tmem_ld_wait();
#pragma unroll
    for (int ch = 0; ch < 4; ++ch) {
      uint4 v;
      v.x = pack_bf16x2(__uint_as_float(o[ch * 8 + 0]) * inv, __uint_as_float(o[ch * 8 + 1]) * inv);
      v.y = pack_bf16x2(__uint_as_float(o[ch * 8 + 2]) * inv, __uint_as_float(o[ch * 8 + 3]) * inv);
      v.z = pack_bf16x2(__uint_as_float(o[ch * 8 + 4]) * inv, __uint_as_float(o[ch * 8 + 5]) * inv);
      v.w = pack_bf16x2(__uint_as_float(o[ch * 8 + 6]) * inv, __uint_as_float(o[ch * 8 + 7]) * inv);
      const int chunk = half * 4 + ch;
      *reinterpret_cast<uint4*>(row_ptr + ((chunk ^ (rr & 7)) << 4)) = v;
    }
  }
  if (half == 0 && p.lse != nullptr && row < p.S) p.lse[((int64_t)b * p.S + row) * p.heads + h] = m_ref * p.scale + __logf(l);
  fence_proxy_async_smem();
  named_bar_sync(pair_bar, 64);  // both halves of the 32 rows are staged
  if (half == 0 && lane == 0 && q_row0 + quad * 32 < p.S) {
    tma_store_3d(tmO, stage_ptr + quad * 32 * 128, h * 64, q_row0 + quad * 32, b);
    tma_store_commit();
    tma_store_wait_all<0>();
  }
}
