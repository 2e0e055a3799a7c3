// Softmax / correction / epilogue role of attn_spatial_tc_kernel (included inside namespace tvi::{anon}).
// One thread per query row; `t` is the 128-row tile this warp group serves.
// The online softmax of src/models/attention_processor.py:598-615 (softmax over keys of
// scale * QK^T) evaluated tile by tile with a shared reference maximum per row.

__device__ __forceinline__ void tc_softmax_role(const TcParams& p, uint32_t tmem_base, int t, int warp, int lane, int b,
                                                int h, int q_row0, uint8_t* stage_ptr, uint64_t* bar_s, uint64_t* bar_p,
                                                uint64_t* bar_o, const CUtensorMap* tmO) {
  const int quad = warp & 3;  // TMEM lane quadrant this warp may touch
  const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
  const uint32_t tS = tmem_base + lane_base + kColS + t * 128;
  const uint32_t tO = tmem_base + lane_base + kColO + t * 64;
  const uint32_t tP = tmem_base + lane_base + kColP + t * 64;
  const float sl2 = p.scale * 1.4426950408889634f;
  const uint64_t sl2x2 = f32x2_pack(sl2, sl2);
  float m_ref = -INFINITY;  // reference maximum of this row, in raw (unscaled) logit units
  float l = 0.f;            // running sum of exp2((s - m_ref) * sl2)

  for (int j = 0; j < p.n_kv; ++j) {
    mbar_wait(bar_s, j & 1);
    tc_fence_after();
    // Pass 1 (row maximum) walks the four 32-column chunks of S two at a time; chunks 2 and 3 stay
    // in registers for pass 2 while chunks 0 and 1 are re-read from TMEM (cheaper than 128 live
    // registers per thread: three warps share an SM sub-partition's register file).
    const bool ragged = (j == p.n_kv - 1) && (p.Skv & (kTcBN - 1)) != 0;
    const int valid = p.Skv - j * kTcBN;
    auto mask_tail = [&](uint32_t* x, int c0) {
      if (ragged) {
#pragma unroll
        for (int c = 0; c < 32; ++c)
          if (c0 + c >= valid) x[c] = 0xff800000u;  // -inf
      }
    };
    auto max32 = [&](const uint32_t* x, float& a, float& bq) {
#pragma unroll
      for (int c = 0; c < 32; c += 4) {
        a = fmax3(a, __uint_as_float(x[c]), __uint_as_float(x[c + 1]));
        bq = fmax3(bq, __uint_as_float(x[c + 2]), __uint_as_float(x[c + 3]));
      }
    };
    uint32_t ra[32], rb[32], rc[32];
    float mx0 = -INFINITY, mx1 = -INFINITY;
    tmem_ld_x32(tS, ra);
    tmem_ld_x32(tS + 32, rb);
    tmem_ld_wait();
    mask_tail(ra, 0);
    mask_tail(rb, 32);
    max32(ra, mx0, mx1);
    max32(rb, mx0, mx1);
    tmem_ld_x32(tS + 64, ra);
    tmem_ld_x32(tS + 96, rb);
    tmem_ld_wait();
    mask_tail(ra, 64);
    mask_tail(rb, 96);
    max32(ra, mx0, mx1);
    max32(rb, mx0, mx1);
    tmem_ld_x32(tS, rc);  // chunk 0 again, in flight under the bookkeeping below
    const float mx = fmaxf(mx0, mx1);
    // advance the reference only when the row maximum outgrew it by more than 2^8
    const bool need = (mx - m_ref) * sl2 > 8.f;
    float alpha = 1.f;
    if (need) {
      alpha = fast_exp2((m_ref - mx) * sl2);  // 0 on the first tile (m_ref = -inf)
      m_ref = mx;
      l *= alpha;
    }
    const float neg = -m_ref * sl2;
    const uint64_t negx2 = f32x2_pack(neg, neg);
    uint64_t sum_a = f32x2_pack(0.f, 0.f), sum_b = sum_a;
    auto exp_store = [&](const uint32_t* x, int cb) {
      uint32_t pk[16];
#pragma unroll
      for (int c = 0; c < 32; c += 4) {
        const uint64_t xa = f32x2_fma(f32x2_pack(__uint_as_float(x[c]), __uint_as_float(x[c + 1])), sl2x2, negx2);
        const uint64_t xb = f32x2_fma(f32x2_pack(__uint_as_float(x[c + 2]), __uint_as_float(x[c + 3])), sl2x2, negx2);
        float a0, a1, b0, b1;
        f32x2_unpack(xa, a0, a1);
        f32x2_unpack(xb, b0, b1);
        a0 = fast_exp2(a0);
        a1 = fast_exp2(a1);
        b0 = fast_exp2(b0);
        b1 = fast_exp2(b1);
        sum_a = f32x2_add(sum_a, f32x2_pack(a0, a1));
        sum_b = f32x2_add(sum_b, f32x2_pack(b0, b1));
        pk[c >> 1] = pack_bf16x2(a0, a1);
        pk[(c >> 1) + 1] = pack_bf16x2(b0, b1);
      }
      tmem_st_x16(tP + (cb >> 1), pk);
    };
    exp_store(ra, 64);
    exp_store(rb, 96);
    tmem_ld_wait();
    tmem_ld_x32(tS + 32, ra);  // chunk 1 again
    mask_tail(rc, 0);
    exp_store(rc, 0);
    tmem_ld_wait();
    mask_tail(ra, 32);
    exp_store(ra, 32);
    if (j > 0 && __any_sync(0xffffffffu, need)) {
      // rescale the O accumulator: PV_{j-1} must have landed, PV_j is not issued before bar_p below
      mbar_wait(bar_o, (j - 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int hcol = 0; hcol < 64; hcol += 32) {
        tmem_ld_x32(tO + hcol, rb);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c) rb[c] = __float_as_uint(__uint_as_float(rb[c]) * alpha);
        tmem_st_x32(tO + hcol, rb);
      }
    }
    {
      float s0, s1, s2, s3;
      f32x2_unpack(sum_a, s0, s1);
      f32x2_unpack(sum_b, s2, s3);
      l += (s0 + s1) + (s2 + s3);
    }
    tmem_st_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_p);
  }

  // ---- epilogue: O / l -> bf16 -> swizzled staging (the dead Q tile) -> TMA store ---------------
  mbar_wait(bar_o, (p.n_kv - 1) & 1);
  tc_fence_after();
  const float inv = 1.f / l;
  const int rr = quad * 32 + lane;  // row inside the tile
  uint8_t* row_ptr = stage_ptr + rr * 128;
#pragma unroll
  for (int hcol = 0; hcol < 64; hcol += 32) {
    uint32_t o[32];
    tmem_ld_x32(tO + hcol, o);
    tmem_ld_wait();
#pragma unroll
    for (int ch = 0; ch < 4; ++ch) {
      uint4 v;
      v.x = pack_bf16x2(__uint_as_float(o[ch * 8 + 0]) * inv, __uint_as_float(o[ch * 8 + 1]) * inv);
      v.y = pack_bf16x2(__uint_as_float(o[ch * 8 + 2]) * inv, __uint_as_float(o[ch * 8 + 3]) * inv);
      v.z = pack_bf16x2(__uint_as_float(o[ch * 8 + 4]) * inv, __uint_as_float(o[ch * 8 + 5]) * inv);
      v.w = pack_bf16x2(__uint_as_float(o[ch * 8 + 6]) * inv, __uint_as_float(o[ch * 8 + 7]) * inv);
      const int chunk = (hcol >> 3) + ch;
      *reinterpret_cast<uint4*>(row_ptr + ((chunk ^ (rr & 7)) << 4)) = v;
    }
  }
  const int row = q_row0 + rr;
  if (p.lse != nullptr && row < p.S) p.lse[((int64_t)b * p.S + row) * p.heads + h] = m_ref * p.scale + __logf(l);
  fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0 && q_row0 + quad * 32 < p.S) {
    tma_store_3d(tmO, stage_ptr + quad * 32 * 128, h * 64, q_row0 + quad * 32, b);
    tma_store_commit();
    tma_store_wait_all<0>();
  }
}
