// GEGLU backward (training through the block-level FeedForward swap).  The forward fuses
// h = value * gelu(gate) into the first GEMM's epilogue and never writes the 2*inner-wide projection; the backward
// recomputes `pre = x W1^T + b1` with the plain GEMM and this kernel turns (pre, dh) into
//   h      = value * gelu(gate)                          (needed by the second Linear's weight gradient)
//   dpre   = [ dh * gelu(gate) | dh * value * gelu'(gate) ],   gelu'(g) = Phi(g) + g * phi(g)   (erf form)
// in one pass over the rows (diffusers GEGLU.forward: `hidden_states, gate = proj(x).chunk(2, dim=-1)`).
#include "tvi_common.cuh"

namespace tvi {
namespace {

__global__ void __launch_bounds__(256)
geglu_bwd_kernel(const tvi_bf16* __restrict__ pre, const tvi_bf16* __restrict__ dh, tvi_bf16* __restrict__ h, tvi_bf16* __restrict__ dpre,
                 int64_t M, int inner8, int64_t ldpre, int64_t lddh, int64_t ldh, int64_t lddpre) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= M * inner8) return;
  const int64_t r = idx / inner8;
  const int c = (int)(idx - r * inner8);
  const int inner = inner8 * 8;
  const uint4 uv = __ldg(reinterpret_cast<const uint4*>(pre + r * ldpre) + c);
  const uint4 ug = __ldg(reinterpret_cast<const uint4*>(pre + r * ldpre + inner) + c);
  const uint4 ud = __ldg(reinterpret_cast<const uint4*>(dh + r * lddh) + c);
  const __nv_bfloat162* hv = reinterpret_cast<const __nv_bfloat162*>(&uv);
  const __nv_bfloat162* hg = reinterpret_cast<const __nv_bfloat162*>(&ug);
  const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&ud);
  uint32_t oh[4], ov[4], og[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float2 v = __bfloat1622float2(hv[q]), g = __bfloat1622float2(hg[q]), d = __bfloat1622float2(hd[q]);
    float hh[2], dv[2], dgt[2];
    const float vv[2] = {v.x, v.y}, gg[2] = {g.x, g.y}, dd[2] = {d.x, d.y};
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float phi_c = 0.5f * (1.0f + erff(gg[k] * 0.70710678118654752f));            // Phi(g)
      const float pdf = 0.3989422804014327f * __expf(-0.5f * gg[k] * gg[k]);               // phi(g)
      const float gelu = gg[k] * phi_c;
      hh[k] = vv[k] * gelu;
      dv[k] = dd[k] * gelu;
      dgt[k] = dd[k] * vv[k] * (phi_c + gg[k] * pdf);
    }
    oh[q] = pack_bf16x2(hh[0], hh[1]);
    ov[q] = pack_bf16x2(dv[0], dv[1]);
    og[q] = pack_bf16x2(dgt[0], dgt[1]);
  }
  if (h != nullptr) *(reinterpret_cast<uint4*>(h + r * ldh) + c) = make_uint4(oh[0], oh[1], oh[2], oh[3]);
  *(reinterpret_cast<uint4*>(dpre + r * lddpre) + c) = make_uint4(ov[0], ov[1], ov[2], ov[3]);
  *(reinterpret_cast<uint4*>(dpre + r * lddpre + inner) + c) = make_uint4(og[0], og[1], og[2], og[3]);
}

}  // namespace
}  // namespace tvi

extern "C" int tvi_geglu_bwd_bf16(const tvi_bf16* pre, const tvi_bf16* dh, tvi_bf16* h, tvi_bf16* dpre, int64_t M, int inner,
                                  int64_t ldpre, int64_t lddh, int64_t ldh, int64_t lddpre, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(M >= 0 && inner >= 0, "tvi_geglu_bwd_bf16: negative size");
  if (M == 0 || inner == 0) return TVI_OK;
  TVI_CHECK_ARG(pre && dh && dpre, "tvi_geglu_bwd_bf16: null pointer");
  TVI_CHECK_ARG(inner % 8 == 0, "tvi_geglu_bwd_bf16: inner=%d must be a multiple of 8", inner);
  TVI_CHECK_ARG(ldpre % 8 == 0 && lddh % 8 == 0 && lddpre % 8 == 0 && ldpre >= 2 * inner && lddpre >= 2 * inner && lddh >= inner &&
                    (h == nullptr || (ldh % 8 == 0 && ldh >= inner)),
                "tvi_geglu_bwd_bf16: bad row strides");
  TVI_CHECK_ARG(((uintptr_t)pre | (uintptr_t)dh | (uintptr_t)h | (uintptr_t)dpre) % 16 == 0, "tvi_geglu_bwd_bf16: pointers must be 16-byte aligned");
  const int64_t total = M * (inner / 8);
  TVI_CHECK_ARG((total + 255) / 256 < ((int64_t)1 << 31), "tvi_geglu_bwd_bf16: grid too large");
  geglu_bwd_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pre, dh, h, dpre, M, inner / 8, ldpre, lddh, ldh,
                                                                                      lddpre);
  TVI_LAUNCH_OK("geglu_bwd_kernel");
  return TVI_OK;
}
