// Per-denoising-step glue of the pipeline loop (row N4 of SURVEY.md 8.6) as ONE kernel per step:
//   src/pipelines/pipeline_stable_video_diffusion_efficient_interp.py
//     :565-566  noise_pred = uncond + guidance_scale[frame] * (cond - uncond)          (per-frame CFG)
//     :569      latents = scheduler.step(noise_pred, t, latents).prev_sample            (EulerDiscreteScheduler.step,
//               diffusers 0.24 schedulers/scheduling_euler_discrete.py, s_churn = 0: un-vendored, restated)
//     :542-546  next iteration: cat([latents] * 2) -> scale_model_input (x / sqrt(sigma^2 + 1)) -> cat with
//               image_latents over channels
// The reference runs these as ~12 elementwise torch kernels over a 2 MB tensor (launch-latency bound).  Here the latents
// stay in an fp32 master copy (updated in place), the Euler arithmetic is fp32 as in the scheduler's upcast, and the
// bf16 model input of the NEXT step is written by the same thread: channels [0, C) of both CFG halves; channels
// [C, Cin) (image latents + mask, constant over the loop) are written once by tvi_model_input_bf16.
#include "tvi_common.cuh"

namespace tvi {
namespace {

// one thread = 4 consecutive spatial positions of one (b, t, c) plane
__global__ void __launch_bounds__(256)
cfg_euler_step_kernel(const tvi_bf16* __restrict__ noise_pred, float* __restrict__ latents, const float* __restrict__ guidance,
                      tvi_bf16* __restrict__ model_input, int64_t B, int T, int C, int Cin, int64_t HW4, float sigma,
                      float sigma_next, int v_prediction, int64_t guidance_stride_b, int round_bf16) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = B * T * C * HW4;
  if (idx >= total) return;
  const int64_t s4 = idx % HW4;
  const int64_t plane = idx / HW4;  // (b * T + t) * C + c
  const int c = (int)(plane % C);
  const int64_t bt = plane / C;
  const int t = (int)(bt % T);
  const int64_t b = bt / T;
  const float gs = __ldg(guidance + b * guidance_stride_b + t);
  const int64_t HW = HW4 * 4;
  const int64_t off = plane * HW + s4 * 4;
  const uint2 un = __ldg(reinterpret_cast<const uint2*>(noise_pred + off));                      // uncond half
  const uint2 co = __ldg(reinterpret_cast<const uint2*>(noise_pred + B * T * C * HW + off));     // cond half
  float4 x = *reinterpret_cast<const float4*>(latents + off);
  const __nv_bfloat162* hu = reinterpret_cast<const __nv_bfloat162*>(&un);
  const __nv_bfloat162* hc = reinterpret_cast<const __nv_bfloat162*>(&co);
  const float2 u0 = __bfloat1622float2(hu[0]), u1 = __bfloat1622float2(hu[1]);
  const float2 c0 = __bfloat1622float2(hc[0]), c1 = __bfloat1622float2(hc[1]);
  float e[4] = {u0.x + gs * (c0.x - u0.x), u0.y + gs * (c0.y - u0.y), u1.x + gs * (c1.x - u1.x), u1.y + gs * (c1.y - u1.y)};
  auto rn = [](float v) { return __bfloat162float(__float2bfloat16_rn(v)); };
  if (round_bf16) {
    // the reference loop's arithmetic types: CFG in the model dtype, every torch op rounded (:565-566)
    const float uu[4] = {u0.x, u0.y, u1.x, u1.y}, cc[4] = {c0.x, c0.y, c1.x, c1.y};
    const float gr = rn(gs);
#pragma unroll
    for (int i = 0; i < 4; ++i) e[i] = rn(uu[i] + rn(gr * rn(cc[i] - uu[i])));
  }
  float xs[4] = {x.x, x.y, x.z, x.w};
  const float s2p1 = sigma * sigma + 1.0f;
  const float dt = sigma_next - sigma;
  const float in_scale = rsqrtf(sigma_next * sigma_next + 1.0f);
  uint32_t packed[2];
  float out[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    // pred_original_sample: epsilon -> x - sigma * e;  v_prediction -> e * (-sigma / sqrt(sigma^2+1)) + x / (sigma^2+1)
    const float x0 = v_prediction ? e[i] * (-sigma / sqrtf(s2p1)) + xs[i] / s2p1 : xs[i] - sigma * e[i];
    const float derivative = (xs[i] - x0) / sigma;
    out[i] = xs[i] + derivative * dt;
    if (round_bf16) out[i] = rn(out[i]);  // EulerDiscreteScheduler.step: prev_sample.to(model_output.dtype)
  }
  *reinterpret_cast<float4*>(latents + off) = make_float4(out[0], out[1], out[2], out[3]);
  if (model_input != nullptr) {
    if (round_bf16) {  // scale_model_input divides (`sample / ((sigma**2 + 1) ** 0.5)`), in the model dtype
      const float den = sqrtf(sigma_next * sigma_next + 1.0f);
      packed[0] = pack_bf16x2(out[0] / den, out[1] / den);
      packed[1] = pack_bf16x2(out[2] / den, out[3] / den);
    } else {
      packed[0] = pack_bf16x2(out[0] * in_scale, out[1] * in_scale);
      packed[1] = pack_bf16x2(out[2] * in_scale, out[3] * in_scale);
    }
    const int64_t moff = (bt * Cin + c) * HW + s4 * 4;
    *reinterpret_cast<uint2*>(model_input + moff) = make_uint2(packed[0], packed[1]);
    *reinterpret_cast<uint2*>(model_input + B * T * Cin * HW + moff) = make_uint2(packed[0], packed[1]);
  }
}

// model_input[h, b, t, c, :] = c < C ? latents[b, t, c, :] * scale : image_latents[h, b, t, c - C, :]   (h = CFG half)
__global__ void __launch_bounds__(256)
model_input_kernel(const float* __restrict__ latents, const tvi_bf16* __restrict__ image_latents, tvi_bf16* __restrict__ out,
                   int64_t BT, int C, int Cin, int64_t HW4, float scale) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = 2 * BT * Cin * HW4;
  if (idx >= total) return;
  const int64_t s4 = idx % HW4;
  const int64_t plane = idx / HW4;  // (h * BT + bt) * Cin + c
  const int c = (int)(plane % Cin);
  const int64_t hbt = plane / Cin;
  const int64_t bt = hbt % BT;
  const int64_t HW = HW4 * 4;
  uint2 v;
  if (c < C) {
    const float4 x = *reinterpret_cast<const float4*>(latents + (bt * C + c) * HW + s4 * 4);
    v = make_uint2(pack_bf16x2(x.x * scale, x.y * scale), pack_bf16x2(x.z * scale, x.w * scale));
  } else {
    v = __ldg(reinterpret_cast<const uint2*>(image_latents + (hbt * (Cin - C) + (c - C)) * HW + s4 * 4));
  }
  *reinterpret_cast<uint2*>(out + plane * HW + s4 * 4) = v;
}

}  // namespace
}  // namespace tvi

extern "C" int tvi_cfg_euler_step_f32(const tvi_bf16* noise_pred, float* latents, const float* guidance, tvi_bf16* model_input,
                                      int64_t B, int T, int C, int Cin, int64_t HW, int64_t guidance_stride_b, float sigma,
                                      float sigma_next, int prediction_type, void* stream) {
  return tvi_cfg_euler_step_round_f32(noise_pred, latents, guidance, model_input, B, T, C, Cin, HW, guidance_stride_b, sigma, sigma_next,
                                      prediction_type, TVI_LATENT_ROUND_NONE, stream);
}

extern "C" int tvi_cfg_euler_step_round_f32(const tvi_bf16* noise_pred, float* latents, const float* guidance, tvi_bf16* model_input,
                                            int64_t B, int T, int C, int Cin, int64_t HW, int64_t guidance_stride_b, float sigma,
                                            float sigma_next, int prediction_type, int latent_rounding, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(latent_rounding == TVI_LATENT_ROUND_NONE || latent_rounding == TVI_LATENT_ROUND_BF16,
                "tvi_cfg_euler_step_f32: unknown latent rounding mode %d", latent_rounding);
  TVI_CHECK_ARG(B >= 0 && T >= 0 && C >= 0 && HW >= 0, "tvi_cfg_euler_step_f32: negative size");
  if (B == 0 || T == 0 || C == 0 || HW == 0) return TVI_OK;
  TVI_CHECK_ARG(noise_pred && latents && guidance, "tvi_cfg_euler_step_f32: null pointer");
  TVI_CHECK_ARG(HW % 4 == 0, "tvi_cfg_euler_step_f32: H*W=%lld must be a multiple of 4", (long long)HW);
  TVI_CHECK_ARG(model_input == nullptr || Cin >= C, "tvi_cfg_euler_step_f32: Cin=%d smaller than C=%d", Cin, C);
  TVI_CHECK_ARG(sigma > 0.f && sigma_next >= 0.f, "tvi_cfg_euler_step_f32: sigma must be positive");
  TVI_CHECK_ARG(prediction_type == TVI_PRED_EPSILON || prediction_type == TVI_PRED_V, "tvi_cfg_euler_step_f32: unknown prediction type %d", prediction_type);
  TVI_CHECK_ARG((uintptr_t)noise_pred % 8 == 0 && (uintptr_t)latents % 16 == 0 && (uintptr_t)model_input % 8 == 0,
                "tvi_cfg_euler_step_f32: misaligned pointer");
  const int64_t total = B * T * C * (HW / 4);
  TVI_CHECK_ARG((total + 255) / 256 < ((int64_t)1 << 31), "tvi_cfg_euler_step_f32: grid too large");
  cfg_euler_step_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      noise_pred, latents, guidance, model_input, B, T, C, Cin, HW / 4, sigma, sigma_next, prediction_type == TVI_PRED_V ? 1 : 0,
      guidance_stride_b, latent_rounding == TVI_LATENT_ROUND_BF16 ? 1 : 0);
  TVI_LAUNCH_OK("cfg_euler_step_kernel");
  return TVI_OK;
}

extern "C" int tvi_model_input_bf16(const float* latents, const tvi_bf16* image_latents, tvi_bf16* model_input, int64_t BT, int C,
                                    int Cin, int64_t HW, float scale, void* stream) {
  using namespace tvi;
  TVI_CHECK_ARG(BT >= 0 && C >= 0 && HW >= 0, "tvi_model_input_bf16: negative size");
  if (BT == 0 || Cin == 0 || HW == 0) return TVI_OK;
  TVI_CHECK_ARG(latents && model_input && (Cin == C || image_latents), "tvi_model_input_bf16: null pointer");
  TVI_CHECK_ARG(Cin >= C && HW % 4 == 0, "tvi_model_input_bf16: Cin >= C and H*W %% 4 == 0 are required");
  TVI_CHECK_ARG((uintptr_t)latents % 16 == 0 && (uintptr_t)image_latents % 8 == 0 && (uintptr_t)model_input % 8 == 0,
                "tvi_model_input_bf16: misaligned pointer");
  const int64_t total = 2 * BT * Cin * (HW / 4);
  TVI_CHECK_ARG((total + 255) / 256 < ((int64_t)1 << 31), "tvi_model_input_bf16: grid too large");
  model_input_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(latents, image_latents, model_input, BT, C, Cin,
                                                                                        HW / 4, scale);
  TVI_LAUNCH_OK("model_input_kernel");
  return TVI_OK;
}
