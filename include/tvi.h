/* libtvi — C ABI of the B200-native (sm_100a) hot path for SVD keyframe interpolation.
 *
 * This is the drop-in boundary: every symbol below is what the reference's
 * Python code would bind through ctypes in place of a torch/xformers/numpy call.  Citations are
 * path:line inside the reference tree (dreamyou070/Timelapse_Video_Interpolation).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the parameter name ends in `_host`;
 *   - bf16 tensors are passed as `const tvi_bf16*` (raw 16-bit storage, torch.bfloat16 layout);
 *   - all leading dimensions / strides are in ELEMENTS, rows are tokens, heads are packed along
 *     the row (head h occupies columns [h*d, (h+1)*d)), so a fused [tokens, 3C] QKV buffer is
 *     consumed in place with ldq = ldk = ldv = 3C and three base pointers;
 *   - `stream` is a cudaStream_t (0 = legacy default stream); nothing synchronises the host;
 *   - nothing allocates: outputs and workspaces are caller-provided;
 *   - return value: 0 on success, a negative TVI_ERR_* otherwise; the message is available from
 *     tvi_last_error() (thread-local).  There is NO CPU fallback anywhere behind this ABI.
 */
#ifndef TVI_H_
#define TVI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint16_t tvi_bf16;

#define TVI_OK 0
#define TVI_ERR_INVALID (-1)     /* bad argument (shape, alignment, null pointer)          */
#define TVI_ERR_CUDA (-2)        /* a CUDA runtime/driver call or a kernel launch failed    */
#define TVI_ERR_UNSUPPORTED (-3) /* valid request this build has no kernel for (e.g. d=96)  */

/* Library identity / diagnostics. */
const char* tvi_version(void);
const char* tvi_last_error(void);
/* Fills SM count and compute capability of the current device. */
int tvi_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---------------------------------------------------------------------------------------------
 * A1  first/last-frame latent slerp.
 * Replaces: frame_slerp_interp.py (0 bytes in the reference; unused flags at
 * src/train_decouples.py:1566-1567, src/test.py:648-649).  Spec: SURVEY.md §8.1 row A1.
 *   v0, v1 : [nclips, n] fp32      out : [nclips, T, n] fp32
 *   c = <v0,v1>/(|v0||v1|) per clip (fp64 accumulation); |c| > dot_threshold -> lerp, else slerp
 *   with t_i = i/(T-1).  T >= 2, n % 4 == 0, pointers 16-byte aligned.
 */
int tvi_slerp_f32(const float* v0, const float* v1, float* out, int64_t nclips, int64_t n, int T,
                  float dot_threshold, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A6  mean per-pixel cosine similarity of two optical-flow fields.
 * Replaces: cosine_similarity(flow_a, flow_b), dataset_construction/step4_flow_check.py:25-34.
 *   fa, fb : [npairs, H, W, 2] fp32 if hwc_layout != 0 (OpenCV layout, step4_flow_check.py:22),
 *            [npairs, 2, H, W] otherwise.     out : [npairs] fp32.
 *   workspace : tvi_flow_cosine_workspace_bytes(npairs, H, W) bytes, 8-byte aligned.
 */
size_t tvi_flow_cosine_workspace_bytes(int64_t npairs, int H, int W);
int tvi_flow_cosine_f32(const float* fa, const float* fb, float* out, void* workspace, int64_t npairs, int H,
                        int W, int hwc_layout, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A7  backward warp + forward/backward consistency check (north-star extension; the reference
 * file dataset_construction/step4_flow_check.py has no warp — spec in SURVEY.md §8.1 row A7,
 * oracle = torch.nn.functional.grid_sample(bilinear, zeros, align_corners=True)).
 *   img2 : [B, C, H, W] fp32   f_fwd, f_bwd : [B, 2, H, W] fp32 (x then y displacement, pixels)
 *   warped : [B, C, H, W] fp32  = bilinear sample of img2 at p + f_fwd(p), zero outside
 *   mask   : [B, H, W] u8, bit0 = occluded: |f_fwd + b~|^2 > alpha(|f_fwd|^2 + |b~|^2) + beta with
 *            b~ = f_bwd sampled at p + f_fwd(p); bit1 = sample point inside the image.
 */
int tvi_backward_warp_f32(const float* img2, const float* f_fwd, const float* f_bwd, float* warped,
                          uint8_t* mask, int64_t B, int C, int H, int W, float alpha, float beta,
                          void* stream);

/* ---------------------------------------------------------------------------------------------
 * A2/A3  projections: Y[M,N] = X[M,K] . W[N,K]^T (+ bias[N]), bf16 in/out, fp32 accumulation
 * (tcgen05.mma, TMEM accumulators, TMA-fed).
 * Replaces: attn.to_q / to_k / to_v / to_out[0] (nn.Linear -> cuBLAS),
 * src/models/attention_processor.py:729,736-737,754 and :1223,1230-1231,1251.
 *   ldx, ldw, ldy: row strides in elements (multiples of 8); K % 8 == 0; N % 8 == 0;
 *   X, W, Y 16-byte aligned; bias nullable.
 */
int tvi_gemm_bf16_tn(const tvi_bf16* X, const tvi_bf16* W, const tvi_bf16* bias, tvi_bf16* Y, int64_t M, int N,
                     int K, int64_t ldx, int64_t ldw, int64_t ldy, void* stream);

/* N1: the same GEMM with a fused epilogue, for FeedForward / GEGLU and residual adds.
 * Replaces: FeedForward.forward (src/models/attention.py:481-535) = GEGLU.proj (nn.Linear dim -> 2*inner),
 * chunk(2), hidden * gelu(gate), net[2] (nn.Linear inner -> dim), and the `ff_output + hidden_states` residual
 * adds of BasicTransformerBlock / TemporalBasicTransformerBlock (:343, :449, :472).
 *   epilogue NONE : Y[M,N]   = X . W^T + bias + residual
 *   epilogue GEGLU: Y[M,N/2] = (X . Wv^T + bv) * gelu(X . Wg^T + bg) + residual, where W[N,K] / bias[N] hold the
 *                   value and gate rows interleaved in blocks of H = tvi_geglu_block(N/2): rows
 *                   [2Hb, 2Hb+H) = value rows [Hb, Hb+H), rows [2Hb+H, 2Hb+2H) = the gate rows of the same
 *                   outputs (the host re-orders GEGLU.proj.weight once; erf-form gelu in fp32).
 *   bias, residual nullable; residual is [M, N or N/2] with row stride ldr.
 */
#define TVI_EPILOGUE_NONE 0
#define TVI_EPILOGUE_GEGLU 1
int tvi_geglu_block(int inner);
int tvi_gemm_bf16_tn_ex(const tvi_bf16* X, const tvi_bf16* W, const tvi_bf16* bias, const tvi_bf16* residual, tvi_bf16* Y,
                        int64_t M, int N, int K, int64_t ldx, int64_t ldw, int64_t ldr, int64_t ldy, int epilogue,
                        void* stream);

/* LayerNorm folded into the GEMM that consumes it (SURVEY 8.6 N1 "LN prologue"; the LayerNorm -> Linear pairs of
 * src/models/attention.py:255+:276 (norm1 -> to_q/k/v), :321+:343 (norm3 -> ff), :441-446 (norm_in -> ff_in), :451, :462):
 *   LN(x) W^T + b = rstd * (x Wg^T - mean * colsum) + bias',  Wg = gamma (.) W (bf16),  colsum_n = sum_k Wg[n, k] (fp32,
 *   summed from the ROUNDED Wg so the mean term cancels exactly),  bias'_n = W_n . beta + b_n.
 * tvi_layernorm_stats_f32 writes stats[M][2] = (mean, rstd) in one read-only pass (C in {320, 640, 1280});
 * tvi_gemm_bf16_tn_ln is tvi_gemm_bf16_tn_ex on the RAW rows X with that affine map in the epilogue (before GEGLU /
 * the residual).  The normalised activations never reach HBM. */
int tvi_layernorm_stats_supported(int C);
int tvi_layernorm_stats_f32(const tvi_bf16* x, float* stats, int64_t M, int C, int64_t ldx, float eps, void* stream);
int tvi_gemm_bf16_tn_ln(const tvi_bf16* X, const tvi_bf16* Wg, const tvi_bf16* bias, const float* colsum, const float* stats,
                        const tvi_bf16* residual, tvi_bf16* Y, int64_t M, int N, int K, int64_t ldx, int64_t ldw, int64_t ldr,
                        int64_t ldy, int epilogue, void* stream);

/* LayerNorm over the channel dimension of token-major activations (fp32 statistics, bf16 in/out).
 * Replaces: nn.LayerNorm norm1/norm2/norm3 and norm_in of the transformer blocks
 * (src/models/attention.py:270, :300, :321 and :441, :446, :462).  gamma, beta nullable; C % 8 == 0, C <= 2048. */
int tvi_layernorm_bf16(const tvi_bf16* x, const tvi_bf16* gamma, const tvi_bf16* beta, tvi_bf16* y, int64_t M, int C,
                       int64_t ldx, int64_t ldy, float eps, void* stream);

/* out[r, :] = v[group(r), :] (+ residual[r, :], nullable) — the one-key cross-attention output (a row broadcast) with
 * the block's `attn_output + hidden_states` (src/models/attention.py:316-317, :459-460) fused.
 * group(r) = (r / (group_inner * rows_per_group)) * group_inner + r % group_inner: group_inner == 1 is the contiguous
 * [group, row] order; group_inner == S addresses the temporal groups (b, s) straight from the [B, T, S] token order,
 * which folds the transposes of src/models/attention.py:436-438 / :474-476 away (row N2). */
int tvi_bcast_rows_add_bf16(const tvi_bf16* v, const tvi_bf16* residual, tvi_bf16* out, int64_t nrows, int64_t rows_per_group,
                            int64_t group_inner, int C, int64_t ldv, int64_t ldr, int64_t ldo, void* stream);

/* Backward of the N1 row (training config 5 through the block-level swap; autograd's native_layer_norm_backward and the
 * backward of diffusers' GEGLU.forward):
 *   layernorm_bwd : dx (nullable), dgamma, dbeta (nullable) from x, dy, gamma; statistics are recomputed from x;
 *       `workspace` = tvi_layernorm_bwd_workspace_bytes(M, C) bytes of fp32 partials, summed in a fixed order.
 *   geglu_bwd     : pre = [value | gate] ([M, 2*inner], the recomputed first projection), dh [M, inner] ->
 *       h = value * gelu(gate) (nullable) and dpre = [dh * gelu(gate) | dh * value * gelu'(gate)] ([M, 2*inner]).
 */
size_t tvi_layernorm_bwd_workspace_bytes(int64_t M, int C);
int tvi_layernorm_bwd_bf16(const tvi_bf16* x, const tvi_bf16* dy, const tvi_bf16* gamma, tvi_bf16* dx, tvi_bf16* dgamma,
                           tvi_bf16* dbeta, void* workspace, int64_t M, int C, int64_t ldx, int64_t lddy, int64_t lddx, float eps,
                           void* stream);
/* The same with the gradient of a residual branch added in the same pass: for `y = x + f(LayerNorm(x))` (every sub-layer
 * of the two blocks, src/models/attention.py:270-287, :300-317, :321-343) dx = LayerNorm'(dy_norm) + dres, where dres is
 * the gradient that reached the `+ x` (nullable: then this is tvi_layernorm_bwd_bf16). */
int tvi_layernorm_bwd_res_bf16(const tvi_bf16* x, const tvi_bf16* dy, const tvi_bf16* gamma, const tvi_bf16* dres, tvi_bf16* dx,
                               tvi_bf16* dgamma, tvi_bf16* dbeta, void* workspace, int64_t M, int C, int64_t ldx, int64_t lddy,
                               int64_t lddres, int64_t lddx, float eps, void* stream);
int tvi_geglu_bwd_bf16(const tvi_bf16* pre, const tvi_bf16* dh, tvi_bf16* h, tvi_bf16* dpre, int64_t M, int inner, int64_t ldpre,
                       int64_t lddh, int64_t ldh, int64_t lddpre, void* stream);
/* The two steps above in one kernel: the backward of `Linear(K -> 2*inner) -> GEGLU` (src/models/attention.py:526-535 under
 * autograd, the first two modules of FeedForward.net) recomputes the projection on the tensor cores and turns the fp32
 * accumulator straight into H = value * gelu(gate) [M, inner] (ldh) and dPre = [dH * gelu(gate) | dH * value * gelu'(gate)]
 * [M, 2*inner] (lddpre, in the un-interleaved [value | gate] column order of the module's weight).  X [M, K] (ldx);
 * Wi / bias_i are the weight [2*inner, K] (ldw) and bias (nullable) with value / gate rows interleaved in blocks of
 * tvi_geglu_block(inner), as tvi_gemm_bf16_tn_ex(..., TVI_EPILOGUE_GEGLU) takes them; dH [M, inner] (lddh) is the gradient
 * that reached H.  H may be NULL (a frozen second projection needs no h).  The 2*inner-wide pre-activation is never
 * written.  inner % 32 == 0, K % 8 == 0. */
int tvi_gemm_bf16_geglu_bwd(const tvi_bf16* X, const tvi_bf16* Wi, const tvi_bf16* bias_i, const tvi_bf16* dH, tvi_bf16* H,
                            tvi_bf16* dPre, int64_t M, int inner, int K, int64_t ldx, int64_t ldw, int64_t lddh, int64_t ldh,
                            int64_t lddpre, void* stream);

/* Layout glue of TransformerSpatioTemporalModel.forward (rows N1 "AlphaBlender" and N2 of SURVEY 8.6).
 *   groupnorm_tokens : y[n, s, c] = GroupNorm(x)[n, c, s] — `self.norm` (nn.GroupNorm, biased variance, fp32/fp64
 *       statistics) fused with `permute(0, 2, 3, 1).reshape(BF, H*W, C)`, src/models/transformer_temporal.py:300-304.
 *       x is NCHW [N, C, HW] contiguous, y token-major [N*HW, C] (ldy); gamma / beta nullable;
 *       `workspace` = tvi_groupnorm_tokens_workspace_bytes(N, C, HW, groups) bytes.  C % 8 == 0, HW % 8 == 0.
 *   tokens_to_nchw_add : out[n, c, s] = y[n, s, c] (+ residual[n, c, s], nullable) — `reshape(BF, H, W, C)
 *       .permute(0, 3, 1, 2).contiguous()` and `output = hidden_states + residual`, :354-357.
 *   blend_rows : out[r, :] = alpha[r / rows_per_group] * x_spatial[r, :] + (1 - alpha[..]) * x_temporal[r, :] —
 *       AlphaBlender.forward (diffusers 0.24 models/resnet.py) called at :349-351; alpha is fp32, one per image.
 */
size_t tvi_groupnorm_tokens_workspace_bytes(int64_t N, int C, int HW, int groups);
int tvi_groupnorm_tokens_bf16(const tvi_bf16* x, const tvi_bf16* gamma, const tvi_bf16* beta, tvi_bf16* y, void* workspace,
                              int64_t N, int C, int HW, int groups, int64_t ldy, float eps, void* stream);
int tvi_tokens_to_nchw_add_bf16(const tvi_bf16* y, const tvi_bf16* residual, tvi_bf16* out, int64_t N, int C, int HW,
                                int64_t ldy, void* stream);
int tvi_blend_rows_bf16(const tvi_bf16* x_spatial, const tvi_bf16* x_temporal, const float* alpha, tvi_bf16* out,
                        int64_t nrows, int64_t rows_per_group, int C, int64_t lds, int64_t ldt, int64_t ldo, void* stream);

/* Backward of the three glue kernels above — what autograd derives for F.group_norm + permute
 * (src/models/transformer_temporal.py:300-304), the permute + residual of :354-357 and AlphaBlender.forward (:349-351)
 * when src/train.py back-propagates through TransformerSpatioTemporalModel.forward.  All deterministic (fixed-order sums).
 *   tvi_nchw_to_tokens_bf16        y[n, s, c] = x[n, c, s]: d tokens of tvi_tokens_to_nchw_add_bf16 (d residual = d out).
 *   tvi_groupnorm_tokens_bwd_bf16  x [N, C, HW] (the forward input), dn [N, HW, C] (lddn) -> dx [N, C, HW] (nullable),
 *                                  dgamma, dbeta [C] bf16 (nullable).  workspace = ..._bwd_workspace_bytes(N, C, HW, groups).
 *   tvi_blend_rows_bwd_bf16        dy -> dx_spatial = alpha dy, dx_temporal = (1 - alpha) dy (row stride ld for all);
 *                                  dalpha_partial (nullable): [frames][tvi_blend_rows_bwd_blocks(rows_per_group, C)] fp32
 *                                  partial sums of dy (x_spatial - x_temporal); the caller adds each frame's row in order. */
int tvi_nchw_to_tokens_bf16(const tvi_bf16* x, tvi_bf16* y, int64_t N, int C, int HW, int64_t ldy, void* stream);
size_t tvi_groupnorm_tokens_bwd_workspace_bytes(int64_t N, int C, int HW, int groups);
int tvi_groupnorm_tokens_bwd_bf16(const tvi_bf16* x, const tvi_bf16* dn, const tvi_bf16* gamma, tvi_bf16* dx, tvi_bf16* dgamma,
                                  tvi_bf16* dbeta, void* workspace, int64_t N, int C, int HW, int groups, int64_t lddn, float eps,
                                  void* stream);
int tvi_blend_rows_bwd_blocks(int64_t rows_per_group, int C);
int tvi_blend_rows_bwd_bf16(const tvi_bf16* dy, const tvi_bf16* x_spatial, const tvi_bf16* x_temporal, const float* alpha,
                            tvi_bf16* dx_spatial, tvi_bf16* dx_temporal, float* dalpha_partial, int64_t frames,
                            int64_t rows_per_group, int C, int64_t ld, void* stream);

/* Per-denoising-step glue of the pipeline loop (row N4 of SURVEY 8.6), one launch per step:
 *   cfg_euler_step : per-frame classifier-free guidance `uncond + guidance[b, t] * (cond - uncond)`
 *       (src/pipelines/pipeline_stable_video_diffusion_efficient_interp.py:565-566), EulerDiscreteScheduler.step with
 *       s_churn = 0 (:569; diffusers 0.24, fp32 as in its upcast) on the fp32 master `latents` [B, T, C, HW] IN PLACE, and
 *       — when `model_input` is given — channels [0, C) of both CFG halves of the NEXT step's UNet input
 *       [2, B, T, Cin, HW] = latents / sqrt(sigma_next^2 + 1) (:542-546).  noise_pred is [2, B, T, C, HW] bf16
 *       (uncond half first, as `noise_pred.chunk(2)`); guidance is fp32, element (b, t) at b * guidance_stride_b + t.
 *   model_input : the whole UNet input of the first step: scaled latents in channels [0, C) and `image_latents`
 *       [2, B*T, Cin - C, HW] (constant over the loop) in channels [C, Cin).
 */
#define TVI_PRED_EPSILON 0
#define TVI_PRED_V 1
int tvi_cfg_euler_step_f32(const tvi_bf16* noise_pred, float* latents, const float* guidance, tvi_bf16* model_input, int64_t B,
                           int T, int C, int Cin, int64_t HW, int64_t guidance_stride_b, float sigma, float sigma_next,
                           int prediction_type, void* stream);
/* The same step with the arithmetic TYPES of the reference loop (latent_rounding = TVI_LATENT_ROUND_BF16): the CFG combine
 * runs in the model dtype with every torch op rounded (:565-566), `prev_sample` is cast back to the model dtype as
 * EulerDiscreteScheduler.step does (`prev_sample.to(model_output.dtype)`), and the next model input is
 * `sample / ((sigma**2 + 1) ** 0.5)` in the model dtype.  `latents` stays an fp32 array whose values are then
 * bf16-representable.  TVI_LATENT_ROUND_NONE is tvi_cfg_euler_step_f32 (fp32 master latents, the default). */
#define TVI_LATENT_ROUND_NONE 0
#define TVI_LATENT_ROUND_BF16 1
int tvi_cfg_euler_step_round_f32(const tvi_bf16* noise_pred, float* latents, const float* guidance, tvi_bf16* model_input, int64_t B,
                                 int T, int C, int Cin, int64_t HW, int64_t guidance_stride_b, float sigma, float sigma_next,
                                 int prediction_type, int latent_rounding, void* stream);
int tvi_model_input_bf16(const float* latents, const tvi_bf16* image_latents, tvi_bf16* model_input, int64_t BT, int C, int Cin,
                         int64_t HW, float scale, void* stream);

/* Dense optical flow after Farneback (row N3 of SURVEY 8.6): the computation of cv2.calcOpticalFlowFarneback(prev, next,
 * None, pyr_scale, levels, winsize, iterations, poly_n, poly_sigma, flags=0) as called at
 * dataset_construction/step4_flow_check.py:13-22 (0.5, 3, 15, 3, 5, 1.2), including the cv2.cvtColor(BGR2GRAY) in front
 * of it when channels == 3.  prev / next: 8-bit device images [H, W, channels] (1 = gray, 3 = BGR); flow: [H, W, 2] fp32
 * (dx, dy); `workspace` = tvi_farneback_workspace_bytes(H, W, pyr_scale, levels) bytes.  ~20 launches per pyramid level, no host sync.
 */
size_t tvi_farneback_workspace_bytes(int H, int W, double pyr_scale, int levels);
int tvi_farneback_f32(const uint8_t* prev, const uint8_t* next, int channels, float* flow, void* workspace, int H, int W,
                      double pyr_scale, int levels, int winsize, int iterations, int poly_n, double poly_sigma, void* stream);
/* The flows of every consecutive pair of a clip (the loop at step4_flow_check.py:55-70): frames [nframes, H, W, channels],
 * flows [nframes - 1, H, W, 2].  Each frame's blur / polynomial-expansion pyramid is built once and serves as "next" of
 * one pair and "prev" of the following one.  `second`: NULL, or a separately stored second frame when nframes == 2. */
int tvi_farneback_clip_f32(const uint8_t* frames, const uint8_t* second, int64_t nframes, int channels, float* flows,
                           void* workspace, int H, int W, double pyr_scale, int levels, int winsize, int iterations, int poly_n,
                           double poly_sigma, void* stream);

/* Backward of the projections (training config 5: src/train_decouples.py back-propagates through the
 * same nn.Linear call sites; autograd's F.linear backward = two cuBLAS GEMMs + a bias column sum).
 *   dgrad : dX[M,K] = dY[M,N] . W[N,K]          (W consumed in place, MN-major tcgen05 operand)
 *   wgrad : dW[N,K] = dY[M,N]^T . X[M,K]        (token reduction split across CTAs; fp32 partials in
 *           `workspace` = tvi_gemm_bf16_wgrad_workspace_bytes(M,N,K) bytes, summed in a fixed order)
 *   colsum: out[g, c] = sum over the rows_per_group rows of group g of x[row, c]  (bias gradient with
 *           one group; also the backward of the one-key cross-attention broadcast).  fp32 accumulation,
 *           `workspace` = tvi_colsum_bf16_workspace_bytes(nrows, rows_per_group, C) bytes.
 */
int tvi_gemm_bf16_dgrad(const tvi_bf16* dY, const tvi_bf16* W, tvi_bf16* dX, int64_t M, int N, int K, int64_t lddy,
                        int64_t ldw, int64_t lddx, void* stream);
size_t tvi_gemm_bf16_wgrad_workspace_bytes(int64_t M, int N, int K);
int tvi_gemm_bf16_wgrad(const tvi_bf16* dY, const tvi_bf16* X, tvi_bf16* dW, void* workspace, int64_t M, int N, int K,
                        int64_t lddy, int64_t ldx, int64_t lddw, void* stream);
size_t tvi_colsum_bf16_workspace_bytes(int64_t nrows, int64_t rows_per_group, int C);
int tvi_colsum_bf16(const tvi_bf16* x, tvi_bf16* out, void* workspace, int64_t nrows, int64_t rows_per_group, int C,
                    int64_t ldx, int64_t ldo, void* stream);
/* The same over interleaved groups: logical row i of group (b, s), s < group_inner, is physical row
 * (b * rows_per_group + i) * group_inner + s (backward of tvi_bcast_rows_add_bf16 with group_inner > 1). */
int tvi_colsum_bf16_interleaved(const tvi_bf16* x, tvi_bf16* out, void* workspace, int64_t nrows, int64_t rows_per_group,
                                int64_t group_inner, int C, int64_t ldx, int64_t ldo, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A3c  short-sequence (temporal) attention core: nseq independent sequences, every head,
 * O = softmax(scale . Q K^T) V with fp32 softmax.  Sequence s owns query rows [s*Tq, (s+1)*Tq)
 * and key/value rows [s*Tk, (s+1)*Tk).  1 <= Tq, Tk <= 64, d in {64, 128}.
 * Replaces: F.scaled_dot_product_attention / xformers / baddbmm+softmax+bmm on the
 * [B*S, T, C] operands, src/models/attention_processor.py:1243, :1156, :598-615+:750, called from
 * src/models/attention.py:451-453.
 *   lse (nullable): [nseq*Tq, heads] fp32 natural-log-sum-exp of the scaled logits (for backward).
 */
int tvi_attn_temporal_fwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse,
                          int64_t nseq, int heads, int Tq, int Tk, int d, int64_t ldq, int64_t ldk,
                          int64_t ldv, int64_t ldo, float scale, void* stream);

/* The same core on INTERLEAVED sequences (row N2 of SURVEY 8.6): token t of sequence s is row
 *   (s / seq_inner) * seq_inner * T + t * seq_inner + s % seq_inner
 * of q / k / v / o (lse is indexed by the same row).  seq_inner == 1 is the layout above; seq_inner == S lets the
 * temporal attention of TemporalBasicTransformerBlock read and write the [B*T, S, C] activations of the spatial
 * block in place, replacing the two full-activation permute copies `[B*T,S,C] -> [B*S,T,C]` and back at
 * src/models/attention.py:436-438 and :474-476 (every other op of that block is row-wise).  nseq % seq_inner == 0. */
int tvi_attn_temporal_fwd_interleaved(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse,
                                      int64_t nseq, int heads, int Tq, int Tk, int d, int64_t seq_inner, int64_t ldq,
                                      int64_t ldk, int64_t ldv, int64_t ldo, float scale, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A3b/A3d  few-key cross-attention core (image-embedding context, 1..8 keys): query row r attends
 * to the L key/value rows of context group (r / rows_per_group).
 * Replaces the same call sites as above for attn2 (src/models/attention.py:311-316, :457-459) and
 * the patched temporal forward src/train_decouples.py:377-495 (additive ramp bias :437-457 and
 * per-frame mean-logit statistics :460-471).
 *   q : [nrows, heads*d] (ldq)   k, v : [ngroups*L, heads*d] (ldk, ldv)   o : [nrows, heads*d] (ldo)
 *   bias  (nullable): [Tq, L] fp32 added to the scaled logits of query row r at index r % Tq
 *   stats (nullable): [Tq, L] fp32, receives the mean over (rows with r % Tq == t, heads) of the
 *                     pre-softmax logits (after bias).  stats_ws: Tq*L doubles (zeroed by the call).
 *   Tq is only used to index bias/stats (pass 1 if both are null).  d in {64, 128}, L <= 8.
 */
int tvi_attn_fewkeys_fwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, int64_t nrows,
                         int64_t rows_per_group, int heads, int L, int d, int64_t ldq, int64_t ldk,
                         int64_t ldv, int64_t ldo, float scale, int Tq, const float* bias, float* stats,
                         void* stats_ws, void* stream);
/* Same over INTERLEAVED groups: logical row i of group g = (b, s), s < group_inner, is physical row
 * (b * rows_per_group + i) * group_inner + s of q / o — the temporal block's cross-attention (rows_per_group = T frames,
 * group_inner = H*W) read straight from the [B*T, S, C] token order, without the permute copies of
 * src/models/attention.py:436-438 / :474-476.  bias / stats index the logical row (i % Tq).  group_inner == 1 is the call above. */
int tvi_attn_fewkeys_fwd_interleaved(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, int64_t nrows,
                                     int64_t rows_per_group, int heads, int L, int d, int64_t group_inner, int64_t ldq,
                                     int64_t ldk, int64_t ldv, int64_t ldo, float scale, int Tq, const float* bias,
                                     float* stats, void* stats_ws, void* stream);

/* No-allocation variant of tvi_attn_fewkeys_fwd with statistics.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments and results as tvi_attn_fewkeys_fwd minus `stats_ws`: the
 * Tq*L fp64 accumulators live in a static device array inside the library (TVI_FEWKEYS_STATIC_MAX_STATS entries =
 * 32 KB per GPU that loads the library, held until exit).  Tq*L beyond that limit -> TVI_ERR_INVALID.  The call
 * allocates nothing (no cudaMalloc / cudaFree / allocating library routine); the caller passes inputs and outputs.
 * All calls on a device share the array, which keeps the last call's contents: the entries a call uses are cleared
 * on `stream` first, and calls that OVERLAP on the GPU (different streams) race on it — serialise them. */
#define TVI_FEWKEYS_STATIC_MAX_STATS 4096
int tvi_attn_fewkeys_fwd_static(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, int64_t nrows,
                                int64_t rows_per_group, int heads, int L, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                                int64_t ldo, float scale, int Tq, const float* bias, float* stats, void* stream);

/* ---------------------------------------------------------------------------------------------
 * A3a  spatial self-attention core (flash-style, tcgen05 QK^T and PV, online fp32 softmax):
 * batch independent sequences of S query rows and Skv key rows, no mask, non-causal.
 * Replaces: F.scaled_dot_product_attention / xformers.ops.memory_efficient_attention on
 * [BF, h, S, 64], src/models/attention_processor.py:1243, :1156 (called from
 * src/models/attention.py:276-281).  d in {64, 128} (128: the mma.sync kernel; TVI_SPATIAL_TCGEN05 is built for d == 64).
 *   lse (nullable): [batch*S, heads] fp32.
 */
int tvi_attn_spatial_fwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse,
                         int batch, int heads, int S, int Skv, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                         int64_t ldo, float scale, void* stream);

/* Same call with an explicit kernel choice (parity tests exercise both; AUTO is what
 * tvi_attn_spatial_fwd uses: tcgen05 when S and Skv are >= 1024, mma.sync otherwise).
 * TVI_SPATIAL_TCGEN05: persistent TMA-fed tcgen05.mma kernel, S/P/O in tensor memory, work items of 256 query rows.
 * TVI_SPATIAL_MMA_SYNC: register-resident flash kernel on mma.sync (short sequences). */
#define TVI_SPATIAL_AUTO 0
#define TVI_SPATIAL_MMA_SYNC 1
#define TVI_SPATIAL_TCGEN05 2
int tvi_attn_spatial_fwd_algo(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, tvi_bf16* o, float* lse,
                              int batch, int heads, int S, int Skv, int d, int64_t ldq, int64_t ldk, int64_t ldv,
                              int64_t ldo, float scale, int algo, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Backward of the attention cores (training config 5; what autograd derives for the call sites of
 * rows A3a-A3d: src/models/attention_processor.py:598-615+:750, :1156, :1243).  With P = softmax(scale QK^T):
 *   dV = P^T dO,  dS = P o (dO V^T - rowsum(dO o O)),  dQ = scale dS K,  dK = scale dS^T Q.
 * All gradients are bf16 with the layout of the tensor they belong to (so dq/dk/dv may be the three
 * column slices of one packed [tokens, 3C] buffer, lddq = lddk = lddv = 3C).
 *
 * spatial : flash-style recomputation from q, k, v and the forward's `lse`; `o` is the forward output.
 *           workspace = tvi_attn_spatial_bwd_workspace_bytes(batch, heads, S) bytes.  d in {64, 128} (128: mma.sync passes).
 * temporal: T <= 64, needs neither o nor lse.  d in {64, 128}.
 * fewkeys : dk, dv are [ngroups*L, heads*d] (sum over the rows of each group); `bias` as in the forward
 *           (a constant: no gradient); q == dq == NULL is allowed when L == 1 (softmax == 1).
 *           workspace = tvi_attn_fewkeys_bwd_workspace_bytes(...) bytes (may be 0).
 */
size_t tvi_attn_spatial_bwd_workspace_bytes(int batch, int heads, int S);
int tvi_attn_spatial_bwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* o, const tvi_bf16* dout,
                         const float* lse, tvi_bf16* dq, tvi_bf16* dk, tvi_bf16* dv, void* workspace, int batch, int heads,
                         int S, int Skv, int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int64_t lddo,
                         int64_t lddq, int64_t lddk, int64_t lddv, float scale, void* stream);
/* Same with an explicit kernel choice (TVI_SPATIAL_AUTO / _MMA_SYNC / _TCGEN05, as for the forward): AUTO takes the
 * tcgen05 / TMEM / TMA passes for S, Skv >= 1024 and the mma.sync passes below. */
int tvi_attn_spatial_bwd_algo(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* o, const tvi_bf16* dout,
                              const float* lse, tvi_bf16* dq, tvi_bf16* dk, tvi_bf16* dv, void* workspace, int batch, int heads,
                              int S, int Skv, int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t ldo, int64_t lddo,
                              int64_t lddq, int64_t lddk, int64_t lddv, float scale, int algo, void* stream);
int tvi_attn_temporal_bwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* dout, tvi_bf16* dq,
                          tvi_bf16* dk, tvi_bf16* dv, int64_t nseq, int heads, int Tq, int Tk, int d, int64_t ldq,
                          int64_t ldk, int64_t ldv, int64_t lddo, int64_t lddq, int64_t lddk, int64_t lddv, float scale,
                          void* stream);
size_t tvi_attn_fewkeys_bwd_workspace_bytes(int64_t nrows, int64_t rows_per_group, int heads, int L, int d);
int tvi_attn_fewkeys_bwd(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* dout, tvi_bf16* dq,
                         tvi_bf16* dk, tvi_bf16* dv, void* workspace, int64_t nrows, int64_t rows_per_group, int heads,
                         int L, int d, int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo, int64_t lddq, int64_t lddk,
                         int64_t lddv, float scale, int Tq, const float* bias, void* stream);
/* The backward kernels over INTERLEAVED sequences / groups (the `[B*T, S, C]` token order read in place, as the
 * `_interleaved` forwards): training through the temporal block without the permute copies of
 * src/models/attention.py:436-438 / :474-476 and without their backward.  seq_inner / group_inner == 1 are the calls above. */
int tvi_attn_temporal_bwd_interleaved(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* dout, tvi_bf16* dq,
                                      tvi_bf16* dk, tvi_bf16* dv, int64_t nseq, int heads, int Tq, int Tk, int d, int64_t seq_inner,
                                      int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo, int64_t lddq, int64_t lddk, int64_t lddv,
                                      float scale, void* stream);
int tvi_attn_fewkeys_bwd_interleaved(const tvi_bf16* q, const tvi_bf16* k, const tvi_bf16* v, const tvi_bf16* dout, tvi_bf16* dq,
                                     tvi_bf16* dk, tvi_bf16* dv, void* workspace, int64_t nrows, int64_t rows_per_group, int heads,
                                     int L, int d, int64_t group_inner, int64_t ldq, int64_t ldk, int64_t ldv, int64_t lddo,
                                     int64_t lddq, int64_t lddk, int64_t lddv, float scale, int Tq, const float* bias, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TVI_H_ */
