"""Per-step glue of the denoising loop on libtvi.

Mirrors src/pipelines/pipeline_stable_video_diffusion_efficient_interp.py:538-569 for `do_classifier_free_guidance=True`:

    stepper = EulerCfgStepper(sigmas, guidance_scale, prediction_type="v_prediction")
    model_input = stepper.begin(latents, image_latents)       # :355 (init_noise_sigma), :542-546
    for i in range(len(sigmas) - 1):
        noise_pred = unet(model_input, ...)                   # [2B, T, C, H, W], uncond half first (:563)
        model_input = stepper.step(noise_pred, i)             # :565-569 and the next iteration's :542-546, ONE launch
    latents = stepper.latents                                 # fp32 master copy

`sigmas` are the scheduler's (EulerDiscreteScheduler.sigmas, trailing 0 included); `guidance_scale` is the per-frame
vector built at :523-530 ([T] or [B, T]).  The latents live in fp32 for the whole loop; the UNet sees bf16.

`latent_rounding="bf16"` reproduces the arithmetic TYPES of the reference loop instead: CFG in the model dtype with every
torch op rounded (:565-566), `prev_sample` cast back to the model dtype after the scheduler's fp32 arithmetic (diffusers
`EulerDiscreteScheduler.step`), the model input divided in the model dtype (`scale_model_input`); `stepper.latents` is then
an fp32 tensor of bf16-representable values.  The default (None) keeps the more accurate fp32 master copy.
"""
from __future__ import annotations

import torch

from . import _lib, ops

__all__ = ["EulerCfgStepper"]


class EulerCfgStepper:
    def __init__(self, sigmas, guidance_scale: torch.Tensor, prediction_type: str = "v_prediction", latent_rounding=None):
        self.sigmas = [float(s) for s in (sigmas.tolist() if isinstance(sigmas, torch.Tensor) else sigmas)]
        if len(self.sigmas) < 2 or any(s <= 0 for s in self.sigmas[:-1]) or self.sigmas[-1] < 0:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "sigmas must be positive with an optional trailing 0")
        if prediction_type not in ("epsilon", "v_prediction"):
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"prediction_type {prediction_type!r}")
        self.pred = _lib.PRED_V if prediction_type == "v_prediction" else _lib.PRED_EPSILON
        if latent_rounding not in (None, "bf16", torch.bfloat16):
            raise _lib.TviError(_lib.TVI_ERR_UNSUPPORTED, f"latent_rounding {latent_rounding!r} (None or 'bf16')")
        self.rounding = _lib.LATENT_ROUND_NONE if latent_rounding is None else _lib.LATENT_ROUND_BF16
        self.guidance = guidance_scale
        self.latents = None
        self.model_input = None

    @property
    def init_noise_sigma(self) -> float:
        """EulerDiscreteScheduler.init_noise_sigma for `timestep_spacing="leading"` (SVD's config): sqrt(sigma_max^2 + 1)."""
        return (max(self.sigmas) ** 2 + 1) ** 0.5

    def begin(self, latents: torch.Tensor, image_latents: torch.Tensor) -> torch.Tensor:
        """latents [B, T, C, H, W] (already multiplied by init_noise_sigma, :355); image_latents [2B, T, Ci, H, W] bf16."""
        if not latents.is_cuda:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "latents are not on a CUDA device (there is no CPU fallback)")
        B, T, C, H, W = latents.shape
        if image_latents.shape[:2] != (2 * B, T) or image_latents.shape[3:] != (H, W) or image_latents.dtype != torch.bfloat16:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "image_latents must be bf16 [2B, T, Ci, H, W]")
        g = self.guidance.to(latents.device, torch.float32).reshape(-1, T)
        if g.shape[0] not in (1, B):
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "guidance_scale must be [T] or [B, T]")
        self._g = g.contiguous()
        self._g_stride = T if g.shape[0] == B and B > 1 else 0
        self.latents = latents.detach().to(torch.float32).contiguous().clone()
        if self.rounding == _lib.LATENT_ROUND_BF16:  # the reference's latents are a model-dtype tensor from the start
            self.latents = self.latents.to(torch.bfloat16).to(torch.float32)
        Ci = image_latents.shape[2]
        self.model_input = torch.empty(2 * B, T, C + Ci, H, W, dtype=torch.bfloat16, device=latents.device)
        self._dims = (B, T, C, C + Ci, H * W)
        scale = (self.sigmas[0] ** 2 + 1) ** -0.5
        ops._hooked("model_input", 1, _lib.load().tvi_model_input_bf16, self.latents.data_ptr(),
                    image_latents.contiguous().data_ptr(), self.model_input.data_ptr(), B * T, C, C + Ci, H * W, scale,
                    ops._stream())
        return self.model_input

    def step(self, noise_pred: torch.Tensor, i: int) -> torch.Tensor:
        if self.latents is None:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "EulerCfgStepper.step before begin")
        B, T, C, Cin, HW = self._dims
        if noise_pred.dtype != torch.bfloat16 or noise_pred.numel() != 2 * B * T * C * HW or not noise_pred.is_cuda:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, "noise_pred must be a CUDA bf16 [2B, T, C, H, W] tensor")
        if not 0 <= i < len(self.sigmas) - 1:
            raise _lib.TviError(_lib.TVI_ERR_INVALID, f"step index {i} outside the sigma schedule")
        last = i == len(self.sigmas) - 2
        ops._hooked("cfg_euler_step", 1, _lib.load().tvi_cfg_euler_step_round_f32, noise_pred.contiguous().data_ptr(),
                    self.latents.data_ptr(), self._g.data_ptr(), None if last else self.model_input.data_ptr(), B, T, C, Cin, HW,
                    self._g_stride, self.sigmas[i], self.sigmas[i + 1], self.pred, self.rounding, ops._stream())
        return self.model_input
